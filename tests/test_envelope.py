"""Byte-level request assembly (``proto/envelope.py``): what goes on the wire must parse to exactly the message the
protobuf object path would have produced, and what arrives must be taken apart without trusting it."""
import numpy as np
import pytest
import torch
from hypothesis import given, settings, strategies as st

from metisfl_b200.proto import controller_pb2, envelope, learner_pb2, metis_pb2, model_pb2
from metisfl_b200.utils.proto_messages_factory import ModelProtoMessages


def _weights():
    g = torch.Generator().manual_seed(0)
    return (["a.weight", "a.bias", "bn.n", "h", "e", "f"], [True, True, False, True, True, True],
            [torch.randn(5, 3, generator=g), torch.randn(5, generator=g).to(torch.bfloat16),
             torch.tensor(7, dtype=torch.int64), np.arange(6, dtype=np.float64).reshape(2, 3),
             np.zeros((0, 4), dtype=np.float32), np.asfortranarray(np.arange(6, dtype=np.int32).reshape(2, 3))])


def test_model_bytes_equal_the_object_path():
    names, trainables, values = _weights()
    want = ModelProtoMessages.construct_model_pb_from_np(values, names, trainables)
    got = model_pb2.Model.FromString(ModelProtoMessages.construct_model_bytes_from_np(values, names, trainables))
    assert got == want
    assert len(got.variables[1].plaintext_tensor.tensor_spec.value) == 10      # bf16: two bytes per element
    assert got.variables[5].plaintext_tensor.tensor_spec.type.fortran_order


def test_encrypted_model_bytes(tmp_path):
    from metisfl_b200.encryption import fhe
    scheme = fhe.CKKS(4096, 52, backend="host")
    scheme.gen_crypto_context_and_keys(str(tmp_path))
    names, trainables, values = _weights()
    m = model_pb2.Model.FromString(ModelProtoMessages.construct_model_bytes_from_np(values[:2], names[:2],
                                                                                   trainables[:2], he_scheme=scheme))
    spec = m.variables[0].ciphertext_tensor.tensor_spec
    assert m.variables[0].HasField("ciphertext_tensor") and list(spec.dimensions) == [5, 3] and spec.length == 15
    np.testing.assert_allclose(scheme.decrypt(spec.value, 15), values[0].numpy().ravel(), atol=1e-6)


def test_run_task_and_evaluate_requests_parse_to_the_same_messages():
    names, trainables, values = _weights()
    model = ModelProtoMessages.construct_model_pb_from_np(values, names, trainables)
    fm = model_pb2.FederatedModel(num_contributors=3, global_iteration=9, model=model)
    task = metis_pb2.LearningTask(global_iteration=9, num_local_updates=12)
    rest = learner_pb2.RunTaskRequest(task=task)
    raw = envelope.embed(learner_pb2.RunTaskRequest, "federated_model", fm.SerializeToString(), rest)
    assert learner_pb2.RunTaskRequest.FromString(raw) == learner_pb2.RunTaskRequest(federated_model=fm, task=task)
    E = learner_pb2.EvaluateModelRequest
    raw = envelope.embed(E, "model", model.SerializeToString(), E(batch_size=50, evaluation_dataset=[E.TEST]))
    assert E.FromString(raw) == E(model=model, batch_size=50, evaluation_dataset=[E.TEST])


def test_mark_task_completed_request_both_task_forms():
    names, trainables, values = _weights()
    model = ModelProtoMessages.construct_model_pb_from_np(values, names, trainables)
    meta = metis_pb2.TaskExecutionMetadata(global_iteration=4, completed_batches=8, batch_size=16)
    task_pb = metis_pb2.CompletedLearningTask(model=model, execution_metadata=meta, aux_metadata="x")
    want = controller_pb2.MarkTaskCompletedRequest(learner_id="l:1", auth_token="tok", task=task_pb)
    wire_task = envelope.WireCompletedTask(model.SerializeToString(), meta, "x")
    for task in (task_pb, wire_task):
        raw = envelope.mark_task_completed_request("l:1", "tok", task)
        assert controller_pb2.MarkTaskCompletedRequest.FromString(raw) == want
    assert wire_task.to_pb() == task_pb and wire_task.model == model
    # the receiving side: the model record comes out as the same bytes, wherever it sits in the request
    R, T = controller_pb2.MarkTaskCompletedRequest, metis_pb2.CompletedLearningTask
    for raw in (envelope.mark_task_completed_request("l:1", "tok", wire_task), want.SerializeToString()):
        outer = envelope.split(raw)
        inner = envelope.split(envelope.merged_bytes(outer, R, "task"))
        assert bytes(envelope.merged_bytes(inner, T, "model")) == model.SerializeToString()
        assert bytes(envelope.last_bytes(outer, R, "learner_id")) == b"l:1"


def test_a_task_split_over_two_records_is_merged():
    R, T = controller_pb2.MarkTaskCompletedRequest, metis_pb2.CompletedLearningTask
    a = T(aux_metadata="first").SerializeToString()
    b = T(execution_metadata=metis_pb2.TaskExecutionMetadata(global_iteration=2)).SerializeToString()
    n = R.DESCRIPTOR.fields_by_name["task"].number
    raw = envelope.record_header(n, len(a)) + a + envelope.record_header(n, len(b)) + b
    merged = T.FromString(bytes(envelope.merged_bytes(envelope.split(raw), R, "task")))
    assert merged == R.FromString(raw).task and merged.aux_metadata == "first"


@settings(max_examples=300, deadline=None)
@given(st.binary(max_size=200))
def test_split_never_reads_outside_its_input(data):
    try:
        rec = envelope.split(data)
    except ValueError:
        return
    total = 0
    for vals in rec.values():
        for wt, v in vals:
            if wt != 0:
                assert len(v) <= len(data)
                total += len(v)
    assert total <= len(data)


@settings(max_examples=200, deadline=None)
@given(st.text(max_size=20), st.text(max_size=20), st.binary(max_size=64), st.integers(0, 2 ** 31 - 1))
def test_split_agrees_with_protobuf_on_valid_messages(lid, tok, payload, it):
    m = model_pb2.Model()
    m.variables.add(name="v").plaintext_tensor.tensor_spec.value = payload
    req = controller_pb2.MarkTaskCompletedRequest(
        learner_id=lid, auth_token=tok, task=metis_pb2.CompletedLearningTask(
            model=m, execution_metadata=metis_pb2.TaskExecutionMetadata(global_iteration=it)))
    R, T = controller_pb2.MarkTaskCompletedRequest, metis_pb2.CompletedLearningTask
    outer = envelope.split(req.SerializeToString())
    got_lid = envelope.last_bytes(outer, R, "learner_id")
    assert (bytes(got_lid).decode() if got_lid is not None else "") == lid
    task = envelope.merged_bytes(outer, R, "task")
    inner = envelope.split(task) if task is not None else {}
    model = envelope.merged_bytes(inner, T, "model")
    assert model_pb2.Model.FromString(bytes(model) if model is not None else b"") == m


def test_controller_takes_mark_task_completed_as_raw_bytes_and_rejects_garbage():
    """Over a real gRPC channel: a request assembled by the envelope, a request serialized by protobuf, and bytes that
    are no message at all (INVALID_ARGUMENT, not a crashed handler); unknown credentials still map to their status."""
    import socket

    import grpc

    from metisfl_b200.controller.service import ControllerServicer
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    p = metis_pb2.ControllerParams()
    p.server_entity.hostname, p.server_entity.port = "127.0.0.1", port
    p.global_model_specs.aggregation_rule.fed_avg.SetInParent()
    p.communication_specs.protocol = metis_pb2.CommunicationSpecs.SYNCHRONOUS
    p.model_store_config.in_memory_store.SetInParent()
    p.model_hyperparams.batch_size, p.model_hyperparams.epochs = 32, 1
    ctrl = ControllerServicer(p)
    ctrl.start()
    try:
        ch = grpc.insecure_channel(f"127.0.0.1:{port}")
        call = ch.unary_unary("/metisfl.ControllerService/MarkTaskCompleted", request_serializer=lambda b: b,
                              response_deserializer=controller_pb2.MarkTaskCompletedResponse.FromString)
        names, trainables, values = _weights()
        model = ModelProtoMessages.construct_model_pb_from_np(values, names, trainables)
        task = envelope.WireCompletedTask(model.SerializeToString(), metis_pb2.TaskExecutionMetadata(global_iteration=1))
        before = ctrl.controller_rx_model_bytes
        for raw in (envelope.mark_task_completed_request("nobody:1", "bad", task),
                    controller_pb2.MarkTaskCompletedRequest(learner_id="nobody:1", auth_token="bad",
                                                            task=task.to_pb()).SerializeToString()):
            with pytest.raises(grpc.RpcError) as e:
                call(raw, timeout=10)
            assert e.value.code() == grpc.StatusCode.UNAUTHENTICATED      # parsed fine; the core refused the identity
        payload = sum(len(v.plaintext_tensor.tensor_spec.value) for v in model.variables)
        assert ctrl.controller_rx_model_bytes - before == 2 * payload
        for garbage in (b"\x1a\xff\xff\xff\xff\x0f" + b"x" * 10, b"\x0a\x05ab", b"\x07"):
            with pytest.raises(grpc.RpcError) as e:
                call(garbage, timeout=10)
            assert e.value.code() == grpc.StatusCode.INVALID_ARGUMENT
    finally:
        ctrl.stop()


def test_split_bounds_the_number_of_records():
    with pytest.raises(ValueError):
        envelope.split(b"\x08\x01" * (envelope.MAX_TOP_LEVEL_RECORDS + 1))
    assert len(envelope.split(b"\x08\x01" * 100)[1]) == 100


def test_mark_task_completed_also_takes_a_parsed_message():
    """In-process callers hand the servicer a message object instead of bytes: same outcome."""
    import grpc

    from metisfl_b200.controller.service import ControllerServicer

    class Ctx:
        code = None

        def abort(self, code, details):
            self.code = code
            raise RuntimeError(details)

    p = metis_pb2.ControllerParams()
    p.server_entity.hostname, p.server_entity.port = "127.0.0.1", 1
    p.global_model_specs.aggregation_rule.fed_avg.SetInParent()
    p.communication_specs.protocol = metis_pb2.CommunicationSpecs.SYNCHRONOUS
    p.model_store_config.in_memory_store.SetInParent()
    p.model_hyperparams.batch_size, p.model_hyperparams.epochs = 32, 1
    ctrl = ControllerServicer(p)
    try:
        names, trainables, values = _weights()
        model = ModelProtoMessages.construct_model_pb_from_np(values, names, trainables)
        req = controller_pb2.MarkTaskCompletedRequest(
            learner_id="nobody:1", auth_token="bad", task=metis_pb2.CompletedLearningTask(model=model))
        for form in (req, req.SerializeToString()):
            ctx = Ctx()
            with pytest.raises(RuntimeError):
                ctrl.MarkTaskCompleted(form, ctx)
            assert ctx.code == grpc.StatusCode.UNAUTHENTICATED
        ctx = Ctx()
        with pytest.raises(RuntimeError):
            ctrl.MarkTaskCompleted(b"\x07", ctx)
        assert ctx.code == grpc.StatusCode.INVALID_ARGUMENT
    finally:
        ctrl.stop()
