"""Wire contract: tensor <-> TensorSpec round trips for every dtype (reference
``proto_messages_factory_test.py:10-52``, ``proto_tensor_serde_test.cc:63-151``) including the
FLOAT16 / BFLOAT16 extensions, native codec == python protobuf, field numbers of the contract."""
import numpy as np
import pytest
import torch

from metisfl_b200 import _core
from metisfl_b200.proto import controller_pb2, learner_pb2, metis_pb2, model_pb2
from metisfl_b200.utils.proto_messages_factory import MetisProtoMessages, ModelProtoMessages

CODEC = ModelProtoMessages.TensorSpecProto
NP_TYPES = [np.int8, np.int16, np.int32, np.int64, np.uint8, np.uint16, np.uint32, np.uint64, np.float32,
            np.float64, np.float16]


@pytest.mark.parametrize("dtype", NP_TYPES)
def test_numpy_round_trip(dtype):
    a = (np.arange(24).reshape(2, 3, 4) % 100).astype(dtype)
    spec = CODEC.numpy_array_to_proto_tensor_spec(a)
    assert spec.length == 24 and list(spec.dimensions) == [2, 3, 4]
    assert spec.value == a.tobytes()
    b = CODEC.proto_tensor_spec_to_numpy_array(model_pb2.TensorSpec.FromString(spec.SerializeToString()))
    assert b.dtype == np.dtype(dtype) and np.array_equal(a, b)


def test_big_endian_and_fortran_order():
    a = np.arange(12, dtype=">i4").reshape(3, 4)
    spec = CODEC.numpy_array_to_proto_tensor_spec(a)
    assert spec.type.byte_order == model_pb2.DType.BIG_ENDIAN_ORDER
    assert np.array_equal(CODEC.proto_tensor_spec_to_numpy_array(spec), a)
    f = np.asfortranarray(np.arange(12, dtype=np.float32).reshape(3, 4))
    spec = CODEC.numpy_array_to_proto_tensor_spec(f)
    assert spec.type.fortran_order
    assert np.array_equal(CODEC.proto_tensor_spec_to_numpy_array(spec), f)
    t = CODEC.proto_tensor_spec_to_torch_tensor(CODEC.numpy_array_to_proto_tensor_spec(a))
    assert t.tolist() == a.tolist()


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32, torch.int64, torch.uint8])
def test_torch_round_trip_including_bfloat16(dtype):
    t = (torch.arange(30).reshape(5, 6) - 7).to(dtype)
    spec = CODEC.torch_tensor_to_proto_tensor_spec(t)
    assert spec.type.type == {torch.bfloat16: model_pb2.DType.BFLOAT16, torch.float16: model_pb2.DType.FLOAT16,
                              torch.float32: model_pb2.DType.FLOAT32, torch.int64: model_pb2.DType.INT64,
                              torch.uint8: model_pb2.DType.UINT8}[dtype]
    back = CODEC.proto_tensor_spec_to_torch_tensor(spec)
    assert back.dtype == dtype and torch.equal(back, t)


def test_native_codec_is_byte_identical_to_protobuf():
    arrays = [np.random.randn(3, 5).astype(np.float32), np.arange(7, dtype=np.int64), np.zeros((0,), np.float64)]
    m = ModelProtoMessages.construct_model_pb_from_np(arrays, ["a", "b/c", "empty"], [True, False, True])
    blob = m.SerializeToString()
    parsed = _core.parse_model(blob)
    assert [p["name"] for p in parsed] == ["a", "b/c", "empty"] and [p["trainable"] for p in parsed] == [True, False, True]
    assert _core.build_model(parsed) == blob
    fm = _core.build_federated_model(3, 9, blob)
    pb = model_pb2.FederatedModel.FromString(fm)
    assert pb.num_contributors == 3 and pb.global_iteration == 9 and pb.model == m
    assert _core.parse_federated_model(pb.SerializeToString()) == (3, 9, blob)


def test_device_slot_extension_round_trips_and_is_ignored_by_old_readers():
    parsed = [{"name": "w", "trainable": True, "length": 4, "dimensions": [4], "dtype": 11, "value": b"",
               "device_slot": (3, 1, 4096, 17)}]
    blob = _core.build_model(parsed)
    v = model_pb2.Model.FromString(blob).variables[0].plaintext_tensor.tensor_spec
    assert (v.device_slot.device_rank, v.device_slot.slot, v.device_slot.offset_bytes, v.device_slot.version) == (3, 1, 4096, 17)
    assert _core.parse_model(blob)[0]["device_slot"] == (3, 1, 4096, 17)


def test_contract_field_numbers():
    def num(msg, field):
        return msg.DESCRIPTOR.fields_by_name[field].number
    assert num(model_pb2.TensorSpec, "value") == 4 and num(model_pb2.Model.Variable, "ciphertext_tensor") == 4
    assert num(metis_pb2.SSLConfig, "ssl_config_files") == 6 and num(metis_pb2.SSLConfig, "ssl_config_stream") == 7
    assert num(metis_pb2.AggregationRule, "aggregation_rule_specs") == 5 and num(metis_pb2.AggregationRule, "fed_async") == 6
    assert num(metis_pb2.FederatedTaskRuntimeMetadata, "model_tensor_quantifiers") == 18
    assert num(learner_pb2.EvaluateModelRequest, "evaluation_dataset") == 3
    assert num(controller_pb2.MarkTaskCompletedRequest, "task") == 3
    assert model_pb2.DType.FLOAT64 == 9 and model_pb2.DType.BFLOAT16 == 11
    svc = controller_pb2.DESCRIPTOR.services_by_name["ControllerService"]
    lsvc = learner_pb2.DESCRIPTOR.services_by_name["LearnerService"]
    # the reference's 4 learner RPCs first and unchanged, then the AggregateOnDevice extension
    assert len(svc.methods) == 12 and [m.name for m in lsvc.methods][:4] == [
        "EvaluateModel", "GetServicesHealthStatus", "RunTask", "ShutDown"]
    assert [m.name for m in lsvc.methods][4:] == ["AggregateOnDevice"]


def test_factories_build_the_oneofs():
    rule = MetisProtoMessages.construct_aggregation_rule_pb("FedStride", "NumParticipants", 3, None)
    assert rule.WhichOneof("rule") == "fed_stride" and rule.fed_stride.stride_length == 3
    rule = MetisProtoMessages.construct_aggregation_rule_pb("FedAsync", "NumTrainingExamples", None, None,
                                                            fed_async_kwargs={"alpha": 0.7, "staleness_fn": "hinge"})
    assert rule.WhichOneof("rule") == "fed_async" and abs(rule.fed_async.alpha - 0.7) < 1e-6
    store = MetisProtoMessages.construct_model_store_config_pb("Redis", "LineageLengthEviction", 2, "h", 6379)
    assert store.redis_db_store.model_store_specs.lineage_length_eviction.lineage_length == 2
    opt = ModelProtoMessages.construct_optimizer_config_pb_from_kwargs({"name": "FedProx", "learning_rate": 0.1,
                                                                        "proximal_term": 0.01})
    assert opt.WhichOneof("config") == "fed_prox"
    with pytest.raises(RuntimeError):
        MetisProtoMessages.construct_aggregation_rule_specs_pb("bogus")
