#!/usr/bin/env python
"""Host baselines of BASELINE.md §3, measured on the same box as the product numbers.

B1 `ref-cpu-agg`  : the reference's aggregation path restated on the CPU - whole models as serialized
                    protobuf `Model` bytes, per-variable OpenMP loop, per-term truncation semantics
                    (`federated_average.cc:15-151`) = `metisfl_b200._core.FedAvg` (the native host core).
B2 `ref-grpc-round`: a reference-style federation round: models travel as protobuf over localhost gRPC
                    (RunTask fan-out, MarkTaskCompleted gather, host aggregation, EvaluateModel fan-out)
                    with no-train echo learners (`test/learner_notrain_noeval.py:170-191`).
B4 `cpu-ckks`     : host RNS-CKKS encrypt / weighted average / decrypt (`ckks_scheme.cc:110-252` semantics).
"""
import argparse
import json
import os
import sys
import threading
import time
from concurrent import futures

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import grpc  # noqa: E402
import numpy as np  # noqa: E402

from metisfl_b200 import _core  # noqa: E402
from metisfl_b200.proto import (controller_pb2, learner_pb2, learner_pb2_grpc, metis_pb2, model_pb2,  # noqa: E402
                                service_common_pb2)
from metisfl_b200.utils.proto_messages_factory import MetisProtoMessages, ModelProtoMessages  # noqa: E402

MODELS = {"mlp": 118282, "cifarcnn": 1610314, "resnet18": 11183582, "gpt2-medium": 354823168}


def synth_model(n_params, n_vars=60, seed=0):
    rng = np.random.default_rng(seed)
    sizes = np.full(n_vars, n_params // n_vars)
    sizes[-1] += n_params - sizes.sum()
    arrays = [rng.standard_normal(int(s), dtype=np.float32) for s in sizes]
    return ModelProtoMessages.construct_model_pb_from_np(arrays, [f"v{i}" for i in range(n_vars)], [True] * n_vars)


def b1(models, learners):
    rows = []
    for name in models:
        n = MODELS[name]
        for L in learners:
            if n * 4 * L > 24e9:
                continue
            blob = synth_model(n).SerializeToString()
            block = [([blob], 1.0 / L)] * L
            agg = _core.FedAvg()
            agg.aggregate(block)
            reps = 3 if n * L < 5e8 else 1
            t0 = time.perf_counter()
            for _ in range(reps):
                agg.aggregate(block)
            ms = (time.perf_counter() - t0) * 1e3 / reps
            rows.append({"model": name, "params": n, "learners": L, "ms": ms, "params_per_s": L * n / (ms * 1e-3)})
    return rows


class _Echo(learner_pb2_grpc.LearnerServiceServicer):
    def __init__(self, ctrl_stub, lid_tok):
        self.ctrl, self.lid_tok = ctrl_stub, lid_tok

    def RunTask(self, request, context):
        def done():
            lid, tok = self.lid_tok
            task = metis_pb2.CompletedLearningTask(
                model=request.federated_model.model,
                execution_metadata=metis_pb2.TaskExecutionMetadata(global_iteration=request.task.global_iteration,
                                                                   completed_batches=1, batch_size=1))
            self.ctrl.MarkTaskCompleted(controller_pb2.MarkTaskCompletedRequest(learner_id=lid, auth_token=tok, task=task))
        threading.Thread(target=done, daemon=True).start()
        return learner_pb2.RunTaskResponse(ack=service_common_pb2.Ack(status=True))

    def EvaluateModel(self, request, context):
        return learner_pb2.EvaluateModelResponse(evaluations=metis_pb2.ModelEvaluations())


def b2(model_name, n_learners, rounds, base_port):
    from metisfl_b200.controller.__main__ import init_controller
    from metisfl_b200.proto import controller_pb2_grpc
    opts = [("grpc.max_send_message_length", -1), ("grpc.max_receive_message_length", -1)]
    se = MetisProtoMessages.construct_server_entity_pb("127.0.0.1", base_port)
    inst = init_controller(se.SerializeToString().hex(), block=False)
    chan = grpc.insecure_channel(f"127.0.0.1:{base_port}", options=opts)
    ctrl = controller_pb2_grpc.ControllerServiceStub(chan)
    servers = []
    model = synth_model(MODELS[model_name])
    ctrl.ReplaceCommunityModel(controller_pb2.ReplaceCommunityModelRequest(
        model=model_pb2.FederatedModel(num_contributors=n_learners, model=model)))
    for i in range(n_learners):
        port = base_port + 1 + i
        holder = [None, None]
        srv = grpc.server(futures.ThreadPoolExecutor(max_workers=4), options=opts)
        learner_pb2_grpc.add_LearnerServiceServicer_to_server(_Echo(ctrl, holder), srv)
        srv.add_insecure_port(f"127.0.0.1:{port}")
        srv.start()
        servers.append(srv)
        r = ctrl.JoinFederation(controller_pb2.JoinFederationRequest(
            server_entity=metis_pb2.ServerEntity(hostname="127.0.0.1", port=port),
            local_dataset_spec=metis_pb2.DatasetSpec(num_training_examples=100)))
        holder[0], holder[1] = r.learner_id, r.auth_token
    core = inst.servicer.core
    t0 = time.time()
    while core.global_iteration() < rounds + 2 and time.time() - t0 < 600:
        time.sleep(0.005)
    md = core.get_runtime_metadata_lineage(0)
    full = [m for m in md if len(m["completed_by_learner_id"]) == n_learners and m["completed_at"] > 0][1:rounds + 1]
    per_round = [m["completed_at"] - m["started_at"] for m in full]
    agg_ms = [m["model_aggregation_total_duration_ms"] for m in full]
    for s in servers:
        s.stop(0)
    inst.shutdown(instantly=True)
    chan.close()
    return {"model": model_name, "params": MODELS[model_name], "learners": n_learners, "rounds": len(per_round),
            "round_time_s": float(np.median(per_round)) if per_round else None,
            "host_aggregation_ms": float(np.median(agg_ms)) if agg_ms else None}


def b4(n_params, learners):
    from metisfl_b200.encryption import fhe
    import tempfile
    s = fhe.CKKS(4096, 52, backend="host")
    s.gen_crypto_context_and_keys(tempfile.mkdtemp(), seed=1)
    x = np.random.default_rng(0).standard_normal(n_params)
    t0 = time.perf_counter(); ct = s.encrypt(x); t_enc = time.perf_counter() - t0
    t0 = time.perf_counter(); avg = _core.ckks_weighted_average([ct] * learners, [1.0 / learners] * learners)
    t_avg = time.perf_counter() - t0
    t0 = time.perf_counter(); y = s.decrypt(avg, n_params); t_dec = time.perf_counter() - t0
    return {"params": n_params, "learners": learners, "ciphertext_bytes": len(ct), "encrypt_s": t_enc,
            "weighted_average_s": t_avg, "decrypt_s": t_dec, "round_s": t_enc + t_avg + t_dec,
            "max_abs_error": float(np.abs(y - x).max())}


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    a = ap.parse_args()
    out = {"cpu_threads": os.cpu_count()}
    out["B1_ref_cpu_agg"] = b1(["mlp", "cifarcnn", "resnet18"] + ([] if a.quick else ["gpt2-medium"]),
                               [2, 8] if a.quick else [2, 8, 64])
    out["B2_ref_grpc_round"] = [b2("mlp", 2, 5, 51500), b2("resnet18", 8, 3, 51600)]
    out["B4_cpu_ckks"] = b4(MODELS["cifarcnn"] if a.quick else MODELS["resnet18"], 8)
    print(json.dumps(out))
