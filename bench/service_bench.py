#!/usr/bin/env python
"""The headline round measured through the SERVICE API: ``DriverSession`` -> controller gRPC -> a learner that hosts a
fabric rank (the call sequence a user of the reference makes, ``driver_session.py:210-239``).

One synchronous round = RunTask RPC -> ``local_steps`` training steps on the GPU (batches copied from the learner's
host dataset every step) -> control-only MarkTaskCompleted (no tensor bytes) -> controller-commanded fused aggregation
(``AggregateOnDevice``) -> next RunTask.  The round time is taken from the controller's own runtime metadata
(``started_at`` .. ``completed_at`` of every round, ``controller.cc:386-393`` in the reference), first rounds dropped
as warm-up.  With ``device_fabric=False`` the same federation runs on the host path (models as protobuf bytes through
the controller, host aggregation) - the reference's architecture on this box.

    python bench/service_bench.py [--rounds 12] [--host]
"""
import argparse
import json
import os
import socket
import statistics
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402
import yaml  # noqa: E402


def _free_ports(n):
    socks = [socket.socket() for _ in range(n)]
    for s in socks:
        s.bind(("127.0.0.1", 0))
    ports = [s.getsockname()[1] for s in socks]
    for s in socks:
        s.close()
    return ports


def _recipe(n, seed, shape):
    def recipe():
        from metisfl_b200.models.model_dataset import ModelDatasetClassification
        rng = np.random.default_rng(seed)
        y = rng.integers(0, 10, size=n)
        x = rng.normal(size=(n,) + shape).astype(np.float32)
        return ModelDatasetClassification(x=x, y=y.astype(np.int64), size=n,
                                          examples_per_class={int(c): int((y == c).sum()) for c in range(10)})
    return recipe


def _ts(t):
    """Timestamp as dict {seconds, nanos}, number, or RFC 3339 string (protobuf JSON) -> seconds."""
    if isinstance(t, dict):
        return t.get("seconds", 0) + t.get("nanos", 0) * 1e-9
    if isinstance(t, str):
        import calendar
        import time as _time
        head, _, frac = t.rstrip("Z").partition(".")
        return calendar.timegm(_time.strptime(head, "%Y-%m-%dT%H:%M:%S")) + (float("0." + frac) if frac else 0.0)
    return float(t)


def service_rounds(device_fabric=True, rounds=12, warmup=4, local_steps=4, batch=128, model="resnet18", cuda_device=0):
    """Returns {"ms_per_round": median, "rounds": n, "aggregation_ms": median block duration, counters...}."""
    from metisfl_b200.driver.driver_session import DriverSession
    from metisfl_b200.models import zoo
    from metisfl_b200.utils.fedenv_parser import FederationEnvironment
    with open(os.path.join(ROOT, "examples", "config", "fashionmnist_2learners_sync_fedavg.yaml")) as f:
        d = yaml.safe_load(f)
    fe = d["FederationEnvironment"]
    ports = _free_ports(2)
    fe["Controller"]["GRPCServicer"]["Port"] = ports[0]
    fe["Learners"] = fe["Learners"][:1]
    fe["Learners"][0]["GRPCServicer"]["Port"] = ports[1]
    fe["Learners"][0]["CudaDevices"] = [cuda_device]
    fe["TerminationSignals"]["FederationRounds"] = rounds + warmup
    fe["TerminationSignals"]["ExecutionCutoffTimeMins"] = 3
    fe["LocalModelConfig"]["BatchSize"] = batch
    fe["LocalModelConfig"]["LocalEpochs"] = 1
    fe["LocalModelConfig"]["OptimizerConfig"] = {"OptimizerName": "MomentumSGD", "LearningRate": 0.02,
                                                 "MomentumFactor": 0.9}
    fe["DeviceFabric"] = bool(device_fabric)
    fe["ModelDType"] = "bf16"
    factory, shape = (zoo.ResNet18, (3, 32, 32)) if model == "resnet18" else (zoo.FashionMnistMLP, (1, 28, 28))
    n = batch * local_steps   # one epoch = local_steps batches
    with tempfile.TemporaryDirectory() as tmp:
        # no snapshot directory: with one, the controller pulls the community model's bytes off the device every round
        os.environ.pop("METISFL_SNAPSHOT_DIR", None)
        torch.manual_seed(0)
        s = DriverSession(FederationEnvironment(d), factory(), _recipe(n, 1, shape), None, _recipe(batch, 3, shape),
                          working_dir=os.path.join(tmp, "w"), launcher="inprocess", banner=False,
                          model_dtype=torch.bfloat16)
        stats = s.run(request_every_secs=0.2)
    md = stats["federation_runtime_metadata"]["metadata"]
    done = [m for m in md if m.get("completed_at") and m.get("started_at")]
    spans = [(_ts(m["completed_at"]) - _ts(m["started_at"])) * 1e3 for m in done][warmup:]
    # round-to-round cadence (start of round i+1 minus start of round i) also counts the controller's scheduling gap
    starts = [_ts(m["started_at"]) for m in done][warmup:]
    cadence = [(b - a) * 1e3 for a, b in zip(starts, starts[1:])]
    agg = [m["model_aggregation_block_duration_ms"][0] for m in done[warmup:]
           if m.get("model_aggregation_block_duration_ms")]
    counters = json.loads(stats["federation_runtime_metadata"].get("json_metadata") or "{}")
    return {"path": "device fabric (control-only messages, controller-commanded K1)" if device_fabric
                    else "host (models as protobuf bytes through the controller, host aggregation)",
            "ms_per_round": statistics.median(cadence) if cadence else (statistics.median(spans) if spans else None),
            "ms_round_span": statistics.median(spans) if spans else None,
            "rounds_timed": len(spans), "local_steps": local_steps, "batch": batch,
            "aggregation_block_ms": statistics.median(agg) if agg else None,
            "controller_rx_model_bytes": counters.get("model_bytes_received"),
            "aggregation_device_timed": counters.get("aggregation_durations_device_timed"),
            "timing": "controller runtime metadata (host clock around each round, gRPC included)"}


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--rounds", type=int, default=12)
    ap.add_argument("--warmup", type=int, default=4)
    ap.add_argument("--local-steps", type=int, default=4)
    ap.add_argument("--batch", type=int, default=128)
    ap.add_argument("--model", default="resnet18", choices=["resnet18", "mlp"])
    ap.add_argument("--host", action="store_true", help="the host (protobuf-through-controller) path instead")
    a = ap.parse_args()
    print(json.dumps(service_rounds(not a.host, a.rounds, a.warmup, a.local_steps, a.batch, a.model)))
