"""Host-path round time: one controller + N in-process learners exchanging whole models over gRPC (the path of
off-box learners and of federations without a device fabric; in the reference it is the only path).

    python bench/host_path_bench.py --learners 1 --width 1060 --depth 10            # real PyTorch learners, lr = 0
    python bench/host_path_bench.py --learners 8 --echo                             # no training: protocol + aggregation

Prints one JSON line: median wall-clock per round from the controller's runtime metadata (started_at -> completed_at),
the median aggregation time, and - for one learner - the median "train leg" (RunTask submitted -> model received back).
CPU only; the model is the reference's stress model (`examples/keras/scalability_testing.py` + `housing_mlp.py:8-43`).
"""
import argparse
import datetime
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402

from examples.scalability_testing import Regressor, recipe  # noqa: E402
from examples.utils.environment_generator import EnvGen  # noqa: E402
from metisfl_b200.driver.driver_session import DriverSession  # noqa: E402


def _ts(x):
    if isinstance(x, dict):
        return x.get("seconds", 0) + x.get("nanos", 0) * 1e-9
    if isinstance(x, str):
        return datetime.datetime.fromisoformat(x.replace("Z", "+00:00")).timestamp()
    return float(x)


if __name__ == "__main__":
    here = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    ap = argparse.ArgumentParser()
    ap.add_argument("--learners", type=int, default=1)
    ap.add_argument("--rounds", type=int, default=8)
    ap.add_argument("--width", type=int, default=1060)
    ap.add_argument("--depth", type=int, default=10)
    ap.add_argument("--echo", action="store_true", help="learners echo the model back without training")
    a = ap.parse_args()
    env = EnvGen(os.path.join(here, "examples", "config", "template.yaml")).generate_localhost(a.rounds, a.learners)
    env.local_model_config.optimizer_config.optimizer_pb_kwargs["learning_rate"] = 0.0
    env.local_model_config.local_epochs = 1
    env.evaluation_metric = "loss"
    model = Regressor(width=a.width, depth=a.depth)
    s = DriverSession(env, model, recipe, None, recipe, launcher="inprocess", banner=False,
                      neural_engine="echo" if a.echo else "pytorch", echo_sleep_secs=0.0)
    stats = s.run(request_every_secs=0.5)
    md = [m for m in stats["federation_runtime_metadata"]["metadata"] if "completed_at" in m][2:]   # skip warm-up rounds
    rounds = [(_ts(m["completed_at"]) - _ts(m["started_at"])) * 1e3 for m in md]
    legs = []
    for m in md:
        sub, rec = m.get("train_task_submitted_at", {}), m.get("train_task_received_at", {})
        legs += [(_ts(rec[k]) - _ts(sub[k])) * 1e3 for k in rec if k in sub]
    params = sum(p.numel() for p in model.parameters())
    print(json.dumps({"learners": a.learners, "echo": a.echo, "params": params, "model_mb": round(params * 4 / 1e6, 1),
                      "rounds_timed": len(rounds), "median_round_ms": round(float(np.median(rounds)), 1),
                      "median_train_leg_ms": round(float(np.median(legs)), 1) if legs else None,
                      "median_aggregation_ms": round(float(np.median(
                          [m.get("model_aggregation_total_duration_ms", 0.0) for m in md])), 1)}))
