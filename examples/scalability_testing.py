"""Controller stress run (reference ``examples/keras/scalability_testing.py:26-115``): N learners x a
HousingMLP of configurable width / depth with learning rate 0, so rounds cost only communication and
aggregation.  ``--launcher inprocess`` keeps everything in one process."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402

from examples.utils.environment_generator import EnvGen  # noqa: E402
from metisfl_b200.driver.driver_session import DriverSession  # noqa: E402
from metisfl_b200.models.model_dataset import ModelDatasetRegression  # noqa: E402
from metisfl_b200.models.zoo import HousingMLP  # noqa: E402

import torch.nn.functional as F  # noqa: E402


def recipe():
    rng = np.random.default_rng(0)
    x = rng.normal(size=(256, 13)).astype(np.float32)
    y = rng.normal(size=(256,)).astype(np.float32)
    return ModelDatasetRegression(x=x, y=y, size=256, min_val=float(y.min()), max_val=float(y.max()),
                                  mean_val=float(y.mean()), median_val=float(np.median(y)), stddev_val=float(y.std()))


class Regressor(HousingMLP):
    def loss_fn(self, out, y):
        return F.mse_loss(out.float(), y.float())


if __name__ == "__main__":
    here = os.path.dirname(os.path.abspath(__file__))
    ap = argparse.ArgumentParser()
    ap.add_argument("--learners", type=int, default=8)
    ap.add_argument("--rounds", type=int, default=5)
    ap.add_argument("--width", type=int, default=100)
    ap.add_argument("--depth", type=int, default=10)
    ap.add_argument("--launcher", default="inprocess")
    ap.add_argument("--echo", action="store_true",
                    help="learners that only sleep and echo the model (the reference's learner_notrain_noeval backend)")
    ap.add_argument("--echo_sleep_secs", type=float, default=0.5)
    a = ap.parse_args()
    env = EnvGen(os.path.join(here, "config", "template.yaml")).generate_localhost(a.rounds, a.learners)
    env.local_model_config.optimizer_config.optimizer_pb_kwargs["learning_rate"] = 0.0  # nothing trains
    env.evaluation_metric = "loss"
    session = DriverSession(env, Regressor(width=a.width, depth=a.depth), recipe, None, recipe, launcher=a.launcher,
                            banner=False, neural_engine="echo" if a.echo else "pytorch",
                            echo_sleep_secs=a.echo_sleep_secs)
    stats = session.run(request_every_secs=1)
    md = [m for m in stats["federation_runtime_metadata"]["metadata"] if "completed_at" in m]
    agg = [m.get("model_aggregation_total_duration_ms", 0.0) for m in md]
    print(json.dumps({"learners": a.learners, "params": sum(p.numel() for p in Regressor(width=a.width, depth=a.depth).parameters()),
                      "rounds": len(md), "median_aggregation_ms": float(np.median(agg)) if agg else None}))
