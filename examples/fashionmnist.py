"""Config #1 (BASELINE.json): FashionMNIST-shaped MLP, 2 learners, synchronous FedAvg - runs on CPU.

The canonical four-call flow of the reference example (``examples/keras/fashionmnist.py:85-97``) with
a PyTorch model and synthetic 28x28 data (no dataset download in this environment).
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402

from metisfl_b200.driver.driver_session import DriverSession  # noqa: E402
from metisfl_b200.models.model_dataset import ModelDatasetClassification  # noqa: E402
from metisfl_b200.models.zoo import FashionMnistMLP  # noqa: E402


def make_recipe(n, seed):
    def recipe():
        rng = np.random.default_rng(seed)
        # class-dependent means so that accuracy actually improves over rounds
        y = rng.integers(0, 10, size=n)
        x = rng.normal(size=(n, 1, 28, 28)).astype(np.float32) + (y[:, None, None, None] / 5.0).astype(np.float32)
        counts = {int(c): int((y == c).sum()) for c in range(10)}
        return ModelDatasetClassification(x=x, y=y.astype(np.int64), size=n, examples_per_class=counts)
    return recipe


if __name__ == "__main__":
    here = os.path.dirname(os.path.abspath(__file__))
    ap = argparse.ArgumentParser()
    ap.add_argument("--env", default=os.path.join(here, "config", "fashionmnist_2learners_sync_fedavg.yaml"))
    ap.add_argument("--launcher", default=None)
    ap.add_argument("--out", default="/tmp/metis/experiment.json")
    args = ap.parse_args()
    session = DriverSession(args.env, FashionMnistMLP(), make_recipe(640, 1), make_recipe(128, 2), make_recipe(128, 3),
                            launcher=args.launcher)
    stats = session.run(request_every_secs=1)
    with open(args.out, "w") as f:
        json.dump(stats, f, indent=1)
    rounds = stats["federation_runtime_metadata"].get("metadata", [])
    print("rounds recorded:", len(rounds), "->", args.out)
