"""Minimal PyTorch federation on a CSV table (counterpart of reference ``examples/pytorch/dummy.py:19-103`` and
``examples/pytorch/models/mlp.py``): a model that owns its own ``fit`` / ``evaluate`` (the ``PyTorchDef`` contract)
next to the framework-owned loop, selected with ``--user_loop``.  The CSV is generated on the fly."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.nn as nn  # noqa: E402

from metisfl_b200.driver.driver_session import DriverSession  # noqa: E402
from metisfl_b200.models.model_dataset import ModelDatasetClassification  # noqa: E402
from metisfl_b200.models.model_def import PyTorchDef  # noqa: E402


class CSVDataset(torch.utils.data.Dataset):
    """Last column is the label, the others are features."""

    def __init__(self, path):
        table = np.loadtxt(path, delimiter=",", dtype=np.float32)
        self.x, self.y = table[:, :-1], table[:, -1].astype(np.int64)

    def __len__(self):
        return len(self.y)

    def __getitem__(self, idx):
        return self.x[idx], self.y[idx]

    def train_test_split(self, n_test=0.1):
        n = int(round(len(self) * n_test))
        return torch.utils.data.random_split(self, [len(self) - n, n], generator=torch.Generator().manual_seed(0))


class MLP(nn.Module):
    def __init__(self, n_inputs=34, hidden=(10, 8), n_classes=2):
        super().__init__()
        dims = [n_inputs] + list(hidden)
        self.body = nn.Sequential(*[m for i in range(len(hidden)) for m in (nn.Linear(dims[i], dims[i + 1]), nn.ReLU())])
        self.out = nn.Linear(dims[-1], n_classes)

    def forward(self, x):
        return self.out(self.body(x))


class SelfTrainingMLP(MLP, PyTorchDef):
    """Owns its loop: the learner calls ``fit(dataset, epochs)`` / ``evaluate(dataset)`` instead of the built-in one."""

    def fit(self, dataset, epochs, *args, **kwargs):
        """``dataset`` iterates ``(x, y)`` mini-batches (the learner wraps the ModelDataset)."""
        opt = torch.optim.SGD(self.parameters(), lr=0.05, momentum=0.9)
        losses = []
        for _ in range(int(epochs)):
            for x, y in dataset:
                opt.zero_grad()
                loss = nn.functional.cross_entropy(self(x), y)
                loss.backward(); opt.step()
            losses.append(float(loss))
        return {"loss": losses}

    def evaluate(self, dataset, *args, **kwargs):
        seen, correct, loss = 0, 0, 0.0
        with torch.no_grad():
            for x, y in dataset:
                out = self(x)
                loss += float(nn.functional.cross_entropy(out, y, reduction="sum"))
                correct += int((out.argmax(1) == y).sum()); seen += len(y)
        return {"loss": loss / max(1, seen), "accuracy": correct / max(1, seen)}


def write_csv(path, n=600, d=34, seed=0):
    rng = np.random.default_rng(seed)
    w = rng.normal(size=d)
    x = rng.normal(size=(n, d)).astype(np.float32)
    y = (x @ w > 0).astype(np.float32)
    np.savetxt(path, np.concatenate([x, y[:, None]], axis=1), delimiter=",", fmt="%.5f")


if __name__ == "__main__":
    here = os.path.dirname(os.path.abspath(__file__))
    ap = argparse.ArgumentParser()
    ap.add_argument("--federation_environment_config_fp",
                    default=os.path.join(here, "..", "config", "fashionmnist_2learners_sync_fedavg.yaml"))
    ap.add_argument("--user_loop", action="store_true")
    ap.add_argument("--launcher", default=None)
    ap.add_argument("--csv", default="/tmp/metis-data/dummy.csv")
    ap.add_argument("--out", default="/tmp/metis/experiment.json")
    args = ap.parse_args()
    os.makedirs(os.path.dirname(args.csv), exist_ok=True)
    write_csv(args.csv)
    csv_path = args.csv

    def _recipe(train):
        def recipe():
            tr, te = CSVDataset(csv_path).train_test_split()
            part = tr if train else te
            x = np.stack([part[i][0] for i in range(len(part))]); y = np.array([part[i][1] for i in range(len(part))])
            return ModelDatasetClassification(x=x, y=y, size=len(y),
                                              examples_per_class={int(c): int((y == c).sum()) for c in np.unique(y)})
        return recipe

    model = SelfTrainingMLP() if args.user_loop else MLP()
    session = DriverSession(args.federation_environment_config_fp, model, _recipe(True), None, _recipe(False),
                            launcher=args.launcher)
    stats = session.run(request_every_secs=1)
    with open(args.out, "w") as f:
        json.dump(stats, f, indent=1)
    print("rounds recorded:", len(stats["federation_runtime_metadata"].get("metadata", [])), "->", args.out)
