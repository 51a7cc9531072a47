"""CIFAR-10-shaped federation through the full Driver / Controller / Learner stack.

Mirrors the flow of the reference example (``examples/keras/cifar10.py:22-111``): partition the training set
IID or non-IID over the learners of the YAML environment, hand one dataset recipe per learner to the driver,
run, dump the statistics JSON.  Data is synthetic CIFAR-shaped (no download here); ``--model`` picks the
reference's ``CifarCNN`` (1.6 M parameters) or the flagship ``ResNet18``.
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402

from examples.utils.data_partitioning import DataPartitioning  # noqa: E402
from metisfl_b200.driver.driver_session import DriverSession  # noqa: E402
from metisfl_b200.models.model_dataset import ModelDatasetClassification  # noqa: E402
from metisfl_b200.models.zoo import CifarCNN, ResNet18  # noqa: E402
from metisfl_b200.utils.fedenv_parser import FederationEnvironment  # noqa: E402


def synthetic_cifar(n, seed):
    rng = np.random.default_rng(seed)
    y = rng.integers(0, 10, size=n)
    x = rng.normal(size=(n, 3, 32, 32)).astype(np.float32)
    x += (y[:, None, None, None] - 4.5).astype(np.float32) / 6.0  # learnable class signal
    return x, y.astype(np.int64)


def dataset_recipe_fn(dataset_fp):
    """Runs inside the learner: one ``.npz`` shard per learner, its path comes from the YAML ``DatasetConfigs``."""
    blob = np.load(dataset_fp)
    x, y = blob["x"], blob["y"]
    counts = {int(c): int((y == c).sum()) for c in np.unique(y)}
    return ModelDatasetClassification(x=x, y=y, size=len(y), examples_per_class=counts)


def save_shards(x_chunks, y_chunks, out_dir, prefix):
    os.makedirs(out_dir, exist_ok=True)
    paths = []
    for i, (xc, yc) in enumerate(zip(x_chunks, y_chunks)):
        paths.append(os.path.join(out_dir, "{}_{}.npz".format(prefix, i)))
        np.savez(paths[-1], x=xc, y=yc)
    return paths


if __name__ == "__main__":
    here = os.path.dirname(os.path.abspath(__file__))
    ap = argparse.ArgumentParser()
    ap.add_argument("--federation_environment_config_fp",
                    default=os.path.join(here, "config", "cifar10", "test_localhost_synchronous_momentumsgd.yaml"))
    ap.add_argument("--generate_iid_partitions", action="store_true")
    ap.add_argument("--generate_noniid_partitions", action="store_true")
    ap.add_argument("--classes_per_partition", type=int, default=2)
    ap.add_argument("--model", default="cifar_cnn", choices=["cifar_cnn", "resnet18"])
    ap.add_argument("--train_examples", type=int, default=2048)
    ap.add_argument("--launcher", default=None)
    ap.add_argument("--data_dir", default="/tmp/metis-data/cifar10")
    ap.add_argument("--out", default="/tmp/metis/experiment.json")
    args = ap.parse_args()

    env = FederationEnvironment(args.federation_environment_config_fp)
    n_learners = len(env.learners.learners)
    x_train, y_train = synthetic_cifar(args.train_examples, 1)
    x_test, y_test = synthetic_cifar(512, 2)
    parts = DataPartitioning(x_train, y_train, n_learners)
    if args.generate_noniid_partitions:
        x_chunks, y_chunks = parts.non_iid_partition(classes_per_partition=args.classes_per_partition)
    else:
        x_chunks, y_chunks = parts.iid_partition()
    print(json.dumps({"partitions": parts.to_json_representation()})[:400])
    train_paths = save_shards(x_chunks, y_chunks, args.data_dir, "train")
    test_path = save_shards([x_test], [y_test], args.data_dir, "test")[0]
    for learner, path in zip(env.learners.learners, train_paths):
        learner.dataset_configs.train_dataset_path = path
        learner.dataset_configs.test_dataset_path = test_path
    model = CifarCNN() if args.model == "cifar_cnn" else ResNet18()
    session = DriverSession(env, model, dataset_recipe_fn, None, dataset_recipe_fn, launcher=args.launcher)
    stats = session.run(request_every_secs=1)
    with open(args.out, "w") as f:
        json.dump(stats, f, indent=1)
    print("rounds recorded:", len(stats["federation_runtime_metadata"].get("metadata", [])), "->", args.out)
