"""Neuro-imaging federations: brain-age regression or Alzheimer's-disease classification with 2-D / 3-D CNNs.

Counterpart of reference ``examples/keras/neuroimaging.py:131-362``: a CSV table maps every subject's scan file to
its label; ``MRIScanGen`` parses the table and packs the volumes into one record file per learner (``.npz`` here,
TFRecords there); ``ModelDatasetHelper`` turns a record file into a regression or classification ``ModelDataset``
inside the learner.  Scans are synthetic (no data in this environment) and small by default; ``--full_volumes``
uses the 91 x 109 x 91 MNI grid of the reference models.
"""
import argparse
import csv
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402

from metisfl_b200.driver.driver_session import DriverSession  # noqa: E402
from metisfl_b200.models.model_dataset import ModelDatasetClassification, ModelDatasetRegression  # noqa: E402
from metisfl_b200.models.zoo import (AlzheimersDisease2DCNN, AlzheimersDisease3DCNN, BrainAge2DCNN,  # noqa: E402
                                     BrainAge3DCNN)
from metisfl_b200.utils.fedenv_parser import FederationEnvironment  # noqa: E402


class MRIScanGen(object):
    """CSV (scan path, label) -> one packed record file."""

    def __init__(self, filepath, image_column, label_column, rescale=True):
        self.filepath, self.image_column, self.label_column, self.rescale = filepath, image_column, label_column, rescale
        self.records_fp = os.path.splitext(filepath)[0] + ".records.npz"

    def parse_csv_table(self):
        with open(self.filepath) as f:
            rows = list(csv.DictReader(f))
        return [r[self.image_column] for r in rows], [float(r[self.label_column]) for r in rows]

    def load_scan(self, scan_path):
        vol = np.load(scan_path).astype(np.float32)
        return vol / max(1e-6, float(vol.max())) if self.rescale else vol

    def generate_records(self):
        paths, labels = self.parse_csv_table()
        np.savez(self.records_fp, images=np.stack([self.load_scan(p) for p in paths]),
                 labels=np.asarray(labels, dtype=np.float32))
        return self.records_fp


class ModelDatasetHelper(object):
    def __init__(self, image_column="scan_path", label_column="label"):
        self.image_column, self.label_column = image_column, label_column

    def _load(self, dataset_fp):
        blob = np.load(MRIScanGen(dataset_fp, self.image_column, self.label_column).generate_records())
        return blob["images"], blob["labels"]

    def regression_task_dataset_recipe_fn(self, dataset_fp):
        x, y = self._load(dataset_fp)
        return ModelDatasetRegression(x=x, y=y, size=len(y), min_val=float(y.min()), max_val=float(y.max()),
                                      mean_val=float(y.mean()), median_val=float(np.median(y)),
                                      mode_val=float(np.round(y).astype(int)[0]), stddev_val=float(y.std()))

    def classification_task_dataset_recipe_fn(self, dataset_fp):
        x, y = self._load(dataset_fp)
        classes = y.astype(int)
        examples_per_class = {int(c): int((classes == c).sum()) for c in set(classes.tolist())}
        return ModelDatasetClassification(x=x, y=y, size=len(y), examples_per_class=examples_per_class)


def write_dummy_scans(out_dir, n, shape, task, seed):
    """Synthetic subjects: the label modulates the volume's mean intensity so the model has something to learn."""
    os.makedirs(out_dir, exist_ok=True)
    rng = np.random.default_rng(seed)
    table = os.path.join(out_dir, "scans.csv")
    with open(table, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["scan_path", "label"])
        for i in range(n):
            label = float(rng.uniform(45, 80)) if task == "brainage" else float(rng.integers(0, 2))
            shift = (label - 62.0) / 40.0 if task == "brainage" else label - 0.5
            vol = np.abs(rng.normal(size=shape) + 2.0 + shift).astype(np.float32)
            path = os.path.join(out_dir, "subject_{:04d}.npy".format(i))
            np.save(path, vol)
            w.writerow([path, label])
    return table


if __name__ == "__main__":
    here = os.path.dirname(os.path.abspath(__file__))
    ap = argparse.ArgumentParser()
    ap.add_argument("--neuroimaging_task", default="brainage", choices=["brainage", "alzheimers"])
    ap.add_argument("--neuroimaging_task_model", default="3dcnn", choices=["2dcnn", "3dcnn"])
    ap.add_argument("--federation_environment_config_fp", default=None)
    ap.add_argument("--dummy_scans_csv_mapping_fp", default=None,
                    help="existing CSV of (scan_path,label); by default synthetic scans are generated per learner")
    ap.add_argument("--full_volumes", action="store_true")
    ap.add_argument("--subjects_per_learner", type=int, default=8)
    ap.add_argument("--launcher", default=None)
    ap.add_argument("--data_dir", default="/tmp/metis-data/neuroimaging")
    ap.add_argument("--out", default="/tmp/metis/experiment.json")
    args = ap.parse_args()

    task, arch = args.neuroimaging_task, args.neuroimaging_task_model
    cfg = args.federation_environment_config_fp or os.path.join(
        here, "config", "brainage" if task == "brainage" else "alzheimers_disease",
        ("brainage" if task == "brainage" else "alzheimers_disease") + "_test_localhost_synchronous.yaml")
    env = FederationEnvironment(cfg)
    shape = (91, 109, 91) if args.full_volumes else ((46, 46, 46) if (task, arch) == ("alzheimers", "3dcnn") else (32, 40, 32))
    for i, learner in enumerate(env.learners.learners):
        table = args.dummy_scans_csv_mapping_fp or write_dummy_scans(
            os.path.join(args.data_dir, "learner_{}".format(i)), args.subjects_per_learner, shape, task, seed=i)
        learner.dataset_configs.train_dataset_path = table
        learner.dataset_configs.test_dataset_path = table
    small = not args.full_volumes
    if task == "brainage":
        model = (BrainAge2DCNN((8, 16, 16, 32, 32)) if small else BrainAge2DCNN()) if arch == "2dcnn" else \
            (BrainAge3DCNN((8, 8, 16, 16, 16)) if small else BrainAge3DCNN())
        recipe = ModelDatasetHelper().regression_task_dataset_recipe_fn
    else:
        model = (AlzheimersDisease2DCNN((8, 16, 16, 32, 32)) if small else AlzheimersDisease2DCNN()) if arch == "2dcnn" else \
            (AlzheimersDisease3DCNN((8, 8, 16, 16)) if small else AlzheimersDisease3DCNN())
        recipe = ModelDatasetHelper().classification_task_dataset_recipe_fn
    session = DriverSession(env, model, recipe, None, recipe, launcher=args.launcher)
    stats = session.run(request_every_secs=1)
    with open(args.out, "w") as f:
        json.dump(stats, f, indent=1)
    print("rounds recorded:", len(stats["federation_runtime_metadata"].get("metadata", [])), "->", args.out)
