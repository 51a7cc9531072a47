"""Example-side utilities: partitioning -> per-learner shard recipes, MRI record packing, convergence series.
Reference: ``examples/utils/{data_partitioning,convergence_plots}.py``, ``examples/keras/{cifar10,neuroimaging}.py``."""
import glob
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def test_every_shipped_yaml_environment_parses():
    from metisfl_b200.utils.fedenv_parser import FederationEnvironment
    files = glob.glob(os.path.join(ROOT, "examples", "config", "**", "*.yaml"), recursive=True)
    assert len(files) >= 16  # the reference ships 16 environments
    for f in files:
        env = FederationEnvironment(f)
        assert env.learners.learners and env.controller.grpc_servicer.port


def test_cifar_shards_round_trip_through_recipe(tmp_path):
    from examples import cifar10
    from examples.utils.data_partitioning import DataPartitioning
    x, y = cifar10.synthetic_cifar(200, 0)
    xs, ys = DataPartitioning(x, y, 4).non_iid_partition(classes_per_partition=2)
    paths = cifar10.save_shards(xs, ys, str(tmp_path), "train")
    sizes = 0
    for p in paths:
        ds = cifar10.dataset_recipe_fn(p)
        sizes += ds.get_size()
        assert sum(ds.get_model_dataset_specifications().values()) == ds.get_size()
    assert sizes == 200


def test_mri_records_and_recipes(tmp_path):
    from examples import neuroimaging as ni
    table = ni.write_dummy_scans(str(tmp_path), 5, (8, 10, 8), "brainage", seed=1)
    ds = ni.ModelDatasetHelper().regression_task_dataset_recipe_fn(table)
    assert ds.get_x().shape == (5, 8, 10, 8) and 45 <= ds.get_model_dataset_specifications()["mean"] <= 80
    table = ni.write_dummy_scans(str(tmp_path / "ad"), 6, (8, 10, 8), "alzheimers", seed=2)
    ds = ni.ModelDatasetHelper().classification_task_dataset_recipe_fn(table)
    assert sum(ds.get_model_dataset_specifications().values()) == 6


def test_convergence_series_and_csv_fallback(tmp_path):
    import json
    from examples.utils import convergence_plots as cp
    assert abs(cp.convert_to_timestamp("1970-01-01T00:00:01.500000000Z") - 1.5) < 1e-6
    exp = {"federation_runtime_metadata": {"metadata": [
        {"global_iteration": 1, "started_at": "2026-01-01T00:00:00Z", "completed_at": "2026-01-01T00:00:02Z"},
        {"global_iteration": 2, "started_at": "2026-01-01T00:00:02Z", "completed_at": "2026-01-01T00:00:05Z"}]},
        "community_model_results": {"community_evaluation": [
            {"global_iteration": 1, "evaluations": {"a": {"test_evaluation": {"metric_values": {"accuracy": "0.5"}}},
                                                    "b": {"test_evaluation": {"metric_values": {"accuracy": "0.7"}}}}},
            {"global_iteration": 2, "evaluations": {"a": {"test_evaluation": {"metric_values": {"accuracy": "0.9"}}}}}]},
        "learners_descriptor": {"learner": [{}, {}]}}
    s = cp.community_series(exp)
    assert [(i, round(v, 3), t) for i, v, t in s] == [(1, 0.6, 2.0), (2, 0.9, 5.0)]
    assert [r[:2] for r in cp.round_durations(exp)] == [(1, 2.0), (2, 3.0)]
    f = tmp_path / "e.json"; f.write_text(json.dumps(exp))
    out = cp.plot_brainage_convergence([str(f)], metric="accuracy", show_global_models_exchanged=True,
                                       out=str(tmp_path / "fig.png"))
    assert os.path.exists(out)
