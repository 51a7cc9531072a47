"""The device-resident federation behind the public API: DriverSession -> controller gRPC -> learners that host a
fabric rank.  After the initial ship of the community model no model bytes reach the controller: MarkTaskCompleted
carries device-slot references, the controller commands the aggregation (AggregateOnDevice) and records the
device-reported durations and quantifier counts (reference flow it replaces:
``controller.cc:201-259,795-950``, ``learner.py:197-206``).  CPU tier: the same control path over the engine's CPU
fallbacks (world 1 in-process, world 2 as gloo subprocesses); the GPU tier runs it on the kernels."""
import os

import numpy as np
import pytest
import torch

from metisfl_b200.driver.driver_session import DriverSession
from metisfl_b200.models.zoo import FashionMnistMLP
from metisfl_b200.proto import model_pb2
from metisfl_b200.utils.fedenv_parser import FederationEnvironment
from test_e2e_cpu import _env, _recipe


def _final_model(snapshot_dir):
    import json
    with open(os.path.join(snapshot_dir, "state.json")) as f:
        state = json.load(f)
    with open(os.path.join(snapshot_dir, state["community_model_file"]), "rb") as f:
        fm = model_pb2.FederatedModel.FromString(f.read())
    out = {}
    for v in fm.model.variables:
        spec = v.plaintext_tensor.tensor_spec
        dt = {8: np.float32, 9: np.float64, 3: np.int64, 11: np.uint16}[spec.type.type]
        out[v.name] = np.frombuffer(spec.value, dtype=dt).copy()
    return out, fm.global_iteration


def _run(tmp_path, tag, fabric, monkeypatch, n_learners=1, launcher="inprocess", rule="FedAvg", rounds=3,
         protocol=None):
    d = _env(**{"GlobalModelConfig.AggregationRule.Name": rule, "TerminationSignals.FederationRounds": rounds})
    fe = d["FederationEnvironment"]
    fe["Learners"] = fe["Learners"][:n_learners]
    fe["DeviceFabric"] = fabric
    fe["ModelDType"] = "fp32"
    if protocol:
        fe["CommunicationProtocol"]["Name"] = protocol
        fe["ModelStoreConfig"]["LineageLength"] = 2
    snap = tmp_path / f"snap_{tag}"
    monkeypatch.setenv("METISFL_SNAPSHOT_DIR", str(snap))
    torch.manual_seed(0)
    s = DriverSession(FederationEnvironment(d), FashionMnistMLP(), _recipe(128, 1), None, _recipe(64, 3),
                      working_dir=str(tmp_path / f"w_{tag}"), launcher=launcher, banner=False)
    stats = s.run(request_every_secs=0.3)
    return s, stats, str(snap)


def _counters(stats):
    import json
    return json.loads(stats["federation_runtime_metadata"]["json_metadata"])


@pytest.mark.timeout(240)
def test_single_learner_device_path_matches_host_path(tmp_path, monkeypatch):
    s_dev, stats_dev, snap_dev = _run(tmp_path, "dev", True, monkeypatch)
    md = stats_dev["federation_runtime_metadata"]["metadata"]
    assert len(md) >= 3
    done = [m for m in md if m.get("model_aggregation_block_duration_ms")]
    assert len(done) >= 3
    # quantifier counts come from the device kernel (K4), one per model variable
    assert len(done[0]["model_tensor_quantifiers"]) == len(list(FashionMnistMLP().state_dict()))
    ev = [e for e in stats_dev["community_model_results"]["community_evaluation"] if e.get("evaluations")]
    assert ev and "accuracy" in list(ev[0]["evaluations"].values())[0]["test_evaluation"]["metric_values"]
    tasks = list(stats_dev["learners_models_results"]["learner_task"].values())[0]["task_metadata"]
    assert tasks[0]["completed_batches"] > 0 and tasks[0]["processing_ms_per_batch"] > 0
    # nothing but references came back from the learner
    c = _counters(stats_dev)
    assert c["model_bytes_received"] == 0 and c["aggregation_durations_device_timed"] and c["fabric_learners"] == 1
    assert c["aggregation_failures"] == 0
    s_host, stats_host, snap_host = _run(tmp_path, "host", False, monkeypatch)
    assert _counters(stats_host)["model_bytes_received"] > 100_000
    a, it_a = _final_model(snap_dev)
    b, it_b = _final_model(snap_host)
    assert set(a) == set(b)
    # both ran at least 3 rounds from the same seed; compare a model of the same round when they stopped together
    if it_a == it_b:
        for k in a:
            np.testing.assert_allclose(a[k], b[k], atol=2e-5, rtol=1e-4)


@pytest.mark.timeout(240)
@pytest.mark.parametrize("rule,protocol", [("FedAvg", None), ("FedStride", None), ("FedAsync", "Asynchronous"),
                                           ("FedRec", "Asynchronous")])
def test_two_gloo_learner_processes_share_a_fabric(tmp_path, monkeypatch, rule, protocol):
    """Controller + 2 learner processes (`Launcher: local`), world-2 fabric over gloo: the control path of every rule."""
    s, stats, snap = _run(tmp_path, "w2", True, monkeypatch, n_learners=2, launcher="local", rule=rule,
                          protocol=protocol, rounds=3)
    md = stats["federation_runtime_metadata"]["metadata"]
    agg = [m for m in md if m.get("model_aggregation_block_duration_ms")]
    assert len(agg) >= 3
    model, it = _final_model(snap)
    assert it >= 2 and all(np.isfinite(v).all() for v in model.values())
    ev = [e for e in stats["community_model_results"]["community_evaluation"] if e.get("evaluations")]
    assert ev
    c = _counters(stats)
    assert c["model_bytes_received"] == 0 and c["fabric_learners"] == 2 and c["aggregation_failures"] == 0
    for name in ("learner0", "learner1"):
        log = open(os.path.join(str(tmp_path / "w_w2"), name + ".log"), errors="replace").read()
        assert "Traceback" not in log, log[-3000:]
