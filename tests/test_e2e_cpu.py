"""End-to-end plumbing on CPU: BASELINE.json config #1 (FashionMNIST-shaped MLP, 2 learners,
synchronous FedAvg) through DriverSession -> controller gRPC -> learners; YAML parsing; the engine
on the gloo backend with world size 2."""
import copy
import os
import socket
import subprocess
import sys

import numpy as np
import pytest
import torch
import yaml

from metisfl_b200.driver.driver_session import DriverSession
from metisfl_b200.models.model_dataset import ModelDatasetClassification
from metisfl_b200.models.zoo import FashionMnistMLP
from metisfl_b200.utils.fedenv_parser import FederationEnvironment

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ENV = os.path.join(ROOT, "examples", "config", "fashionmnist_2learners_sync_fedavg.yaml")


def _free_ports(n):
    socks = [socket.socket() for _ in range(n)]
    for s in socks:
        s.bind(("127.0.0.1", 0))
    ports = [s.getsockname()[1] for s in socks]
    for s in socks:
        s.close()
    return ports


def _env(**over):
    with open(ENV) as f:
        d = yaml.safe_load(f)
    fe = d["FederationEnvironment"]
    ports = _free_ports(1 + len(fe["Learners"]))
    fe["Controller"]["GRPCServicer"]["Port"] = ports[0]
    for l, p in zip(fe["Learners"], ports[1:]):
        l["GRPCServicer"]["Port"] = p
    for k, v in over.items():
        node = fe
        *path, last = k.split(".")
        for part in path:
            node = node[part]
        node[last] = v
    return d


def _recipe(n, seed):
    def recipe():
        rng = np.random.default_rng(seed)
        y = rng.integers(0, 10, size=n)
        x = rng.normal(size=(n, 1, 28, 28)).astype(np.float32) + (y[:, None, None, None] / 4.0).astype(np.float32)
        return ModelDatasetClassification(x=x, y=y.astype(np.int64), size=n,
                                          examples_per_class={int(c): int((y == c).sum()) for c in range(10)})
    return recipe


def test_fedenv_parser_reads_the_reference_schema():
    env = FederationEnvironment(ENV)
    assert env.communication_protocol.is_synchronous and not env.communication_protocol.enable_ssl
    assert env.termination_signals.federation_rounds == 3
    assert env.global_model_config.aggregation_rule.aggregation_rule_name == "FedAvg"
    assert env.global_model_config.aggregation_rule.aggregation_rule_scaling_factor == "NumTrainingExamples"
    assert env.local_model_config.optimizer_config.optimizer_pb_kwargs == {
        "learning_rate": 0.05, "name": "VanillaSGD", "l2_reg": 0.0}
    assert env.model_store_config.name == "InMemory" and env.model_store_config.eviction_lineage_length == 1
    assert len(env.learners) == 2 and env.learners[0].grpc_servicer.port == 50152
    d = _env()
    d["FederationEnvironment"]["HomomorphicEncryption"] = {"Scheme": "CKKS", "BatchSize": 4096, "ScalingFactorBits": 52}
    with pytest.raises(AssertionError):  # HE requires the PWA rule (fedenv_parser.py:302-309)
        FederationEnvironment(d)
    d["FederationEnvironment"]["GlobalModelConfig"]["AggregationRule"]["Name"] = "PWA"
    assert FederationEnvironment(d).homomorphic_encryption.batch_size == 4096


def _check_statistics(stats, n_learners=2, min_rounds=3):
    assert set(stats) == {"learners_descriptor", "learners_models_results", "federation_runtime_metadata",
                          "community_model_results"}
    rounds = stats["federation_runtime_metadata"]["metadata"]
    assert len(rounds) >= min_rounds
    # learners join at their own pace: early rounds may have run with a smaller federation
    assert max(len(r.get("completed_by_learner_id", [])) for r in rounds) == n_learners
    assert len(stats["learners_descriptor"]["learner"]) == n_learners
    evals = stats["community_model_results"]["community_evaluation"]
    assert len(evals) >= min_rounds - 1
    assert max(len(e.get("evaluations", {})) for e in evals) == n_learners
    first = list(evals[0]["evaluations"].values())[0]
    assert "accuracy" in first["test_evaluation"]["metric_values"]
    tasks = list(stats["learners_models_results"]["learner_task"].values())[0]["task_metadata"]
    assert tasks[0]["completed_batches"] > 0 and tasks[0]["processing_ms_per_batch"] > 0


def test_config1_inprocess(tmp_path):
    torch.manual_seed(0)
    s = DriverSession(FederationEnvironment(_env()), FashionMnistMLP(), _recipe(320, 1), _recipe(64, 2),
                      _recipe(64, 3), working_dir=str(tmp_path / "w"), launcher="inprocess", banner=False)
    stats = s.run(request_every_secs=0.5)
    _check_statistics(stats)
    losses = [float(list(e["evaluations"].values())[0]["test_evaluation"]["metric_values"]["loss"])
              for e in stats["community_model_results"]["community_evaluation"] if e.get("evaluations")]
    assert min(losses[1:]) < losses[0]  # the federation actually learns


@pytest.mark.parametrize("rule,protocol", [("FedRec", "Asynchronous"), ("FedAsync", "Asynchronous"),
                                           ("FedStride", "SEMI_SYNCHRONOUS")])
def test_other_protocols_inprocess(tmp_path, rule, protocol):
    over = {"GlobalModelConfig.AggregationRule.Name": rule, "CommunicationProtocol.Name": protocol,
            "ModelStoreConfig.LineageLength": 2}
    d = _env(**over)
    if protocol == "SEMI_SYNCHRONOUS":
        d["FederationEnvironment"]["CommunicationProtocol"]["Specifications"] = {
            "SemiSynchronousLambda": 1, "SemiSynchronousRecomputeSteps": True}
        d["FederationEnvironment"]["GlobalModelConfig"]["AggregationRule"]["RuleSpecifications"]["StrideLength"] = 1
    s = DriverSession(FederationEnvironment(d), FashionMnistMLP(), _recipe(160, 1), None, _recipe(64, 3),
                      working_dir=str(tmp_path / "w"), launcher="inprocess", banner=False)
    stats = s.run(request_every_secs=0.5)
    assert len(stats["federation_runtime_metadata"]["metadata"]) >= 3
    assert stats["community_model_results"]["community_evaluation"]


def test_config1_secure_aggregation_inprocess(tmp_path):
    """PWA + CKKS end to end: encrypted initial model, learners decrypt / train / re-encrypt."""
    d = _env(**{"GlobalModelConfig.AggregationRule.Name": "PWA", "TerminationSignals.FederationRounds": 2,
                "LocalModelConfig.OptimizerConfig.LearningRate": 0.0005})
    d["FederationEnvironment"]["HomomorphicEncryption"] = {"Scheme": "CKKS", "BatchSize": 4096, "ScalingFactorBits": 52}
    model = torch.nn.Sequential(torch.nn.Flatten(), torch.nn.Linear(784, 10))
    s = DriverSession(FederationEnvironment(d), model, _recipe(96, 1), None, _recipe(64, 3),
                      working_dir=str(tmp_path / "w"), launcher="inprocess", banner=False)
    stats = s.run(request_every_secs=0.5)
    assert len(stats["federation_runtime_metadata"]["metadata"]) >= 2
    ev = [e for e in stats["community_model_results"]["community_evaluation"] if e.get("evaluations")]
    loss = float(list(ev[-1]["evaluations"].values())[0]["test_evaluation"]["metric_values"]["loss"])
    assert np.isfinite(loss) and loss < 5.0


@pytest.mark.multiprocess
def test_config1_local_subprocess_launcher(tmp_path):
    """Controller and learners as real child processes (`python -m metisfl_b200.controller|learner`)."""
    s = DriverSession(FederationEnvironment(_env()), FashionMnistMLP(), _recipe(160, 1), None, _recipe(64, 3),
                      working_dir=str(tmp_path / "w"), launcher="local", banner=False)
    stats = s.run(request_every_secs=1)
    _check_statistics(stats)


_GLOO_WORKER = r"""
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, {root!r})
from metisfl_b200.parallel.engine import EngineConfig, FederationEngine
from metisfl_b200.models.zoo import FashionMnistMLP
from metisfl_b200.ops.optim import OptimHyper
dist.init_process_group("gloo")
rank = dist.get_rank()
cfg = EngineConfig(learners_per_gpu=2, model_dtype=torch.float32, optimizer=OptimHyper("VanillaSGD", lr=0.1))
eng = FederationEngine(FashionMnistMLP, cfg, device="cpu", examples_per_learner=[100, 200, 300, 400])
g = torch.Generator().manual_seed(rank)
data = [[(torch.randn(16, 1, 28, 28, generator=g), torch.randint(0, 10, (16,), generator=g)) for _ in range(2)] for _ in range(2)]
for r in range(2):
    out = eng.run_round(data)
sd = eng.community_state_dict()
flat = torch.cat([v.float().flatten() for v in sd.values()])
gathered = [torch.empty_like(flat) for _ in range(2)]
dist.all_gather(gathered, flat)
assert torch.equal(gathered[0], gathered[1]), "ranks disagree on the community model"
if rank == 0:
    print("GLOO_OK", float(flat.abs().sum()), out["loss"])
dist.destroy_process_group()
"""


def test_engine_world_size_2_gloo(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(_GLOO_WORKER.format(root=ROOT))
    port = _free_ports(1)[0]
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", str(port), str(script)],
                       capture_output=True, text=True, timeout=300)
    assert "GLOO_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-3000:]


def test_echo_learners_drive_the_controller_without_training(tmp_path):
    """The reference's fake backend (``test/learner_notrain_noeval.py``): learners sleep and echo the model, so the
    rounds measure only the controller path; the community model stays equal to the initial one under FedAvg."""
    torch.manual_seed(0)
    model = FashionMnistMLP()
    want = {k: v.clone() for k, v in model.state_dict().items()}
    d = _env()
    d["FederationEnvironment"]["TerminationSignals"]["FederationRounds"] = 3
    s = DriverSession(FederationEnvironment(d), model, _recipe(16, 1), None, None, working_dir=str(tmp_path / "w"),
                      launcher="inprocess", banner=False, neural_engine="echo", echo_sleep_secs=0.2)
    stats = s.run(request_every_secs=0.3)
    md = stats["federation_runtime_metadata"]["metadata"]
    assert len(md) >= 3 and all(len(m["completed_by_learner_id"]) == 2 for m in md[:3])
    tasks = stats["learners_models_results"]["learner_task"]
    assert len(tasks) == 2


@pytest.mark.multiprocess
@pytest.mark.timeout(300)
def test_ssh_launcher_over_a_local_exec_transport(tmp_path, monkeypatch):
    """`Launcher: ssh` (the reference's only launcher, driver_session.py:225-262) end to end: the driver builds the
    remote commands, ships the model archive and the dataset recipes, and the controller / learners it starts that way
    run a federation.  There is no sshd in this image, so `fabric.Connection` is replaced by a transport that runs the
    very same command lines with `bash -c` and copies the very same files - everything above the socket is exercised."""
    import shutil
    import fabric

    calls = {"run": [], "put": [], "hosts": set()}

    class LocalExecConnection:
        def __init__(self, host=None, port=None, user=None, connect_kwargs=None, **kw):
            calls["hosts"].add((host, user))
            self.env = dict(os.environ)
            self.env["PYTHONPATH"] = ROOT + os.pathsep + self.env.get("PYTHONPATH", "")

        def run(self, cmd, **kw):
            calls["run"].append(cmd)
            return subprocess.run(["bash", "-c", cmd], env=self.env, stdout=subprocess.DEVNULL,
                                  stderr=subprocess.DEVNULL, check=False)

        def put(self, local, remote):
            calls["put"].append((os.path.basename(local), remote))
            shutil.copy(local, remote)

        def close(self):
            pass

    monkeypatch.setattr(fabric, "Connection", LocalExecConnection)
    d = _env(**{"TerminationSignals.FederationRounds": 2, "TerminationSignals.ExecutionCutoffTimeMins": 3})
    d["FederationEnvironment"]["Learners"] = d["FederationEnvironment"]["Learners"][:1]
    s = DriverSession(FederationEnvironment(d), FashionMnistMLP(), _recipe(96, 1), None, _recipe(32, 3),
                      working_dir=str(tmp_path / "w"), launcher="ssh", banner=False)
    stats = s.run(request_every_secs=0.5)
    assert len(stats["federation_runtime_metadata"]["metadata"]) >= 2
    assert stats["community_model_results"]["community_evaluation"]
    # one remote command for the controller, mkdir + one command for the learner; the learner got its files
    assert any("metisfl" in c and "controller" in c for c in calls["run"])
    learner_cmds = [c for c in calls["run"] if "learner" in c and "tar -xzf" in c]
    assert len(learner_cmds) == 1 and "mkdir -p /tmp/metis/workdir_learner_" in learner_cmds[0]
    shipped = [name for name, _ in calls["put"]]
    assert any(name.endswith(".tar.gz") for name in shipped) and len(shipped) >= 3   # model archive + 2 recipes
    assert calls["hosts"] == {("localhost", "root")}


def test_participation_ratio_through_the_driver(tmp_path):
    """`GlobalModelConfig.ParticipationRatio: 0.5` with two learners: after the first round (every joiner holds the
    initial task) each round trains and aggregates ONE sampled learner (the reference parses the field and never reads
    it, `fedenv_parser.py` / `metis.proto:307`, quirk Q7)."""
    d = _env(**{"GlobalModelConfig.ParticipationRatio": 0.5, "TerminationSignals.FederationRounds": 5})
    s = DriverSession(FederationEnvironment(d), FashionMnistMLP(), _recipe(96, 1), None, _recipe(32, 3),
                      working_dir=str(tmp_path / "w"), launcher="inprocess", banner=False)
    stats = s.run(request_every_secs=0.3)
    md = stats["federation_runtime_metadata"]["metadata"]
    closed = [m for m in md[1:] if m.get("model_aggregation_block_size")]
    assert len(closed) >= 3
    # (a learner that joins while a round is open gets the initial task and widens that round: only the rounds after
    # both have joined are required to be cohorts of one)
    ones = [m for m in closed if len(m["assigned_to_learner_id"]) == 1 and m["model_aggregation_block_size"] == [1.0]]
    assert len(ones) >= 3 and all(m in ones for m in closed[-2:])
    # a cohort of one publishes that learner's model (not example-count x model): the community model stays sane
    ev = [e for e in stats["community_model_results"]["community_evaluation"] if e.get("evaluations")]
    losses = [float(list(e["evaluations"].values())[0]["test_evaluation"]["metric_values"]["loss"]) for e in ev]
    assert losses and all(np.isfinite(l) and l < 10 for l in losses), losses


def test_validation_percentage_holds_out_a_stratified_split(tmp_path):
    """`LocalModelConfig.ValidationPercentage: 0.25` with learners that bring no validation set: each learner trains on
    75 % of its examples and reports the held-out 25 % (per class) as its validation evaluation (the reference sends the
    fraction with every task, `controller.cc:730-731`, and no learner reads it - quirk Q7)."""
    d = _env(**{"LocalModelConfig.ValidationPercentage": 0.25, "TerminationSignals.FederationRounds": 2})
    s = DriverSession(FederationEnvironment(d), FashionMnistMLP(), _recipe(96, 1), None, _recipe(32, 3),
                      working_dir=str(tmp_path / "w"), launcher="inprocess", banner=False)
    stats = s.run(request_every_secs=0.3)
    ev = [e for e in stats["community_model_results"]["community_evaluation"] if e.get("evaluations")]
    assert ev
    per_learner = list(ev[-1]["evaluations"].values())
    assert all(p["validation_evaluation"]["metric_values"].get("loss") for p in per_learner)
    learners = [svc.learner for name, svc in s._inprocess if name.startswith("learner")]
    assert len(learners) == 2
    for l in learners:
        train, valid, _ = l._datasets
        assert train.get_size() + valid.get_size() == 96 and 20 <= valid.get_size() <= 28
        assert set(valid.examples_per_class) <= set(train.examples_per_class)


def test_holdout_split_unit():
    from metisfl_b200.learner.learner import Learner
    from metisfl_b200.models.model_dataset import ModelDatasetClassification
    rng = np.random.default_rng(0)
    y = np.repeat(np.arange(4), [40, 20, 12, 8])
    x = rng.normal(size=(80, 3)).astype(np.float32)
    full = ModelDatasetClassification(x=x, y=y, size=80, examples_per_class={0: 40, 1: 20, 2: 12, 3: 8})
    empty = ModelDatasetClassification()
    l = Learner.__new__(Learner)
    l._datasets, l._holdout_done = (full, empty, empty), False
    l._apply_validation_holdout(25)          # percent form
    train, valid, _ = l._datasets
    assert valid.examples_per_class == {0: 10, 1: 5, 2: 3, 3: 2} and train.get_size() == 60
    assert train.examples_per_class == {0: 30, 1: 15, 2: 9, 3: 6}
    both = np.concatenate([train.get_x(), valid.get_x()])
    assert sorted(map(tuple, both.tolist())) == sorted(map(tuple, x.tolist()))   # a partition of the examples
    l._apply_validation_holdout(0.5)         # once per learner
    assert l._datasets[0] is train
    # a learner that brought its own validation set keeps it
    l2 = Learner.__new__(Learner)
    l2._datasets, l2._holdout_done = (full, full, empty), False
    l2._apply_validation_holdout(0.25)
    assert l2._datasets[0] is full


def test_metric_cutoff_ends_the_federation(tmp_path):
    """`TerminationSignals.MetricCutoffScore`: the driver stops as soon as the learners' mean test metric reaches it
    (`driver_session.py:423-480`); here the bar is 0, so the first filed evaluation ends the run."""
    d = _env(**{"TerminationSignals.FederationRounds": 500, "TerminationSignals.MetricCutoffScore": 0.0,
                "TerminationSignals.ExecutionCutoffTimeMins": 3})
    s = DriverSession(FederationEnvironment(d), FashionMnistMLP(), _recipe(96, 1), None, _recipe(32, 3),
                      working_dir=str(tmp_path / "w"), launcher="inprocess", banner=False)
    import time
    t0 = time.time()
    stats = s.run(request_every_secs=0.3)
    assert time.time() - t0 < 60
    md = stats["federation_runtime_metadata"]["metadata"]
    assert 1 <= max(m["global_iteration"] for m in md) < 500
    assert [e for e in stats["community_model_results"]["community_evaluation"] if e.get("evaluations")]
