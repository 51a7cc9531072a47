"""Federation state machine of the native core, driven with in-process echo learners (the
reference's manual load-test learner, ``test/learner_notrain_noeval.py:16-198``, turned into an
automated protocol test)."""
import threading
import time

import numpy as np
import pytest

from metisfl_b200 import _core
from helpers import model_arrays, model_bytes


class Federation:
    """Controller + N fake learners that answer RunTask by adding ``delta`` to the model."""

    def __init__(self, rule="FedAvg", protocol=1, n=3, examples=None, stride=0, deadline=0.0, lineage=None,
                 semi_lambda=0, auto=True, participation=1.0, sampling_seed=0):
        p = _core.ControllerParams()
        p.rule, p.protocol, p.stride_length = rule, protocol, stride
        p.participation_ratio, p.sampling_seed = participation, sampling_seed
        p.batch_size, p.epochs = 10, 2
        p.task_deadline_secs = deadline
        p.semi_sync_lambda = semi_lambda
        if lineage:
            p.lineage_length_eviction, p.lineage_length = True, lineage
        self.run_log, self.eval_log = [], []
        self.lock = threading.Lock()
        self.auto = auto
        self.tokens = {}
        self.delta = {}
        self.ms_per_batch = {}
        self.mute = set()
        self.core = _core.Controller(p, self._run, self._eval)
        self.examples = examples or [100] * n
        self.n = n

    def _run(self, items, fm):
        c, it, model = _core.parse_federated_model(fm)
        with self.lock:
            self.run_log.append(([i[0] for i in items], items[0][4], [i[5] for i in items]))
        if not self.auto:
            return
        for (lid, host, port, cert, iteration, updates, pct) in items:
            if lid in self.mute:
                continue
            threading.Thread(target=self._complete, args=(lid, model, iteration, updates), daemon=True).start()

    def _complete(self, lid, model, iteration, updates):
        t0 = time.time()
        while lid not in self.tokens and time.time() - t0 < 5:   # the first task can arrive before join() has returned
            time.sleep(0.002)
        x = model_arrays(model)[0]
        new = model_bytes([x + self.delta.get(lid, 0.0)])
        self.core.learner_completed_task(lid, self.tokens[lid], new, iteration, 1.0, updates, 10,
                                         self.ms_per_batch.get(lid, 1.0) * 10, self.ms_per_batch.get(lid, 1.0), b"")

    def _eval(self, learners, model, batch_size, idx, iteration):
        with self.lock:
            self.eval_log.append(([l[0] for l in learners], idx, iteration))
        for l in learners:
            self.core.evaluation_completed(l[0], idx, b"")

    def release(self):
        """Switch to automatic echo mode and answer the tasks that were issued while it was off."""
        self.auto = True
        c, it, model = _core.parse_federated_model(self.core.community_model())
        with self.lock:
            pending = list(self.run_log)
        for ids, iteration, updates in pending:
            for lid, u in zip(ids, updates):
                threading.Thread(target=self._complete, args=(lid, model, iteration, u), daemon=True).start()

    def join(self, i):
        lid, tok = self.core.add_learner("127.0.0.1", 9000 + i, b"", b"", self.examples[i], 0, 0)
        self.tokens[lid] = tok
        return lid

    def wait_rounds(self, n, timeout=20):
        t0 = time.time()
        while time.time() - t0 < timeout:
            if self.core.global_iteration() > n:
                return
            time.sleep(0.01)
        raise AssertionError(f"only reached iteration {self.core.global_iteration()}")

    def community(self):
        c, it, model = _core.parse_federated_model(self.core.community_model())
        return c, it, model_arrays(model)[0]


def test_join_validation_and_tokens():
    f = Federation(auto=False)
    f.core.replace_community_model(_core.build_federated_model(0, 0, model_bytes([np.zeros(4, np.float32)])))
    lid = f.join(0)
    assert lid == "127.0.0.1:9000" and len(f.tokens[lid]) == 64
    with pytest.raises(_core.ControllerError) as e:
        f.join(0)
    assert e.value.code == int(_core.Status.ALREADY_EXISTS)
    with pytest.raises(_core.ControllerError) as e:
        f.core.add_learner("", 1, b"", b"", 10, 0, 0)
    assert e.value.code == int(_core.Status.INVALID_ARGUMENT)
    with pytest.raises(_core.ControllerError) as e:
        f.core.add_learner("h", 1, b"", b"", 0, 0, 0)
    assert e.value.code == int(_core.Status.INVALID_ARGUMENT)
    with pytest.raises(_core.ControllerError) as e:
        f.core.learner_completed_task(lid, "wrong", b"", 1, 0, 0, 0, 0, 0, b"")
    assert e.value.code == int(_core.Status.UNAUTHENTICATED)
    f.core.wait_idle()
    # task template: epochs * ceil(n_train / batch) = 2 * ceil(100 / 10)
    assert f.run_log[0] == ([lid], 1, [20])
    f.core.remove_learner(lid, f.tokens[lid])
    assert f.core.get_learners() == []
    f.core.shutdown()


def test_learner_joining_before_the_model_waits_for_it():
    f = Federation(auto=False)
    lid = f.join(0)
    f.core.wait_idle()
    assert f.run_log == []
    f.core.replace_community_model(_core.build_federated_model(0, 0, model_bytes([np.ones(4, np.float32)])))
    f.core.wait_idle()
    assert f.run_log and f.run_log[0][0] == [lid]
    f.core.shutdown()


def test_synchronous_fedavg_rounds_and_metadata():
    f = Federation(n=3, examples=[100, 300, 600], auto=False)
    f.core.replace_community_model(_core.build_federated_model(0, 0, model_bytes([np.zeros(8, np.float32)])))
    ids = [f.join(i) for i in range(3)]
    for lid, d in zip(ids, (1.0, 2.0, 3.0)):
        f.delta[lid] = d
    f.core.wait_idle()
    f.release()  # everybody has joined before the first task completes
    f.wait_rounds(3)
    f.mute.update(ids)
    time.sleep(0.1)
    f.core.wait_idle()
    md = f.core.get_runtime_metadata_lineage(0)
    r1 = md[0]
    assert r1["global_iteration"] == 1 and sorted(r1["completed_by_learner_id"]) == sorted(ids)
    assert r1["model_aggregation_block_size"] == [3.0] and len(r1["model_tensor_quantifiers"]) == 1
    assert r1["model_tensor_quantifiers"][0] == (8, 0, 32)
    assert set(r1["train_task_received_at"]) == set(ids) and set(r1["eval_task_received_at"]) == set(ids)
    # community after round 1 = weighted mean of deltas: 0.1*1 + 0.3*2 + 0.6*3 = 2.5; then +2.5 per round
    lineage = f.core.get_community_model_lineage(0)
    vals = [model_arrays(_core.parse_federated_model(b)[2])[0][0] for b in lineage]
    assert len(vals) >= 2 and all(abs(v / 2.5 - round(v / 2.5)) < 1e-4 for v in vals) and max(vals) >= 7.5
    c, it, x = f.community()
    assert c == 3 and abs(x[0] / 2.5 - round(x[0] / 2.5)) < 1e-5
    ev = f.core.get_community_evaluation_lineage(0)
    assert ev[0][0] == 1 and set(ev[0][1]) == set(ids)
    tasks = f.core.get_local_task_lineage(ids[:1], 2)
    assert len(tasks[ids[0]]) == 2
    assert len(f.core.get_runtime_metadata_lineage(2)) == 2
    f.core.shutdown()


def test_fedstride_blocks_give_the_same_average():
    f = Federation(rule="FedStride", stride=2, n=4, auto=False)
    f.core.replace_community_model(_core.build_federated_model(0, 0, model_bytes([np.zeros(4, np.float32)])))
    ids = [f.join(i) for i in range(4)]
    f.core.wait_idle()
    for lid, d in zip(ids, (1.0, 2.0, 3.0, 4.0)):
        f.core.learner_completed_task(lid, f.tokens[lid], model_bytes([np.full(4, d, np.float32)]), 1, 1, 5, 10, 1, 1, b"")
    f.core.wait_idle()
    md = f.core.get_runtime_metadata_lineage(0)[0]
    assert md["model_aggregation_block_size"] == [2.0, 2.0]
    c, it, x = f.community()
    assert c == 4 and it == 1
    np.testing.assert_allclose(x, 2.5, rtol=1e-6)
    f.core.shutdown()


@pytest.mark.parametrize("rule", ["FedRec", "FedAsync"])
def test_asynchronous_rules_fold_the_completing_learner(rule):
    f = Federation(rule=rule, protocol=2, n=2, lineage=2, auto=False)
    f.core.replace_community_model(_core.build_federated_model(0, 0, model_bytes([np.zeros(4, np.float32)])))
    a, b = f.join(0), f.join(1)
    f.core.wait_idle()
    assert sorted(r[0][0] for r in f.run_log) == sorted([a, b])
    # b (lexicographically second) completes first: quirk Q2 would have folded `a` instead
    f.core.learner_completed_task(b, f.tokens[b], model_bytes([np.full(4, 8.0, np.float32)]), 1, 1, 5, 10, 1, 1, b"")
    f.core.wait_idle()
    c, it, x = f.community()
    if rule == "FedRec":
        np.testing.assert_allclose(x, 8.0)
    else:  # C = (1 - a) * 0 + a * 8 with a = 0.5 * s(staleness 0)
        np.testing.assert_allclose(x, 4.0)
    assert f.run_log[-1][0] == [b]          # asynchronous: only the caller is rescheduled
    f.core.learner_completed_task(a, f.tokens[a], model_bytes([np.full(4, 2.0, np.float32)]), 1, 1, 5, 10, 1, 1, b"")
    f.core.wait_idle()
    c, it, x = f.community()
    if rule == "FedRec":
        np.testing.assert_allclose(x, 5.0)  # (100*8 + 100*2) / 200
        f.core.learner_completed_task(b, f.tokens[b], model_bytes([np.full(4, 4.0, np.float32)]), 2, 1, 5, 10, 1, 1, b"")
        f.core.wait_idle()
        np.testing.assert_allclose(f.community()[2], 3.0)  # b's 8 swapped for 4: z stays 200
    else:
        s = _core.FedAsync.staleness(1, 1.0, 0.5, 4.0)  # a started from iteration 1, community is at 2
        np.testing.assert_allclose(x, (1 - 0.5 * s) * 4.0 + 0.5 * s * 2.0, rtol=1e-6)
    f.core.shutdown()


def test_semi_synchronous_resizes_tasks_after_first_round():
    f = Federation(protocol=3, n=2, semi_lambda=2, auto=False)
    f.core.replace_community_model(_core.build_federated_model(0, 0, model_bytes([np.zeros(4, np.float32)])))
    ids = [f.join(i) for i in range(2)]
    f.ms_per_batch = {ids[0]: 1.0, ids[1]: 4.0}   # ms/epoch = 10x: slowest epoch = 40 ms -> t_max = 80 ms
    f.core.wait_idle()
    f.release()
    f.wait_rounds(1)
    f.mute.update(ids)
    time.sleep(0.1)
    f.core.wait_idle()
    round2 = [r for r in f.run_log if r[1] == 2][0]
    assert dict(zip(round2[0], round2[2])) == {ids[0]: 80, ids[1]: 20}
    f.core.shutdown()


def test_dead_learner_is_evicted_and_the_round_completes():
    f = Federation(n=2, deadline=0.3, auto=False)
    f.core.replace_community_model(_core.build_federated_model(0, 0, model_bytes([np.zeros(4, np.float32)])))
    ids = [f.join(i) for i in range(2)]
    f.mute.add(ids[1])  # never answers (reference quirk Q9: the round would hang forever)
    f.delta[ids[0]] = 1.0
    f.core.wait_idle()
    f.auto = True
    f._complete(ids[0], _core.parse_federated_model(f.core.community_model())[2], 1, 20)
    f.wait_rounds(2, timeout=10)
    assert [l["id"] for l in f.core.get_learners()] == [ids[0]]
    f.core.shutdown()


def test_clean_leave_shrinks_the_barrier():
    f = Federation(n=2, auto=False)
    f.core.replace_community_model(_core.build_federated_model(0, 0, model_bytes([np.zeros(4, np.float32)])))
    a, b = f.join(0), f.join(1)
    f.core.wait_idle()
    f.core.learner_completed_task(a, f.tokens[a], model_bytes([np.ones(4, np.float32)]), 1, 1, 5, 10, 1, 1, b"")
    f.core.wait_idle()
    assert f.core.global_iteration() == 1
    f.core.remove_learner(b, f.tokens[b])
    f.core.wait_idle()
    time.sleep(0.05)
    f.core.wait_idle()
    assert f.core.global_iteration() == 2
    f.core.shutdown()


# ---- device aggregation commands (no GPU: a fake fabric that answers with references) ---------------------------
def _ref_model(rank, slot, version, numel=8, ciphertext=False):
    """A control-only model as a fabric learner sends it: one fp32 variable, no bytes, `TensorSpec.device_slot` set
    (under `ciphertext_tensor` when the learner of a secure federation reports an encrypted slot)."""
    from metisfl_b200.proto import model_pb2
    m = model_pb2.Model()
    var = m.variables.add(name="w", trainable=True)
    spec = (var.ciphertext_tensor if ciphertext else var.plaintext_tensor).tensor_spec
    spec.length = numel
    spec.dimensions.append(numel)
    spec.type.type = model_pb2.DType.FLOAT32
    spec.type.byte_order = model_pb2.DType.LITTLE_ENDIAN_ORDER
    spec.device_slot.device_rank = rank
    spec.device_slot.slot = slot
    spec.device_slot.version = version
    return m.SerializeToString()


class FabricFederation:
    """Controller whose learners answer RunTask with device-slot references; the aggregation callback records the
    command it is given and answers like an executor rank would (a reference to the community buffer)."""
    COMMUNITY = 0xFFFFFFFF

    def __init__(self, rule="FedAvg", protocol=1, n=2, examples=None, stride=0, fail_first=False, he_context="",
                 bytes_from=(), participation=1.0):
        p = _core.ControllerParams()
        p.rule, p.protocol, p.stride_length = rule, protocol, stride
        p.participation_ratio, p.sampling_seed = participation, 11
        if he_context:
            p.he_crypto_context_file = he_context
        self.ciphertext = rule == "PWA"
        self.bytes_from = set(bytes_from)     # ranks whose learners answer with (ciphertext) bytes instead of references
        p.batch_size, p.epochs = 10, 1
        if rule in ("FedRec", "FedAsync"):
            p.lineage_length_eviction, p.lineage_length = True, 2
        self.lock = threading.Lock()
        self.commands, self.tokens, self.rank_of, self.task_no = [], {}, {}, {}
        self.fail_first = fail_first
        self.core = _core.Controller(p, self._run, self._eval, self._aggregate)
        self.examples = examples or [100] * n

    def _run(self, items, fm, ref=None):
        for (lid, host, port, cert, iteration, updates, pct) in items:
            threading.Thread(target=self._complete, args=(lid, iteration, updates), daemon=True).start()

    def _complete(self, lid, iteration, updates):
        with self.lock:
            self.task_no[lid] = self.task_no.get(lid, 0) + 1
            n = self.task_no[lid]
        rank = int(lid.rsplit(":", 1)[1]) - 9100      # the first task can arrive before join() has returned
        model = _ref_model(rank, slot=n % 2, version=n, ciphertext=self.ciphertext)
        if rank in self.bytes_from:
            model = self.bytes_model
        t0 = time.time()
        while lid not in self.tokens and time.time() - t0 < 5:
            time.sleep(0.005)
        try:
            self.core.learner_completed_task(lid, self.tokens[lid], model, iteration, 1.0, updates, 10, 10.0, 1.0, b"")
        except Exception:
            pass  # shutdown races

    def _eval(self, learners, model, batch_size, idx, iteration, ref=None):
        for l in learners:
            self.core.evaluation_completed(l[0], idx, b"")

    def _aggregate(self, req):
        with self.lock:
            self.commands.append(req)
            first = len(self.commands) == 1
        if self.fail_first and first:
            return {"error": "injected failure"}
        return {"model_ref": _ref_model(0, self.COMMUNITY, req["global_iteration"] + 1, ciphertext=self.ciphertext),
                "device_ms": 0.125,
                "quantifiers": [(7, 1, 32)], "model_bytes": None, "error": None}

    def join(self, i):
        lid, tok = self.core.add_learner("127.0.0.1", 9100 + i, b"", b"", self.examples[i], 0, 0)
        self.tokens[lid], self.rank_of[lid] = tok, i
        return lid

    def wait_commands(self, n, timeout=20):
        t0 = time.time()
        while time.time() - t0 < timeout:
            with self.lock:
                if len(self.commands) >= n:
                    return
            time.sleep(0.01)
        raise AssertionError(f"only {len(self.commands)} aggregation commands")


def test_synchronous_round_is_commanded_with_slots_and_scaling_factors():
    f = FabricFederation(n=2, examples=[100, 300])
    f.core.replace_community_model(_core.build_federated_model(0, 0, model_bytes([np.zeros(8, np.float32)])))
    ids = [f.join(i) for i in range(2)]
    f.wait_commands(2)
    f.core.wait_idle()
    full = [c for c in f.commands if len(c["sources"]) == 2]
    assert full, [len(c["sources"]) for c in f.commands]   # a round with both learners (the first may close with one)
    c = full[0]
    assert c["rule"] == "FedAvg" and c["first_block"] and c["last_block"] and c["quantify"]
    by_id = {s["learner_id"]: s for s in c["sources"]}
    assert set(by_id) == set(ids)
    assert by_id[ids[0]]["device_rank"] == 0 and by_id[ids[1]]["device_rank"] == 1
    assert all(s["staged_model"] is None for s in c["sources"])           # references only: nothing to stage
    assert abs(by_id[ids[0]]["weight"] - 0.25) < 1e-9 and abs(by_id[ids[1]]["weight"] - 0.75) < 1e-9
    md = [m for m in f.core.get_runtime_metadata_lineage(0) if m["model_aggregation_block_duration_ms"]]
    assert md and md[0]["device_timed"] and md[0]["model_aggregation_block_duration_ms"][0] == 0.125
    assert md[0]["model_tensor_quantifiers"][0] == (7, 1, 32)            # the fabric's counts, not a host pass
    assert f.core.aggregation_failures() == 0
    f.core.shutdown()


def test_fedrec_command_carries_the_previous_contribution():
    f = FabricFederation(rule="FedRec", protocol=2, n=1, examples=[100])
    f.core.replace_community_model(_core.build_federated_model(0, 0, model_bytes([np.zeros(8, np.float32)])))
    f.join(0)
    f.wait_commands(3)
    f.core.wait_idle()
    first, second = f.commands[0], f.commands[1]
    assert first["rule"] == "FedRec" and len(first["sources"]) == 1 and not first["sources"][0]["has_previous"]
    s = second["sources"][0]
    assert s["has_previous"] and s["previous_slot"] == first["sources"][0]["slot"] and s["slot"] != s["previous_slot"]
    assert abs(s["previous_weight"] - first["sources"][0]["weight"]) < 1e-12 and second["z"] > 0
    f.core.shutdown()


def test_failed_device_aggregation_closes_the_round_and_is_counted():
    f = FabricFederation(n=1, examples=[100], fail_first=True)
    f.core.replace_community_model(_core.build_federated_model(0, 0, model_bytes([np.zeros(8, np.float32)])))
    f.join(0)
    f.wait_commands(3)          # the federation keeps going after the failed command
    f.core.wait_idle()
    assert f.core.aggregation_failures() == 1
    f.core.shutdown()


def test_fedstride_round_is_commanded_block_by_block():
    """Stride 1 with two learners: one command per block; the first one starts the running result, the last one
    finishes it (and is the only one that quantifies); the metadata lists both blocks with their device times."""
    f = FabricFederation(rule="FedStride", n=2, examples=[100, 100], stride=1)
    f.core.replace_community_model(_core.build_federated_model(0, 0, model_bytes([np.zeros(8, np.float32)])))
    ids = [f.join(i) for i in range(2)]
    f.wait_commands(6)
    f.core.wait_idle()
    md = [m for m in f.core.get_runtime_metadata_lineage(0) if len(m["model_aggregation_block_size"]) == 2]
    assert md, [m["model_aggregation_block_size"] for m in f.core.get_runtime_metadata_lineage(0)]
    assert md[0]["model_aggregation_block_size"] == [1.0, 1.0] and md[0]["model_aggregation_block_duration_ms"] == [0.125, 0.125]
    # find two consecutive commands of one round
    pairs = [(a, b) for a, b in zip(f.commands, f.commands[1:])
             if a["global_iteration"] == b["global_iteration"] and a["first_block"] and not a["last_block"]]
    assert pairs
    a, b = pairs[0]
    assert not b["first_block"] and b["last_block"] and len(a["sources"]) == len(b["sources"]) == 1
    assert {a["sources"][0]["learner_id"], b["sources"][0]["learner_id"]} == set(ids)
    assert abs(a["sources"][0]["weight"] - 0.5) < 1e-9 and b["quantify"]
    f.core.shutdown()


def _secure_setup(tmp_path):
    from metisfl_b200.encryption import fhe
    from metisfl_b200.proto import model_pb2
    scheme = fhe.CKKS(4096, 52, backend="host")
    scheme.gen_crypto_context_and_keys(str(tmp_path / "keys"))
    m = model_pb2.Model()
    var = m.variables.add(name="w", trainable=True)
    spec = var.ciphertext_tensor.tensor_spec
    spec.length = 8
    spec.dimensions.append(8)
    spec.type.type = model_pb2.DType.FLOAT32
    spec.type.byte_order = model_pb2.DType.LITTLE_ENDIAN_ORDER
    spec.value = scheme.encrypt(np.arange(8, dtype=np.float64))
    return scheme.get_crypto_params_files()["crypto_context_file"], m.SerializeToString()


def test_secure_round_is_commanded_on_the_fabric(tmp_path):
    """PWA with learners that report ciphertext REFERENCES: the controller commands the reduction (rule, ranks, its
    scaling factors) instead of calling the host PWA (`private_weighted_average.cc:24-82`), which has nothing to add."""
    ctx, initial = _secure_setup(tmp_path)
    f = FabricFederation(rule="PWA", n=2, examples=[100, 300], he_context=ctx)
    f.core.replace_community_model(_core.build_federated_model(0, 0, initial))
    ids = [f.join(i) for i in range(2)]
    f.wait_commands(2)
    f.core.wait_idle()
    full = [c for c in f.commands if len(c["sources"]) == 2]
    assert full
    c = full[0]
    assert c["rule"] == "PWA" and c["first_block"] and c["last_block"]
    by_id = {s["learner_id"]: s for s in c["sources"]}
    assert by_id[ids[0]]["device_rank"] == 0 and by_id[ids[1]]["device_rank"] == 1
    assert all(s["staged_model"] is None for s in c["sources"])
    assert abs(by_id[ids[0]]["weight"] - 0.25) < 1e-9 and abs(by_id[ids[1]]["weight"] - 0.75) < 1e-9
    assert f.core.aggregation_failures() == 0
    f.core.shutdown()


def test_secure_round_refuses_to_mix_references_and_ciphertext_bytes(tmp_path):
    ctx, initial = _secure_setup(tmp_path)
    f = FabricFederation(rule="PWA", n=2, examples=[100, 100], he_context=ctx, bytes_from=(1,))
    f.bytes_model = initial
    f.core.replace_community_model(_core.build_federated_model(0, 0, initial))
    [f.join(i) for i in range(2)]
    t0 = time.time()
    while f.core.aggregation_failures() == 0 and time.time() - t0 < 20:
        time.sleep(0.02)
    assert f.core.aggregation_failures() >= 1
    with f.lock:
        assert not [c for c in f.commands if len(c["sources"]) == 2]   # no command ever mixed the two kinds
    f.core.shutdown()


# ---- client sampling (GlobalModelConfig.ParticipationRatio; plumbed but never read in the reference, quirk Q7) ------
def test_participation_ratio_draws_a_cohort_per_round():
    f = Federation(n=4, auto=False, participation=0.5, sampling_seed=7)
    f.core.replace_community_model(_core.build_federated_model(0, 0, model_bytes([np.zeros(4, np.float32)])))
    ids = [f.join(i) for i in range(4)]
    for lid, d in zip(ids, (1.0, 2.0, 4.0, 8.0)):
        f.delta[lid] = d
    f.core.wait_idle()
    f.release()          # round 1: everybody who joined holds the initial task
    f.wait_rounds(8)
    f.mute.update(ids)
    time.sleep(0.1)
    f.core.wait_idle()
    md = f.core.get_runtime_metadata_lineage(0)
    assert sorted(md[0]["assigned_to_learner_id"]) == sorted(ids)
    later = md[1:-1]                                                  # the last round was still open when muted
    assert len(later) >= 6
    cohorts = [tuple(sorted(m["assigned_to_learner_id"])) for m in later]
    assert all(len(c) == 2 for c in cohorts), cohorts                 # ceil(0.5 * 4) learners per round
    assert all(sorted(m["completed_by_learner_id"]) == sorted(m["assigned_to_learner_id"]) for m in later)
    assert all(m["model_aggregation_block_size"] == [2.0] for m in later)   # only the cohort is averaged
    assert len(set(cohorts)) > 1                                      # not the same pair every round
    # evaluation follows the cohort that trains next
    assert all(len(e[0]) == 2 for e in f.eval_log[1:])
    # every round adds the mean of its cohort's deltas (equal example counts): the increments identify the cohort
    lineage = f.core.get_community_model_lineage(0)
    vals = [float(model_arrays(_core.parse_federated_model(b)[2])[0][0]) for b in lineage]
    delta_of = dict(zip(ids, (1.0, 2.0, 4.0, 8.0)))
    its = [_core.parse_federated_model(b)[1] for b in lineage]
    by_it = {m["global_iteration"]: m for m in md}
    for (a, b, it) in zip(vals, vals[1:], its[1:]):
        cohort = by_it[it]["completed_by_learner_id"]
        assert abs((b - a) - sum(delta_of[c] for c in cohort) / len(cohort)) < 1e-4
    f.core.shutdown()


def test_round_whose_whole_cohort_leaves_is_closed_and_a_new_cohort_is_drawn():
    f = Federation(n=3, auto=False, participation=0.34, sampling_seed=3)   # cohorts of ceil(0.34 * 3) = 2 ... of 1 below
    f.core.replace_community_model(_core.build_federated_model(0, 0, model_bytes([np.zeros(4, np.float32)])))
    ids = [f.join(i) for i in range(3)]
    f.core.wait_idle()
    f.release()
    f.wait_rounds(1)
    f.auto = False       # tasks of the next round stay unanswered
    time.sleep(0.2)
    f.core.wait_idle()
    md = f.core.get_runtime_metadata_lineage(0)
    open_round = md[-1]
    cohort = list(open_round["assigned_to_learner_id"])
    assert 1 <= len(cohort) <= 2 and not open_round["completed_by_learner_id"]
    it = f.core.global_iteration()
    for lid in cohort:
        assert f.core.evict_learner(lid, "test: cohort member gone")
    t0 = time.time()
    while f.core.global_iteration() == it and time.time() - t0 < 10:
        time.sleep(0.01)
    assert f.core.global_iteration() == it + 1           # closed without contributions, next cohort drawn
    f.core.wait_idle()
    md = f.core.get_runtime_metadata_lineage(0)
    survivors = set(ids) - set(cohort)
    assert md[-1]["assigned_to_learner_id"] and set(md[-1]["assigned_to_learner_id"]) <= survivors
    assert f.core.aggregation_failures() == 0
    f.core.shutdown()


def test_participation_ratio_is_validated():
    p = _core.ControllerParams()
    p.rule, p.protocol, p.batch_size, p.epochs, p.participation_ratio = "FedAvg", 1, 10, 1, 1.5
    with pytest.raises(Exception):
        _core.Controller(p, lambda *a: None, lambda *a: None)


def test_sampled_cohorts_are_commanded_on_the_fabric():
    """Client sampling + device fabric: the aggregation command names the cohort's slots only, with scaling factors
    normalised over the cohort (every rank still launches - the fused kernel is collective - but reads two sources)."""
    f = FabricFederation(n=4, examples=[100, 100, 300, 300], participation=0.5)
    f.core.replace_community_model(_core.build_federated_model(0, 0, model_bytes([np.zeros(8, np.float32)])))
    ids = [f.join(i) for i in range(4)]
    with f.lock:
        joined_at = len(f.commands)      # rounds already run while the others were still joining
    f.wait_commands(joined_at + 7)
    f.core.wait_idle()
    with f.lock:
        later = f.commands[joined_at + 2:]   # (a learner that joined while a round was open widened that round)
    assert later and all(len(c["sources"]) == 2 for c in later), [len(c["sources"]) for c in f.commands]
    for c in later:
        assert abs(sum(s["weight"] for s in c["sources"]) - 1.0) < 1e-9
        assert {s["learner_id"] for s in c["sources"]} <= set(ids)
    assert f.core.aggregation_failures() == 0
    f.core.shutdown()


def test_a_round_with_one_contributor_publishes_that_model_not_a_multiple_of_it():
    """Stateless FedAvg + a lone contributor in a federation of two (a sampled cohort of one): the scalers hand out the
    RAW example count for a single participant (meant for the rolling rules, `train_dataset_size_scaler.cc:7-49`); the
    community model must still be that learner's model, i.e. grow by its delta per round."""
    f = Federation(n=2, auto=False, participation=0.5, sampling_seed=5, examples=[100, 300])
    f.core.replace_community_model(_core.build_federated_model(0, 0, model_bytes([np.zeros(4, np.float32)])))
    ids = [f.join(i) for i in range(2)]
    f.delta[ids[0]], f.delta[ids[1]] = 1.0, 2.0
    f.core.wait_idle()
    f.release()
    f.wait_rounds(7)
    f.mute.update(ids)
    time.sleep(0.1)
    f.core.wait_idle()
    md = {m["global_iteration"]: m for m in f.core.get_runtime_metadata_lineage(0)}
    lineage = [_core.parse_federated_model(b) for b in f.core.get_community_model_lineage(0)]
    vals = [(it, float(model_arrays(m)[0][0])) for _, it, m in lineage]
    assert all(np.isfinite(v) and v < 100 for _, v in vals), vals
    delta_of = {ids[0]: 1.0, ids[1]: 2.0}
    for (_, a), (it, b) in zip(vals, vals[1:]):
        who = md[it]["completed_by_learner_id"]
        assert len(who) == 1 and abs((b - a) - delta_of[who[0]]) < 1e-5, (it, a, b, who)
    f.core.shutdown()


@pytest.mark.parametrize("rule", ["FedAvg", "FedStride"])
def test_device_command_for_a_lone_contributor_carries_weight_one(rule):
    """On the device path the blocks are plain weighted sums (no division by z), so the lone contributor of a round
    must be commanded with factor 1, not with its raw example count."""
    f = FabricFederation(rule=rule, n=2, examples=[100, 300], participation=0.5)
    f.core.replace_community_model(_core.build_federated_model(0, 0, model_bytes([np.zeros(8, np.float32)])))
    [f.join(i) for i in range(2)]
    with f.lock:
        joined_at = len(f.commands)
    f.wait_commands(joined_at + 6)
    f.core.wait_idle()
    with f.lock:
        later = f.commands[joined_at + 2:]
    assert later and all(len(c["sources"]) == 1 for c in later)
    assert all(abs(c["sources"][0]["weight"] - 1.0) < 1e-12 for c in later), [c["sources"][0]["weight"] for c in later]
    f.core.shutdown()


def test_first_round_that_a_late_joiner_missed_publishes_the_early_learners_model():
    """No sampling: learner 0 finishes round 1 before learner 1 joins.  The barrier closes with one contributor in a
    federation that by then may count two learners; the community model is that contributor's model (+1), and from
    the round both take part in it moves by the weighted mean (0.25 * 1 + 0.75 * 3 = 2.5)."""
    f = Federation(n=2, examples=[100, 300])
    f.core.replace_community_model(_core.build_federated_model(0, 0, model_bytes([np.zeros(4, np.float32)])))
    f.delta["127.0.0.1:9000"], f.delta["127.0.0.1:9001"] = 1.0, 3.0   # known before the first task can complete
    a = f.join(0)
    f.wait_rounds(2)
    b = f.join(1)
    assert (a, b) == ("127.0.0.1:9000", "127.0.0.1:9001")
    f.wait_rounds(8)
    f.mute.update([a, b])
    time.sleep(0.1)
    f.core.wait_idle()
    md = {m["global_iteration"]: m for m in f.core.get_runtime_metadata_lineage(0)}
    lineage = [_core.parse_federated_model(x) for x in f.core.get_community_model_lineage(0)]
    vals = [(it, float(model_arrays(m)[0][0])) for _, it, m in lineage]
    assert all(np.isfinite(v) and v < 100 for _, v in vals), vals
    for (_, x0), (it, x1) in zip(vals, vals[1:]):
        who = sorted(md[it]["completed_by_learner_id"])
        want = 1.0 if who == [a] else 3.0 if who == [b] else 2.5
        assert abs((x1 - x0) - want) < 1e-4, (it, x0, x1, who)
    f.core.shutdown()


def test_semi_synchronous_with_sampling_resizes_every_learners_task():
    """Semi-synchronous + client sampling: the task sizes are recomputed from ALL active learners' timings (a learner
    outside this round's cohort keeps a template that matches the slowest learner), and rounds keep closing."""
    f = Federation(protocol=3, n=4, semi_lambda=2, auto=False, participation=0.5, sampling_seed=9)
    f.core.replace_community_model(_core.build_federated_model(0, 0, model_bytes([np.zeros(4, np.float32)])))
    ids = [f.join(i) for i in range(4)]
    f.ms_per_batch = {ids[0]: 1.0, ids[1]: 4.0, ids[2]: 2.0, ids[3]: 1.0}   # slowest epoch 40 ms -> t_max = 80 ms
    f.core.wait_idle()
    f.release()
    f.wait_rounds(5)
    f.mute.update(ids)
    time.sleep(0.1)
    f.core.wait_idle()
    want = {ids[0]: 80, ids[1]: 20, ids[2]: 40, ids[3]: 80}
    later = [r for r in f.run_log if r[1] >= 2]
    assert len(later) >= 4 and all(len(r[0]) == 2 for r in later)
    for lids, _, updates in later:
        assert dict(zip(lids, updates)) == {l: want[l] for l in lids}
    f.core.shutdown()
