"""Engine modes beyond the synchronous round: asynchronous FedAsync / FedRec with simulated
heterogeneous learners (config #3) and the streaming many-learner round (config #5) - CPU tier."""
import pytest
import torch

from metisfl_b200.models.zoo import FashionMnistMLP
from metisfl_b200.ops.optim import OptimHyper
from metisfl_b200.parallel.engine import EngineConfig, FederationEngine, scaling_factors, staleness_factor
from metisfl_b200.parallel.simulate import AsyncFederation


def _data(n, seed=0):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(n, 32, 1, 28, 28, generator=g), torch.randint(0, 10, (n, 32), generator=g)


def test_staleness_and_scaling_tables():
    assert staleness_factor("constant", 9) == 1.0
    assert abs(staleness_factor("polynomial", 3, a=0.5) - 0.5) < 1e-12
    assert staleness_factor("hinge", 3, a=0.5, b=4) == 1.0 and abs(staleness_factor("hinge", 6, a=0.5, b=4) - 0.5) < 1e-12
    assert scaling_factors("NumTrainingExamples", [1, 3], [5, 5], 2) == [0.25, 0.75]
    assert scaling_factors("NumParticipants", [1, 3], [5, 5], 2) == [0.5, 0.5]
    assert scaling_factors("NumCompletedBatches", [1, 3], [2, 6], 4) == [0.25, 0.75]
    assert scaling_factors("NumTrainingExamples", [7], [1], 1) == [1.0]
    assert scaling_factors("NumTrainingExamples", [7], [1], 3) == [7.0]


@pytest.mark.parametrize("rule", ["FedAsync", "FedRec"])
def test_async_simulation_orders_events_and_tracks_staleness(rule):
    X, Y = _data(4)
    cfg = EngineConfig(learners_per_gpu=4, model_dtype=torch.float32, rule=rule,
                       optimizer=OptimHyper("VanillaSGD", lr=0.05))
    e = FederationEngine(FashionMnistMLP, cfg, device="cpu")
    before = {k: v.clone() for k, v in e.community_state_dict().items()}
    sim = AsyncFederation(e, [1.0, 2.0, 4.0, 8.0], lambda g, n: [(X[g], Y[g])])
    out = sim.run(15)
    assert out["updates"] == 15 and e.global_iteration == 15
    # the fastest learner contributes most often; the slowest one is the most stale
    assert sim.task_no[0] > sim.task_no[3] and out["max_staleness"] >= 4
    after = e.community_state_dict()
    assert any(not torch.equal(before[k], after[k]) for k in before)
    assert all(torch.isfinite(v.float()).all() for v in after.values())


def test_streaming_round_equals_plain_fedavg():
    X, Y = _data(6, seed=3)
    mk = lambda k, **kw: EngineConfig(learners_per_gpu=k, model_dtype=torch.float32,
                                      optimizer=OptimHyper("VanillaSGD", lr=0.05), **kw)
    stream = FederationEngine(FashionMnistMLP, mk(2), device="cpu")
    stream.run_round_streaming(6, lambda g: [(X[g], Y[g])] * 2, examples_fn=lambda g: 100.0 * (g + 1))
    plain = FederationEngine(FashionMnistMLP, mk(6), device="cpu", examples_per_learner=[100 * (g + 1) for g in range(6)])
    plain.run_round([[(X[g], Y[g])] * 2 for g in range(6)])
    a, b = stream.community_state_dict(), plain.community_state_dict()
    for k in a:
        torch.testing.assert_close(a[k], b[k], atol=1e-6, rtol=1e-6)


def test_fedstride_blocks_equal_fedavg_in_engine():
    X, Y = _data(4, seed=5)
    data = [[(X[g], Y[g])] for g in range(4)]
    a = FederationEngine(FashionMnistMLP, EngineConfig(learners_per_gpu=4, model_dtype=torch.float32, rule="FedStride",
                                                       stride_length=2), device="cpu")
    b = FederationEngine(FashionMnistMLP, EngineConfig(learners_per_gpu=4, model_dtype=torch.float32), device="cpu")
    a.run_round(data)
    b.run_round(data)
    for k, v in a.community_state_dict().items():
        torch.testing.assert_close(v, b.community_state_dict()[k], atol=1e-6, rtol=1e-6)


def test_lone_participant_publishes_its_model():
    """`aggregate_sync(participants=[j])` in an engine of several learners: the community model is learner j's model,
    not example-count x model (the scalers return the raw count for a single participant)."""
    import torch
    cfg = EngineConfig(learners_per_gpu=3, model_dtype=torch.float32, use_cuda_graph=False)
    e = FederationEngine(FashionMnistMLP, cfg, device="cpu", examples_per_learner=[100, 200, 300])
    g = torch.Generator().manual_seed(0)
    for j in range(3):
        for v in e.slot_views[j]:
            if v.is_floating_point():
                v.copy_(torch.randn(v.shape, generator=g))
    e.aggregate_sync(participants=[1])
    for got, want in zip(e.community_views, e.slot_views[e.publish_slot(1)]):
        if got.is_floating_point():
            torch.testing.assert_close(got, want, atol=1e-6, rtol=0)
