"""Numerics of the sm_100a kernels against plain PyTorch fp32 references of the same op.

Golden expectations ported from the reference's gtest suites: integer truncation
of FedAvg (``federated_average_test.cc:95-208``), rolling S/z updates
(``federated_stride_test.cc``, ``federated_recency_test.cc``).
"""
import pytest
import torch

from metisfl_b200.ops import aggregation as agg
from metisfl_b200.ops.optim import OptimHyper, optimizer_step

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _rand(n, dtype, seed):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return torch.randn(n, generator=g).to(dtype).to(DEV)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
@pytest.mark.parametrize("n", [7, 8, 4099, 1 << 20])
@pytest.mark.parametrize("L", [1, 3, 8, 13])
def test_weighted_sum_fast(dtype, n, L):
    xs = [_rand(n, dtype, 10 + l) for l in range(L)]
    w = [(l + 1) / sum(range(1, L + 1)) for l in range(L)]
    out = agg.weighted_sum(xs, w)
    ref = sum(wi * x.float() for wi, x in zip(w, xs))
    tol = 1e-6 if dtype == torch.float32 else 1e-2
    torch.testing.assert_close(out.float(), ref.to(dtype).float(), atol=tol, rtol=tol)
    out32 = agg.weighted_sum(xs, w, out_dtype=torch.float32)
    torch.testing.assert_close(out32, ref, atol=1e-5, rtol=1e-5)


def test_weighted_sum_many_sources_streaming():
    n, L = 10007, 150  # > MFL_MAX_SRC: chunked through an fp32 running sum
    xs = [_rand(n, torch.bfloat16, l) for l in range(L)]
    w = [1.0 / L] * L
    out = agg.weighted_sum(xs, w)
    ref = sum(x.float() for x in xs) / L
    torch.testing.assert_close(out.float(), ref, atol=2e-2, rtol=2e-2)
    acc = torch.zeros(n, dtype=torch.float32, device=DEV)
    for i in range(0, L, 50):
        agg.weighted_sum(xs[i:i + 50], w[i:i + 50], acc=acc, accumulate=True, write_out=False)
    torch.testing.assert_close(acc, ref, atol=1e-4, rtol=1e-4)


@pytest.mark.parametrize("dtype,expect", [
    (torch.int32, [0, 2, 2, 4, 4, 6, 6, 8, 8, 10]),
    (torch.int16, [0, 2, 2, 4, 4, 6, 6, 8, 8, 10]),
    (torch.float32, [1, 2, 3, 4, 5, 6, 7, 8, 9, 10]),
    (torch.float64, [1, 2, 3, 4, 5, 6, 7, 8, 9, 10]),
])
def test_fedavg_reference_truncation(dtype, expect):
    """Two identical models x 0.5: integers truncate per term (0.5*1 -> 0), floats are exact."""
    x = torch.arange(1, 11).to(dtype).to(DEV)
    out = agg.weighted_sum([x, x.clone()], [0.5, 0.5], exact=True)
    assert out.dtype == dtype
    assert out.cpu().tolist() == expect


def test_exact_matches_cpu_oracle_int64():
    g = torch.Generator().manual_seed(3)
    xs = [torch.randint(-1000, 1000, (1001,), generator=g) for _ in range(5)]
    w = [0.1, 0.2, 0.3, 0.15, 0.25]
    cpu = agg.weighted_sum(xs, w, exact=True)
    gpu = agg.weighted_sum([x.to(DEV) for x in xs], w, exact=True)
    assert torch.equal(cpu, gpu.cpu())


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_rolling_update_fedrec_sequence(dtype):
    """3 first-time committers (weights 1,2,3) then learner 0 re-commits: z stays 6, S swaps."""
    n = 5003
    S = torch.zeros(n, dtype=torch.float32, device=DEV)
    C = torch.zeros(n, dtype=dtype, device=DEV)
    xs = [_rand(n, dtype, 40 + i) for i in range(4)]
    z = 0.0
    for i, wgt in enumerate([1.0, 2.0, 3.0]):
        z += wgt
        agg.rolling_update(C, S, xs[i], wgt, None, 0.0, z)
    ref = (xs[0].float() + 2 * xs[1].float() + 3 * xs[2].float()) / 6
    torch.testing.assert_close(C.float(), ref.to(dtype).float(), atol=1e-2, rtol=1e-2)
    agg.rolling_update(C, S, xs[3], 1.0, xs[0], 1.0, z)
    ref = (xs[3].float() + 2 * xs[1].float() + 3 * xs[2].float()) / 6
    torch.testing.assert_close(C.float(), ref.to(dtype).float(), atol=1e-2, rtol=1e-2)


@pytest.mark.parametrize("name", ["VanillaSGD", "MomentumSGD", "FedProx", "Adam", "AdamW"])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_optimizer_step_matches_cpu_reference(name, dtype):
    n = 40013
    h = OptimHyper(name, lr=0.05, l1=1e-4, l2=1e-3, momentum=0.9, mu=0.1, weight_decay=0.01, eps=1e-7)
    master = _rand(n, torch.float32, 1)
    anchor = _rand(n, torch.float32, 2)
    s1 = torch.zeros(n, device=DEV); s2 = torch.zeros(n, device=DEV)
    model = master.to(dtype) if dtype != torch.float32 else None
    pub = torch.zeros(n, dtype=dtype, device=DEV)
    c_master, c_anchor = master.cpu().clone(), anchor.cpu().clone()
    c_s1, c_s2 = torch.zeros(n), torch.zeros(n)
    c_model = c_master.to(dtype) if dtype != torch.float32 else None
    c_pub = torch.zeros(n, dtype=dtype)
    for step in range(1, 4):
        g = _rand(n, dtype, 100 + step)
        cg = g.cpu().clone()
        optimizer_step(master, model, g, h, s1, s2, anchor, publish=pub, step=step, publish_scale=2.0, zero_grad=True)
        optimizer_step(c_master, c_model, cg, h, c_s1, c_s2, c_anchor, publish=c_pub, step=step, publish_scale=2.0,
                       zero_grad=True)
        assert float(g.float().abs().sum()) == 0.0  # grads zeroed in the same pass
    torch.testing.assert_close(master.cpu(), c_master, atol=2e-5, rtol=2e-5)
    tol = 1e-5 if dtype == torch.float32 else 2e-2
    torch.testing.assert_close(pub.cpu().float(), c_pub.float(), atol=tol, rtol=tol)
    if model is not None:
        torch.testing.assert_close(model.cpu().float(), c_model.float(), atol=tol, rtol=tol)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32, torch.int64, torch.uint8])
def test_quantifier(dtype):
    sizes = [3, 0, 100000, 17, 64, 1, 250000, 5]
    offs = [0]
    for s in sizes:
        offs.append(offs[-1] + s)
    g = torch.Generator().manual_seed(5)
    flat = (torch.randint(0, 3, (offs[-1],), generator=g) - 1).to(dtype)  # values in {-1,0,1} (uint8: 255,0,1)
    if dtype.is_floating_point:
        flat[5] = -0.0
    q = agg.Quantifier(offs, DEV)
    got = q(flat.to(DEV)).cpu()
    ref = agg.Quantifier(offs, "cpu")(flat)
    assert torch.equal(got, ref)
    got2 = q(flat.to(DEV)).cpu()  # scratch reuse
    assert torch.equal(got2, ref)


def test_fused_aggregate_single_rank_matches_weighted_sum():
    """world=1 exercise of the fused kernel: barriers against itself, multi-segment, install."""
    from metisfl_b200.models.zoo import ResNet18
    from metisfl_b200.parallel.engine import EngineConfig, FederationEngine
    eng = FederationEngine(ResNet18, EngineConfig(learners_per_gpu=3, use_cuda_graph=False), device=DEV,
                           examples_per_learner=[100, 200, 300])
    nb = eng.layout.total_bytes
    g = torch.Generator().manual_seed(9)
    for j in range(3):  # fill slots with random models of the right per-segment dtype
        for si, s in enumerate(eng.layout.segments):
            v = eng.slot_views[j][si]
            if s.dtype.is_floating_point:
                v.copy_(torch.randn(v.numel(), generator=g).to(s.dtype))
            else:
                v.copy_(torch.randint(0, 50, (v.numel(),), generator=g).to(s.dtype))
    eng.aggregate_sync()
    torch.cuda.synchronize()
    w = [1 / 6, 2 / 6, 3 / 6]
    for si, s in enumerate(eng.layout.segments):
        srcs = [eng.slot_views[j][si].cpu() for j in range(3)]
        ref = agg.weighted_sum(srcs, w, exact=s.exact)
        got = eng.community_views[si].cpu()
        if s.exact:
            assert torch.equal(got, ref)
        else:
            torch.testing.assert_close(got.float(), ref.float(), atol=2e-2, rtol=2e-2)


def test_engine_round_gpu_matches_cpu_math():
    """One FedAvg round on the GPU (CUDA graph + K2 + K1) tracks the same round on the CPU path."""
    from metisfl_b200.models.zoo import FashionMnistMLP
    from metisfl_b200.parallel.engine import EngineConfig, FederationEngine
    mk = lambda: EngineConfig(learners_per_gpu=2, model_dtype=torch.float32,
                              optimizer=OptimHyper("MomentumSGD", lr=0.05, momentum=0.9))
    eg = FederationEngine(FashionMnistMLP, mk(), device=DEV)
    ec = FederationEngine(FashionMnistMLP, mk(), device="cpu")
    g = torch.Generator().manual_seed(0)
    data = [[(torch.randn(64, 1, 28, 28, generator=g), torch.randint(0, 10, (64,), generator=g)) for _ in range(3)]
            for _ in range(2)]
    for r in range(2):
        og = eg.run_round(data)
        oc = ec.run_round(data)
        for a, b in zip(og["loss"], oc["loss"]):
            assert abs(a - b) < 5e-3, (og, oc)
    sg, sc = eg.community_state_dict(), ec.community_state_dict()
    for k in sc:
        torch.testing.assert_close(sg[k].cpu(), sc[k], atol=2e-3, rtol=2e-3)
