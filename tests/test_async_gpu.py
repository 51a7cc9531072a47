"""K3 fused asynchronous fold, version counters and the rotating community buffers on one GPU (world 1; the
cross-GPU pushes are exercised by tests/test_multigpu.py).  Oracle: the same arithmetic in fp32 PyTorch
(``state = decay * state + w_new x_new - w_old x_old; C = cast(state * post)``; reference semantics
``federated_rolling_average_base.cc:175-293``, ``federated_recency.cc:27-94``)."""
import pytest
import torch

from metisfl_b200.models.zoo import FashionMnistMLP
from metisfl_b200.ops import aggregation as agg
from metisfl_b200.ops.optim import OptimHyper
from metisfl_b200.parallel.engine import EngineConfig, FederationEngine
from metisfl_b200.parallel.simulate import AsyncFederation

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _segments(n_bf16, n_f32, n_i64):
    """One arena-like buffer: a bf16 segment, an fp32 segment, an int64 (exact) segment, 256-byte aligned."""
    off, segs = 0, []
    for n, dt, exact in ((n_bf16, torch.bfloat16, False), (n_f32, torch.float32, False), (n_i64, torch.int64, True)):
        off = (off + 255) // 256 * 256
        segs.append((off, n, dt, exact))
        off += n * dt.itemsize
    return segs, (off + 255) // 256 * 256


def _fill(buf, segs, gen):
    views = []
    for off, n, dt, exact in segs:
        v = buf[off:off + n * dt.itemsize].view(dt)
        if exact:
            v.copy_(torch.randint(0, 1000, (n,), generator=gen).to(DEV))
        else:
            v.copy_(torch.randn(n, generator=gen).to(DEV).to(dt))
        views.append(v)
    return views


@pytest.mark.parametrize("with_old", [False, True])
def test_async_fold_matches_fp32_reference(with_old):
    gen = torch.Generator().manual_seed(3)
    segs, nb = _segments(100_003, 5_001, 37)   # odd sizes: exercises the < 8-element tails
    x_new = torch.zeros(nb, dtype=torch.uint8, device=DEV)
    x_old = torch.zeros(nb, dtype=torch.uint8, device=DEV)
    out = torch.zeros(nb, dtype=torch.uint8, device=DEV)
    vn, vo = _fill(x_new, segs, gen), _fill(x_old, segs, gen)
    offs, total = [], 0
    for off, n, dt, exact in segs:
        offs.append(total)
        if not exact:
            total += (n + 3) // 4 * 4
    state = torch.randn(total, generator=gen).to(DEV)
    ref_state = state.clone()
    flags = torch.zeros(8, dtype=torch.int64, device=DEV)
    done = torch.zeros(4, dtype=torch.int32, device=DEV)
    decay, w_new, w_old, post = 0.75, 0.4, 0.15, 1.0 / 3.0
    for launch in range(2):   # two launches in a row: the completion counter resets itself
        agg.async_fold(x_new.data_ptr(), w_new, x_old.data_ptr() if with_old else 0, w_old, decay, post, state, offs,
                       segs, [out.data_ptr()], 0, [flags.data_ptr()], new_version=launch + 1, done_counter=done,
                       n_blocks=32)
        for si, (off, n, dt, exact) in enumerate(segs):
            if exact:
                continue
            st = ref_state[offs[si]:offs[si] + n]
            st.mul_(decay).add_(vn[si].float(), alpha=w_new)
            if with_old:
                st.add_(vo[si].float(), alpha=-w_old)
    torch.cuda.synchronize()
    assert int(flags[0]) == 2 and int(done[0]) == 0
    for si, (off, n, dt, exact) in enumerate(segs):
        got = out[off:off + n * dt.itemsize].view(dt)
        if exact:
            assert torch.equal(got, vn[si])
            continue
        want = ref_state[offs[si]:offs[si] + n] * post
        torch.testing.assert_close(state[offs[si]:offs[si] + n], ref_state[offs[si]:offs[si] + n], atol=1e-5, rtol=1e-5)
        tol = 1e-5 if dt == torch.float32 else 2e-2
        torch.testing.assert_close(got.float(), want, atol=tol, rtol=tol)
    assert agg.spin_timeouts() == 0


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32, torch.float16])
def test_rolling_update_vector_path(dtype):
    gen = torch.Generator().manual_seed(5)
    n = 1_000_003
    x_new = torch.randn(n, generator=gen).to(DEV).to(dtype)
    x_old = torch.randn(n, generator=gen).to(DEV).to(dtype)
    S = torch.randn(n, generator=gen).to(DEV)
    want_S = S + 2.0 * x_new.float() - 0.5 * x_old.float()
    c = torch.empty(n, dtype=dtype, device=DEV)
    agg.rolling_update(c, S, x_new, 2.0, x_old, 0.5, z=4.0)
    torch.testing.assert_close(S, want_S, atol=1e-5, rtol=1e-5)
    tol = 1e-5 if dtype == torch.float32 else 2e-2
    torch.testing.assert_close(c.float(), want_S / 4.0, atol=tol, rtol=tol)


def test_snapshot_install_picks_the_buffer_of_the_version():
    gen = torch.Generator().manual_seed(7)
    segs, nb = _segments(40_003, 1_001, 9)
    bufs = torch.zeros(3 * nb, dtype=torch.uint8, device=DEV)
    views = [_fill(bufs[b * nb:(b + 1) * nb], segs, gen) for b in range(3)]
    flag = torch.zeros(8, dtype=torch.int64, device=DEV)
    status = torch.zeros(2, dtype=torch.int64, device=DEV)
    dst = torch.zeros(nb, dtype=torch.uint8, device=DEV)
    master = torch.zeros(segs[0][1], dtype=torch.float32, device=DEV)
    for v in (0, 1, 2, 7):
        flag[0] = v
        agg.snapshot_install([bufs.data_ptr() + b * nb for b in range(3)], flag.data_ptr(), status, dst, nb, master,
                             segs[0])
        torch.cuda.synchronize()
        assert int(status[0]) == v and int(status[1]) == 0
        assert torch.equal(dst, bufs[(v % 3) * nb:(v % 3 + 1) * nb])
        torch.testing.assert_close(master, views[v % 3][0].float(), atol=0, rtol=0)


@pytest.mark.parametrize("rule", ["FedAsync", "FedRec"])
def test_async_federation_on_one_gpu_matches_the_cpu_engine(rule):
    """Same simulated schedule on the GPU (K3 on the fold stream, snapshot installs, events) and on the CPU engine."""
    g = torch.Generator().manual_seed(11)
    X = torch.randn(4, 32, 1, 28, 28, generator=g)
    Y = torch.randint(0, 10, (4, 32), generator=g)
    outs = {}
    for dev in ("cpu", DEV):
        cfg = EngineConfig(learners_per_gpu=4, model_dtype=torch.float32, rule=rule,
                           optimizer=OptimHyper("VanillaSGD", lr=0.05), use_cuda_graph=False, use_b200_linear=False)
        e = FederationEngine(FashionMnistMLP, cfg, device=dev)
        sim = AsyncFederation(e, [1.0, 2.0, 4.0, 8.0], lambda gl, n: [(X[gl], Y[gl])])
        out = sim.run(12)
        e.sync_folds()
        assert e.version == 12 and out["max_staleness"] >= 3
        if dev != "cpu":
            h = e.async_health()
            assert h == {"torn_reads": 0, "spin_timeouts": 0}, h
        outs[dev] = {k: v.float().cpu() for k, v in e.community_state_dict().items()}
    for k in outs["cpu"]:
        torch.testing.assert_close(outs[DEV][k], outs["cpu"][k], atol=2e-4, rtol=2e-3)


def test_engine_phase_timers_are_device_timed():
    """``EngineConfig.trace``: CUDA-event phase timers (utils/tracing.py) around training and aggregation."""
    e = FederationEngine(FashionMnistMLP, EngineConfig(trace=True, use_cuda_graph=False), device=DEV)
    x = torch.randn(32, 1, 28, 28).pin_memory()
    y = torch.randint(0, 10, (32,)).pin_memory()
    for _ in range(2):
        e.run_round([[(x, y)] * 2])
    t = e.phase_times()
    assert set(t) == {"train_learner", "aggregate"} and all(0.0 < v < 5e3 for v in t.values()), t
    assert e.phase_times() == {}
