"""Fused cross-GPU aggregation (K1 over NVLink peer memory) at world size 2/4/8: every rank must end
with the same community model, equal to the CPU oracle's weighted average of all learners' slots
(SURVEY §4.4 item 7).  Skipped when fewer GPUs are visible."""
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r"""
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, {root!r})
rank = int(os.environ["RANK"]); lr = int(os.environ["LOCAL_RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
from metisfl_b200.parallel.engine import EngineConfig, FederationEngine
from metisfl_b200.models.zoo import ResNet18
from metisfl_b200.ops import aggregation as agg
k = {k}
ex = [100 * (i + 1) for i in range(world * k)]
import torch.distributed._symmetric_memory as _symm
try:  # NVLS needs multicast objects (NVSwitch + fabric manager); the peer-pointer path is tested either way
    have_mc = bool(_symm._SymmetricMemory.has_multicast_support(_symm.DeviceType.CUDA, lr))
except Exception:
    have_mc = False
eng = FederationEngine(ResNet18, EngineConfig(learners_per_gpu=k, use_cuda_graph=False, nvls=True if have_mc else False),
                       device=f"cuda:{{lr}}", examples_per_learner=ex)
g = torch.Generator().manual_seed(100 + rank)
for j in range(k):
    for si, s in enumerate(eng.layout.segments):
        v = eng.slot_views[j][si]
        if s.dtype.is_floating_point:
            v.copy_(torch.randn(v.numel(), generator=g).to(s.dtype))
        else:
            v.copy_(torch.randint(0, 50, (v.numel(),), generator=g).to(s.dtype))
torch.cuda.synchronize(); dist.barrier()
for it in range(3):          # back-to-back launches reuse the self-resetting signal pads
    eng.aggregate_sync()
torch.cuda.synchronize(); dist.barrier()
# oracle: gather all slots to every rank through NCCL and average on the CPU
nb = eng.layout.total_bytes
allslots = [torch.empty_like(eng.slots.local) for _ in range(world)]
dist.all_gather(allslots, eng.slots.local)
w = [e / sum(ex) for e in ex]
ok = True
for si, s in enumerate(eng.layout.segments):
    srcs = []
    for i in range(world * k):
        r, j = divmod(i, k)
        buf = allslots[r][j * nb:(j + 1) * nb]
        srcs.append(buf[s.byte_offset:s.byte_offset + s.nbytes].view(s.dtype).cpu())
    ref = agg.weighted_sum(srcs, w, exact=s.exact)
    got = eng.community_views[si].cpu()
    if s.exact:
        ok &= bool(torch.equal(got, ref))
    else:
        ok &= bool(torch.allclose(got.float(), ref.float(), atol=2e-2, rtol=2e-2))
if eng.direct_install:
    ok &= bool(torch.allclose(eng.arena.master[0].cpu(), eng.community_views[0].cpu().float()))
# NVLS path: the same aggregation with the reduction done inside the NVSwitch (slots pre-scaled by w_l)
nvls_ok = True
if eng.nvls:
    for j in range(k):
        wl = eng._static_weight(eng.learner_id(j))
        for si, s in enumerate(eng.layout.segments):
            if not s.exact:
                eng.slot_views[j][si].mul_(wl)
        eng._slot_scale[j] = wl
    torch.cuda.synchronize(); dist.barrier()
    for it in range(2):
        eng.aggregate_sync()
    torch.cuda.synchronize(); dist.barrier()
    for si, s in enumerate(eng.layout.segments):
        srcs = []
        for i in range(world * k):
            r, j = divmod(i, k)
            buf = allslots[r][j * nb:(j + 1) * nb]
            srcs.append(buf[s.byte_offset:s.byte_offset + s.nbytes].view(s.dtype).cpu())
        ref = agg.weighted_sum(srcs, w, exact=s.exact)
        got = eng.community_views[si].cpu()
        if s.exact:
            nvls_ok &= bool(torch.equal(got, ref))
        else:
            nvls_ok &= bool(torch.allclose(got.float(), ref.float(), atol=3e-2, rtol=3e-2))
    ok &= nvls_ok
    eng._slot_scale = [1.0] * k
t = torch.tensor([1.0 if ok else 0.0], device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MIN)
# one real round as well: train + publish + fused aggregation
data = [[(torch.randn(32, 3, 32, 32), torch.randint(0, 10, (32,))) for _ in range(2)] for _ in range(k)]
out = eng.run_round(data)
flat = eng.community.local[:nb].clone()
gathered = [torch.empty_like(flat) for _ in range(world)]
dist.all_gather(gathered, flat)
same = all(torch.equal(gathered[0], x) for x in gathered)
if rank == 0:
    print("MULTIGPU_OK" if (t.item() == 1.0 and same) else f"MULTIGPU_BAD agg={{t.item()}} same={{same}}", out["loss"],
          "nvls" if eng.nvls else "p2p-only")
eng.close()
dist.destroy_process_group()
"""


def _run(world, k, tmp_path):
    script = tmp_path / f"worker_{world}_{k}.py"
    script.write_text(WORKER.format(root=ROOT, k=k))
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                        "--master-addr", "127.0.0.1", "--master-port", str(port), str(script)],
                       capture_output=True, text=True, timeout=420)
    assert "MULTIGPU_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


@pytest.mark.parametrize("world,k", [(2, 1), (2, 2), (4, 1), (8, 1)])
def test_fused_aggregation_across_gpus(world, k, tmp_path):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    _run(world, k, tmp_path)


# ---- the fp32 master install under inter-rank skew (phase 3 reads only what its own flag partners wrote) ------------
STRESS_WORKER = r"""
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, {root!r})
rank = int(os.environ["RANK"]); lr = int(os.environ["LOCAL_RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
from metisfl_b200.parallel.engine import EngineConfig, FederationEngine
from metisfl_b200.models.zoo import ResNet18
eng = FederationEngine(ResNet18, EngineConfig(use_cuda_graph=False, nvls={nvls}), device=f"cuda:{{lr}}")
assert eng.direct_install and eng.arena.master[0] is not eng.arena.seg[0]
g = torch.Generator().manual_seed(7 + rank)
bad = 0
for it in range({iters}):
    # new slot contents every launch, so a stale read cannot hide behind identical values
    v = eng.slot_views[0][0]
    v.copy_(torch.randn(v.numel(), generator=g).to(v.dtype).to(v.device))
    if eng.nvls:
        wl = eng._static_weight(eng.learner_id(0)); v.mul_(wl); eng._slot_scale[0] = wl
    if rank == it % world:
        torch.cuda._sleep(int(2e6))     # ~1 ms of skew: this rank reaches the kernel late
    eng.aggregate_sync()
    torch.cuda.synchronize()
    if not torch.equal(eng.arena.master[0], eng.community_views[0].float()):
        bad += 1
t = torch.tensor([float(bad)], device="cuda"); dist.all_reduce(t)
from metisfl_b200.ops import aggregation as agg
to = agg.spin_timeouts()
if rank == 0:
    print("STRESS_OK" if (t.item() == 0 and to == 0) else f"STRESS_BAD stale_installs={{t.item()}} timeouts={{to}}",
          "nvls" if eng.nvls else "p2p")
eng.close()
dist.destroy_process_group()
"""

ASYNC_WORKER = r"""
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, {root!r})
rank = int(os.environ["RANK"]); lr = int(os.environ["LOCAL_RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
from metisfl_b200.parallel.engine import EngineConfig, FederationEngine
from metisfl_b200.parallel.simulate import AsyncFederation
from metisfl_b200.models.zoo import FashionMnistMLP
from metisfl_b200.ops.optim import OptimHyper
k = 2
res = {{}}
for rule in ("FedAsync", "FedRec"):
    cfg = EngineConfig(learners_per_gpu=k, model_dtype=torch.float32, rule=rule, use_cuda_graph=False,
                       use_b200_linear=False, optimizer=OptimHyper("VanillaSGD", lr=0.05))
    eng = FederationEngine(FashionMnistMLP, cfg, device=f"cuda:{{lr}}")
    g = torch.Generator().manual_seed(5)
    X = torch.randn(world * k, 32, 1, 28, 28, generator=g); Y = torch.randint(0, 10, (world * k, 32), generator=g)
    durations = [1.0 + 0.7 * i for i in range(world * k)]
    sim = AsyncFederation(eng, durations, lambda gl, n: [(X[gl], Y[gl])])
    sim.run(6 * world * k)
    eng.sync_folds()
    h = eng.async_health()
    flat = torch.cat([v.float().flatten() for v in eng.community_state_dict().values()])
    gathered = [torch.empty_like(flat) for _ in range(world)]
    dist.all_gather(gathered, flat)
    same = all(torch.equal(gathered[0], x) for x in gathered)
    # CPU oracle: the same schedule on one CPU engine hosting all learners (fp32 both sides)
    ok = True
    if rank == 0:
        ce = FederationEngine(FashionMnistMLP, EngineConfig(learners_per_gpu=world * k, model_dtype=torch.float32, rule=rule,
                              optimizer=OptimHyper("VanillaSGD", lr=0.05), standalone=True), device="cpu")
        cs = AsyncFederation(ce, durations, lambda gl, n: [(X[gl], Y[gl])])
        cs.run(6 * world * k)
        ref = torch.cat([v.float().flatten() for v in ce.community_state_dict().values()])
        # a task may have started from a NEWER community version on the GPUs than the simulated order promises, so
        # this is a sanity bound, not bit-equality
        ok = bool(torch.isfinite(flat).all()) and float((flat.cpu() - ref).abs().max()) < 0.05
    res[rule] = (same, h["torn_reads"], h["spin_timeouts"], ok, eng.version)
    eng.close()
if rank == 0:
    good = all(v[0] and v[1] == 0 and v[2] == 0 and v[3] and v[4] == 6 * world * k for v in res.values())
    print("ASYNC_OK" if good else "ASYNC_BAD", res)
dist.destroy_process_group()
"""


def _torchrun(script, world, timeout=420):
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    return subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                           "--master-addr", "127.0.0.1", "--master-port", str(port), str(script)],
                          capture_output=True, text=True, timeout=timeout)


@pytest.mark.parametrize("world,nvls", [(2, False), (4, False), (4, True), (8, False), (8, True)])
def test_master_install_is_never_stale_under_skew(world, nvls, tmp_path):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    script = tmp_path / "stress.py"
    script.write_text(STRESS_WORKER.format(root=ROOT, nvls=nvls, iters=200))
    r = _torchrun(script, world)
    assert "STRESS_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


@pytest.mark.parametrize("world", [2, 4, 8])
def test_async_folds_across_gpus(world, tmp_path):
    """K3 with peer loads, ready / version counters and rotating buffers: all ranks end with one community model."""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    script = tmp_path / "async.py"
    script.write_text(ASYNC_WORKER.format(root=ROOT))
    r = _torchrun(script, world)
    assert "ASYNC_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]
