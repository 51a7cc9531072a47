#!/usr/bin/env python
"""Aggregation-only benchmark: fused K1 kernel vs the `nccl-fedavg` baseline.

Per model size (ResNet-18 = 11.2 M, GPT-2-medium = 355 M parameters, bf16) and world size:
  * ours : ONE launch of mfl_fused_aggregate per rank (weighted reduce-scatter + all-gather over NVLink
           peer loads/stores, in-kernel barriers);
  * nvls : the same launch with the reduction done inside the NVSwitch (multimem.ld_reduce on pre-scaled slots)
           and the result multicast-stored to every GPU (multimem.st);
  * nccl : per rank `buf *= w` then `dist.all_reduce(SUM)` on the same flat bf16 buffer.
Timed with CUDA events over `--iters` launches after warm-up, max over ranks; a buffer larger than
L2 is written between iterations when `--flush` is given.  Reports params/s, per-rank NVLink bytes
and the fraction of the link roofline (bytes over NVLink / 770 GB/s measured peer-copy peak), or of
the measured HBM copy peak for the single-GPU case.

Launch: `python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 bench/agg_bench.py`.
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

NVLINK_GBS = 770.0  # measured peer copy per direction (B200_PROFILING.md)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--learners-per-gpu", type=int, default=1)
    ap.add_argument("--flush", action="store_true")
    ap.add_argument("--sizes", default="11183582,354823168")
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    lr = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    from metisfl_b200.ops import aggregation as agg
    from metisfl_b200.parallel.fabric import Fabric
    fab = Fabric(dev)
    hbm = 6553.9
    try:
        with open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")) as f:
            hbm = json.load(f)["hbm_gbs"]
    except Exception:
        pass
    k = args.learners_per_gpu
    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev) if args.flush else None
    results = []
    for n in [int(s) for s in args.sizes.split(",")]:
        n = (n + 7) // 8 * 8
        nbytes = n * 2
        slots = fab.alloc(nbytes * k, multicast=True)
        comm = fab.alloc(nbytes, multicast=True)
        pads = fab.signal_pads(296)
        slots.local.view(torch.bfloat16)[:n * k].copy_(torch.randn(n * k, device=dev).to(torch.bfloat16))
        L = world * k
        w = [1.0 / L] * L
        srcs = [slots.ptrs[r] + j * nbytes for r in range(world) for j in range(k)]
        seg = [(0, n, torch.bfloat16, False)]
        fab.host_barrier()

        def ours():
            agg.fused_aggregate(srcs, w, comm.ptrs, pads.ptrs, rank, world, seg, n_blocks=296)

        have_mc = bool(world > 1 and slots.mc_ptr and comm.mc_ptr and k <= 8)

        def nvls():  # slots are taken as already scaled by w (what K2's publish_scale produces)
            agg.fused_aggregate(srcs, w, comm.ptrs, pads.ptrs, rank, world, seg, n_blocks=296,
                                mc_src_ptrs=[slots.mc_ptr + j * nbytes for j in range(k)], mc_out_ptr=comm.mc_ptr)

        flat = slots.local.view(torch.bfloat16)[:n].clone()

        def nccl():
            flat.mul_(1.0 / L)
            if world > 1:
                dist.all_reduce(flat)

        def timeit(fn):
            for _ in range(args.warmup):
                fn()
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            tot = 0.0
            for _ in range(args.iters):
                if flush_buf is not None:
                    flush_buf.fill_(1)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                fn()
                e1.record()
                e1.synchronize()
                tot += e0.elapsed_time(e1)
            t = torch.tensor([tot / args.iters], device=dev, dtype=torch.float64)
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t)

        t_ours = timeit(ours)
        # correctness spot check against a local reference of shard 0
        torch.cuda.synchronize()
        ref = None
        if world == 1:
            ref = sum(slots.local.view(torch.bfloat16)[j * n:(j + 1) * n][:4096].float() for j in range(k)) / L
            got = comm.local.view(torch.bfloat16)[:4096].float()
            assert torch.allclose(got, ref, atol=2e-2, rtol=2e-2)
        t_nvls, nvls_err = None, None
        if have_mc:
            ours(); torch.cuda.synchronize(); dist.barrier()
            want = comm.local.view(torch.bfloat16)[:n].float().clone() * L   # unweighted sum = L x the average
            t_nvls = timeit(nvls)
            torch.cuda.synchronize(); dist.barrier()
            nvls_err = float((comm.local.view(torch.bfloat16)[:n].float() - want).abs().max())
        t_nccl = timeit(nccl)
        # bytes that must cross NVLink per rank: read (L - k) remote shards of n/world, write (world-1) shards
        shard = nbytes / world
        link_bytes = (L - k) * shard + (world - 1) * shard
        hbm_bytes = L * shard + world * shard if world == 1 else None
        roof_ms = (link_bytes / (NVLINK_GBS * 1e9) * 1e3) if world > 1 else (hbm_bytes / (hbm * 1e9) * 1e3)
        results.append({
            "params": n, "learners": L, "world": world, "ours_ms": t_ours, "nccl_fedavg_ms": t_nccl,
            "speedup_vs_nccl": t_nccl / t_ours, "ours_params_per_s": L * n / (t_ours * 1e-3),
            "roofline_ms": roof_ms, "fraction_of_roofline": roof_ms / t_ours,
            "nvls_ms": t_nvls, "nvls_max_abs_err_vs_p2p": nvls_err,
            # per direction: the switch pulls (world-1)/world of each of my k slots, I push my 1/world shard once
            "nvls_link_bytes_per_rank": ((world - 1) * k + 1) * shard if have_mc else None,
            "nvls_fraction_of_link_roofline": (((world - 1) * k + 1) * shard / (NVLINK_GBS * 1e9) * 1e3 / t_nvls)
            if t_nvls else None,
            "roofline_kind": "NVLink 770 GB/s measured peer copy" if world > 1 else f"HBM copy {hbm} GB/s of measured",
        })
        del slots, comm
    if rank == 0:
        print(json.dumps({"bench": "agg_bench", "flush_l2": args.flush, "results": results}))
    fab.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
