#!/usr/bin/env python
"""Headline benchmark: federation round time + aggregated params/sec (BASELINE.json).

Config #2 of BASELINE.json: CIFAR-10-shaped ResNet-18, one learner per B200, synchronous FedAvg,
bf16.  One *step* = one full federation round on every GPU:

    local training (``--local-steps`` SGD steps of ``--batch`` synthetic 32x32 images: forward,
    backward, fused optimizer step; the last step publishes into the peer-visible slot)
    -> fused cross-GPU weighted aggregation (reduce-scatter + all-gather + fp32 install, one launch)

``value`` is learners x parameters aggregated per second over the whole job (weak scaling: one
learner per GPU).  Two timings are taken, each over exactly K rounds after W warm-up rounds,
bracketed by barrier + synchronize, max over ranks:

* device: inputs already resident, CUDA events around the K rounds;
* e2e:    ``FederationEngine.run_round`` (the public API) with every step's batch copied from
          pinned host memory and the round's losses read back to the host.

``--impl reference`` would run the unmodified reference; it cannot be installed offline (needs
Bazel + network, Python <= 3.10), so that arm reports ``unavailable``.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

REF_WHY = ("reference needs Bazel 4.2.1 + network fetches (protobuf/gRPC/PALISADE) and Python<=3.10; "
           "pip install from /root/reference fails at setup.py (see DESIGN.md)")


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--local-steps", type=int, default=4)
    ap.add_argument("--batch", type=int, default=128)
    ap.add_argument("--model", default="resnet18", choices=["resnet18", "mlp", "gpt2-medium", "gpt2-small"])
    ap.add_argument("--seq-len", type=int, default=1024)
    ap.add_argument("--learners-per-gpu", type=int, default=1)
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--own-conv", action="store_true", help="convolutions on the hand-written kernels instead of cuDNN")
    ap.add_argument("--no-stock", action="store_true", help="skip the stock PyTorch/cuDNN + NCCL arm (vs_baseline)")
    ap.add_argument("--no-service", action="store_true", help="skip the DriverSession/gRPC sub-record (1 GPU only)")
    ap.add_argument("--config", type=int, default=2, choices=[2, 3, 4, 5],
                    help="BASELINE.json north-star config: 2 = the headline (default), 3 = asynchronous ResNet-18 with "
                         "8 heterogeneous learners per GPU, 4 = ResNet-18 + CKKS secure aggregation, 5 = GPT-2 medium "
                         "with 128 streamed learners per GPU")
    return ap.parse_args()


def run_other_config(args):
    """Configs #3 / #4 / #5 live in ``examples/``; this runs the script in-process on every rank (same launch
    contract as the headline), samples clocks around it and re-emits rank 0's result in the bench.py line format."""
    import contextlib
    import io
    import runpy
    here = os.path.dirname(os.path.abspath(__file__))
    script, argv, metric, pick = {
        3: ("resnet_async.py", ["--rule", "FedAsync", "--updates", str(max(32, args.steps * 8))],
            "community_updates_per_sec (asynchronous FedAsync, 8 heterogeneous learners per GPU)",
            lambda r: (r["updates_per_s"], "updates/s", r["ms_per_update"], "weak")),
        4: ("resnet_secure.py", ["--rounds", str(args.steps)],
            "secure_round_time_s (ResNet-18, CKKS encrypt -> fused ciphertext reduce over NVLink -> decrypt)",
            lambda r: (r["round_time_s"], "s", r["round_time_s"] * 1e3, "weak")),
        5: ("gpt2_many_learners.py", ["--rounds", "1"],
            "aggregated_params_per_sec (GPT-2 medium, 128 streamed learners per GPU)",
            lambda r: (r["aggregated_params_per_s"], "params/s", r["round_time_s"] * 1e3, "weak")),
    }[args.config]
    rank = int(os.environ.get("RANK", "0"))
    sampler = ClockSampler(int(os.environ.get("LOCAL_RANK", "0"))) if rank == 0 else None
    sys.path.insert(0, os.path.join(here, "examples"))
    old_argv, buf = sys.argv, io.StringIO()
    sys.argv = [script] + argv
    try:
        with contextlib.redirect_stdout(buf):
            runpy.run_path(os.path.join(here, "examples", script), run_name="__main__")
    finally:
        sys.argv = old_argv
    clocks = sampler.stop() if sampler else None
    if rank != 0:
        return 0
    lines = [l for l in buf.getvalue().splitlines() if l.startswith("{")]
    res = json.loads(lines[-1])
    value, unit, ms, scaling = pick(res)
    print(json.dumps({"metric": metric, "value": value, "unit": unit, "n_gpus": int(os.environ.get("WORLD_SIZE", "1")),
                      "steps": args.steps, "warmup": 1, "ms_per_step": ms, "higher_is_better": args.config != 4,
                      "scaling": scaling, "vs_baseline": None, "dtype": "bf16",
                      "data": "synthetic (random batches of the named shape, random-init weights)", "impl": "ours",
                      "config": {"north_star_config": args.config, "script": "examples/" + script, **res},
                      "clocks": clocks, "e2e": None, "gpu_launches": None}))
    return 0


class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap,utilization.gpu")

    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "20", "-i", str(index)], stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            pass

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons, busy = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f:
            parts = [c.strip() for c in line.split(",")]
            if len(parts) < 9:
                continue
            try:
                sm.append(float(parts[1])); mx.append(float(parts[2]))
            except ValueError:
                continue
            try:
                busy.append(float(parts[9]) >= 10.0 if len(parts) > 9 else True)
            except ValueError:
                busy.append(True)
            for n, v in zip(names, parts[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        os.unlink(self.f.name)
        if sm:
            # "under load": samples taken while the GPU reported work (a sampler that brackets a whole script also
            # sees its idle set-up phases at the idle clock)
            loaded = [v for v, b in zip(sm, busy) if b] or sm
            out = {"sm_mhz": statistics.median(loaded), "sm_max_mhz": max(mx), "reasons": sorted(reasons),
                   "samples": len(sm), "samples_under_load": len(loaded)}
        return out


def main():
    args = parse()
    if args.impl == "reference":
        print(json.dumps({"impl": "reference", "unavailable": REF_WHY}))
        return 0
    if args.config != 2:
        return run_other_config(args)

    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        print(json.dumps({"error": "bench.py needs a CUDA device (B200)"}))
        return 1
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    assert world == args.gpus or world == 1, f"--gpus {args.gpus} but WORLD_SIZE={world}"

    from metisfl_b200.models import zoo
    from metisfl_b200.ops import _lib
    from metisfl_b200.ops.optim import OptimHyper
    from metisfl_b200.parallel.engine import EngineConfig, FederationEngine

    torch.backends.cudnn.benchmark = True
    cfg = EngineConfig(learners_per_gpu=args.learners_per_gpu, model_dtype=torch.bfloat16,
                       optimizer=OptimHyper("MomentumSGD", lr=0.02, momentum=0.9), rule="FedAvg",
                       scaling="NumTrainingExamples", use_cuda_graph=not args.no_graph, own_conv=args.own_conv)
    B, S = args.batch, args.local_steps
    if args.model == "resnet18":
        factory, loss_fn = zoo.ResNet18, None
        def make_batch(g):
            x = torch.randn(B, 3, 32, 32, generator=g).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
            return x.pin_memory(), torch.randint(0, 10, (B,), generator=g).pin_memory()
    elif args.model == "mlp":
        factory, loss_fn = zoo.FashionMnistMLP, None
        def make_batch(g):
            return (torch.randn(B, 1, 28, 28, generator=g).to(torch.bfloat16).pin_memory(),
                    torch.randint(0, 10, (B,), generator=g).pin_memory())
    else:
        factory = zoo.gpt2_medium if args.model == "gpt2-medium" else zoo.gpt2_small
        loss_fn = zoo.lm_loss
        cfg.channels_last = False
        def make_batch(g):
            t = torch.randint(0, 50257, (B, args.seq_len), generator=g).pin_memory()
            return t, t
    kw = {"loss_fn": loss_fn} if loss_fn is not None else {}
    eng = FederationEngine(factory, cfg, device=device, **kw)
    k = cfg.learners_per_gpu
    g = torch.Generator().manual_seed(1000 + rank)
    # a small pool of distinct pinned batches per local learner, cycled across rounds
    pool = [[make_batch(g) for _ in range(S * 2)] for _ in range(k)]
    def host_batches(r):
        return [[pool[j][(r * S + s) % len(pool[j])] for s in range(S)] for j in range(k)]
    dev_pool = [[(x.to(device), y.to(device)) for x, y in pj] for pj in pool]
    def dev_batches(r):
        return [[dev_pool[j][(r * S + s) % len(dev_pool[j])] for s in range(S)] for j in range(k)]

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def device_round(r):
        for j in range(k):
            eng.train_learner(j, dev_batches(r)[j])
        eng.aggregate_sync()

    # ---- warm-up (also captures the CUDA graph) -------------------------------------------------
    for r in range(max(3, args.warmup)):
        eng.run_round(host_batches(r))
    for r in range(2):
        device_round(r)
    barrier()

    sampler = ClockSampler(local_rank) if rank == 0 else None
    # ---- e2e: public API, H2D every step, loss D2H every round ----------------------------------
    barrier()
    t0 = time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    last = None
    for r in range(args.steps):
        last = eng.run_round(host_batches(r))
    e1.record()
    barrier()
    e2e_ms_wall = (time.perf_counter() - t0) * 1e3 / args.steps
    e2e_ms = e0.elapsed_time(e1) / args.steps
    e2e_ms = max(e2e_ms, e2e_ms_wall)  # the host read-back is part of the contract

    # ---- device-timed: inputs resident, CUDA events --------------------------------------------
    barrier()
    launches0 = _lib.LAUNCHES
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for r in range(args.steps):
        device_round(r)
    e1.record()
    barrier()
    dev_ms = e0.elapsed_time(e1) / args.steps
    launches = _lib.LAUNCHES - launches0
    clocks = sampler.stop() if sampler else None

    # ---- the product by itself: K1 only vs NCCL `buf *= w; all_reduce`, same buffers, same clocks ----------
    nb = eng.layout.total_bytes
    def time_loop(fn, iters=30):
        barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        fn(); fn()
        barrier()
        a.record()
        for _ in range(iters):
            fn()
        b.record()
        barrier()
        return a.elapsed_time(b) / iters
    k1_ms = time_loop(eng.aggregate_sync)
    flat = torch.empty(eng.layout.segments[0].numel, dtype=torch.bfloat16, device=device).normal_()
    def nccl_fedavg():
        flat.mul_(1.0 / max(1, world))
        if world > 1:
            dist.all_reduce(flat)
    nccl_ms = time_loop(nccl_fedavg)

    # ---- check what the timed rounds produced: every rank holds the same community model, and it is the weighted
    # average of the published slots (fp32 oracle through NCCL) within bf16 rounding --------------------------------
    eng.aggregate_sync()
    torch.cuda.synchronize()
    comm = eng.community_views[0].float()
    mine = eng.slot_views[0][0].float() if k == 1 else None
    check = {"identical_across_ranks": True, "max_abs_err_vs_fp32_oracle": None}
    digest = torch.stack([eng.community_bytes()[:nb].view(torch.int32).to(torch.int64).sum(),
                          eng.community_bytes()[:nb].view(torch.int32)[::7].to(torch.int64).sum()])
    if world > 1:
        all_d = [torch.empty_like(digest) for _ in range(world)]
        dist.all_gather(all_d, digest)
        check["identical_across_ranks"] = all(torch.equal(all_d[0], d) for d in all_d)
    if mine is not None:
        scale = eng._slot_scale[0]
        w_me = eng._static_weight(eng.learner_id(0))
        oracle = mine * (w_me / scale)      # NVLS mode publishes w_l * theta_l already
        if world > 1:
            dist.all_reduce(oracle)
        err = (oracle - comm).abs().max().item()
        ref = oracle.abs().max().item()
        check["max_abs_err_vs_fp32_oracle"] = err
        check["oracle_abs_max"] = ref
        assert err <= 2.0 ** -7 * max(ref, 1.0), f"community model is off the fp32 oracle by {err} (|max| {ref})"
    assert check["identical_across_ranks"], "ranks disagree on the community model"

    # ---- stock arm: plain PyTorch / cuDNN learner + NCCL FedAvg, the whole round -----------------------------
    stock = None
    if not args.no_stock and args.model == "resnet18":
        sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "bench"))
        from stock import StockLearner
        res = {}
        for mode, graph in (("eager", False), ("cuda_graph", True)):
            sl = StockLearner(device, (B, 3, 32, 32), graph=graph)
            if graph and sl.graph is None:
                res[mode] = None
                continue
            def stock_round(r):
                for s_ in range(S):
                    x, y = dev_pool[0][(r * S + s_) % len(dev_pool[0])]
                    sl.step(x, y)
                sl.fedavg(1.0 / world, world)
            for r in range(3):
                stock_round(r)
            barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for r in range(args.steps):
                stock_round(r)
            b.record()
            barrier()
            res[mode] = a.elapsed_time(b) / args.steps
            del sl
        stock = res

    t = torch.tensor([dev_ms, e2e_ms, k1_ms, nccl_ms] + [stock[m] if stock and stock[m] else 0.0
                                                         for m in ("eager", "cuda_graph")],
                     device=device, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, e2e_ms, k1_ms, nccl_ms, stock_eager_ms, stock_graph_ms = t.tolist()
    n_params = eng.layout.num_params
    learners = eng.num_learners
    value = learners * n_params / (dev_ms * 1e-3)
    e2e_value = learners * n_params / (e2e_ms * 1e-3)
    stock_best_ms = min([v for v in (stock_eager_ms, stock_graph_ms) if v > 0], default=0.0)
    vs_baseline = (stock_best_ms / dev_ms) if stock_best_ms > 0 else None   # = value / stock value (same numerator)
    link_gbs = 770.0  # measured peer-copy bandwidth per direction (B200_PROFILING.md)
    # roofline of the aggregation alone = INBOUND NVLink bytes per GPU / measured link bandwidth.  Peer loads / stores:
    # this GPU reads its shard from world-1 peers and receives world-1 result shards = 2 (world-1)/world of a model.
    # In-switch reduction: the reduced shard (1/world) plus the multicast all-gather (a whole model, own shard
    # included).  At world 1 it is an HBM copy of the slots.
    seg_bytes = sum(sg.nbytes for sg in eng.layout.segments)
    if world > 1:
        inbound = seg_bytes * (1.0 / world + 1.0) if eng.nvls else seg_bytes * 2.0 * (world - 1) / world
        roof_ms = inbound / (link_gbs * 1e9) * 1e3
    else:
        roof_ms = seg_bytes * (k + 1) / (6553.9e9) * 1e3
    x0, y0 = pool[0][0]
    h2d = k * S * (x0.numel() * x0.element_size() + (y0.numel() * y0.element_size() if y0 is not x0 else 0))
    d2h = 4 * k
    peak_mb = torch.cuda.max_memory_allocated(device) / 2**20
    agg_mode = "nvls" if eng.nvls else ("p2p" if world > 1 else "local")
    # ---- the same round through DriverSession -> controller gRPC -> learner (the reference user's call sequence) ----
    service = None
    if world == 1 and not args.no_service and args.model == "resnet18":
        eng.close()
        eng = None
        try:
            sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "bench"))
            import contextlib
            import io
            from service_bench import service_rounds
            with contextlib.redirect_stdout(io.StringIO()):
                service = service_rounds(True, rounds=max(8, min(args.steps, 20)), warmup=4, local_steps=S, batch=B,
                                         cuda_device=local_rank)
                host = service_rounds(False, rounds=6, warmup=2, local_steps=S, batch=B, cuda_device=local_rank)
            if service.get("ms_per_round"):
                service["value"] = n_params / (service["ms_per_round"] * 1e-3)
                service["unit"] = "params/s"
            # the reference's architecture on this box: same learner, same GPU, but models travel as protobuf bytes
            # through the controller and are averaged on the host
            service["host_path_ms_per_round"] = host.get("ms_per_round")
            service["host_path_controller_rx_model_bytes"] = host.get("controller_rx_model_bytes")
            if service.get("ms_per_round") and host.get("ms_per_round"):
                service["device_path_speedup_over_host_path"] = host["ms_per_round"] / service["ms_per_round"]
        except Exception as e:  # the headline must not die with the side measurement
            service = {"error": repr(e)[:300]}
    if rank == 0:
        print(json.dumps({
            "metric": "aggregated_params_per_sec (federation round: local training + fused FedAvg)",
            "value": value, "unit": "params/s", "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup),
            "ms_per_step": dev_ms, "round_time_s": dev_ms * 1e-3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": vs_baseline, "dtype": "bf16", "data": "synthetic (random CIFAR-shaped batches, random-init weights)",
            "impl": "ours",
            "config": {"model": args.model, "params": n_params, "learners": learners,
                       "learners_per_gpu": k, "protocol": "synchronous", "rule": "FedAvg",
                       "scaling_factor": "NumTrainingExamples", "local_steps": S, "global_batch": B * learners,
                       "batch_per_learner": B, "seq_len": args.seq_len if args.model.startswith("gpt2") else None,
                       "parallelism": f"fed-dp{world}", "optimizer": "MomentumSGD(fused K2)",
                       "cuda_graph": not args.no_graph,
                       "convolutions": "own tcgen05 kernels for every convolution (K11/K11s/K15/K15t/K15w)" if args.own_conv
                                       else ("per-shape dispatch: own tcgen05 gradients where measured faster "
                                             "(stride-2 / 1x1 dgrad, K15w + 1x1 wgrad), cuDNN elsewhere"
                                             if os.environ.get("MFL_CONV_HYBRID", "1") != "0" else "cuDNN"),
                       "l2": f"no flush: per-round working set (peak {peak_mb:.0f} MiB allocated) exceeds the 126 MB L2"},
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "params/s", "ms_per_step": e2e_ms,
                    "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "api": "FederationEngine.run_round(pinned host batches) -> host losses"},
            "baseline_arm": {"what": "stock PyTorch/cuDNN ResNet-18 (bf16 channels_last, fused SGD) + NCCL FedAvg "
                                     "(buf *= w; all_reduce), same batch / local steps / GPUs, measured in this process; "
                                     "vs_baseline = its best round time / ours (the reference publishes no number and "
                                     "cannot be installed offline)",
                             "ms_per_round_eager": stock_eager_ms or None,
                             "ms_per_round_cuda_graph": stock_graph_ms or None},
            "aggregation": {"k1_fused_ms": k1_ms, "nccl_scale_allreduce_ms": nccl_ms,
                            "k1_vs_nccl": nccl_ms / k1_ms if (k1_ms > 0 and world > 1) else None,  # world 1: no collective to compare with
                            "mode": agg_mode,
                            "roofline_ms": roof_ms, "fraction_of_roofline": roof_ms / k1_ms if k1_ms > 0 else None,
                            "roofline_def": "inbound NVLink bytes per GPU (shard reads + result shards; in-switch "
                                            "reduction: reduced shard + multicast all-gather) / 770 GB/s measured "
                                            "(world 1: slot reads + community write / measured HBM copy)"},
            "community_check": check,
            "e2e_service": service,
            "gpu_launches": launches, "gpu_launches_per_round": launches / args.steps,
            "last_loss": last["loss"] if last else None,
        }))
    if eng is not None:
        eng.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
