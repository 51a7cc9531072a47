#!/usr/bin/env python
"""Fused BatchNorm(+residual)(+ReLU) (K14) vs the ATen sequence, forward + backward, ResNet-18 layer shapes, batch 128.
Bytes counted algorithmically: forward reads x (twice) [+ residual] and writes y; backward reads x, dy, y twice and
writes dx [+ dres]."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.nn as nn  # noqa: E402
import torch.nn.functional as F  # noqa: E402

from metisfl_b200.ops import bn as fbn  # noqa: E402

fbn.FUSED_FINALIZE = os.environ.get("BN_COOP", "1") == "1"
ONE = os.environ.get("BN_BENCH_ONE") == "1"  # single shape, few iterations: the ncu target


def timeit(fn, iters=20, warmup=5):
    """Capture ``fn`` (forward + backward) in a CUDA graph so the host-side autograd overhead is not measured."""
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(warmup):
            fn()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    if not ONE:
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            fn()
        fn = g.replay
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record(); e1.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3


rows = []
shapes = [(128, 64, 32, 32)] if ONE else [(128, 64, 32, 32), (128, 128, 16, 16), (128, 256, 8, 8), (128, 512, 4, 4)]
for shape in shapes:
    for use_res in (False, True):
        c = shape[1]
        mod = nn.BatchNorm2d(c).cuda()
        mk = lambda: torch.randn(shape, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        x, res, gy = mk().requires_grad_(), (mk().requires_grad_() if use_res else None), mk()

        def run(enabled):
            fbn.ENABLED = enabled
            y = fbn.bn_act(mod, x, res, relu=True)
            y.backward(gy)
            x.grad = None
            if res is not None:
                res.grad = None

        fbn.set_cluster_mode(0)
        t_fused = timeit(lambda: run(True), 3 if ONE else 20, 2 if ONE else 5)     # three launches per direction
        fbn.set_cluster_mode(1)
        t_cluster = timeit(lambda: run(True), 3 if ONE else 20, 2 if ONE else 5)   # one launch on thread-block clusters
        fbn.set_cluster_mode(-1)
        t_aten = None if ONE else timeit(lambda: run(False))
        nbytes = x.numel() * 2
        moved = nbytes * (3 + 7 + (2 if use_res else 0))
        rows.append({"shape": shape, "residual": use_res, "fused_us": t_fused, "cluster_us": t_cluster, "aten_us": t_aten,
                     "algorithmic_GBps": moved / t_fused / 1e3})
fbn.ENABLED = True
print(json.dumps({"bench": "bn_bench", "results": rows}))
