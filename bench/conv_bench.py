#!/usr/bin/env python
"""Implicit-GEMM conv family (K11: forward, data gradient, weight gradient) vs cuDNN on the CIFAR ResNet-18 layer
shapes, batch 128, L2 flushed between timed iterations."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402

from metisfl_b200.ops import conv  # noqa: E402

torch.backends.cudnn.benchmark = True


def timeit(fn, iters=20, warmup=5):
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(iters):
        flush.fill_(0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); e1.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / iters


rows = []
for (c, h, k, r, st, p) in [(64, 32, 64, 3, 1, 1), (64, 32, 128, 3, 2, 1), (128, 16, 128, 3, 1, 1), (128, 16, 256, 3, 2, 1),
                            (256, 8, 256, 3, 1, 1), (256, 8, 512, 3, 2, 1), (512, 4, 512, 3, 1, 1), (64, 32, 128, 1, 2, 0)]:
    n = 128
    x = torch.randn(n, c, h, h, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    w = torch.randn(k, c, r, r, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    t_ours = timeit(lambda: conv.conv2d_fprop(x, w, st, p))
    t_halo = timeit(lambda: conv.conv3x3(x, w)) if (r == 3 and st == 1 and p == 1 and conv.conv3x3_supported(x, w)) else None
    t_cudnn = timeit(lambda: F.conv2d(x, w, None, st, p))
    tt_f = tt_d = None
    if r == 3 and st == 1:
        tt_f = timeit(lambda: conv.conv3x3_t(x, w))
    oh = (h + 2 * p - r) // st + 1
    gy = torch.randn(n, k, oh, oh, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    cb = lambda mask: torch.ops.aten.convolution_backward(gy, x, w, None, [st, st], [p, p], [1, 1], False, [0, 0], 1, mask)
    t_dgrad = timeit(lambda: conv.conv2d_dgrad(gy, w, x.shape, st, p)) if conv.dgrad_supported(x.shape, w.shape, st, p) else None
    t_dgrad_cudnn = timeit(lambda: cb([True, False, False]))
    if r == 3 and st == 1:
        tt_d = timeit(lambda: conv.conv3x3_t_dgrad(gy, w))
    t_wgrad = timeit(lambda: conv.conv2d_wgrad(gy, x, w.shape, st, p))
    t_wgrad_cudnn = timeit(lambda: cb([False, True, False]))
    fl = 2.0 * n * oh * oh * k * c * r * r
    rows.append({"C": c, "HW": h, "K": k, "R": r, "stride": st, "ours_us": t_ours * 1e3, "cudnn_us": t_cudnn * 1e3,
                 "k15_us": t_halo * 1e3 if t_halo else None, "k15_tflops": fl / t_halo / 1e9 if t_halo else None,
                 "ours_tflops": fl / t_ours / 1e9, "cudnn_tflops": fl / t_cudnn / 1e9,
                 "dgrad_us": t_dgrad * 1e3 if t_dgrad else None, "dgrad_cudnn_us": t_dgrad_cudnn * 1e3,
                 "wgrad_us": t_wgrad * 1e3, "wgrad_cudnn_us": t_wgrad_cudnn * 1e3,
                 "k15t_fprop_us": tt_f * 1e3 if tt_f else None, "k15t_dgrad_us": tt_d * 1e3 if tt_d else None,
                 "fprop_vs_cudnn": t_cudnn / t_ours, "dgrad_vs_cudnn": t_dgrad_cudnn / t_dgrad if t_dgrad else None,
                 "wgrad_vs_cudnn": t_wgrad_cudnn / t_wgrad})
print(json.dumps({"bench": "conv_bench", "batch": 128, "results": rows}))
