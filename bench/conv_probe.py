#!/usr/bin/env python
"""One convolution op of one ResNet-18 layer shape, a few launches - the target of `ncu -k regex:...` captures.

    python bench/conv_probe.py --op wgrad --c 512 --hw 4 --k 512 [--stride 1] [--r 3] [--iters 3]
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from metisfl_b200.ops import conv  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--op", default="wgrad", choices=["fprop", "dgrad", "wgrad", "halo_fprop", "halo_dgrad"])
ap.add_argument("--n", type=int, default=128)
ap.add_argument("--c", type=int, default=512)
ap.add_argument("--hw", type=int, default=4)
ap.add_argument("--k", type=int, default=512)
ap.add_argument("--r", type=int, default=3)
ap.add_argument("--stride", type=int, default=1)
ap.add_argument("--iters", type=int, default=3)
a = ap.parse_args()
pad = 1 if a.r == 3 else 0
x = torch.randn(a.n, a.c, a.hw, a.hw, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
w = torch.randn(a.k, a.c, a.r, a.r, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
oh = (a.hw + 2 * pad - a.r) // a.stride + 1
gy = torch.randn(a.n, a.k, oh, oh, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
fn = {"fprop": lambda: conv.conv2d_fprop(x, w, a.stride, pad),
      "dgrad": lambda: conv.conv2d_dgrad(gy, w, x.shape, a.stride, pad),
      "wgrad": lambda: conv.conv2d_wgrad(gy, x, w.shape, a.stride, pad),
      "halo_fprop": lambda: conv.conv3x3(x, w),
      "halo_dgrad": lambda: conv.conv3x3_dgrad(gy, w)}[a.op]
for _ in range(a.iters):
    fn()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20):
    fn()
e1.record()
e1.synchronize()
print(f"{a.op} c={a.c} hw={a.hw} k={a.k} r={a.r} s={a.stride}: {e0.elapsed_time(e1) / 20 * 1e3:.1f} us per call (warm L2)")
