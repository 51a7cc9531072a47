#!/usr/bin/env python
"""The network stem (3 -> 64 channels, 3x3, 32x32 images, batch 128): K11s forward / weight gradient against cuDNN.
CUDA-event timing, L2 flushed between iterations."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from metisfl_b200.ops import conv as C  # noqa: E402


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
    return tot / iters * 1e3


torch.backends.cudnn.benchmark = True
n, c, h, k = 128, 3, 32, 64
x = torch.randn(n, c, h, h, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
w = torch.randn(k, c, 3, 3, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
gy = torch.randn(n, k, h, h, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
res = {"shape": [n, c, h, h, k]}
res["fprop_cudnn_us"] = timeit(lambda: torch.nn.functional.conv2d(x, w, padding=1))
res["fprop_own_us"] = timeit(lambda: C._StemFn.apply(x, w))            # FMA kernel (K11s)
res["fprop_im2col_gemm_us"] = timeit(lambda: C._StemGemmFn.apply(x, w))  # patch matrix + K10
res["wgrad_cudnn_us"] = timeit(lambda: torch.ops.aten.convolution_backward(
    gy, x, w, None, [1, 1], [1, 1], [1, 1], False, [0, 0], 1, [False, True, False]))
# the weight-gradient kernel pair called directly (through autograd the Python overhead would dominate the timing)
from metisfl_b200.ops import _lib  # noqa: E402
gw = torch.empty_like(w)
def own_wgrad():
    st = torch.cuda.current_stream()
    ws = C._workspace(x.device, 0, st)
    rc = C._bind().mfl_conv_stem_wgrad_bf16(gw.data_ptr(), gy.data_ptr(), x.data_ptr(), ws.data_ptr(), n, h, h, c, k, 0,
                                            st.cuda_stream)
    _lib.check(rc, "mfl_conv_stem_wgrad_bf16", launches=2)
res["wgrad_own_us"] = timeit(own_wgrad)
from metisfl_b200.ops import gemm as G  # noqa: E402
patches = torch.empty((n * h * h, 32), dtype=torch.bfloat16, device="cuda")
g2 = gy.permute(0, 2, 3, 1).reshape(n * h * h, k)
res["wgrad_im2col_gemm_us"] = timeit(lambda: G.gemm_ex(g2, patches, a_mn=True, b_mn=True))   # split-k K10 + finalize
ref = torch.nn.functional.conv2d(x.float(), w.float(), padding=1)
res["fprop_max_err"] = float((C._StemFn.apply(x, w).float() - ref).abs().max())
print(json.dumps(res))
