#!/usr/bin/env python
"""Launch the implicit-GEMM convolution (K11) a few times on one ResNet-18 shape - the command `ncu` wraps."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from metisfl_b200.ops import conv  # noqa: E402

n, c, hw, k, r, stride, pad = [int(v) for v in (sys.argv[1:8] if len(sys.argv) > 7 else (128, 128, 16, 128, 3, 1, 1))]
x = torch.randn(n, c, hw, hw, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
w = torch.randn(k, c, r, r, device="cuda").to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
for _ in range(4):
    y = conv.conv3x3(x, w) if os.environ.get("K15") == "1" else conv.conv2d_fprop(x, w, stride, pad)
torch.cuda.synchronize()
print("profile_conv ok", tuple(y.shape))
