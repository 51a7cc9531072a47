#!/usr/bin/env python
"""Launch the tcgen05 GEMM a few times on one shape - the command `ncu` wraps for K10."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from metisfl_b200.ops import gemm  # noqa: E402

m, n, k = [int(x) for x in (sys.argv[1:4] if len(sys.argv) > 3 else (8192, 3072, 1024))]
a = torch.randn(m, k, device="cuda").to(torch.bfloat16)
b = torch.randn(n, k, device="cuda").to(torch.bfloat16)
out = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
for _ in range(4):
    gemm.gemm_nt(a, b, out=out)
torch.cuda.synchronize()
print("profile_gemm ok", m, n, k)
