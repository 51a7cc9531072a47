#!/usr/bin/env python
"""Small driver that launches each hand-written kernel a few times on one GPU - the command that
`ncu` wraps (profiles/README.md).  Sizes = ResNet-18 arena (11.2 M bf16 parameters, 8 learner slots)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from metisfl_b200.ops import aggregation as agg  # noqa: E402
from metisfl_b200.ops.optim import OptimHyper, optimizer_step  # noqa: E402


def main():
    dev = torch.device("cuda", 0)
    n, L = 11183584, 8
    slots = [torch.randn(n, device=dev).to(torch.bfloat16) for _ in range(L)]
    out = torch.empty(n, dtype=torch.bfloat16, device=dev)
    pads = torch.zeros(1 << 16, dtype=torch.uint8, device=dev)
    master = torch.randn(n, device=dev)
    model = master.to(torch.bfloat16)
    grad = torch.randn(n, device=dev).to(torch.bfloat16)
    mom = torch.zeros(n, device=dev)
    pub = torch.empty(n, dtype=torch.bfloat16, device=dev)
    install = torch.empty(n, device=dev)
    h = OptimHyper("MomentumSGD", lr=0.01, momentum=0.9)
    q = agg.Quantifier([0, 1000, 5000000, n], dev)
    iters = int(sys.argv[1]) if len(sys.argv) > 1 else 5
    for _ in range(iters):
        agg.fused_aggregate([s.data_ptr() for s in slots], [1.0 / L] * L, [out.data_ptr()], [pads.data_ptr()], 0, 1,
                            [(0, n, torch.bfloat16, False)], install_f32=install)
        agg.weighted_sum(slots, [1.0 / L] * L, out=out)
        optimizer_step(master, model, grad, h, mom, publish=pub, zero_grad=True)
        q(out)
    torch.cuda.synchronize()
    print("profile_kernels ok")


if __name__ == "__main__":
    main()
