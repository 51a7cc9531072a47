"""Python front-end of the fused optimizer-step + publish kernel (K2, ``csrc/kernels/optim.cu``)."""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional

import torch

from metisfl_b200.ops import _lib

KINDS = {"VanillaSGD": 0, "MomentumSGD": 1, "FedProx": 2, "Adam": 3, "AdamW": 4}


@dataclass
class OptimHyper:
    """Hyper-parameters of ``metisfl.OptimizerConfig`` (reference ``model.proto:116-152``)."""
    name: str = "VanillaSGD"
    lr: float = 0.01
    l1: float = 0.0
    l2: float = 0.0
    momentum: float = 0.0
    mu: float = 0.0
    beta1: float = 0.9
    beta2: float = 0.999
    eps: float = 1e-7
    weight_decay: float = 0.0

    @property
    def kind(self) -> int:
        return KINDS[self.name]


def _cpu_step(master, model, grad, s1, s2, anchor, publish, h: OptimHyper, step, grad_scale, publish_scale,
              zero_grad):
    g = grad.to(torch.float32) * grad_scale
    p = master
    k = h.kind
    if k == 0:
        g = g + h.l2 * p + h.l1 * torch.sign(p)
        p.sub_(h.lr * g)
    elif k == 1:
        s1.mul_(h.momentum).sub_(h.lr * g)
        p.add_(s1)
    elif k == 2:
        p.sub_(h.lr * (g + h.mu * (p - anchor)))
    else:
        if k == 4:
            p.mul_(1.0 - h.lr * h.weight_decay)
        s1.mul_(h.beta1).add_((1 - h.beta1) * g)
        s2.mul_(h.beta2).add_((1 - h.beta2) * g * g)
        mhat = s1 / (1 - h.beta1 ** step)
        vhat = s2 / (1 - h.beta2 ** step)
        p.sub_(h.lr * mhat / (vhat.sqrt() + h.eps))
    if model is not None:
        model.copy_(p.to(model.dtype))
    if publish is not None:
        publish.copy_((p * publish_scale).to(publish.dtype))
    if zero_grad:
        grad.zero_()


def optimizer_step(master: torch.Tensor, model: Optional[torch.Tensor], grad: torch.Tensor, hyper: OptimHyper,
                   state1: Optional[torch.Tensor] = None, state2: Optional[torch.Tensor] = None,
                   anchor: Optional[torch.Tensor] = None, publish: Optional[torch.Tensor] = None,
                   step: int = 1, grad_scale: float = 1.0, publish_scale: float = 1.0,
                   zero_grad: bool = False, stream=None):
    """One optimizer step over a flat arena; optionally writes the published copy in the same pass."""
    n = master.numel()
    assert master.dtype == torch.float32 and grad.numel() == n
    if not master.is_cuda:
        return _cpu_step(master, model, grad, state1, state2, anchor, publish, hyper, step, grad_scale,
                         publish_scale, zero_grad)
    cfg = _lib.OptimCfg(hyper.kind, hyper.lr, hyper.l1, hyper.l2, hyper.momentum, hyper.mu, hyper.beta1,
                        hyper.beta2, hyper.eps, hyper.weight_decay, int(step), float(grad_scale),
                        float(publish_scale))
    import ctypes as C
    rc = _lib.lib().mfl_optimizer_step(
        master.data_ptr(), _lib.ptr(model), _lib.wire(model) if model is not None else 0,
        grad.data_ptr(), _lib.wire(grad), _lib.ptr(state1), _lib.ptr(state2), _lib.ptr(anchor),
        _lib.ptr(publish), _lib.wire(publish) if publish is not None else 0, n, C.byref(cfg),
        int(zero_grad), _lib.stream_ptr(stream))
    _lib.check(rc, "mfl_optimizer_step")
