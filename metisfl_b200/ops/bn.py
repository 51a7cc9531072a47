"""Fused BatchNorm2d (+ residual add) (+ ReLU) on NHWC bf16 activations (K14, ``csrc/kernels/bn_relu.cu``).

``bn_act(bn_module, x, residual=None, relu=True)`` is the functional entry the model zoo calls; on CUDA with bf16
``channels_last`` activations it runs three launches forward (partial statistics, finalize, apply) and three backward,
replacing the four ATen batch-norm kernels, the ReLU / ReLU-backward / residual-add kernels and the running-statistics
launches of the eager path; everywhere else it is ``F.relu(bn(x) + residual)``.  ``use_fused_bn(model)`` re-classes
plain ``nn.BatchNorm2d`` layers (no activation fusion) for models that were not written against ``bn_act``.
Reference counterpart: the framework normalisation layers under ``model.fit`` (``keras_model_ops.py:156``).
"""
from __future__ import annotations

import ctypes as C

import torch
import torch.nn as nn
import torch.nn.functional as F

from metisfl_b200.ops import _lib

_BOUND = False
ENABLED = True  # flipped by tests / benchmarks to compare against the ATen path


def _bind():
    global _BOUND
    L = _lib.lib()
    if not _BOUND:
        vp, i, f, sz = C.c_void_p, C.c_int, C.c_float, C.c_size_t
        L.mfl_bn_workspace_bytes.restype = sz
        L.mfl_bn_workspace_bytes.argtypes = [i, i]
        L.mfl_bn_set_cluster_mode.argtypes = [i]
        L.mfl_bn_fwd.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, sz, vp, i, i, f, f, i, i, vp]
        L.mfl_bn_bwd.argtypes = [vp, vp, vp, vp, i, vp, vp, vp, vp, vp, vp, vp, sz, vp, i, i, i, vp]
        _BOUND = True
    return L


def set_cluster_mode(mode: int) -> int:
    """Back-end C of K14 (one launch per direction on thread-block clusters, ``csrc/kernels/bn_cluster.cuh``):
    -1 automatic (layers up to ~4.5 M elements), 0 never, 1 wherever a plan exists.  Returns the previous mode."""
    return _bind().mfl_bn_set_cluster_mode(int(mode))


def _ptr(t):
    return None if t is None else t.data_ptr()


_BAR_WORDS = 64 + 32 * 32  # kBarWords in csrc/kernels/bn_relu.cu
# One cooperative launch per direction (grid barriers between the phases) instead of three.  Measured on B200
# (bench/bn_bench.py, CUDA-graph replays, fwd+bwd): 51.8 vs 45.4 us at 128x64x32x32, 24.8 vs 22.8 us at 128x512x4x4 -
# the two grid barriers cost as much as the two launch boundaries they replace, so three launches stay the default.
FUSED_FINALIZE = False


def _tickets(bn: nn.Module, device) -> torch.Tensor:
    """Grid-barrier state {count, generation} x {forward, backward}, owned by the module: a module never runs
    concurrently with itself, which is the sharing rule of the barrier words (``include/metisfl_b200.h``)."""
    t = bn.__dict__.get("_mfl_tickets")
    if t is None or t.device != device:
        t = torch.zeros(2 * _BAR_WORDS, dtype=torch.int32, device=device)  # forward barrier, backward barrier
        bn.__dict__["_mfl_tickets"] = t
    return t


def supported(bn: nn.Module, x: torch.Tensor) -> bool:
    c = x.shape[1] if x.dim() == 4 else 0
    return (ENABLED and x.is_cuda and x.dtype == torch.bfloat16 and x.dim() == 4 and 8 <= c <= 2048 and c & (c - 1) == 0
            and (bn.weight is None or bn.weight.dtype == torch.float32)
            and (bn.momentum is not None or not bn.training)
            and (bn.running_mean is not None or bn.training)
            and (bn.running_mean is None or bn.running_mean.dtype == torch.float32) and _lib.available())


class _BNActFn(torch.autograd.Function):
    """x, residual: NHWC bf16.  ``state`` carries the module's buffers (updated in place, never differentiated)."""

    @staticmethod
    def forward(ctx, x, residual, weight, bias, state, training, momentum, eps, relu, tickets):
        L = _bind()
        n, c, h, w = x.shape
        m = n * h * w
        x = x.contiguous(memory_format=torch.channels_last)
        if residual is not None:
            residual = residual.contiguous(memory_format=torch.channels_last)
        y = torch.empty_like(x, memory_format=torch.channels_last)
        ws_bytes = L.mfl_bn_workspace_bytes(m, c)
        ws = torch.empty(ws_bytes // 4, dtype=torch.float32, device=x.device)
        rm, rv, nbt = state
        save_mean = save_invstd = None
        if training:
            save_mean = torch.empty(c, dtype=torch.float32, device=x.device)
            save_invstd = torch.empty(c, dtype=torch.float32, device=x.device)
        rc = L.mfl_bn_fwd(y.data_ptr(), x.data_ptr(), _ptr(residual), _ptr(weight), _ptr(bias), _ptr(rm), _ptr(rv),
                          _ptr(nbt) if training else None, _ptr(save_mean), _ptr(save_invstd), ws.data_ptr(), ws_bytes,
                          _ptr(tickets), m, c, float(eps), float(momentum or 0.0), int(training), int(relu),
                          _lib.stream_ptr())
        _lib.check(rc, "mfl_bn_fwd", 1 if (tickets is not None and training) else (3 if training else 2))
        if training:
            ctx.save_for_backward(x, y if relu else None, weight, save_mean, save_invstd)
            ctx.relu, ctx.has_res, ctx.has_bias = relu, residual is not None, bias is not None
            ctx.bias_ref, ctx.tickets = bias, tickets
        return y

    @staticmethod
    def backward(ctx, gy):
        L = _bind()
        x, y, weight, save_mean, save_invstd = ctx.saved_tensors
        n, c, h, w = x.shape
        m = n * h * w
        gy = gy.contiguous(memory_format=torch.channels_last)
        dx = torch.empty_like(x, memory_format=torch.channels_last)
        dres = torch.empty_like(x, memory_format=torch.channels_last) if ctx.has_res and ctx.needs_input_grad[1] else None
        ws_bytes = L.mfl_bn_workspace_bytes(m, c)
        ws = torch.empty(ws_bytes // 4, dtype=torch.float32, device=x.device)
        bias = ctx.bias_ref
        # accumulate straight into pre-allocated flat gradients (the arena's) when they exist: saves two
        # fp32 add launches per layer; otherwise hand the gradients back to autograd
        direct = (weight is not None and weight.grad is not None and weight.grad.dtype == torch.float32
                  and weight.grad.is_contiguous() and bias is not None and bias.grad is not None
                  and bias.grad.dtype == torch.float32 and bias.grad.is_contiguous())
        dgamma = dbeta = None
        if direct:
            pg, pb, acc = weight.grad.data_ptr(), bias.grad.data_ptr(), 1
        elif weight is not None:
            dgamma = torch.empty(c, dtype=torch.float32, device=x.device)
            dbeta = torch.empty(c, dtype=torch.float32, device=x.device)
            pg, pb, acc = dgamma.data_ptr(), dbeta.data_ptr(), 0
        else:
            pg = pb = None
            acc = 0
        rc = L.mfl_bn_bwd(dx.data_ptr(), _ptr(dres), pg, pb, acc, gy.data_ptr(), _ptr(y), x.data_ptr(), _ptr(weight),
                          save_mean.data_ptr(), save_invstd.data_ptr(), ws.data_ptr(), ws_bytes,
                          None if ctx.tickets is None else ctx.tickets.data_ptr() + 4 * _BAR_WORDS, m, c, int(ctx.relu),
                          _lib.stream_ptr())
        _lib.check(rc, "mfl_bn_bwd", 1 if ctx.tickets is not None else 3)
        if dgamma is not None:
            dgamma, dbeta = dgamma.to(weight.dtype), (dbeta.to(bias.dtype) if bias is not None else None)
        return dx, dres, dgamma, dbeta, None, None, None, None, None, None


def bn_act(bn: nn.Module, x: torch.Tensor, residual: torch.Tensor = None, relu: bool = True) -> torch.Tensor:
    """``relu(bn(x) + residual)`` with everything after the convolution in one pass over the activations."""
    if supported(bn, x) and (residual is None or (residual.dtype == x.dtype and residual.shape == x.shape)):
        use_batch_stats = bn.training or bn.running_mean is None
        state = (bn.running_mean, bn.running_var, bn.num_batches_tracked) if bn.track_running_stats else (None, None, None)
        return _BNActFn.apply(x, residual, bn.weight, bn.bias, state, use_batch_stats, bn.momentum, bn.eps, relu,
                              _tickets(bn, x.device) if FUSED_FINALIZE else None)
    y = bn(x)
    if residual is not None:
        y = y + residual
    return F.relu(y, inplace=True) if relu else y


class FusedBatchNorm2d(nn.BatchNorm2d):
    def forward(self, x):
        return bn_act(self, x, None, relu=False) if supported(self, x) else super().forward(x)


def use_fused_bn(module: nn.Module) -> nn.Module:
    for m in module.modules():
        if type(m) is nn.BatchNorm2d:
            m.__class__ = FusedBatchNorm2d
    return module
