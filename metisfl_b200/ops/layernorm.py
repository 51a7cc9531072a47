"""Fused residual-add + LayerNorm (K18, ``csrc/kernels/layernorm.cu``).

``add_layer_norm(a, b, ln)`` returns ``(r, h)`` with ``r = a + b`` (the new residual stream) and ``h = LN(r)``;
``layer_norm(x, ln)`` is the variant without the add.  In backward one kernel produces the gradient of the residual
stream - ``dr = dr_in + LN'(dh)``, which autograd then hands to BOTH ``a`` and ``b`` without any add - and per-CTA
partial rows of ``dgamma`` / ``dbeta``.  Off the GPU (or for unsupported widths) it is plain ``F.layer_norm``.
Reference counterpart: the framework normalisation layers under ``model.fit`` (``keras_model_ops.py:156``).
"""
from __future__ import annotations

import ctypes as C

import torch
import torch.nn as nn
import torch.nn.functional as F

from metisfl_b200.ops import _lib

_BOUND = False
ENABLED = True


def _bind():
    global _BOUND
    L = _lib.lib()
    if not _BOUND:
        vp, i, f = C.c_void_p, C.c_int, C.c_float
        L.mfl_ln_supported.argtypes = [i]
        L.mfl_ln_partial_rows.argtypes = [i]
        L.mfl_add_ln_fwd.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, i, i, i, f, vp]
        L.mfl_add_ln_bwd.argtypes = [vp, vp, vp, vp, vp, vp, i, vp, vp, i, i, vp]
        _BOUND = True
    return L


def _ptr(t):
    return None if t is None else t.data_ptr()


def supported(x: torch.Tensor, ln: nn.LayerNorm) -> bool:
    d = x.shape[-1]
    if not (ENABLED and x.is_cuda and x.dtype == torch.bfloat16 and _lib.available() and len(ln.normalized_shape) == 1
            and ln.normalized_shape[0] == d):
        return False
    w, b = ln.weight, ln.bias
    if w is not None and w.dtype not in (torch.bfloat16, torch.float32):
        return False
    if w is not None and b is not None and w.dtype != b.dtype:
        return False
    return bool(_bind().mfl_ln_supported(d))


class _AddLayerNormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, a, b, weight, bias, eps):
        L = _bind()
        d = a.shape[-1]
        a2 = a.reshape(-1, d).contiguous()
        b2 = b.reshape(-1, d).contiguous() if b is not None else None
        rows = a2.shape[0]
        r = torch.empty_like(a2) if b2 is not None else a2
        h = torch.empty_like(a2)
        mean = torch.empty(rows, dtype=torch.float32, device=a.device)
        rstd = torch.empty(rows, dtype=torch.float32, device=a.device)
        pbf16 = int(weight is not None and weight.dtype == torch.bfloat16)
        rc = L.mfl_add_ln_fwd(r.data_ptr() if b2 is not None else None, h.data_ptr(), mean.data_ptr(), rstd.data_ptr(),
                              a2.data_ptr(), _ptr(b2), _ptr(weight), _ptr(bias), pbf16, rows, d, float(eps),
                              _lib.stream_ptr())
        _lib.check(rc, "mfl_add_ln_fwd")
        ctx.save_for_backward(r, weight, mean, rstd)
        ctx.has_b, ctx.has_bias, ctx.shape, ctx.pbf16 = b is not None, bias is not None, a.shape, pbf16
        ctx.bias_dtype = bias.dtype if bias is not None else None
        # leaf inputs may both adopt the returned gradient tensor as their .grad: give them separate storage then
        ctx.leafy = b is not None and (a.is_leaf or b.is_leaf)
        if b is not None:
            return r.view(a.shape), h.view(a.shape)
        dummy = h.new_empty(0)  # no residual output in the plain-LayerNorm form
        ctx.mark_non_differentiable(dummy)
        return dummy, h.view(a.shape)

    @staticmethod
    def backward(ctx, dr_in, dh):
        L = _bind()
        r, weight, mean, rstd = ctx.saved_tensors
        rows, d = r.shape
        dh2 = dh.reshape(rows, d).contiguous()
        dr2 = dr_in.reshape(rows, d).contiguous() if (dr_in is not None and ctx.has_b) else None
        out = torch.empty_like(r)
        partial = torch.empty(L.mfl_ln_partial_rows(rows), 2, d, dtype=torch.float32, device=r.device)
        rc = L.mfl_add_ln_bwd(out.data_ptr(), partial.data_ptr(), dh2.data_ptr(), _ptr(dr2), r.data_ptr(), _ptr(weight),
                              ctx.pbf16, mean.data_ptr(), rstd.data_ptr(), rows, d, _lib.stream_ptr())
        _lib.check(rc, "mfl_add_ln_bwd")
        dgamma = dbeta = None
        if weight is not None:
            p = partial.sum(0)
            dgamma = p[0].to(weight.dtype)
            dbeta = p[1].to(ctx.bias_dtype) if ctx.has_bias else None
        dr = out.view(ctx.shape)
        return dr, ((dr.clone() if ctx.leafy else dr) if ctx.has_b else None), dgamma, dbeta, None


def add_layer_norm(a: torch.Tensor, b: torch.Tensor, ln: nn.LayerNorm):
    """``r = a + b``; ``h = ln(r)``; returns ``(r, h)``."""
    if supported(a, ln) and b.shape == a.shape and b.dtype == a.dtype:
        return _AddLayerNormFn.apply(a, b, ln.weight, ln.bias, ln.eps)
    r = a + b
    return r, ln(r)


def layer_norm(x: torch.Tensor, ln: nn.LayerNorm) -> torch.Tensor:
    if supported(x, ln):
        return _AddLayerNormFn.apply(x, None, ln.weight, ln.bias, ln.eps)[1]
    return ln(x)
