"""Fused language-model cross-entropy (K16, ``csrc/kernels/cross_entropy.cu``).

``lm_cross_entropy(logits, targets)`` returns the mean loss over the rows whose target is not ``ignore_index``; the
gradient with respect to the logits is produced by the same kernel in the forward pass (bf16, already divided by the
number of contributing rows), so backward is one in-place multiply by the upstream gradient.  Falls back to
``F.cross_entropy`` on fp32 logits off the GPU.  Reference counterpart: the loss inside ``model.fit``
(``keras_model_ops.py:156``).
"""
from __future__ import annotations

import ctypes as C

import torch
import torch.nn.functional as F

from metisfl_b200.ops import _lib

_BOUND = False


def _bind():
    global _BOUND
    L = _lib.lib()
    if not _BOUND:
        vp, i, i64, f, ll = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_longlong
        L.mfl_lm_cross_entropy.argtypes = [vp, i64, vp, vp, i64, vp, i, i, f, ll, vp]
        _BOUND = True
    return L


class _LMCrossEntropyFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits2d, targets, n_valid, ignore_index):
        rows, vocab = logits2d.shape
        dlogits = torch.empty_like(logits2d)
        row_loss = torch.empty(rows, dtype=torch.float32, device=logits2d.device)
        rc = _bind().mfl_lm_cross_entropy(dlogits.data_ptr(), dlogits.stride(0), row_loss.data_ptr(), logits2d.data_ptr(),
                                          logits2d.stride(0), targets.data_ptr(), rows, vocab, 1.0 / float(n_valid),
                                          int(ignore_index), _lib.stream_ptr())
        _lib.check(rc, "mfl_lm_cross_entropy")
        ctx.save_for_backward(dlogits)
        return row_loss.sum() / float(n_valid)

    @staticmethod
    def backward(ctx, g):
        (dlogits,) = ctx.saved_tensors
        return dlogits.mul_(g.to(dlogits.dtype)), None, None, None


def lm_cross_entropy(logits: torch.Tensor, targets: torch.Tensor, n_valid: int = None, ignore_index: int = -100):
    """``logits [..., V]``, ``targets [...]`` (int64).  ``n_valid``: number of rows with a real target (pass it when it
    is known statically - it keeps the call free of host synchronisation)."""
    v = logits.shape[-1]
    if (logits.is_cuda and logits.dtype == torch.bfloat16 and _lib.available() and v % 8 == 0):
        l2 = logits.reshape(-1, v)
        if l2.stride(1) != 1 or l2.stride(0) % 8:
            l2 = l2.contiguous()
        t = targets.reshape(-1).to(torch.int64).contiguous()
        if n_valid is None:
            n_valid = int((t != ignore_index).sum())
        return _LMCrossEntropyFn.apply(l2, t, max(1, n_valid), ignore_index)
    return F.cross_entropy(logits.reshape(-1, v).float(), targets.reshape(-1), ignore_index=ignore_index)
