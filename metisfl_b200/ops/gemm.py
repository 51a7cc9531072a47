"""Front-end of the tcgen05 GEMM (K10, ``csrc/kernels/gemm_tcgen05.cu``) and an autograd-aware
``linear`` built on it.

``y = act(x @ W^T + b)`` runs as ONE kernel (bias and activation fused in the TMEM epilogue).  The
backward pass runs on the same kernel with MN-major operand descriptors (``gemm_ex``): ``dX = dY @ W`` reads ``W``
and ``dW = dY^T @ X`` reads ``dY`` and ``X`` exactly as they are stored - no transposed copies.
Reference counterpart: the framework matmuls under ``model.fit`` (``keras_model_ops.py:156``).
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch
import torch.nn as nn
import torch.nn.functional as F

from metisfl_b200.ops import _lib

SPLIT_K = True              # gemm_ex may hand the kernel a workspace for split-k (small outputs, long reductions)
SPLIT_K_MAX_OUT = 1 << 21   # outputs up to 2 M elements (an 8 MiB fp32 workspace)
ACTS = {None: 0, "none": 0, "relu": 1, "gelu": 2, "gelu_tanh": 2}
_BOUND = False


def _bind():
    global _BOUND
    L = _lib.lib()
    if not _BOUND:
        vp, i, i64 = C.c_void_p, C.c_int, C.c_int64
        L.mfl_gemm_bf16.argtypes = [vp, i, i64, vp, i64, vp, i64, vp, i, i, i, i, i, vp]
        L.mfl_gemm_bf16_ex.argtypes = [vp, i, i64, vp, i64, i, vp, i64, i, vp, i, i, i, i, i, vp]
        L.mfl_gemm_set_pair_mode.argtypes = [i]
        L.mfl_gemm_bf16_ws.argtypes = [vp, i, i64, vp, i64, i, vp, i64, i, i, i, i, vp, C.c_size_t, vp]
        L.mfl_colsum_slices.argtypes = [i]
        L.mfl_colsum_bf16.argtypes = [vp, vp, i64, i, i, vp]
        _BOUND = True
    return L


def set_pair_mode(mode: int) -> int:
    """CTA-pair kernel (``tcgen05.mma.cta_group::2``) selection: -1 automatic, 0 never, 1 wherever the shape allows.
    Returns the previous mode."""
    return _bind().mfl_gemm_set_pair_mode(int(mode))


def supported(x: torch.Tensor, w: torch.Tensor) -> bool:
    return (x.is_cuda and x.dtype == torch.bfloat16 and w.dtype == torch.bfloat16 and x.shape[-1] % 8 == 0
            and _lib.available())


def gemm_nt(a: torch.Tensor, b: torch.Tensor, bias: Optional[torch.Tensor] = None, act: Optional[str] = None,
            out_dtype: torch.dtype = torch.bfloat16, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``act(a @ b.T + bias)`` for 2-D bf16 ``a [M,K]`` and ``b [N,K]`` (K contiguous in both)."""
    assert a.dim() == 2 and b.dim() == 2 and a.shape[1] == b.shape[1], (a.shape, b.shape)
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16
    if a.stride(1) != 1 or a.stride(0) % 8 or a.data_ptr() % 16:
        a = a.contiguous()
    if b.stride(1) != 1 or b.stride(0) % 8 or b.data_ptr() % 16:
        b = b.contiguous()
    m, k = a.shape
    n = b.shape[0]
    if out is None:
        out = torch.empty(m, n, dtype=out_dtype, device=a.device)
    if bias is not None:
        assert bias.numel() == n and bias.dtype in (torch.bfloat16, torch.float32) and bias.is_contiguous()
    rc = _bind().mfl_gemm_bf16(out.data_ptr(), _lib.wire(out), out.stride(0), a.data_ptr(), a.stride(0), b.data_ptr(),
                               b.stride(0), _lib.ptr(bias), _lib.wire(bias) if bias is not None else 0, m, n, k,
                               ACTS[act], _lib.stream_ptr())
    _lib.check(rc, "mfl_gemm_bf16")
    return out


def _ok2d(t: torch.Tensor) -> bool:
    return t.dim() == 2 and t.stride(1) == 1 and t.stride(0) % 8 == 0 and t.data_ptr() % 16 == 0


def gemm_ex(a: torch.Tensor, b: torch.Tensor, a_mn: bool = False, b_mn: bool = False,
            out_dtype: torch.dtype = torch.bfloat16) -> torch.Tensor:
    """General-major product on the tcgen05 kernel, no transposed copies.

    ``a_mn=False``: ``a`` is ``[M, K]``; ``a_mn=True``: ``a`` is ``[K, M]`` (the transpose of the logical operand).
    ``b_mn=False``: ``b`` is ``[N, K]``; ``b_mn=True``: ``b`` is ``[K, N]``.  Returns ``[M, N]``.
    """
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16
    if not _ok2d(a):
        a = a.contiguous()
    if not _ok2d(b):
        b = b.contiguous()
    m, k = (a.shape[1], a.shape[0]) if a_mn else (a.shape[0], a.shape[1])
    n, k2 = (b.shape[1], b.shape[0]) if b_mn else (b.shape[0], b.shape[1])
    assert k == k2, (a.shape, b.shape, a_mn, b_mn)
    out = torch.empty(m, n, dtype=out_dtype, device=a.device)
    if SPLIT_K and out_dtype == torch.bfloat16 and m * n <= SPLIT_K_MAX_OUT and k >= 8 * max(m, n, 512) // 4:
        # few output tiles, long reduction (dW of a square layer): let the kernel split k over the idle SMs
        from metisfl_b200.ops.conv import _workspace
        ws = _workspace(a.device, m * n * 4, torch.cuda.current_stream(a.device))
        rc = _bind().mfl_gemm_bf16_ws(out.data_ptr(), _lib.wire(out), out.stride(0), a.data_ptr(), a.stride(0),
                                      int(a_mn), b.data_ptr(), b.stride(0), int(b_mn), m, n, k, ws.data_ptr(),
                                      ws.numel() * 4, _lib.stream_ptr())
        _lib.check(rc, "mfl_gemm_bf16_ws")
        return out
    rc = _bind().mfl_gemm_bf16_ex(out.data_ptr(), _lib.wire(out), out.stride(0), a.data_ptr(), a.stride(0), int(a_mn),
                                  b.data_ptr(), b.stride(0), int(b_mn), None, 0, m, n, k, 0, _lib.stream_ptr())
    _lib.check(rc, "mfl_gemm_bf16_ex")
    return out


def colsum(x: torch.Tensor) -> torch.Tensor:
    """fp32 column sums of a bf16 ``[rows, cols]`` matrix (K17): the bias gradient, without an fp32 copy of ``x``."""
    rows, cols = x.shape
    if not (x.is_cuda and x.dtype == torch.bfloat16 and cols % 8 == 0 and _ok2d(x) and _lib.available()):
        return x.sum(0, dtype=torch.float32)
    L = _bind()
    partial = torch.empty(L.mfl_colsum_slices(rows), cols, dtype=torch.float32, device=x.device)
    _lib.check(L.mfl_colsum_bf16(partial.data_ptr(), x.data_ptr(), x.stride(0), rows, cols, _lib.stream_ptr()),
               "mfl_colsum_bf16")
    return partial.sum(0) if partial.shape[0] > 1 else partial[0]


MN_MAJOR_BACKWARD = True  # dX / dW straight from the stored layouts (no .t().contiguous() copies)


class _LinearFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, bias):
        x2 = x.reshape(-1, x.shape[-1])
        y = gemm_nt(x2, w, bias)
        ctx.save_for_backward(x2, w)
        ctx.has_bias = bias is not None
        ctx.bias_dtype = bias.dtype if bias is not None else None
        ctx.x_shape = x.shape
        return y.view(*x.shape[:-1], w.shape[0])

    @staticmethod
    def backward(ctx, gy):
        x2, w = ctx.saved_tensors
        g2 = gy.reshape(-1, gy.shape[-1])
        if g2.dtype != torch.bfloat16:
            g2 = g2.to(torch.bfloat16)
        gx = gw = gb = None
        n, k = w.shape
        if ctx.needs_input_grad[0]:
            if MN_MAJOR_BACKWARD and n % 8 == 0 and k % 8 == 0:
                gx = gemm_ex(g2, w, False, True).view(ctx.x_shape)          # dY [M,N] . W [N,K]: W read MN-major
            elif n % 8 == 0:
                gx = gemm_nt(g2, w.t().contiguous()).view(ctx.x_shape)      # [M,N] x [K,N]^T
            else:
                gx = (g2 @ w).view(ctx.x_shape)
        if ctx.needs_input_grad[1]:
            m = g2.shape[0]
            if MN_MAJOR_BACKWARD and n % 8 == 0 and k % 8 == 0:
                gw = gemm_ex(g2, x2, True, True)                            # dY^T . X: both read MN-major
            elif m % 8 == 0:
                gw = gemm_nt(g2.t().contiguous(), x2.t().contiguous())      # [N,M] x [K,M]^T
            else:
                gw = g2.t() @ x2
        if ctx.has_bias and ctx.needs_input_grad[2]:
            gb = colsum(g2).to(ctx.bias_dtype)  # K17: fp32 accumulation straight from the bf16 dY
        return gx, gw, gb


def linear(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Drop-in for ``F.linear`` that uses the tcgen05 kernel when the operands allow it."""
    if supported(x, w):
        return _LinearFn.apply(x, w, bias)
    return F.linear(x, w, bias)


class B200Linear(nn.Linear):
    def forward(self, x):
        return linear(x, self.weight, self.bias)


def use_b200_linear(module: nn.Module) -> nn.Module:
    """Re-class every ``nn.Linear`` in ``module`` so its forward/backward run on the tcgen05 kernel."""
    for m in module.modules():
        if type(m) is nn.Linear:
            m.__class__ = B200Linear
    return module
