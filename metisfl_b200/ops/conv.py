"""Front-end of the implicit-GEMM convolution (K11) and an autograd-aware ``conv2d`` built on it.

Forward and the stride-1 data gradient run on the tcgen05 kernel (the data gradient is the same
convolution with spatially flipped, channel-transposed filters and padding ``R-1-p``); the weight
gradient and stride-2 data gradients still go through ATen (cuDNN) - see DESIGN.md for the gap.
Tensors are NHWC (``torch.channels_last``) bf16.  Reference counterpart: the framework convolutions
under ``model.fit`` (``keras_model_ops.py:156``).
"""
from __future__ import annotations

import ctypes as C

import torch
import torch.nn as nn
import torch.nn.functional as F

from metisfl_b200.ops import _lib

_BOUND = False


def _bind():
    global _BOUND
    L = _lib.lib()
    if not _BOUND:
        vp, i = C.c_void_p, C.c_int
        L.mfl_conv2d_fprop_bf16.argtypes = [vp, vp, vp, i, i, i, i, i, i, i, i, i, vp]
        L.mfl_conv3x3_supported.argtypes = [i, i, i, i, i]
        L.mfl_conv3x3_bf16.argtypes = [vp, vp, vp, vp, i, i, i, i, i, vp]
        _BOUND = True
    return L


def supported(x: torch.Tensor, w: torch.Tensor, stride, padding, dilation=(1, 1), groups=1) -> bool:
    return (x.is_cuda and x.dtype == torch.bfloat16 and w.dtype == torch.bfloat16 and x.dim() == 4
            and groups == 1 and tuple(dilation) == (1, 1) and stride[0] == stride[1] and padding[0] == padding[1]
            and x.shape[1] % 64 == 0 and w.shape[0] % 8 == 0 and _lib.available())


def conv2d_fprop(x: torch.Tensor, w: torch.Tensor, stride: int, pad: int) -> torch.Tensor:
    """x: [N,C,H,W] channels_last bf16, w: [K,C,R,S] channels_last bf16 -> y: [N,K,OH,OW] channels_last."""
    n, c, h, wd = x.shape
    k, c2, r, s = w.shape
    assert c == c2
    x = x.contiguous(memory_format=torch.channels_last)
    w = w.contiguous(memory_format=torch.channels_last)
    oh, ow = (h + 2 * pad - r) // stride + 1, (wd + 2 * pad - s) // stride + 1
    y = torch.empty((n, k, oh, ow), dtype=torch.bfloat16, device=x.device, memory_format=torch.channels_last)
    rc = _bind().mfl_conv2d_fprop_bf16(y.data_ptr(), x.data_ptr(), w.data_ptr(), n, h, wd, c, k, r, s, stride, pad,
                                       _lib.stream_ptr())
    _lib.check(rc, "mfl_conv2d_fprop_bf16")
    return y


def conv3x3_supported(x: torch.Tensor, w: torch.Tensor) -> bool:
    """Shapes the weight-stationary 3x3 kernel (K15) takes: stride 1, padding 1, C in {64, 128}, K % 64 == 0."""
    if not (x.is_cuda and x.dtype == torch.bfloat16 and w.dtype == torch.bfloat16 and x.dim() == 4
            and tuple(w.shape[2:]) == (3, 3) and _lib.available()):
        return False
    n, c, h, wd = x.shape
    return bool(_bind().mfl_conv3x3_supported(n, h, wd, c, w.shape[0]))


def conv3x3(x: torch.Tensor, w: torch.Tensor, stats: torch.Tensor = None) -> torch.Tensor:
    """K15: y = conv2d(x, w, stride 1, padding 1); x [N,C,H,W] and w [K,C,3,3] channels_last bf16.
    ``stats`` ([2, K] fp32) receives += the per-channel sum and sum of squares of y."""
    n, c, h, wd = x.shape
    k = w.shape[0]
    x = x.contiguous(memory_format=torch.channels_last)
    w = w.contiguous(memory_format=torch.channels_last)
    y = torch.empty((n, k, h, wd), dtype=torch.bfloat16, device=x.device, memory_format=torch.channels_last)
    rc = _bind().mfl_conv3x3_bf16(y.data_ptr(), x.data_ptr(), w.data_ptr(), None if stats is None else stats.data_ptr(),
                                  n, h, wd, c, k, _lib.stream_ptr())
    _lib.check(rc, "mfl_conv3x3_bf16")
    return y


class _Conv2dFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, stride, pad):
        ctx.save_for_backward(x, w)
        ctx.stride, ctx.pad = stride, pad
        return conv2d_fprop(x, w, stride, pad)

    @staticmethod
    def backward(ctx, gy):
        x, w = ctx.saved_tensors
        stride, pad = ctx.stride, ctx.pad
        gy = gy.contiguous(memory_format=torch.channels_last)
        k, c, r, s = w.shape
        gx = gw = None
        need_x, need_w = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        if need_x and stride == 1 and k % 64 == 0 and c % 8 == 0 and r - 1 - pad >= 0:
            # dX = conv(dY, flip(W)^T): filters [C, K, R, S] with both spatial axes reversed
            wt = w.flip(2, 3).permute(1, 0, 2, 3).contiguous(memory_format=torch.channels_last)
            gx = conv2d_fprop(gy, wt, 1, r - 1 - pad)
            need_x = False
        if need_x or need_w:
            gx2, gw, _ = torch.ops.aten.convolution_backward(
                gy, x, w, None, [stride, stride], [pad, pad], [1, 1], False, [0, 0], 1, [need_x, need_w, False])
            if need_x:
                gx = gx2
        return gx, gw, None, None


def conv2d(x: torch.Tensor, w: torch.Tensor, bias=None, stride=(1, 1), padding=(0, 0), dilation=(1, 1), groups=1):
    """Drop-in for ``F.conv2d`` that uses the tcgen05 implicit-GEMM kernel when the operands allow it."""
    stride = (stride, stride) if isinstance(stride, int) else tuple(stride)
    padding = (padding, padding) if isinstance(padding, int) else tuple(padding)
    if isinstance(padding, tuple) and supported(x, w, stride, padding, dilation if not isinstance(dilation, int)
                                                else (dilation, dilation), groups):
        y = _Conv2dFn.apply(x, w, stride[0], padding[0])
        return y if bias is None else y + bias.view(1, -1, 1, 1)
    return F.conv2d(x, w, bias, stride, padding, dilation, groups)


class B200Conv2d(nn.Conv2d):
    def forward(self, x):
        if isinstance(self.padding, str) or self.padding_mode != "zeros":
            return super().forward(x)
        return conv2d(x, self.weight, self.bias, self.stride, self.padding, self.dilation, self.groups)


def use_b200_conv(module: nn.Module) -> nn.Module:
    """Re-class every ``nn.Conv2d`` so eligible layers run on the tcgen05 implicit-GEMM kernel."""
    for m in module.modules():
        if type(m) is nn.Conv2d:
            m.__class__ = B200Conv2d
    return module


# ---- weight gradients on a side stream -------------------------------------------------------------------------
# In backward the data gradient of a convolution is on the critical path (the previous layer waits for it) while the
# weight gradient is only needed by the optimizer at the very end.  Autograd would still run both back to back on
# one stream; here the weight-gradient convolution is issued on a side stream and accumulates straight into the flat
# gradient buffer, so it overlaps the dgrad -> BatchNorm-backward chain of the following layers.  Inside a CUDA
# graph capture the fork / join become parallel branches of the graph.  ``join()`` must run before the gradients
# are consumed (``LocalTrainer`` does it right after ``loss.backward()``).
class _SideStream:
    def __init__(self):
        self.streams = {}
        self.pending = []   # tensors the side stream still reads
        self.dirty = set()

    def stream(self, device) -> "torch.cuda.Stream":
        s = self.streams.get(device)
        if s is None:
            s = self.streams[device] = torch.cuda.Stream(device=device)
        return s


_SIDE = _SideStream()
OVERLAP_WGRAD = True


def join():
    """Make the current stream wait for every weight gradient issued on the side stream."""
    for device in list(_SIDE.dirty):
        torch.cuda.current_stream(device).wait_stream(_SIDE.streams[device])
    _SIDE.dirty.clear()
    _SIDE.pending.clear()


class _OverlapConvFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, stride, padding, dilation, groups):
        ctx.save_for_backward(x, w)
        ctx.cfg = (list(stride), list(padding), list(dilation), groups)
        return torch.ops.aten.convolution(x, w, None, stride, padding, dilation, False, [0, 0], groups)

    @staticmethod
    def backward(ctx, gy):
        x, w = ctx.saved_tensors
        stride, padding, dilation, groups = ctx.cfg
        need_x, need_w = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        gw = None
        direct = need_w and OVERLAP_WGRAD and w.grad is not None and w.grad.dtype == w.dtype
        if direct:
            main = torch.cuda.current_stream(x.device)
            side = _SIDE.stream(x.device)
            side.wait_stream(main)                       # gy (and x) are ready on the main stream
            with torch.cuda.stream(side):
                g = torch.ops.aten.convolution_backward(gy, x, w, None, stride, padding, dilation, False, [0, 0], groups,
                                                        [False, True, False])[1]
                w.grad.add_(g)
            _SIDE.pending.append((gy, x, g))             # keep the operands alive until join()
            _SIDE.dirty.add(x.device)
            need_w = False
        gx = None
        if need_x or need_w:
            gx, gw, _ = torch.ops.aten.convolution_backward(gy, x, w, None, stride, padding, dilation, False, [0, 0],
                                                            groups, [need_x, need_w, False])
        return gx, gw, None, None, None, None


class OverlappedConv2d(nn.Conv2d):
    def forward(self, x):
        if (x.is_cuda and self.bias is None and self.padding_mode == "zeros" and not isinstance(self.padding, str)
                and torch.is_grad_enabled()):
            return _OverlapConvFn.apply(x, self.weight, self.stride, self.padding, self.dilation, self.groups)
        return super().forward(x)


def use_overlapped_wgrad(module: nn.Module) -> nn.Module:
    """Re-class bias-free ``nn.Conv2d`` layers so their weight gradients run on the side stream."""
    for m in module.modules():
        if type(m) is nn.Conv2d and m.bias is None:
            m.__class__ = OverlappedConv2d
    return module
