"""Front-end of the implicit-GEMM convolution family (K11) and an autograd-aware ``conv2d`` built on it.

Forward, data gradient (stride 1 and 2) and weight gradient all run on the persistent tcgen05 pipeline of
``csrc/kernels/gemm_tcgen05.cu``:

* forward: A = im2col TMA of the activations, B = filters (K-major);
* data gradient: A = im2col TMA of dY, B = the SAME filter tensor read MN-major through the tensor map (no flipped /
  transposed copy is made); stride 2 is split into its four (h mod 2, w mod 2) classes, each a stride-1 correlation
  with the taps of matching parity, written through a strided 4-D tensor map;
* weight gradient: A = dY read MN-major, B = im2col TMA boxes of the activations read MN-major, the reduction over
  the N*OH*OW output pixels split over the SMs and combined with TMA reduce-add stores into an fp32 workspace; a
  finalize kernel adds it into the (bf16) gradient arena.

The weight gradient is issued on a side stream (it is only needed by the optimizer), see below.
Tensors are NHWC (``torch.channels_last``) bf16.  Reference counterpart: the framework convolutions
under ``model.fit`` (``keras_model_ops.py:156``).
"""
from __future__ import annotations

import ctypes as C
import os

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
        L.mfl_conv3x3_dgrad_bf16.argtypes = [vp, vp, vp, i, i, i, i, i, vp]
        L.mfl_conv3x3_wgrad_supported.argtypes = [i, i, i, i, i]
        L.mfl_conv3x3_wgrad_bf16.argtypes = [vp, vp, vp, vp, i, i, i, i, i, i, vp]
        L.mfl_conv3x3_t_supported.argtypes = [i, i, i, i, i]
        L.mfl_conv3x3_t_fprop_bf16.argtypes = [vp, vp, vp, i, i, i, i, i, vp]
        L.mfl_conv3x3_t_dgrad_bf16.argtypes = [vp, vp, vp, i, i, i, i, i, vp]
        L.mfl_conv_stem_supported.argtypes = [i, i]
        L.mfl_conv_stem_fprop_bf16.argtypes = [vp, vp, vp, i, i, i, i, i, vp]
        L.mfl_conv_stem_wgrad_bf16.argtypes = [vp, vp, vp, vp, i, i, i, i, i, i, vp]
        L.mfl_conv_stem_im2col_bf16.argtypes = [vp, vp, i, i, i, i, vp]
        L.mfl_conv2d_wgrad_bf16.argtypes = [vp, vp, vp, vp, i, i, i, i, i, i, i, i, i, i, vp]
        L.mfl_conv2d_dgrad_bf16.argtypes = [vp, vp, vp, i, i, i, i, i, i, i, i, i, vp]
        L.mfl_conv2d_fprop_bf16_ws.argtypes = [vp, vp, vp, vp, C.c_size_t, i, i, i, i, i, i, i, i, i, vp]
        L.mfl_conv2d_dgrad_bf16_ws.argtypes = [vp, vp, vp, vp, C.c_size_t, i, i, i, i, i, i, i, i, i, vp]
        L.mfl_conv2d_wgrad_workspace_bytes.restype = C.c_size_t
        L.mfl_conv2d_wgrad_workspace_bytes.argtypes = [i, i, i, i]
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
    stream = torch.cuda.current_stream(x.device)
    ws = _workspace(x.device, 0, stream)
    rc = _bind().mfl_conv2d_fprop_bf16_ws(y.data_ptr(), x.data_ptr(), w.data_ptr(), ws.data_ptr(), ws.numel() * 4, n, h, wd,
                                          c, k, r, s, stride, pad, stream.cuda_stream)
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


def _pow2(v: int) -> bool:
    return v > 0 and (v & (v - 1)) == 0


def dgrad_supported(x_shape, w_shape, stride: int, pad: int) -> bool:
    n, c, h, wd = x_shape
    k, _, r, s = w_shape
    if k % 64 != 0 or c % 8 != 0 or pad >= r or pad >= s or r * s > 16:
        return False
    if stride == 1:
        return True
    if stride != 2 or c % 64 != 0 or h % 2 or wd % 2:
        return False
    oh, ow = (h + 2 * pad - r) // 2 + 1, (wd + 2 * pad - s) // 2 + 1
    return oh == h // 2 and ow == wd // 2 and _pow2(oh) and _pow2(ow)


def conv2d_dgrad(gy: torch.Tensor, w: torch.Tensor, x_shape, stride: int, pad: int) -> torch.Tensor:
    """dX of ``conv2d(x, w, stride, pad)``: gy [N,K,OH,OW] and w [K,C,R,S] channels_last bf16 -> [N,C,H,W]."""
    n, c, h, wd = x_shape
    k, _, r, s = w.shape
    gy = gy.contiguous(memory_format=torch.channels_last)
    w = w.contiguous(memory_format=torch.channels_last)
    gx = torch.empty((n, c, h, wd), dtype=torch.bfloat16, device=gy.device, memory_format=torch.channels_last)
    stream = torch.cuda.current_stream(gy.device)
    ws = _workspace(gy.device, 0, stream)
    rc = _bind().mfl_conv2d_dgrad_bf16_ws(gx.data_ptr(), gy.data_ptr(), w.data_ptr(), ws.data_ptr(), ws.numel() * 4, n, h,
                                          wd, c, k, r, s, stride, pad, stream.cuda_stream)
    _lib.check(rc, "mfl_conv2d_dgrad_bf16", launches=stride * stride)
    return gx


USE_T = True      # 3x3 / stride-1 layers on the pixels-as-N kernel (K15t) where the image is wide enough
T_MIN_WIDTH = 16  # below this the zero-padded pixel grid wastes too much of every tile (4x4 images: 16 of 36 columns)
USE_HALO = True   # route 3x3 / stride-1 layers with <= 128 reduction channels to the weight-stationary kernel (K15)
_WS = {}   # (device, stream id) -> zeroed fp32 split-K workspace, handed back zeroed by every call


def _workspace(device, nbytes: int, stream) -> torch.Tensor:
    key = (device, stream.cuda_stream)
    ws = _WS.get(key)
    if ws is None or ws.numel() * 4 < nbytes:
        with torch.cuda.stream(stream):
            ws = torch.zeros(max(nbytes // 4, 4 << 20), dtype=torch.float32, device=device)  # >= 16 MiB
        _WS[key] = ws
    return ws


def conv2d_wgrad(gy: torch.Tensor, x: torch.Tensor, w_shape, stride: int, pad: int, out: torch.Tensor = None,
                 accumulate: bool = False) -> torch.Tensor:
    """dW of ``conv2d(x, w, stride, pad)``: [K,C,R,S] channels_last bf16 (``out += dW`` when ``accumulate``)."""
    n, c, h, wd = x.shape
    k, _, r, s = w_shape
    gy = gy.contiguous(memory_format=torch.channels_last)
    x = x.contiguous(memory_format=torch.channels_last)
    if out is None:
        out = torch.empty((k, c, r, s), dtype=torch.bfloat16, device=x.device, memory_format=torch.channels_last)
        accumulate = False
    assert out.is_contiguous(memory_format=torch.channels_last) and out.dtype == torch.bfloat16
    L = _bind()
    stream = torch.cuda.current_stream(x.device)
    ws = _workspace(x.device, L.mfl_conv2d_wgrad_workspace_bytes(c, k, r, s), stream)
    if USE_HALO and r == 3 and s == 3 and stride == 1 and pad == 1 and L.mfl_conv3x3_wgrad_supported(n, h, wd, c, k):
        # K15w: one halo tile per filter row, the three taps of the row in one N = 192 instruction
        rc = L.mfl_conv3x3_wgrad_bf16(out.data_ptr(), gy.data_ptr(), x.data_ptr(), ws.data_ptr(), n, h, wd, c, k,
                                      int(accumulate), stream.cuda_stream)
        _lib.check(rc, "mfl_conv3x3_wgrad_bf16", launches=2)
        return out
    rc = L.mfl_conv2d_wgrad_bf16(out.data_ptr(), gy.data_ptr(), x.data_ptr(), ws.data_ptr(), n, h, wd, c, k, r, s, stride,
                                 pad, int(accumulate), stream.cuda_stream)
    _lib.check(rc, "mfl_conv2d_wgrad_bf16", launches=2)
    return out


def stem_supported(x: torch.Tensor, w: torch.Tensor, stride, padding) -> bool:
    """The network stem: 3x3 / stride 1 / padding 1 over an image with at most 4 channels (K11s)."""
    return (x.is_cuda and x.dtype == torch.bfloat16 and w.dtype == torch.bfloat16 and x.dim() == 4
            and tuple(w.shape[2:]) == (3, 3) and tuple(stride) == (1, 1) and tuple(padding) == (1, 1)
            and _lib.available() and bool(_bind().mfl_conv_stem_supported(x.shape[1], w.shape[0])))


class _StemFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w):
        n, c, h, wd = x.shape
        k = w.shape[0]
        x = x.contiguous(memory_format=torch.channels_last)
        wc = w.contiguous(memory_format=torch.channels_last)
        ctx.save_for_backward(x, w)
        y = torch.empty((n, k, h, wd), dtype=torch.bfloat16, device=x.device, memory_format=torch.channels_last)
        rc = _bind().mfl_conv_stem_fprop_bf16(y.data_ptr(), x.data_ptr(), wc.data_ptr(), n, h, wd, c, k, _lib.stream_ptr())
        _lib.check(rc, "mfl_conv_stem_fprop_bf16")
        return y

    @staticmethod
    def backward(ctx, gy):
        x, w = ctx.saved_tensors
        n, c, h, wd = x.shape
        k = w.shape[0]
        gy = gy.contiguous(memory_format=torch.channels_last)
        gx = None
        if ctx.needs_input_grad[0]:  # an image that wants a gradient: rare, leave it to the library
            gx = torch.ops.aten.convolution_backward(gy, x, w, None, [1, 1], [1, 1], [1, 1], False, [0, 0], 1,
                                                     [True, False, False])[0]
        if not ctx.needs_input_grad[1]:
            return gx, None
        L = _bind()

        def run(out, accumulate):
            stream = torch.cuda.current_stream(x.device)
            ws = _workspace(x.device, 0, stream)
            rc = L.mfl_conv_stem_wgrad_bf16(out.data_ptr(), gy.data_ptr(), x.data_ptr(), ws.data_ptr(), n, h, wd, c, k,
                                            int(accumulate), stream.cuda_stream)
            _lib.check(rc, "mfl_conv_stem_wgrad_bf16", launches=2)

        direct = OVERLAP_WGRAD and w.grad is not None and w.grad.dtype == torch.bfloat16 and \
            w.grad.is_contiguous(memory_format=torch.channels_last)
        if direct:
            main = torch.cuda.current_stream(x.device)
            side = _SIDE.stream(x.device)
            side.wait_stream(main)
            with torch.cuda.stream(side):
                run(w.grad, True)
            _SIDE.pending.append((gy, x))
            _SIDE.dirty.add(x.device)
            return gx, None
        gw = torch.empty(w.shape, dtype=torch.bfloat16, device=x.device, memory_format=torch.channels_last)
        run(gw, False)
        return gx, gw


STEM_GEMM = os.environ.get("MFL_STEM_GEMM", "1") != "0"   # default path: the stem as im2col + K10 GEMMs
_PATCH_COLS = 32


def stem_gemm_supported(x: torch.Tensor, w: torch.Tensor, stride, padding, dilation=(1, 1), groups=1) -> bool:
    """3x3 / stride 1 / padding 1 over an image with at most 3 channels (27 patch values fit a 32-wide row)."""
    return (x.is_cuda and x.dtype == torch.bfloat16 and w.dtype == torch.bfloat16 and x.dim() == 4
            and x.shape[1] <= 3 and tuple(w.shape[2:]) == (3, 3) and tuple(stride) == (1, 1)
            and tuple(padding) == (1, 1) and tuple(dilation) == (1, 1) and groups == 1 and w.shape[0] % 8 == 0
            and _lib.available())


class _StemGemmFn(torch.autograd.Function):
    """The stem on the tensor cores (``csrc/kernels/conv_stem.cu: stem_im2col_kernel``): one small kernel writes the
    [N*H*W, 32] patch matrix, the forward is ``patches . W32^T`` on K10, the weight gradient ``dY^T . patches`` on K10
    with split k (both operands read MN-major, no transposed copies).  The patch matrix (64 bytes per pixel) is kept
    for the backward pass."""

    @staticmethod
    def forward(ctx, x, w):
        from metisfl_b200.ops import gemm as G
        n, c, h, wd = x.shape
        k = w.shape[0]
        x = x.contiguous(memory_format=torch.channels_last)
        patches = torch.empty((n * h * wd, _PATCH_COLS), dtype=torch.bfloat16, device=x.device)
        rc = _bind().mfl_conv_stem_im2col_bf16(patches.data_ptr(), x.data_ptr(), n, h, wd, c, _lib.stream_ptr())
        _lib.check(rc, "mfl_conv_stem_im2col_bf16")
        w_rows = w.permute(0, 2, 3, 1).reshape(k, 9 * c)            # [K, R, S, C]: a view of NHWC-stored filters
        w32 = torch.nn.functional.pad(w_rows, (0, _PATCH_COLS - 9 * c))
        y = G.gemm_nt(patches, w32)                                  # [N*H*W, K]
        ctx.save_for_backward(patches, x, w)
        return y.view(n, h, wd, k).permute(0, 3, 1, 2)               # NCHW view of the NHWC result

    @staticmethod
    def backward(ctx, gy):
        from metisfl_b200.ops import gemm as G
        patches, x, w = ctx.saved_tensors
        n, c, h, wd = x.shape
        k = w.shape[0]
        gx = gw = None
        if ctx.needs_input_grad[0]:  # an image that wants a gradient: rare, leave it to the library
            gx = torch.ops.aten.convolution_backward(gy, x, w, None, [1, 1], [1, 1], [1, 1], False, [0, 0], 1,
                                                     [True, False, False])[0]
        if ctx.needs_input_grad[1]:
            g2 = gy.contiguous(memory_format=torch.channels_last).permute(0, 2, 3, 1).reshape(n * h * wd, k)
            gw32 = G.gemm_ex(g2, patches, a_mn=True, b_mn=True)      # [K, 32] = dY^T . patches, reduction over pixels
            gw = gw32[:, :9 * c].reshape(k, 3, 3, c).permute(0, 3, 1, 2)
        return gx, gw


def conv3x3_dgrad(gy: torch.Tensor, w: torch.Tensor) -> torch.Tensor:
    """K15 in data-gradient mode: dX of ``conv3x3(x, w)`` from gy [N,K,H,W]; w is the FORWARD filter [K,C,3,3]."""
    n, k, h, wd = gy.shape
    c = w.shape[1]
    gy = gy.contiguous(memory_format=torch.channels_last)
    w = w.contiguous(memory_format=torch.channels_last)
    gx = torch.empty((n, c, h, wd), dtype=torch.bfloat16, device=gy.device, memory_format=torch.channels_last)
    rc = _bind().mfl_conv3x3_dgrad_bf16(gx.data_ptr(), gy.data_ptr(), w.data_ptr(), n, h, wd, c, k, _lib.stream_ptr())
    _lib.check(rc, "mfl_conv3x3_dgrad_bf16")
    return gx


def conv3x3_t(x: torch.Tensor, w: torch.Tensor) -> torch.Tensor:
    """K15t forward: y = conv2d(x, w, stride 1, padding 1) with the pixels as the MMA's N dimension."""
    n, c, h, wd = x.shape
    k = w.shape[0]
    x = x.contiguous(memory_format=torch.channels_last)
    w = w.contiguous(memory_format=torch.channels_last)
    y = torch.empty((n, k, h, wd), dtype=torch.bfloat16, device=x.device, memory_format=torch.channels_last)
    rc = _bind().mfl_conv3x3_t_fprop_bf16(y.data_ptr(), x.data_ptr(), w.data_ptr(), n, h, wd, c, k, _lib.stream_ptr())
    _lib.check(rc, "mfl_conv3x3_t_fprop_bf16")
    return y


def conv3x3_t_dgrad(gy: torch.Tensor, w: torch.Tensor) -> torch.Tensor:
    """K15t data gradient: dX from gy [N,K,H,W] and the FORWARD filter w [K,C,3,3] (read MN-major, taps mirrored)."""
    n, k, h, wd = gy.shape
    c = w.shape[1]
    gy = gy.contiguous(memory_format=torch.channels_last)
    w = w.contiguous(memory_format=torch.channels_last)
    gx = torch.empty((n, c, h, wd), dtype=torch.bfloat16, device=gy.device, memory_format=torch.channels_last)
    rc = _bind().mfl_conv3x3_t_dgrad_bf16(gx.data_ptr(), gy.data_ptr(), w.data_ptr(), n, h, wd, c, k, _lib.stream_ptr())
    _lib.check(rc, "mfl_conv3x3_t_dgrad_bf16")
    return gx


def _t_ok(n, c_red, h, wd, c_out, r, s, stride, pad, dgrad: bool) -> bool:
    # measured (profiles/README.md): K15t wins among the own kernels for 128-channel layers on 16x16 images; 64-channel
    # layers stay on K15 (their 1024 small tiles are bound by K15t's transposed epilogue), 8x8 and 4x4 images on K11
    return (USE_T and r == 3 and s == 3 and stride == 1 and pad == 1 and (not dgrad or c_out % 64 == 0)
            and wd >= T_MIN_WIDTH and c_red >= 128 and c_out >= 128
            and bool(_bind().mfl_conv3x3_t_supported(n, h, wd, c_red, c_out)))


def _halo_ok(n, c_red, h, wd, k_out, r, s, stride, pad) -> bool:
    """Shapes where the weight-stationary halo kernel (K15) is the faster formulation: 3x3 / stride 1 / padding 1 with
    at most 128 reduction channels (all taps of 64 output channels stay resident in shared memory)."""
    return (r == 3 and s == 3 and stride == 1 and pad == 1 and
            bool(_bind().mfl_conv3x3_supported(n, h, wd, c_red, k_out)))


class _Conv2dFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, stride, pad):
        ctx.save_for_backward(x, w)
        ctx.stride, ctx.pad = stride, pad
        n, c, h, wd = x.shape
        if _t_ok(n, c, h, wd, w.shape[0], w.shape[2], w.shape[3], stride, pad, False):
            return conv3x3_t(x, w)
        if USE_HALO and _halo_ok(n, c, h, wd, w.shape[0], w.shape[2], w.shape[3], stride, pad):
            return conv3x3(x, w)
        return conv2d_fprop(x, w, stride, pad)

    @staticmethod
    def backward(ctx, gy):
        x, w = ctx.saved_tensors
        stride, pad = ctx.stride, ctx.pad
        gy = gy.contiguous(memory_format=torch.channels_last)
        gx = gw = None
        need_x, need_w = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        own_w = need_w and x.shape[1] % 64 == 0 and w.shape[0] % 8 == 0
        if own_w:
            # the weight gradient is off the critical path: side stream, accumulated straight into the flat arena
            direct = OVERLAP_WGRAD and w.grad is not None and w.grad.dtype == torch.bfloat16 and \
                w.grad.is_contiguous(memory_format=torch.channels_last)
            if direct:
                main = torch.cuda.current_stream(x.device)
                side = _SIDE.stream(x.device)
                side.wait_stream(main)
                with torch.cuda.stream(side):
                    conv2d_wgrad(gy, x, w.shape, stride, pad, out=w.grad, accumulate=True)
                _SIDE.pending.append((gy, x))
                _SIDE.dirty.add(x.device)
            else:
                gw = conv2d_wgrad(gy, x, w.shape, stride, pad)
            need_w = False
        if need_x and _t_ok(x.shape[0], w.shape[0], x.shape[2], x.shape[3], w.shape[1], w.shape[2], w.shape[3], stride,
                            pad, True):
            gx = conv3x3_t_dgrad(gy, w)
            need_x = False
        if need_x and USE_HALO and _halo_ok(x.shape[0], w.shape[0], x.shape[2], x.shape[3], w.shape[1], w.shape[2],
                                            w.shape[3], stride, pad):
            gx = conv3x3_dgrad(gy, w)
            need_x = False
        if need_x and dgrad_supported(x.shape, w.shape, stride, pad):
            gx = conv2d_dgrad(gy, w, x.shape, stride, pad)
            need_x = False
        if need_x or need_w:
            gx2, gw2, _ = torch.ops.aten.convolution_backward(
                gy, x, w, None, [stride, stride], [pad, pad], [1, 1], False, [0, 0], 1, [need_x, need_w, False])
            gx = gx2 if need_x else gx
            gw = gw2 if need_w else gw
        return gx, gw, None, None


def conv2d(x: torch.Tensor, w: torch.Tensor, bias=None, stride=(1, 1), padding=(0, 0), dilation=(1, 1), groups=1):
    """Drop-in for ``F.conv2d`` that uses the tcgen05 implicit-GEMM kernel when the operands allow it."""
    stride = (stride, stride) if isinstance(stride, int) else tuple(stride)
    padding = (padding, padding) if isinstance(padding, int) else tuple(padding)
    if isinstance(padding, tuple) and groups == 1 and stem_supported(x, w, stride, padding) and \
            (dilation == 1 or tuple(dilation) == (1, 1)):
        y = _StemGemmFn.apply(x, w) if (STEM_GEMM and stem_gemm_supported(x, w, stride, padding)) else _StemFn.apply(x, w)
        return y if bias is None else y + bias.view(1, -1, 1, 1)
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
        own_w = direct and HYBRID and _own_wgrad_wins(x, w, stride, padding, dilation, groups) and \
            w.grad.is_contiguous(memory_format=torch.channels_last)
        if direct:
            main = torch.cuda.current_stream(x.device)
            side = _SIDE.stream(x.device)
            side.wait_stream(main)                       # gy (and x) are ready on the main stream
            with torch.cuda.stream(side):
                if own_w:   # hand-written weight gradient, accumulated straight into the flat gradient arena
                    conv2d_wgrad(gy.contiguous(memory_format=torch.channels_last), x, w.shape, stride[0], padding[0],
                                 out=w.grad, accumulate=True)
                    g = None
                else:
                    g = torch.ops.aten.convolution_backward(gy, x, w, None, stride, padding, dilation, False, [0, 0],
                                                            groups, [False, True, False])[1]
                    w.grad.add_(g)
            _SIDE.pending.append((gy, x, g))             # keep the operands alive until join()
            _SIDE.dirty.add(x.device)
            need_w = False
        gx = None
        if need_x and HYBRID and _own_dgrad_wins(x, w, stride, padding, dilation, groups):
            gx = conv2d_dgrad(gy, w, x.shape, stride[0], padding[0])
            need_x = False
        if need_x or need_w:
            gx, gw, _ = torch.ops.aten.convolution_backward(gy, x, w, None, stride, padding, dilation, False, [0, 0],
                                                            groups, [need_x, need_w, False])
        return gx, gw, None, None, None, None


# Per-shape dispatch of the cuDNN-based default path: where a hand-written gradient kernel measured faster than cuDNN on
# the ResNet-18 shapes (bench/conv_bench.py, profiles/README.md) it takes that gradient; the rest stays on the library.
HYBRID = os.environ.get("MFL_CONV_HYBRID", "1") != "0"


def _bf16_nhwc(x, w, dilation, groups) -> bool:
    return (x.is_cuda and x.dtype == torch.bfloat16 and w.dtype == torch.bfloat16 and groups == 1
            and tuple(dilation) == (1, 1) and _lib.available())


def _own_dgrad_wins(x, w, stride, padding, dilation, groups) -> bool:
    """Stride-2 data gradients (one launch over the four parity classes) from 128 input channels up, and every 1x1
    stride-2 shortcut: 18.9 vs 59.4 us, 21.9 vs 31.1 us, 17.4 vs 21.2 us against cuDNN."""
    if not _bf16_nhwc(x, w, dilation, groups) or tuple(stride) != (2, 2) or stride[0] != stride[1]:
        return False
    k, c, r, s = w.shape
    if not dgrad_supported(x.shape, w.shape, 2, padding[0]) or padding[0] != padding[1]:
        return False
    return (r == 1 and s == 1) or (r == 3 and s == 3 and c >= 128)


def _own_wgrad_wins(x, w, stride, padding, dilation, groups) -> bool:
    """K15w on 3x3 stride-1 layers with images of 16x16 and more (29.6 vs 32.8 us, 23.5 vs 28.3 us), and the 1x1
    shortcuts (17.4 vs 22.5 us)."""
    if not _bf16_nhwc(x, w, dilation, groups) or x.shape[1] % 64 != 0 or w.shape[0] % 8 != 0:
        return False
    k, c, r, s = w.shape
    if padding[0] != padding[1] or stride[0] != stride[1]:
        return False
    if r == 3 and s == 3 and stride[0] == 1 and padding[0] == 1:
        return x.shape[3] >= 16 and bool(_bind().mfl_conv3x3_wgrad_supported(x.shape[0], x.shape[2], x.shape[3], c, k))
    return r == 1 and s == 1 and padding[0] == 0


class OverlappedConv2d(nn.Conv2d):
    def forward(self, x):
        if (x.is_cuda and self.bias is None and self.padding_mode == "zeros" and not isinstance(self.padding, str)
                and torch.is_grad_enabled()):
            if HYBRID and STEM_GEMM and stem_gemm_supported(x, self.weight, self.stride, self.padding, self.dilation,
                                                              self.groups):
                return _StemGemmFn.apply(x, self.weight)   # the 3-channel stem: im2col + K10 (fwd and dW)
            return _OverlapConvFn.apply(x, self.weight, self.stride, self.padding, self.dilation, self.groups)
        return super().forward(x)


def use_overlapped_wgrad(module: nn.Module) -> nn.Module:
    """Re-class bias-free ``nn.Conv2d`` layers so their weight gradients run on the side stream."""
    for m in module.modules():
        if type(m) is nn.Conv2d and m.bias is None:
            m.__class__ = OverlappedConv2d
    return module
