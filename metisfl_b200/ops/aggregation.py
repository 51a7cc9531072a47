"""Python front-ends of the aggregation kernels (K1, K3, K4).

CUDA tensors go to the sm_100a kernels in ``csrc/kernels/aggregate.cu`` / ``quantify.cu``; CPU
tensors take a plain PyTorch restatement of the same math, which is what the CPU-only test tier
and the controller's host path use.  Reference semantics: ``federated_average.cc:15-59`` (FedAvg),
``federated_rolling_average_base.cc:175-293`` (rolling S/z update), ``proto_tensor_serde.h:35-50``
(quantifier).
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import torch

from metisfl_b200.ops import _lib

_INT_DTYPES = (torch.int8, torch.int16, torch.int32, torch.int64, torch.uint8, torch.uint16,
               torch.uint32, torch.uint64)


def _cpu_exact(srcs, weights, out, accumulate):
    """Reference semantics: every term is computed in double, truncated to T, summed in T."""
    T = srcs[0].dtype
    acc = out.clone() if accumulate else torch.zeros_like(srcs[0])
    for x, w in zip(srcs, weights):
        term = (x.to(torch.float64) * float(w))
        if T in _INT_DTYPES:
            term = term.trunc().to(torch.int64).to(T) if T not in (torch.uint64,) else term.trunc().to(T)
        else:
            term = term.to(T)
        acc = (acc + term).to(T)
    out.copy_(acc)
    return out


def weighted_sum(srcs: Sequence[torch.Tensor], weights: Sequence[float], out: Optional[torch.Tensor] = None,
                 out_dtype: Optional[torch.dtype] = None, acc: Optional[torch.Tensor] = None,
                 accumulate: bool = False, post_scale: float = 1.0, exact: bool = False,
                 write_out: bool = True, stream=None) -> Optional[torch.Tensor]:
    """``out = cast((accumulate ? prev : 0) + sum_l w_l * x_l) * post_scale`` over flat tensors.

    ``acc`` is an optional fp32 running sum (streaming aggregation); ``exact`` selects the reference's
    per-term-truncation semantics (needed bit-for-bit for integer variables).
    """
    assert len(srcs) == len(weights) and len(srcs) > 0
    x0 = srcs[0]
    n = x0.numel()
    for x in srcs:
        assert x.dtype == x0.dtype and x.numel() == n and x.is_contiguous() and x.device == x0.device
    if out is None and write_out:
        out = torch.empty(n, dtype=out_dtype or x0.dtype, device=x0.device)
    if not x0.is_cuda:
        if exact:
            return _cpu_exact(srcs, weights, out, accumulate)
        s = torch.zeros(n, dtype=torch.float32)
        if accumulate:
            s = acc.clone() if acc is not None else out.to(torch.float32)
        for x, w in zip(srcs, weights):
            s = s + float(w) * x.to(torch.float32)
        if acc is not None:
            acc.copy_(s)
        if write_out:
            out.copy_((s * post_scale).to(out.dtype))
        return out
    L = _lib.lib()
    st = _lib.stream_ptr(stream)
    chunks = [(srcs[i:i + _lib.MAX_SRC], weights[i:i + _lib.MAX_SRC]) for i in range(0, len(srcs), _lib.MAX_SRC)]
    if not exact and len(chunks) > 1 and acc is None:
        acc = torch.empty(n, dtype=torch.float32, device=x0.device)  # running sum across chunks
    for ci, (chunk, wch) in enumerate(chunks):
        last = ci == len(chunks) - 1
        dst = out if (exact or (last and write_out)) else None
        rc = L.mfl_weighted_sum(
            _lib.ptr(dst), _lib.wire(dst) if dst is not None else 0,
            _lib.voidp_array([x.data_ptr() for x in chunk]), _lib.double_array(wch), len(chunk), _lib.wire(x0), n,
            None if exact else _lib.ptr(acc) or None, int(accumulate or ci > 0),
            float(post_scale if last else 1.0), int(exact), st)
        _lib.check(rc, "mfl_weighted_sum")
    return out


def rolling_update(community: Optional[torch.Tensor], scaled_sum: torch.Tensor, x_new: torch.Tensor,
                   w_new: float, x_old: Optional[torch.Tensor] = None, w_old: float = 0.0,
                   z: float = 1.0, stream=None):
    """``S += w_new*x_new - w_old*x_old`` (fp32 state) and ``community = cast(S / z)``."""
    n = x_new.numel()
    assert scaled_sum.dtype == torch.float32 and scaled_sum.numel() == n
    if not x_new.is_cuda:
        scaled_sum.add_(x_new.to(torch.float32), alpha=float(w_new))
        if x_old is not None:
            scaled_sum.add_(x_old.to(torch.float32), alpha=-float(w_old))
        if community is not None:
            community.copy_((scaled_sum / float(z)).to(community.dtype))
        return community
    rc = _lib.lib().mfl_rolling_update(_lib.ptr(community), _lib.wire(x_new), _lib.ptr(scaled_sum),
                                       _lib.ptr(x_new), float(w_new), _lib.ptr(x_old), float(w_old),
                                       float(z), n, _lib.stream_ptr(stream))
    _lib.check(rc, "mfl_rolling_update")
    return community


class Quantifier:
    """Per-variable zero / non-zero counts of a flat arena (K4); owns its scratch and tables."""

    def __init__(self, offsets: Sequence[int], device):
        self.n_vars = len(offsets) - 1
        self.device = torch.device(device)
        self.offsets_host = list(int(o) for o in offsets)
        if self.device.type == "cuda":
            self.offsets = torch.tensor(self.offsets_host, dtype=torch.int64, device=self.device)
            self.counts = torch.zeros(2 * self.n_vars, dtype=torch.int64, device=self.device)
            nbytes = _lib.lib().mfl_quantify_scratch_bytes(self.n_vars)
            self.scratch = torch.empty(nbytes, dtype=torch.uint8, device=self.device)

    def __call__(self, flat: torch.Tensor, stream=None) -> torch.Tensor:
        """Returns an int64 tensor ``[n_vars, 2]`` of (non_zeros, zeros); asynchronous on CUDA."""
        assert flat.numel() >= self.offsets_host[-1]
        if not flat.is_cuda:
            out = torch.empty(self.n_vars, 2, dtype=torch.int64)
            for v in range(self.n_vars):
                seg = flat[self.offsets_host[v]:self.offsets_host[v + 1]]
                nz = int((seg != 0).sum())
                out[v, 0], out[v, 1] = nz, seg.numel() - nz
            return out
        rc = _lib.lib().mfl_quantify(flat.data_ptr(), _lib.wire(flat), self.offsets.data_ptr(), self.n_vars,
                                     self.counts.data_ptr(), self.scratch.data_ptr(), _lib.stream_ptr(stream))
        _lib.check(rc, "mfl_quantify", launches=2)
        return self.counts.view(self.n_vars, 2)


def fused_aggregate(src_slot_ptrs: Sequence[int], weights: Sequence[float], out_ptrs: Sequence[int],
                    signal_pad_ptrs: Sequence[int], rank: int, world: int, segments, install_f32=None,
                    n_blocks: Optional[int] = None, stream=None, mc_src_ptrs: Optional[Sequence[int]] = None,
                    mc_out_ptr: int = 0):
    """Launch the fused reduce-scatter + all-gather kernel (every rank must make the same call).

    ``segments`` is a list of ``(offset_bytes, numel, torch_dtype_or_wire, exact)``.  With ``mc_src_ptrs`` (multicast
    addresses of the local slot indices) and ``mc_out_ptr`` the floating-point segments are reduced inside the
    NVSwitch and multicast back (NVLS); the slots must then hold pre-scaled models and all learners participate.
    """
    L = _lib.lib()
    segs = (_lib.Segment * len(segments))()
    for s, (off, numel, dt, exact) in zip(segs, segments):
        s.offset_bytes, s.numel = int(off), int(numel)
        s.dtype = dt if isinstance(dt, int) else _lib.wire(dt)
        s.exact = int(bool(exact))
    if n_blocks is None:
        n_blocks = default_fused_blocks()
    if mc_src_ptrs:
        rc = L.mfl_fused_aggregate_nvls(_lib.voidp_array(src_slot_ptrs), _lib.double_array(weights),
                                        len(src_slot_ptrs), _lib.voidp_array(out_ptrs),
                                        _lib.voidp_array(signal_pad_ptrs), rank, world, segs, len(segments),
                                        _lib.ptr(install_f32), int(n_blocks), _lib.stream_ptr(stream),
                                        _lib.voidp_array(mc_src_ptrs), len(mc_src_ptrs), int(mc_out_ptr))
        _lib.check(rc, "mfl_fused_aggregate_nvls")
        return
    rc = L.mfl_fused_aggregate(_lib.voidp_array(src_slot_ptrs), _lib.double_array(weights), len(src_slot_ptrs),
                               _lib.voidp_array(out_ptrs), _lib.voidp_array(signal_pad_ptrs), rank, world,
                               segs, len(segments), _lib.ptr(install_f32), int(n_blocks), _lib.stream_ptr(stream))
    _lib.check(rc, "mfl_fused_aggregate")


def default_fused_blocks() -> int:
    return 296
