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
    segs = _lib.segment_array(segments)
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


# ---- K3 fused over NVLink (csrc/kernels/async_fold.cu) ----------------------------------------------------
def async_fold(x_new_ptr: int, w_new: float, x_old_ptr: int, w_old: float, decay: float, post_scale: float,
               state: torch.Tensor, state_elem_offsets: Sequence[int], segments, out_ptrs: Sequence[int],
               mc_out_ptr: int = 0, version_flag_ptrs: Optional[Sequence[int]] = None, new_version: int = 0,
               wait_flag_ptr: int = 0, wait_value: int = 0, done_counter: Optional[torch.Tensor] = None,
               n_blocks: int = 64, stream=None):
    """One asynchronous community update on the calling GPU (FedRec: ``decay=1, post_scale=1/z``; FedAsync:
    ``decay=1-a, w_new=a``): reads the finisher's slot(s) through (peer) pointers, updates the fp32 ``state`` and
    stores the new community model into every rank's buffer, then releases ``new_version`` to every rank."""
    import ctypes as C
    L = _lib.lib()
    segs = _lib.segment_array(segments)
    offs = (C.c_uint64 * len(state_elem_offsets))(*[int(o) for o in state_elem_offsets])
    flags = _lib.voidp_array(version_flag_ptrs) if version_flag_ptrs else None
    rc = L.mfl_async_fold(int(x_new_ptr), float(w_new), int(x_old_ptr) or None, float(w_old), float(decay),
                          float(post_scale), state.data_ptr(), offs, segs, len(segments),
                          _lib.voidp_array(out_ptrs), int(mc_out_ptr) or None, len(out_ptrs), flags,
                          int(new_version), int(wait_flag_ptr) or None, int(wait_value),
                          _lib.ptr(done_counter) or None, int(n_blocks), _lib.stream_ptr(stream))
    _lib.check(rc, "mfl_async_fold")


def flag_signal(flag_ptr: int, value: int, stream=None):
    _lib.check(_lib.lib().mfl_flag_signal(int(flag_ptr), int(value), _lib.stream_ptr(stream)), "mfl_flag_signal")


def flag_wait(flag_ptr: int, value: int, stream=None):
    _lib.check(_lib.lib().mfl_flag_wait(int(flag_ptr), int(value), _lib.stream_ptr(stream)), "mfl_flag_wait")


def snapshot_install(community_ptrs: Sequence[int], version_flag_ptr: int, status: torch.Tensor, dst: torch.Tensor,
                     nbytes: int, master: Optional[torch.Tensor], segment0, n_blocks: int = 0, stream=None):
    """Install the newest community model (rotating buffers, picked by the version counter) into ``dst`` and,
    if given, segment 0 into the fp32 ``master``; ``status`` (2 x uint64 on the device) gets the version read and
    the torn-read count."""
    seg = _lib.segment_array([segment0]) if segment0 is not None else None
    rc = _lib.lib().mfl_snapshot_install(_lib.voidp_array(community_ptrs), len(community_ptrs),
                                         int(version_flag_ptr), status.data_ptr(), dst.data_ptr(), int(nbytes),
                                         _lib.ptr(master) or None, seg, int(n_blocks), _lib.stream_ptr(stream))
    _lib.check(rc, "mfl_snapshot_install", launches=2)


def spin_timeouts() -> int:
    return int(_lib.lib().mfl_spin_timeouts())
