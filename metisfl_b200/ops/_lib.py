"""ctypes binding of ``libmetisfl_b200_kernels.so`` (C ABI in ``include/metisfl_b200.h``).

The library is built in-tree by ``build.py``; it is loaded lazily so that CPU-only
code paths (protocol, controller core, CPU oracles) never need it.  On a machine with a
GPU a missing library is a hard error: the ops never fall back to eager PyTorch silently.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

from metisfl_b200.utils.dtypes import torch_to_wire

_LIB = None
LIB_PATH = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                        "lib", "libmetisfl_b200_kernels.so")

MAX_SRC = 64
MAX_PEERS = 16
MAX_SEGMENTS = 8


class Segment(C.Structure):
    _fields_ = [("offset_bytes", C.c_uint64), ("numel", C.c_uint64),
                ("dtype", C.c_int32), ("exact", C.c_int32)]


class OptimCfg(C.Structure):
    _fields_ = [("kind", C.c_int32), ("lr", C.c_float), ("l1", C.c_float), ("l2", C.c_float),
                ("momentum", C.c_float), ("mu", C.c_float), ("beta1", C.c_float),
                ("beta2", C.c_float), ("eps", C.c_float), ("weight_decay", C.c_float),
                ("step", C.c_int32), ("grad_scale", C.c_float), ("publish_scale", C.c_float)]


class KernelLibraryError(RuntimeError):
    pass


def lib():
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise KernelLibraryError(
                f"{LIB_PATH} is missing: run `python build.py` (or __graft_entry__.build()) first")
        L = C.CDLL(LIB_PATH)
        L.mfl_last_error.restype = C.c_char_p
        vp, i, sz, d = C.c_void_p, C.c_int, C.c_size_t, C.c_double
        L.mfl_weighted_sum.argtypes = [vp, i, C.POINTER(vp), C.POINTER(d), i, i, sz, vp, i, d, i, vp]
        L.mfl_rolling_update.argtypes = [vp, i, vp, vp, d, vp, d, d, sz, vp]
        L.mfl_fused_aggregate.argtypes = [C.POINTER(vp), C.POINTER(d), i, C.POINTER(vp), C.POINTER(vp),
                                          i, i, C.POINTER(Segment), i, vp, i, vp]
        L.mfl_fused_aggregate_nvls.argtypes = [C.POINTER(vp), C.POINTER(d), i, C.POINTER(vp), C.POINTER(vp),
                                               i, i, C.POINTER(Segment), i, vp, i, vp, C.POINTER(vp), i, vp]
        L.mfl_signal_pad_bytes.restype = sz
        L.mfl_signal_pad_bytes.argtypes = [i]
        L.mfl_optimizer_step.argtypes = [vp, vp, i, vp, i, vp, vp, vp, vp, i, sz, C.POINTER(OptimCfg), i, vp]
        L.mfl_quantify_scratch_bytes.restype = sz
        L.mfl_quantify_scratch_bytes.argtypes = [i]
        L.mfl_quantify.argtypes = [vp, i, vp, i, vp, vp, vp]
        L.mfl_pack_cast.argtypes = [vp, i, vp, i, vp, vp, vp, i, i, vp]
        L.mfl_fabric_alloc.argtypes = [C.POINTER(vp), sz, vp]
        L.mfl_fabric_open.argtypes = [C.POINTER(vp), vp]
        L.mfl_fabric_close.argtypes = [vp]
        L.mfl_fabric_free.argtypes = [vp]
        L.mfl_fabric_enable_peers.argtypes = [i, i]
        L.mfl_fabric_barrier.argtypes = [C.POINTER(vp), i, i, i, vp]
        L.mfl_sm_count.argtypes = [i]
        ull = C.c_ulonglong
        L.mfl_async_fold.argtypes = [vp, d, vp, d, d, d, vp, C.POINTER(C.c_uint64), C.POINTER(Segment), i,
                                     C.POINTER(vp), vp, i, C.POINTER(vp), ull, vp, ull, vp, i, vp]
        L.mfl_flag_signal.argtypes = [vp, ull, vp]
        L.mfl_flag_wait.argtypes = [vp, ull, vp]
        L.mfl_snapshot_install.argtypes = [C.POINTER(vp), i, vp, vp, vp, sz, vp, C.POINTER(Segment), i, vp]
        L.mfl_spin_timeouts.argtypes = []
        _LIB = L
    return _LIB


def available() -> bool:
    return os.path.exists(LIB_PATH) and torch.cuda.is_available()


LAUNCHES = 0  # kernels of this library enqueued so far (bench.py reports the delta)


def check(rc: int, what: str, launches: int = 1):
    global LAUNCHES
    LAUNCHES += launches
    if rc != 0:
        msg = lib().mfl_last_error().decode(errors="replace")
        raise KernelLibraryError(f"{what} failed (rc={rc}): {msg}")


def stream_ptr(stream=None) -> int:
    s = stream if stream is not None else torch.cuda.current_stream()
    return s.cuda_stream


def ptr(t) -> int:
    """Device address of a tensor (or 0 for None)."""
    return 0 if t is None else t.data_ptr()


def wire(t_or_dtype) -> int:
    dt = t_or_dtype.dtype if isinstance(t_or_dtype, torch.Tensor) else t_or_dtype
    return torch_to_wire(dt)


def voidp_array(ptrs):
    return (C.c_void_p * len(ptrs))(*[C.c_void_p(p) for p in ptrs])


def segment_array(segments):
    """``(offset_bytes, numel, torch dtype or wire code, exact)`` rows -> ``mfl_segment[]``."""
    segs = (Segment * len(segments))()
    for s, (off, numel, dt, exact) in zip(segs, segments):
        s.offset_bytes, s.numel = int(off), int(numel)
        s.dtype = dt if isinstance(dt, int) else wire(dt)
        s.exact = int(bool(exact))
    return segs


def double_array(vals):
    return (C.c_double * len(vals))(*[float(v) for v in vals])
