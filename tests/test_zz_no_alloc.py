"""The allocation-free quantifier entry point, driven through the C ABI exactly as a non-Python
host would drive it.

``mfl_quantify_noalloc`` exists for hosts that cannot get device scratch of their own (plain-C
controller plug-ins, processes whose GPU memory belongs to another allocator, latency-critical
paths that must not hit the synchronisation inside cudaMalloc).  It must give the same counts as
``mfl_quantify`` (which takes caller scratch) - counts are integers, so the comparison is exact,
like the other quantifier tests - and it must do so on consecutive calls, because all calls share
one built-in scratch array that still holds the previous call's records.

This file sorts last on purpose: a run that stops at the first failure has then already executed
every other test.
"""
import ctypes as C

import pytest
import torch

from metisfl_b200.ops import _lib

pytestmark = pytest.mark.gpu

NOALLOC_SCRATCH_BYTES = 16384  # MFL_QUANTIFY_NOALLOC_SCRATCH_BYTES in include/metisfl_b200.h


def _abi():
    L = C.CDLL(_lib.LIB_PATH)
    vp = C.c_void_p
    L.mfl_last_error.restype = C.c_char_p
    L.mfl_quantify_scratch_bytes.restype = C.c_size_t
    L.mfl_quantify_scratch_bytes.argtypes = [C.c_int]
    L.mfl_quantify.argtypes = [vp, C.c_int, vp, C.c_int, vp, vp, vp]
    L.mfl_quantify.restype = C.c_int
    L.mfl_quantify_noalloc.argtypes = [vp, C.c_int, vp, C.c_int, vp, vp]
    L.mfl_quantify_noalloc.restype = C.c_int
    return L


def _case(sizes, dtype, seed, dev):
    offs = [0]
    for s in sizes:
        offs.append(offs[-1] + s)
    g = torch.Generator().manual_seed(seed)
    flat = (torch.randint(0, 3, (offs[-1],), generator=g) - 1).to(dtype).to(dev)
    offsets = torch.tensor(offs, dtype=torch.int64, device=dev)
    return flat, offsets, len(sizes)


def test_noalloc_quantifier_matches_scratch_version_twice():
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    L = _abi()
    stream = torch.cuda.current_stream(dev).cuda_stream
    cases = [
        # many CTAs see partial variables -> every scratch record is populated
        _case([3, 0, 1_000_003, 17, 64, 1, 2_500_000, 5], torch.bfloat16, 11, dev),
        # much smaller arena, other dtype: most CTAs get an empty range and must overwrite the
        # records the first call left behind
        _case([5, 129, 0, 7, 4096], torch.float32, 12, dev),
    ]
    for flat, offsets, n_vars in cases:
        need = L.mfl_quantify_scratch_bytes(n_vars)
        assert need <= NOALLOC_SCRATCH_BYTES
        scratch = torch.empty(need, dtype=torch.uint8, device=dev)
        want = torch.full((2 * n_vars,), -1, dtype=torch.int64, device=dev)
        got = torch.full((2 * n_vars,), -2, dtype=torch.int64, device=dev)
        code = _lib.wire(flat)
        rc = L.mfl_quantify(flat.data_ptr(), code, offsets.data_ptr(), n_vars, want.data_ptr(),
                            scratch.data_ptr(), stream)
        assert rc == 0, L.mfl_last_error()
        rc = L.mfl_quantify_noalloc(flat.data_ptr(), code, offsets.data_ptr(), n_vars, got.data_ptr(), stream)
        assert rc == 0, L.mfl_last_error()
        torch.cuda.synchronize(dev)
        assert torch.equal(got, want), (got.tolist(), want.tolist())
        # and against a straightforward count on the host
        host = flat.cpu()
        offs = offsets.cpu().tolist()
        for v in range(n_vars):
            seg = host[offs[v]:offs[v + 1]]
            nz = int((seg != 0).sum())
            assert got[2 * v].item() == nz and got[2 * v + 1].item() == seg.numel() - nz


def test_noalloc_quantifier_rejects_bad_arguments():
    L = _abi()
    rc = L.mfl_quantify_noalloc(None, 8, None, 1, None, None)
    assert rc != 0 and b"required" in L.mfl_last_error()
