"""Peer-memory fabric: symmetric device buffers shared by the GPU processes of one NVSwitch box.

Reference counterpart: none — the reference ships every model as protobuf bytes over gRPC
(``controller.cc:696-759``, ``learner.py:197-206``).  Here every rank allocates the same set of
buffers, exports them with cudaIpc (``csrc/kernels/runtime.cu``), and maps all peers' copies, so the
aggregation kernels can ``ld.global`` / ``st.global`` any learner's memory over NVLink.

``torch.distributed`` (NCCL on GPUs, gloo on CPU) is used only for bootstrap: exchanging the
64-byte IPC handles and small metadata.  With ``world == 1`` or on CPU the same API degrades to
plain local tensors, which is what the CPU test tier runs.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional

import torch
import torch.distributed as dist

from metisfl_b200.ops import _lib


class _CudaMem:
    """``__cuda_array_interface__`` shim to view raw device memory as a torch tensor."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False),
                                         "version": 2, "strides": None}


class SymmetricBuffer:
    """One buffer per rank, all mapped into this process. ``ptrs[r]`` is rank r's base address here."""

    def __init__(self, local: torch.Tensor, ptrs: List[int], owner=None, mc_ptr: int = 0):
        self.local = local
        self.ptrs = ptrs
        self._owner = owner
        self.mc_ptr = mc_ptr  # multicast (NVLS) address of the same buffers, 0 when unavailable

    @property
    def nbytes(self) -> int:
        return self.local.numel()


class Fabric:
    def __init__(self, device=None, group=None, standalone: bool = False):
        self.group = group
        # standalone: a world-1 fabric inside a process that is part of a larger job (oracles, evaluation copies)
        self.distributed = dist.is_available() and dist.is_initialized() and not standalone
        self.rank = dist.get_rank(group) if self.distributed else 0
        self.world = dist.get_world_size(group) if self.distributed else 1
        if device is None:
            device = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() \
                else torch.device("cpu")
        self.device = torch.device(device)
        self.peer_mapped = self.device.type == "cuda"
        self._allocs = []
        self._symm_handles = []
        self._opened = []
        if self.peer_mapped and self.world > 1:
            torch.cuda.set_device(self.device)
        self._pads: Optional[SymmetricBuffer] = None

    # -- collective helpers (bootstrap only) -------------------------------------------------
    def all_gather_object(self, obj):
        if not self.distributed or self.world == 1:
            return [obj]
        out = [None] * self.world
        dist.all_gather_object(out, obj, group=self.group)
        return out

    def host_barrier(self):
        if self.distributed and self.world > 1 and dist.is_initialized():
            if self.device.type == "cuda":
                torch.cuda.synchronize(self.device)
            dist.barrier(group=self.group)

    # -- allocation --------------------------------------------------------------------------
    def alloc(self, nbytes: int, multicast: bool = False) -> SymmetricBuffer:
        """Collective: every rank must call with the same size, in the same order.

        ``multicast=True`` asks for a buffer that also has an NVLS multicast address (``mc_ptr``): the allocation and
        the cuMulticast binding are done by ``torch.distributed._symmetric_memory`` (plumbing only - the kernels that
        use the addresses are ours).  Falls back to the cudaIpc allocation (``mc_ptr == 0``) when unsupported."""
        nbytes = (int(nbytes) + 255) // 256 * 256
        if multicast and self.peer_mapped and self.world > 1:
            buf = self._alloc_multicast(nbytes)
            if buf is not None:
                return buf
        if not self.peer_mapped:
            t = torch.zeros(nbytes, dtype=torch.uint8, device=self.device)
            return SymmetricBuffer(t, [t.data_ptr()] * 1 if self.world == 1 else [0] * self.world)
        if self.world == 1:
            t = torch.zeros(nbytes, dtype=torch.uint8, device=self.device)
            return SymmetricBuffer(t, [t.data_ptr()])
        L = _lib.lib()
        p = C.c_void_p()
        handle = (C.c_ubyte * 64)()
        with torch.cuda.device(self.device):
            _lib.check(L.mfl_fabric_alloc(C.byref(p), nbytes, handle), "mfl_fabric_alloc")
        self._allocs.append(p.value)
        handles = self.all_gather_object(bytes(handle))
        ptrs = []
        for r, h in enumerate(handles):
            if r == self.rank:
                ptrs.append(p.value)
                continue
            q = C.c_void_p()
            hb = (C.c_ubyte * 64).from_buffer_copy(h)
            with torch.cuda.device(self.device):
                _lib.check(L.mfl_fabric_open(C.byref(q), hb), "mfl_fabric_open")
            self._opened.append(q.value)
            ptrs.append(q.value)
        local = torch.as_tensor(_CudaMem(p.value, nbytes), device=self.device)
        self.host_barrier()
        return SymmetricBuffer(local, ptrs, owner=self)

    def _alloc_multicast(self, nbytes: int) -> Optional[SymmetricBuffer]:
        ok, t, hdl = 1, None, None
        try:
            import torch.distributed._symmetric_memory as symm
            t = symm.empty(nbytes, dtype=torch.uint8, device=self.device)
            hdl = symm.rendezvous(t, self.group if self.group is not None else dist.group.WORLD)
            if not hdl.multicast_ptr or len(hdl.buffer_ptrs) != self.world:
                ok = 0
        except Exception:  # older driver / no fabric manager / API moved
            ok = 0
        flags = self.all_gather_object(ok)
        if not all(flags):
            return None
        t.zero_()
        self._symm_handles.append((t, hdl))
        self.host_barrier()
        return SymmetricBuffer(t, [int(p) for p in hdl.buffer_ptrs], owner=hdl, mc_ptr=int(hdl.multicast_ptr))

    def signal_pads(self, n_blocks: int = 296) -> SymmetricBuffer:
        if self._pads is None:
            nbytes = _lib.lib().mfl_signal_pad_bytes(n_blocks) if self.peer_mapped else 256
            self._pads = self.alloc(nbytes)
        return self._pads

    def device_barrier(self, channel: int = 0, stream=None):
        """Cross-GPU barrier enqueued on the stream (K13); a host barrier on CPU."""
        if not self.peer_mapped:
            return self.host_barrier()
        pads = self.signal_pads()
        rc = _lib.lib().mfl_fabric_barrier(_lib.voidp_array(pads.ptrs), self.rank, self.world, channel,
                                           _lib.stream_ptr(stream))
        _lib.check(rc, "mfl_fabric_barrier")

    def close(self):
        if not self.peer_mapped or self.world == 1:
            return
        torch.cuda.synchronize(self.device)
        self.host_barrier()
        L = _lib.lib()
        for q in self._opened:
            L.mfl_fabric_close(q)
        self._opened.clear()
        self.host_barrier()
        for p in self._allocs:
            L.mfl_fabric_free(p)
        self._allocs.clear()
