"""K12 front-end: move a whole wire-format model into / out of a flat device arena in bulk.

Reference: every tensor is (de)serialised on the host one by one (``proto_messages_factory.py:419-494``,
``proto_tensor_serde.h:14-32``).  Here the payload bytes of all variables are concatenated into one pinned
staging buffer, copied to the device ONCE, and scattered / cast into the arena segments by one
``mfl_pack_cast`` launch per (wire dtype -> storage dtype) pair (``csrc/kernels/pack.cu``); the way
back is a single device->host copy of the arena followed by zero-copy slicing.
"""
from __future__ import annotations

from collections import defaultdict
from typing import List

import numpy as np
import torch

from metisfl_b200.ops import _lib
from metisfl_b200.proto import model_pb2
from metisfl_b200.utils import dtypes


def pack_cast(dst: torch.Tensor, dst_wire: int, src: torch.Tensor, src_wire: int, dst_off: List[int],
              src_off: List[int], numel: List[int], byte_swap: bool = False):
    """``dst[dst_off[t]:] <- cast(src[src_off[t]:])`` for every tensor t (byte offsets, uint8 buffers)."""
    n = len(numel)
    if n == 0:
        return
    if not dst.is_cuda:
        for d, s, k in zip(dst_off, src_off, numel):
            sv = src[s:s + k * dtypes.WIRE_SIZE[src_wire]].view(dtypes.wire_to_torch(src_wire))
            if byte_swap:
                sv = torch.from_numpy(sv.view(torch.uint8).numpy().reshape(k, -1)[:, ::-1].copy().reshape(-1)).view(sv.dtype)
            dst[d:d + k * dtypes.WIRE_SIZE[dst_wire]].view(dtypes.wire_to_torch(dst_wire)).copy_(sv)
        return
    table = torch.tensor([dst_off, src_off, numel], dtype=torch.int64).to(dst.device, non_blocking=True)
    rc = _lib.lib().mfl_pack_cast(dst.data_ptr(), dst_wire, src.data_ptr(), src_wire, table[0].data_ptr(),
                                  table[1].data_ptr(), table[2].data_ptr(), n, int(byte_swap), _lib.stream_ptr())
    _lib.check(rc, "mfl_pack_cast")


def unpack_model_into_arena(model_pb: model_pb2.Model, arena) -> None:
    """Install the plaintext variables of ``model_pb`` (matched by name, else by position) into ``arena``."""
    layout = arena.layout
    by_name = {v.name: v for v in layout.variables}
    positional = len(model_pb.variables) == len(layout.variables) and \
        not all(v.name in by_name for v in model_pb.variables)
    chunks, groups, cursor = [], defaultdict(lambda: ([], [], [])), 0
    for i, var in enumerate(model_pb.variables):
        if not var.HasField("plaintext_tensor"):
            raise RuntimeError("unpack_model_into_arena handles plaintext tensors; decrypt ciphertexts first")
        if not positional and var.name not in by_name:
            raise ValueError(f"variable {var.name!r} is not part of this model")
        info = layout.variables[i] if positional else by_name[var.name]
        spec = var.plaintext_tensor.tensor_spec
        if spec.length != info.numel:
            raise ValueError(f"variable {var.name}: {spec.length} values on the wire, {info.numel} in the model")
        # the payload is copied / cast on the device by element count: it must hold exactly that many elements of the
        # type it declares (a short or mistyped value would make the kernel read past the staging buffer)
        wire_type = int(spec.type.type)
        if not 0 <= wire_type < len(dtypes.WIRE_SIZE):
            raise ValueError(f"variable {var.name}: unknown wire data type {wire_type}")
        if len(spec.value) != info.numel * dtypes.WIRE_SIZE[wire_type]:
            raise ValueError(f"variable {var.name}: {len(spec.value)} payload bytes for {info.numel} "
                             f"{dtypes.WIRE_NAMES[wire_type]} values")
        seg = layout.segments[info.segment]
        swap = spec.type.byte_order == model_pb2.DType.BIG_ENDIAN_ORDER
        key = (int(spec.type.type), dtypes.torch_to_wire(seg.dtype), swap)
        pad = (-cursor) % 16
        cursor += pad
        chunks.append((pad, spec.value))
        d, s, k = groups[key]
        d.append(seg.byte_offset + info.elem_offset * seg.dtype.itemsize)
        s.append(cursor)
        k.append(info.numel)
        cursor += len(spec.value)
    staging = torch.empty(max(cursor, 16), dtype=torch.uint8, pin_memory=arena.device.type == "cuda")
    buf = staging.numpy()
    pos = 0
    for pad, value in chunks:
        pos += pad
        buf[pos:pos + len(value)] = np.frombuffer(value, dtype=np.uint8)
        pos += len(value)
    dev = staging.to(arena.device, non_blocking=True)
    for (src_wire, dst_wire, swap), (d, s, k) in groups.items():
        pack_cast(arena.state, dst_wire, dev, src_wire, d, s, k, swap)
    # convolution filters live in [K,R,S,C] order inside the arena; the wire carries logical [K,C,R,S]
    seg_views = layout.views(arena.state)
    for i, var in enumerate(model_pb.variables):
        info = layout.variables[i] if positional else by_name[var.name]
        if info.channels_last:
            flat = seg_views[info.segment][info.elem_offset:info.elem_offset + info.numel]
            k_, c_, r_, s_ = info.shape
            flat.copy_(flat.view(k_, c_, r_, s_).permute(0, 2, 3, 1).contiguous().view(-1))
    arena.sync_master_from_model()


def arena_to_model_pb(arena) -> model_pb2.Model:
    """One device->host copy of the arena, then slicing: the wire dtype is the storage dtype."""
    host = arena.state.detach().to("cpu")
    raw = host.numpy()
    m = model_pb2.Model()
    for info in arena.layout.variables:
        seg = arena.layout.segments[info.segment]
        lo = seg.byte_offset + info.elem_offset * seg.dtype.itemsize
        var = m.variables.add(name=info.name, trainable=info.trainable)
        spec = var.plaintext_tensor.tensor_spec
        spec.length = info.numel
        spec.dimensions.extend(info.shape)
        spec.type.type = dtypes.torch_to_wire(seg.dtype)
        spec.type.byte_order = model_pb2.DType.NA if seg.dtype.itemsize == 1 else model_pb2.DType.LITTLE_ENDIAN_ORDER
        chunk = raw[lo:lo + info.numel * seg.dtype.itemsize]
        if info.channels_last:  # stored [K,R,S,C] -> wire order [K,C,R,S]
            k_, c_, r_, s_ = info.shape
            chunk = np.ascontiguousarray(chunk.reshape(k_, r_, s_, c_, seg.dtype.itemsize).transpose(0, 3, 1, 2, 4))
        spec.value = chunk.tobytes()
    return m


def arena_to_model_bytes(arena) -> bytes:
    """``arena_to_model_pb(arena).SerializeToString()`` without the message objects: one device->host copy of the arena,
    then the core's wire writer copies each variable once, straight out of that host buffer (``_core.build_model``)."""
    from metisfl_b200 import _core
    raw = arena.state.detach().to("cpu").numpy()
    D = model_pb2.DType
    variables = []
    for info in arena.layout.variables:
        seg = arena.layout.segments[info.segment]
        lo = seg.byte_offset + info.elem_offset * seg.dtype.itemsize
        chunk = raw[lo:lo + info.numel * seg.dtype.itemsize]
        if info.channels_last:  # stored [K,R,S,C] -> wire order [K,C,R,S]
            k_, c_, r_, s_ = info.shape
            chunk = np.ascontiguousarray(
                chunk.reshape(k_, r_, s_, c_, seg.dtype.itemsize).transpose(0, 3, 1, 2, 4)).reshape(-1)
        variables.append({"name": info.name, "trainable": bool(info.trainable), "length": int(info.numel),
                          "dimensions": [int(d) for d in info.shape], "dtype": int(dtypes.torch_to_wire(seg.dtype)),
                          "byte_order": int(D.NA if seg.dtype.itemsize == 1 else D.LITTLE_ENDIAN_ORDER),
                          "value": chunk})
    return _core.build_model(variables)
