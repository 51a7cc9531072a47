"""Flat, dtype-segmented model arenas.

The reference moves a model as a list of independently serialised tensors
(``Model.variables``, ``model.proto:79-99``) and touches them one at a time
(``federated_average.cc:101-145``).  Here a model's whole state (parameters AND buffers, fixing
reference quirk Q10) lives in ONE contiguous device buffer with a tiny table of dtype-homogeneous
segments, so optimizer steps, publishing, aggregation and packing are each a single launch, and
the same byte layout is used by the working copy, the peer-visible publish slots and the
community buffers.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Callable, Dict, List, Optional, Tuple

import torch
import torch.nn as nn

from metisfl_b200.utils.dtypes import torch_to_wire

_ALIGN_VAR = 16      # every variable starts on a 16-byte boundary (vector loads)
_ALIGN_SEG = 256     # segments start on 256-byte boundaries

_NORM_TYPES = (nn.modules.batchnorm._BatchNorm,)
_INT_DTYPES = (torch.int8, torch.int16, torch.int32, torch.int64, torch.uint8, torch.bool)


@dataclass
class VarInfo:
    name: str
    shape: Tuple[int, ...]
    dtype: torch.dtype        # storage dtype inside the arena
    trainable: bool
    segment: int
    elem_offset: int          # element offset inside the segment (in units of dtype)
    numel: int
    channels_last: bool = False  # 4-D convolution filters are stored [K,R,S,C] (what NHWC convolutions consume)

    def shaped(self, flat: torch.Tensor) -> torch.Tensor:
        """Logical-shape view of this variable's flat storage slice."""
        if self.channels_last:
            k, c, r, s = self.shape
            return flat.view(k, r, s, c).permute(0, 3, 1, 2)
        return flat.view(self.shape)


@dataclass
class SegmentInfo:
    dtype: torch.dtype
    trainable: bool
    byte_offset: int
    numel: int                # padded element count
    exact: bool               # aggregate with reference truncation semantics (integer types)

    @property
    def nbytes(self) -> int:
        return self.numel * self.dtype.itemsize


def default_dtype_policy(model_dtype: torch.dtype) -> Callable[[nn.Module, str, torch.Tensor, bool], torch.dtype]:
    """Low-precision storage for matmul/conv/embedding/LayerNorm parameters; BatchNorm affine
    parameters and running statistics stay fp32 (cuDNN's mixed mode); integers keep their type."""

    def policy(owner: nn.Module, name: str, t: torch.Tensor, is_param: bool) -> torch.dtype:
        if not t.is_floating_point():
            return t.dtype
        if isinstance(owner, _NORM_TYPES):
            return torch.float32
        return model_dtype if is_param else torch.float32

    return policy


class ArenaLayout:
    """Variable table + segment table of a model; independent of any storage."""

    def __init__(self, variables: List[VarInfo], segments: List[SegmentInfo]):
        self.variables = variables
        self.segments = segments
        self.total_bytes = 0
        for s in segments:
            self.total_bytes = max(self.total_bytes, s.byte_offset + s.nbytes)
        self.total_bytes = (self.total_bytes + _ALIGN_SEG - 1) // _ALIGN_SEG * _ALIGN_SEG

    @property
    def num_params(self) -> int:
        return sum(v.numel for v in self.variables)

    def segment_table(self):
        """``(offset_bytes, numel, torch dtype, exact)`` rows for the fused aggregation kernel."""
        return [(s.byte_offset, s.numel, s.dtype, s.exact) for s in self.segments]

    def views(self, buf: torch.Tensor) -> List[torch.Tensor]:
        """Typed flat view of every segment of ``buf`` (a uint8 tensor with this layout)."""
        out = []
        for s in self.segments:
            out.append(buf[s.byte_offset:s.byte_offset + s.nbytes].view(s.dtype))
        return out

    def var_view(self, seg_views: List[torch.Tensor], v: VarInfo) -> torch.Tensor:
        return v.shaped(seg_views[v.segment][v.elem_offset:v.elem_offset + v.numel])

    @staticmethod
    def from_module(module: nn.Module, model_dtype: torch.dtype, policy=None,
                    conv_channels_last: bool = True) -> "ArenaLayout":
        """``conv_channels_last``: keep Conv2d filters in NHWC order inside the arena, so that the NHWC convolution
        kernels read them in place and write weight gradients straight in the layout of the flat gradient buffer
        (otherwise every step pays a layout copy per filter in forward and backward plus a strided accumulate; 225 us
        per ResNet-18 step, profiles/r1_step_breakdown.md).  The wire format and ``state_dict`` keep the logical
        [K,C,R,S] order - ``VarInfo.shaped`` returns a permuted view."""
        policy = policy or default_dtype_policy(model_dtype)
        cl_ids = set()
        if conv_channels_last:
            for mod in module.modules():
                if isinstance(mod, nn.Conv2d) and mod.weight is not None and mod.weight.dim() == 4:
                    cl_ids.add(id(mod.weight))
        entries = []  # (name, tensor, trainable, dtype)
        seen = set()
        for mod_name, mod in module.named_modules():
            for pname, p in mod.named_parameters(recurse=False):
                if id(p) in seen:
                    continue  # tied weights are stored once
                seen.add(id(p))
                full = f"{mod_name}.{pname}" if mod_name else pname
                entries.append((full, p, bool(p.requires_grad), policy(mod, pname, p, True)))
            for bname, b in mod.named_buffers(recurse=False):
                if b is None or id(b) in seen or bname in mod._non_persistent_buffers_set:
                    continue
                seen.add(id(b))
                full = f"{mod_name}.{bname}" if mod_name else bname
                entries.append((full, b, False, policy(mod, bname, b, False)))
        # segment key: (trainable, dtype); trainable low-precision first so it is segment 0
        keys: List[Tuple[bool, torch.dtype]] = []
        for _, _, tr, dt in entries:
            if (tr, dt) not in keys:
                keys.append((tr, dt))
        keys.sort(key=lambda k: (not k[0], 0 if k[1] == model_dtype else 1, str(k[1])))
        cursors: Dict[Tuple[bool, torch.dtype], int] = {k: 0 for k in keys}
        variables = []
        for name, t, tr, dt in entries:
            k = (tr, dt)
            per = _ALIGN_VAR // dt.itemsize if dt.itemsize <= _ALIGN_VAR else 1
            off = (cursors[k] + per - 1) // per * per
            variables.append(VarInfo(name, tuple(t.shape), dt, tr, keys.index(k), off, t.numel(), id(t) in cl_ids))
            cursors[k] = off + t.numel()
        segments, byte = [], 0
        for k in keys:
            tr, dt = k
            per = _ALIGN_VAR // dt.itemsize if dt.itemsize <= _ALIGN_VAR else 1
            numel = (cursors[k] + per - 1) // per * per
            byte = (byte + _ALIGN_SEG - 1) // _ALIGN_SEG * _ALIGN_SEG
            segments.append(SegmentInfo(dt, tr, byte, numel, exact=dt in _INT_DTYPES))
            byte += numel * dt.itemsize
        return ArenaLayout(variables, segments)


class ModelArena:
    """A module whose parameters/buffers are views into one flat buffer with an ``ArenaLayout``."""

    def __init__(self, module: nn.Module, device, model_dtype: torch.dtype = torch.bfloat16, policy=None,
                 storage: Optional[torch.Tensor] = None):
        self.module = module
        self.device = torch.device(device)
        self.model_dtype = model_dtype
        self.layout = ArenaLayout.from_module(module, model_dtype, policy)
        self.state = storage if storage is not None else torch.zeros(
            self.layout.total_bytes, dtype=torch.uint8, device=self.device)
        assert self.state.numel() >= self.layout.total_bytes and self.state.dtype == torch.uint8
        self.seg = self.layout.views(self.state)
        # fp32 master copy + flat grads for every trainable segment
        self.master: List[Optional[torch.Tensor]] = []
        self.grad: List[Optional[torch.Tensor]] = []
        for s, view in zip(self.layout.segments, self.seg):
            if s.trainable:
                self.grad.append(torch.zeros(s.numel, dtype=s.dtype, device=self.device))
                self.master.append(view if s.dtype == torch.float32 else
                                   torch.zeros(s.numel, dtype=torch.float32, device=self.device))
            else:
                self.grad.append(None)
                self.master.append(None)
        self._bind()

    def _bind(self):
        """Copy the module's current values into the arena and re-point its tensors at the views."""
        tensors = dict(self.module.named_parameters(remove_duplicate=True))
        tensors.update(dict(self.module.named_buffers(remove_duplicate=True)))
        with torch.no_grad():
            for v in self.layout.variables:
                t = tensors[v.name]
                view = self.layout.var_view(self.seg, v)
                view.copy_(t.detach().to(device=self.device, dtype=v.dtype))
                t.data = view
                if v.trainable:
                    g = v.shaped(self.grad[v.segment][v.elem_offset:v.elem_offset + v.numel])
                    t.grad = g
        # non-persistent buffers and anything else still on the old device
        for mod in self.module.modules():
            for bname in list(mod._non_persistent_buffers_set):
                b = mod._buffers.get(bname)
                if b is not None:
                    mod._buffers[bname] = b.to(self.device)
        self.sync_master_from_model()

    def sync_master_from_model(self):
        for s, view, m in zip(self.layout.segments, self.seg, self.master):
            if s.trainable and m is not view:
                m.copy_(view)

    def load_state_bytes(self, src: torch.Tensor):
        """Install a community model (same layout) as the working copy and refresh the fp32 masters."""
        self.state.copy_(src[:self.state.numel()], non_blocking=True)
        self.sync_master_from_model()

    def named_variables(self):
        """``(name, tensor view, trainable)`` in layout order — what goes on the wire."""
        for v in self.layout.variables:
            yield v.name, self.layout.var_view(self.seg, v), v.trainable
