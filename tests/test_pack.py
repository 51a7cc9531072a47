"""K12: bulk wire <-> arena packing (CPU path here; the CUDA kernel is covered in test_kernels_gpu)."""
import numpy as np
import pytest
import torch

from metisfl_b200.models.zoo import ResNet18
from metisfl_b200.ops import pack
from metisfl_b200.parallel.arena import ModelArena
from metisfl_b200.proto import model_pb2


def _roundtrip(device):
    torch.manual_seed(0)
    src = ModelArena(ResNet18(), device, torch.bfloat16)
    for seg in src.seg:
        if seg.dtype.is_floating_point:
            seg.copy_(torch.randn(seg.numel()).to(seg.dtype))
        else:
            seg.copy_(torch.randint(0, 99, (seg.numel(),)).to(seg.dtype))
    m = pack.arena_to_model_pb(src)
    assert len(m.variables) == 122 and m.variables[0].plaintext_tensor.tensor_spec.type.type == model_pb2.DType.BFLOAT16
    dst = ModelArena(ResNet18(), device, torch.bfloat16)
    pack.unpack_model_into_arena(model_pb2.Model.FromString(m.SerializeToString()), dst)
    for a, b in zip(src.named_variables(), dst.named_variables()):
        assert a[0] == b[0] and torch.equal(a[1].cpu(), b[1].cpu()), a[0]
    assert torch.equal(dst.master[0].cpu(), dst.seg[0].float().cpu())
    # a float32 / big-endian sender is converted on the way in
    for v in m.variables:
        spec = v.plaintext_tensor.tensor_spec
        if spec.type.type == model_pb2.DType.BFLOAT16:
            t = torch.frombuffer(bytearray(spec.value), dtype=torch.bfloat16).float().numpy()
            spec.value = t.astype(">f4").tobytes()
            spec.type.type, spec.type.byte_order = model_pb2.DType.FLOAT32, model_pb2.DType.BIG_ENDIAN_ORDER
    dst2 = ModelArena(ResNet18(), device, torch.bfloat16)
    pack.unpack_model_into_arena(m, dst2)
    for a, b in zip(src.named_variables(), dst2.named_variables()):
        assert torch.equal(a[1].cpu(), b[1].cpu()), a[0]


def test_pack_roundtrip_cpu():
    _roundtrip("cpu")


@pytest.mark.gpu
def test_pack_roundtrip_cuda_kernel():
    _roundtrip("cuda")
