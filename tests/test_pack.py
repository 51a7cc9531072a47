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


def test_conv_filters_keep_logical_order_on_the_wire():
    """Conv2d filters are stored [K,R,S,C] inside the arena (NHWC kernels read them in place); the protobuf model and
    ``state_dict`` still see logical [K,C,R,S]."""
    import torch.nn as nn
    from metisfl_b200.ops import pack
    from metisfl_b200.parallel.arena import ModelArena
    from metisfl_b200.utils.proto_messages_factory import ModelProtoMessages
    torch.manual_seed(0)
    make = lambda: nn.Sequential(nn.Conv2d(3, 8, 3, padding=1), nn.BatchNorm2d(8), nn.ReLU(), nn.Conv2d(8, 16, 1),
                                 nn.Flatten(), nn.Linear(16 * 4 * 4, 5))
    src = make()
    want = {k: v.detach().clone() for k, v in src.state_dict().items()}
    a = ModelArena(src, "cpu", torch.float32)
    info = {v.name: v for v in a.layout.variables}
    assert info["0.weight"].channels_last and info["3.weight"].channels_last and not info["5.weight"].channels_last
    assert src[0].weight.is_contiguous(memory_format=torch.channels_last)
    assert src[0].weight.grad.stride() == src[0].weight.stride()
    for k, v in src.state_dict().items():          # binding did not change any logical value
        torch.testing.assert_close(v, want[k])
    pb = pack.arena_to_model_pb(a)
    for var in pb.variables:
        got = ModelProtoMessages.TensorSpecProto.proto_tensor_spec_to_numpy_array(var.plaintext_tensor.tensor_spec)
        assert tuple(got.shape) == tuple(want[var.name].shape)
        torch.testing.assert_close(torch.from_numpy(got.copy()).to(want[var.name].dtype), want[var.name])
    dst = make()
    b = ModelArena(dst, "cpu", torch.float32)
    pack.unpack_model_into_arena(pb, b)
    for k, v in dst.state_dict().items():
        torch.testing.assert_close(v, want[k])
    # the forward pass agrees too
    x = torch.randn(2, 3, 4, 4)
    src.eval(); dst.eval()
    torch.testing.assert_close(src(x), dst(x))


def test_unpack_refuses_payloads_that_do_not_match_their_declaration():
    """A wire model from outside (an off-box learner staged by the executor rank) is data, not code: a value shorter
    than its declared length, an unknown dtype code or an unknown variable is refused before anything is copied."""
    import pytest
    from metisfl_b200.models.zoo import FashionMnistMLP
    from metisfl_b200.ops.pack import arena_to_model_pb, unpack_model_into_arena
    from metisfl_b200.parallel.arena import ModelArena
    arena = ModelArena(FashionMnistMLP(), "cpu", torch.float32)
    good = arena_to_model_pb(arena)
    unpack_model_into_arena(good, arena)                      # the well-formed model installs
    short = type(good)(); short.CopyFrom(good)
    spec = short.variables[0].plaintext_tensor.tensor_spec
    spec.value = spec.value[:-4]
    with pytest.raises(ValueError):
        unpack_model_into_arena(short, arena)
    typed = type(good)(); typed.CopyFrom(good)
    typed.variables[0].plaintext_tensor.tensor_spec.type.type = 9     # declares FLOAT64, carries fp32 bytes
    with pytest.raises(ValueError):
        unpack_model_into_arena(typed, arena)
    named = type(good)(); named.CopyFrom(good)
    named.variables[0].name = "no.such.variable"      # (same count + unknown names = positional matching, by design)
    del named.variables[-1]                            # a different count: names must resolve
    with pytest.raises(ValueError):
        unpack_model_into_arena(named, arena)
    wrong_len = type(good)(); wrong_len.CopyFrom(good)
    wrong_len.variables[0].plaintext_tensor.tensor_spec.length += 1
    with pytest.raises(ValueError):
        unpack_model_into_arena(wrong_len, arena)


def test_arena_to_model_bytes_equals_the_message_path():
    """Serialized straight out of the arena == the message built variable by variable (channels-last filters included)."""
    import torch

    from metisfl_b200.ops import pack
    from metisfl_b200.parallel.arena import ModelArena
    from metisfl_b200.proto import model_pb2
    torch.manual_seed(0)
    net = torch.nn.Sequential(torch.nn.Conv2d(3, 8, 3), torch.nn.BatchNorm2d(8), torch.nn.Flatten(),
                              torch.nn.Linear(8, 4)).to(memory_format=torch.channels_last)
    arena = ModelArena(net, torch.device("cpu"), torch.float32)
    want = pack.arena_to_model_pb(arena)
    got = model_pb2.Model.FromString(pack.arena_to_model_bytes(arena))
    assert got == want and len(got.variables) == len(list(net.state_dict()))
