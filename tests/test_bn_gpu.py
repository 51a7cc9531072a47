"""Fused BatchNorm (+ residual) (+ ReLU) kernels (K14) against a plain fp32 PyTorch reference of the same op."""
import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

pytestmark = pytest.mark.gpu



@pytest.fixture(params=["cooperative", "three_launches", "clusters"], autouse=True)
def finalize_mode(request):
    """All back-ends: one cooperative launch with grid barriers, the three-launch sequence, and one launch on
    thread-block clusters (channel slices, partial sums combined through distributed shared memory)."""
    from metisfl_b200.ops import bn as fbn
    old = fbn.FUSED_FINALIZE
    fbn.FUSED_FINALIZE = request.param == "cooperative"
    old_cluster = fbn.set_cluster_mode(1 if request.param == "clusters" else 0)
    yield
    fbn.FUSED_FINALIZE = old
    fbn.set_cluster_mode(old_cluster)


SHAPES = [(128, 64, 32, 32), (32, 128, 16, 16), (7, 256, 5, 3), (16, 512, 4, 4), (3, 8, 9, 11), (2, 2048, 2, 2),
          (2, 64, 1, 1), (5, 32, 17, 13), (128, 128, 16, 16), (128, 256, 8, 8), (128, 512, 4, 4), (64, 16, 24, 24)]


def _inputs(shape, seed, device="cuda"):
    g = torch.Generator().manual_seed(seed)
    mk = lambda: (torch.randn(shape, generator=g) * 1.7 + 0.3).to(torch.bfloat16).to(device) \
        .contiguous(memory_format=torch.channels_last)
    return mk(), mk(), mk()


def _bn(c, affine=True):
    bn = nn.BatchNorm2d(c, affine=affine).cuda()
    if affine:
        with torch.no_grad():
            bn.weight.copy_(torch.rand(c) + 0.5)
            bn.bias.copy_(torch.randn(c) * 0.2)
    return bn


def _reference(bn_ref, x, res, relu):
    y = F.batch_norm(x, bn_ref.running_mean, bn_ref.running_var, bn_ref.weight, bn_ref.bias, bn_ref.training,
                     bn_ref.momentum, bn_ref.eps)
    if res is not None:
        y = y + res
    return F.relu(y) if relu else y


@pytest.mark.parametrize("shape", SHAPES)
@pytest.mark.parametrize("relu,use_res", [(True, True), (True, False), (False, True), (False, False)])
def test_training_forward_backward_match_fp32_reference(shape, relu, use_res):
    from metisfl_b200.ops import bn as fbn
    x, res, gy = _inputs(shape, 1)
    c = shape[1]
    mod = _bn(c)
    ref = _bn(c)
    ref.load_state_dict(mod.state_dict())
    x1, r1 = x.clone().requires_grad_(), (res.clone().requires_grad_() if use_res else None)
    y = fbn.bn_act(mod, x1, r1, relu=relu)
    assert y.dtype == torch.bfloat16 and y.is_contiguous(memory_format=torch.channels_last)
    y.backward(gy)
    xr, rr = x.float().requires_grad_(), (res.float().requires_grad_() if use_res else None)
    pre = _reference(ref, xr, rr, False)
    yr = F.relu(pre) if relu else pre
    # the kernel's ReLU mask comes from its own bf16 output; the reference gradient uses the same mask so that
    # elements within rounding distance of zero do not flip the comparison
    (pre * (y.float() > 0) if relu else pre).backward(gy.float())
    torch.testing.assert_close(y.float(), yr, atol=3e-2, rtol=2e-2)
    torch.testing.assert_close(mod.running_mean, ref.running_mean, atol=1e-4, rtol=1e-4)
    torch.testing.assert_close(mod.running_var, ref.running_var, atol=1e-4, rtol=1e-3)
    assert int(mod.num_batches_tracked) == 1
    m = shape[0] * shape[2] * shape[3]
    if m > 1:
        scale = max(1.0, float(xr.grad.abs().max()))
        torch.testing.assert_close(x1.grad.float(), xr.grad, atol=4e-2 * scale, rtol=4e-2)
        torch.testing.assert_close(mod.weight.grad, ref.weight.grad, atol=2e-2 * max(1.0, m ** 0.5), rtol=2e-2)
        torch.testing.assert_close(mod.bias.grad, ref.bias.grad, atol=2e-2 * max(1.0, m ** 0.5), rtol=2e-2)
        if use_res:
            torch.testing.assert_close(r1.grad.float(), rr.grad, atol=2e-2, rtol=2e-2)


def test_eval_mode_uses_running_statistics():
    from metisfl_b200.ops import bn as fbn
    x, res, _ = _inputs((8, 64, 6, 6), 3)
    mod = _bn(64)
    with torch.no_grad():
        mod.running_mean.copy_(torch.randn(64) * 0.1); mod.running_var.copy_(torch.rand(64) + 0.5)
    mod.eval()
    y = fbn.bn_act(mod, x, res, relu=True)
    yr = _reference(mod, x.float(), res.float(), True)
    torch.testing.assert_close(y.float(), yr, atol=3e-2, rtol=2e-2)
    assert int(mod.num_batches_tracked) == 0


def test_affine_free_and_direct_gradient_accumulation():
    from metisfl_b200.ops import bn as fbn
    x, _, gy = _inputs((16, 128, 8, 8), 5)
    plain = nn.BatchNorm2d(128, affine=False).cuda()
    x1 = x.clone().requires_grad_()
    fbn.bn_act(plain, x1, None, relu=False).backward(gy)
    xr = x.float().requires_grad_()
    F.batch_norm(xr, None, None, None, None, True).backward(gy.float())
    torch.testing.assert_close(x1.grad.float(), xr.grad, atol=4e-2, rtol=4e-2)
    # pre-allocated .grad (the arena case): the kernel accumulates into it, autograd gets nothing to add
    mod = _bn(128)
    mod.weight.grad = torch.full((128,), 2.0, device="cuda"); mod.bias.grad = torch.full((128,), -1.0, device="cuda")
    ref = _bn(128); ref.load_state_dict(mod.state_dict())
    fbn.bn_act(mod, x.clone().requires_grad_(), None, relu=False).backward(gy)
    F.batch_norm(x.float(), None, None, ref.weight, ref.bias, True).backward(gy.float())
    torch.testing.assert_close(mod.weight.grad - 2.0, ref.weight.grad, atol=0.5, rtol=2e-2)
    torch.testing.assert_close(mod.bias.grad + 1.0, ref.bias.grad, atol=0.5, rtol=2e-2)


def test_resnet18_fused_bn_is_as_close_to_fp32_as_the_aten_path():
    """bf16 noise through 20 layers makes two bf16 paths differ; judge both against an fp32 run of the same net."""
    from metisfl_b200.models.zoo import ResNet18
    from metisfl_b200.ops import bn as fbn
    torch.manual_seed(0)
    nets = [ResNet18().cuda().to(memory_format=torch.channels_last) for _ in range(3)]
    for n in nets[1:]:
        n.load_state_dict(nets[0].state_dict())
    for n in nets[:2]:
        for mod in n.modules():
            if isinstance(mod, (nn.Conv2d, nn.Linear)):
                mod.to(torch.bfloat16)
    x = torch.randn(64, 3, 32, 32, device="cuda").contiguous(memory_format=torch.channels_last)
    t = torch.randint(0, 10, (64,), device="cuda")
    losses, first, last = [], [], []
    for model, enabled, dt in ((nets[0], True, torch.bfloat16), (nets[1], False, torch.bfloat16),
                               (nets[2], False, torch.float32)):
        fbn.ENABLED = enabled
        try:
            loss = F.cross_entropy(model(x.to(dt)).float(), t)
            loss.backward()
        finally:
            fbn.ENABLED = True
        losses.append(float(loss.detach()))
        first.append(model.layers[0].conv1.weight.grad.float().flatten())
        last.append(model.layers[7].conv2.weight.grad.float().flatten())
    cos = lambda a, b: float(F.cosine_similarity(a, b, dim=0))
    assert abs(losses[0] - losses[2]) < 5e-2 and abs(losses[1] - losses[2]) < 5e-2, losses
    assert cos(last[0], last[2]) > max(0.95, cos(last[1], last[2]) - 0.02), (cos(last[0], last[2]), cos(last[1], last[2]))
    assert cos(first[0], first[2]) > cos(first[1], first[2]) - 0.03, (cos(first[0], first[2]), cos(first[1], first[2]))
    torch.testing.assert_close(nets[0].bn1.running_var, nets[2].bn1.running_var, atol=1e-3, rtol=1e-2)
