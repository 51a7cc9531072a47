"""tcgen05 GEMM (K10) against an fp32 PyTorch reference of the same op, including ragged shapes,
fused bias/activation epilogues, fp32 output and the autograd path."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _ref(a, b, bias, act):
    y = a.float() @ b.float().t()
    if bias is not None:
        y = y + bias.float()
    if act == "relu":
        y = F.relu(y)
    elif act == "gelu":
        y = F.gelu(y, approximate="tanh")
    return y


@pytest.mark.parametrize("m,n,k", [(128, 128, 64), (128, 128, 512), (256, 384, 1024), (100, 10, 784), (77, 200, 72),
                                   (1000, 130, 40), (2048, 4096, 1024)])
@pytest.mark.parametrize("act", [None, "relu", "gelu"])
def test_gemm_matches_fp32_reference(m, n, k, act):
    from metisfl_b200.ops import gemm
    g = torch.Generator(device="cpu").manual_seed(m * 7 + n * 3 + k)
    a = (torch.randn(m, k, generator=g) / k ** 0.5).to(torch.bfloat16).cuda()
    b = torch.randn(n, k, generator=g).to(torch.bfloat16).cuda()
    bias = torch.randn(n, generator=g).to(torch.bfloat16).cuda()
    y = gemm.gemm_nt(a, b, bias, act)
    ref = _ref(a, b, bias, act)
    torch.testing.assert_close(y.float(), ref, atol=3e-2, rtol=3e-2)
    y32 = gemm.gemm_nt(a, b, None, None, out_dtype=torch.float32)
    torch.testing.assert_close(y32, _ref(a, b, None, None), atol=2e-3, rtol=2e-3)


@pytest.fixture
def pair_kernel():
    from metisfl_b200.ops import gemm
    old = gemm.set_pair_mode(1)
    yield
    gemm.set_pair_mode(old)


@pytest.mark.parametrize("m,n,k", [(256, 256, 64), (256, 128, 512), (512, 384, 1024), (300, 200, 72), (1000, 130, 40),
                                   (2048, 3072, 1024), (4096, 1024, 4096), (640, 1000, 264)])
@pytest.mark.parametrize("act", [None, "gelu"])
def test_pair_kernel_matches_fp32_reference(pair_kernel, m, n, k, act):
    """The CTA-pair kernel (cta_group::2, M = 256 per instruction) forced on: ragged M / N / K, all three tile widths
    (256, 192 for N = 3072, 128), bias + activation, bf16 and fp32 output, and the same result as the single-CTA kernel."""
    from metisfl_b200.ops import gemm
    g = torch.Generator(device="cpu").manual_seed(m * 5 + n * 3 + k)
    a = (torch.randn(m, k, generator=g) / k ** 0.5).to(torch.bfloat16).cuda()
    b = torch.randn(n, k, generator=g).to(torch.bfloat16).cuda()
    bias = torch.randn(n, generator=g).to(torch.bfloat16).cuda()
    y = gemm.gemm_nt(a, b, bias, act)
    torch.testing.assert_close(y.float(), _ref(a, b, bias, act), atol=3e-2, rtol=3e-2)
    y32 = gemm.gemm_nt(a, b, None, None, out_dtype=torch.float32)
    torch.testing.assert_close(y32, _ref(a, b, None, None), atol=2e-3, rtol=2e-3)
    gemm.set_pair_mode(0)
    y1 = gemm.gemm_nt(a, b, None, None, out_dtype=torch.float32)
    gemm.set_pair_mode(1)
    torch.testing.assert_close(y1, y32, atol=1e-4, rtol=1e-4)   # same reduction order per element


@pytest.mark.parametrize("m,n,k", [(256, 128, 64), (384, 320, 192), (4096, 1024, 1024), (1024, 3072, 1024)])
@pytest.mark.parametrize("a_mn,b_mn", [(False, True), (True, True), (True, False)])
def test_pair_kernel_mn_major_operands(pair_kernel, m, n, k, a_mn, b_mn):
    from metisfl_b200.ops import gemm
    g = torch.Generator().manual_seed(m + n + k)
    a = torch.randn(m, k, generator=g).to(torch.bfloat16).cuda()
    b = torch.randn(n, k, generator=g).to(torch.bfloat16).cuda()
    ref = a.float() @ b.float().t()
    out = gemm.gemm_ex(a.t().contiguous() if a_mn else a, b.t().contiguous() if b_mn else b, a_mn, b_mn)
    # fp32 reference from the same bf16-rounded inputs: what is left is the rounding of the bf16 output (relative
    # 2^-8 per element at most, 1.7e-3 in the Frobenius norm) and the fp32 accumulation order
    err = (out.float() - ref).norm() / ref.norm()
    assert err < 5e-3, float(err)
    torch.testing.assert_close(out.float(), ref, atol=k ** 0.5 * 8e-3, rtol=1.5e-2)


def test_gemm_fp32_bias_and_strided_inputs():
    from metisfl_b200.ops import gemm
    big = torch.randn(300, 512, device="cuda").to(torch.bfloat16)
    a = big[:, :256]                      # row pitch 512, still K-contiguous: no copy needed
    b = torch.randn(64, 256, device="cuda").to(torch.bfloat16)
    bias = torch.randn(64, device="cuda")
    y = gemm.gemm_nt(a, b, bias)
    torch.testing.assert_close(y.float(), _ref(a, b, bias, None), atol=5e-2, rtol=3e-2)


def test_linear_autograd_matches_torch():
    from metisfl_b200.ops import gemm
    torch.manual_seed(0)
    x = torch.randn(4, 32, 256, device="cuda").to(torch.bfloat16).requires_grad_()
    w = (torch.randn(384, 256, device="cuda") / 16).to(torch.bfloat16).requires_grad_()
    b = torch.randn(384, device="cuda").to(torch.bfloat16).requires_grad_()
    y = gemm.linear(x, w, b)
    gy = torch.randn_like(y)
    y.backward(gy)
    xr, wr, br = [t.detach().float().requires_grad_() for t in (x, w, b)]
    yr = F.linear(xr, wr, br)
    yr.backward(gy.float())
    torch.testing.assert_close(y.float(), yr, atol=5e-2, rtol=3e-2)
    torch.testing.assert_close(x.grad.float(), xr.grad, atol=5e-2, rtol=5e-2)
    torch.testing.assert_close(w.grad.float(), wr.grad, atol=1e-1, rtol=5e-2)
    torch.testing.assert_close(b.grad.float(), br.grad, atol=1e-1, rtol=5e-2)


def test_mlp_trains_with_b200_linear():
    """The FashionMNIST MLP with every Linear on the tcgen05 kernel tracks the cuBLAS-backed model."""
    from metisfl_b200.models.zoo import FashionMnistMLP
    from metisfl_b200.ops import gemm
    torch.manual_seed(1)
    ref = FashionMnistMLP().cuda().to(torch.bfloat16)
    ours = FashionMnistMLP().cuda().to(torch.bfloat16)
    ours.load_state_dict(ref.state_dict())
    gemm.use_b200_linear(ours)
    x = torch.randn(256, 1, 28, 28, device="cuda").to(torch.bfloat16)
    y = torch.randint(0, 10, (256,), device="cuda")
    for model in (ref, ours):
        F.cross_entropy(model(x).float(), y).backward()
    for (n1, p1), (_, p2) in zip(ref.named_parameters(), ours.named_parameters()):
        torch.testing.assert_close(p2.grad.float(), p1.grad.float(), atol=2e-2, rtol=1e-1), n1


@pytest.mark.parametrize("m,n,k", [(256, 128, 64), (384, 320, 192), (4096, 1024, 1024), (200, 72, 136), (1024, 3072, 1024),
                                   (1024, 1024, 4096), (512, 768, 8192)])   # the last two take the split-k path
@pytest.mark.parametrize("a_mn,b_mn", [(False, True), (True, True), (True, False)])
def test_mn_major_operands_match_reference(m, n, k, a_mn, b_mn):
    """General-major GEMM: operands handed over as their transposes, transposed inside the tensor core."""
    from metisfl_b200.ops import gemm
    g = torch.Generator().manual_seed(m + n + k)
    a = torch.randn(m, k, generator=g).to(torch.bfloat16).cuda()
    b = torch.randn(n, k, generator=g).to(torch.bfloat16).cuda()
    ref = a.float() @ b.float().t()
    out = gemm.gemm_ex(a.t().contiguous() if a_mn else a, b.t().contiguous() if b_mn else b, a_mn, b_mn)
    # fp32 reference from the same bf16-rounded inputs: what is left is the rounding of the bf16 output (relative
    # 2^-8 per element at most, 1.7e-3 in the Frobenius norm) and the fp32 accumulation order
    err = (out.float() - ref).norm() / ref.norm()
    assert err < 5e-3, float(err)
    torch.testing.assert_close(out.float(), ref, atol=k ** 0.5 * 8e-3, rtol=1.5e-2)


def test_linear_backward_without_transposed_copies_matches_autograd():
    from metisfl_b200.ops import gemm
    torch.manual_seed(0)
    x = torch.randn(6, 64, 256, device="cuda").to(torch.bfloat16).requires_grad_()
    w = (torch.randn(512, 256, device="cuda") / 16).to(torch.bfloat16).requires_grad_()
    b = torch.randn(512, device="cuda").to(torch.bfloat16).requires_grad_()
    gy = torch.randn(6, 64, 512, device="cuda").to(torch.bfloat16)
    gemm.linear(x, w, b).backward(gy)
    xr, wr, br = (t.detach().float().requires_grad_() for t in (x, w, b))
    torch.nn.functional.linear(xr, wr, br).backward(gy.float())
    torch.testing.assert_close(x.grad.float(), xr.grad, atol=0.3, rtol=3e-2)
    torch.testing.assert_close(w.grad.float(), wr.grad, atol=0.6, rtol=3e-2)


@pytest.mark.parametrize("rows,cols", [(4096, 1024), (4096, 3072), (100, 72), (7, 8), (5000, 4096)])
def test_colsum_matches_reference(rows, cols):
    from metisfl_b200.ops import gemm
    x = torch.randn(rows, cols, device="cuda").to(torch.bfloat16)
    got = gemm.colsum(x)
    ref = x.float().sum(0)
    assert got.dtype == torch.float32
    torch.testing.assert_close(got, ref, atol=1e-2 * rows ** 0.5, rtol=1e-3)
