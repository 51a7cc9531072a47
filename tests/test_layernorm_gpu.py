"""Fused residual-add + LayerNorm (K18) against fp32 PyTorch of the same op, forward and backward."""
import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("rows,d", [(4096, 1024), (77, 768), (5, 256), (300, 512)])
@pytest.mark.parametrize("param_dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("with_add", [True, False])
def test_add_layer_norm_matches_reference(rows, d, param_dtype, with_add):
    from metisfl_b200.ops import layernorm as fln
    g = torch.Generator().manual_seed(rows + d)
    mk = lambda: (torch.randn(rows, d, generator=g) * 1.5 + 0.2).to(torch.bfloat16).cuda()
    a, b, gh, gr = mk(), mk(), mk(), mk()
    ln = nn.LayerNorm(d).cuda()
    with torch.no_grad():
        ln.weight.copy_(torch.rand(d) + 0.5); ln.bias.copy_(torch.randn(d) * 0.1)
    ln = ln.to(param_dtype)
    a1, b1 = a.clone().requires_grad_(), b.clone().requires_grad_()
    if with_add:
        r, h = fln.add_layer_norm(a1, b1, ln)
        (r.float() * gr.float()).sum().backward(retain_graph=True)
        (h.float() * gh.float()).sum().backward()
    else:
        h = fln.layer_norm(a1, ln)
        (h.float() * gh.float()).sum().backward()
    ar, br = a.float().requires_grad_(), b.float().requires_grad_()
    w32, b32 = ln.weight.detach().float().requires_grad_(), ln.bias.detach().float().requires_grad_()
    # value: the bf16-rounded sum the kernel stores; gradient: straight through to both addends (counted once)
    rr = (ar + br).detach().to(torch.bfloat16).float() + ((ar + br) - (ar + br).detach()) if with_add else ar
    hr = F.layer_norm(rr, (d,), w32, b32, ln.eps)
    loss = (hr * gh.float()).sum() + ((rr * gr.float()).sum() if with_add else 0.0)
    loss.backward()
    torch.testing.assert_close(h.float(), hr, atol=4e-2, rtol=3e-2)
    if with_add:
        torch.testing.assert_close(r.float(), rr, atol=2e-2, rtol=2e-2)
        torch.testing.assert_close(b1.grad.float(), br.grad, atol=6e-2, rtol=5e-2)
    torch.testing.assert_close(a1.grad.float(), ar.grad, atol=6e-2, rtol=5e-2)
    tol = 0.05 * max(1.0, rows ** 0.5)
    torch.testing.assert_close(ln.weight.grad.float(), w32.grad, atol=tol, rtol=5e-2)
    torch.testing.assert_close(ln.bias.grad.float(), b32.grad, atol=tol, rtol=5e-2)
