"""Fused LM cross-entropy (K16) against fp32 ``F.cross_entropy`` on the same bf16 logits: loss and logit gradients."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("rows,vocab", [(64, 50304), (33, 1000), (7, 264), (128, 8)])
def test_lm_cross_entropy_matches_reference(rows, vocab):
    from metisfl_b200.ops.loss import lm_cross_entropy
    g = torch.Generator().manual_seed(rows + vocab)
    logits = (torch.randn(rows, vocab, generator=g) * 3).to(torch.bfloat16).cuda().requires_grad_()
    targets = torch.randint(0, vocab, (rows,), generator=g).cuda()
    targets[::5] = -100
    loss = lm_cross_entropy(logits, targets)
    (loss * 2.0).backward()
    ref_in = logits.detach().float().requires_grad_()
    ref = F.cross_entropy(ref_in, targets, ignore_index=-100)
    (ref * 2.0).backward()
    torch.testing.assert_close(loss.float(), ref, atol=2e-3, rtol=2e-3)
    torch.testing.assert_close(logits.grad.float(), ref_in.grad, atol=2e-4, rtol=2e-2)
    assert float(logits.grad[0].abs().sum()) == 0.0  # ignored row


def test_gpt2_lm_loss_uses_fused_kernel_and_matches():
    from metisfl_b200.models import zoo
    torch.manual_seed(0)
    tokens = torch.randint(0, 50257, (2, 64), device="cuda")
    logits = torch.randn(2, 64, 50304, device="cuda").to(torch.bfloat16).requires_grad_()
    loss = zoo.lm_loss(logits, tokens)
    loss.backward()
    ref_in = logits.detach().float().requires_grad_()
    ref = F.cross_entropy(ref_in[:, :-1].reshape(-1, 50304), tokens[:, 1:].reshape(-1))
    ref.backward()
    torch.testing.assert_close(loss.float(), ref, atol=2e-3, rtol=2e-3)
    torch.testing.assert_close(logits.grad.float(), ref_in.grad, atol=2e-4, rtol=2e-2)
