"""GPT-2 decoder (medium = 24 layers, d=1024, 16 heads, vocab 50257, ctx 1024: ~355 M parameters with
tied embeddings) — north-star config #5.  The reference ships no transformer (largest sequence model:
a 2-layer BiLSTM, ``examples/keras/models/imdb_lstm.py:21-26``).  Attention uses PyTorch SDPA
(flash kernels); the linears are the GEMM hot spots targeted by the tcgen05 kernel (K10).
"""
import math

import torch
import torch.nn as nn
import torch.nn.functional as F

from metisfl_b200.ops.gemm import linear
from metisfl_b200.ops.layernorm import add_layer_norm, layer_norm


class _SplitQKV(torch.autograd.Function):
    """[B, T, 3, H, hd] -> three [B, H, T, hd] views for the attention kernel.  Autograd's own backward for the
    equivalent view / permute / unbind chain stacks dq, dk, dv and then permutes the stack back (two passes over the
    3 * B * T * D tensor per layer, 2.3 ms per GPT-2-medium step); here the three gradients are written straight into
    the packed layout the QKV projection's backward GEMMs read."""

    @staticmethod
    def forward(ctx, qkv):
        ctx.shape = qkv.shape
        return tuple(qkv[:, :, i].transpose(1, 2) for i in range(3))

    @staticmethod
    def backward(ctx, dq, dk, dv):
        out = torch.empty(ctx.shape, dtype=dq.dtype, device=dq.device)
        for i, g in enumerate((dq, dk, dv)):
            out[:, :, i].copy_(g.transpose(1, 2))
        return out


class Block(nn.Module):
    def __init__(self, d, heads):
        super().__init__()
        self.ln1 = nn.LayerNorm(d)
        self.qkv = nn.Linear(d, 3 * d)
        self.proj = nn.Linear(d, d)
        self.ln2 = nn.LayerNorm(d)
        self.fc = nn.Linear(d, 4 * d)
        self.out = nn.Linear(4 * d, d)
        self.heads = heads

    def attn_branch(self, h):
        """proj(attention(qkv(h))) for an already normalised ``h``."""
        B, T, D = h.shape
        q, k, v = _SplitQKV.apply(self.qkv(h).view(B, T, 3, self.heads, D // self.heads))
        a = F.scaled_dot_product_attention(q, k, v, is_causal=True)
        return self.proj(a.transpose(1, 2).reshape(B, T, D))

    def mlp_branch(self, h):
        return self.out(F.gelu(self.fc(h), approximate="tanh"))

    def forward(self, x):
        x = x + self.attn_branch(self.ln1(x))
        return x + self.mlp_branch(self.ln2(x))


class GPT2(nn.Module):
    def __init__(self, vocab=50257, ctx=1024, d=1024, layers=24, heads=16, pad_vocab_to=64):
        super().__init__()
        # the embedding table / tied LM head is padded to a multiple of 64 rows (50257 -> 50304): an odd leading
        # dimension forces unaligned legacy GEMM kernels for the head (8.7 of 42 ms per step, profiles/README.md);
        # the extra rows are never indexed and only add 47 * d unused parameters
        self.vocab = vocab
        self.vocab_padded = (vocab + pad_vocab_to - 1) // pad_vocab_to * pad_vocab_to if pad_vocab_to else vocab
        self.wte = nn.Embedding(self.vocab_padded, d)
        self.wpe = nn.Embedding(ctx, d)
        self.blocks = nn.ModuleList([Block(d, heads) for _ in range(layers)])
        self.ln_f = nn.LayerNorm(d)
        self.apply(self._init)
        for b in self.blocks:  # GPT-2 residual scaling
            for lin in (b.proj, b.out):
                nn.init.normal_(lin.weight, std=0.02 / math.sqrt(2 * layers))

    @staticmethod
    def _init(m):
        if isinstance(m, (nn.Linear, nn.Embedding)):
            nn.init.normal_(m.weight, std=0.02)
            if isinstance(m, nn.Linear) and m.bias is not None:
                nn.init.zeros_(m.bias)

    def forward(self, idx):
        B, T = idx.shape
        r = self.wte(idx) + self.wpe(torch.arange(T, device=idx.device))
        # Pre-LN blocks with every "residual add + next LayerNorm" pair fused (K18, ops/layernorm.py): r is the
        # residual stream, h its normalised view for the next branch.  Off the GPU the two ops run separately.
        h = layer_norm(r, self.blocks[0].ln1)
        n = len(self.blocks)
        for i, blk in enumerate(self.blocks):
            r, h = add_layer_norm(r, blk.attn_branch(h), blk.ln2)
            r, h = add_layer_norm(r, blk.mlp_branch(h), self.blocks[i + 1].ln1 if i + 1 < n else self.ln_f)
        return linear(h, self.wte.weight)  # tied LM head (tcgen05 GEMM on CUDA bf16, F.linear elsewhere)


def gpt2_medium(**kw):
    return GPT2(d=1024, layers=24, heads=16, **kw)


def gpt2_small(**kw):
    return GPT2(d=768, layers=12, heads=12, **kw)


def lm_loss(logits, tokens):
    """Next-token cross entropy; ``tokens`` are the inputs themselves.  On CUDA bf16 logits the fused kernel K16
    (``ops/loss.py``) computes loss and logit gradients without an fp32 copy of the [B*T, V] logits: the last
    position of every sequence gets the ignore label instead of slicing the logits."""
    if logits.is_cuda and logits.dtype == torch.bfloat16:
        from metisfl_b200.ops.loss import lm_cross_entropy
        b, t = tokens.shape
        targets = torch.full((b, t), -100, dtype=torch.int64, device=tokens.device)
        targets[:, :-1] = tokens[:, 1:]
        return lm_cross_entropy(logits, targets, n_valid=b * (t - 1))
    return F.cross_entropy(logits[:, :-1].reshape(-1, logits.size(-1)).float(), tokens[:, 1:].reshape(-1))
