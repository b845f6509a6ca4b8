"""A small LLaMA (RMSNorm, RoPE attention, SwiGLU MLP) with pipeline-stage variants.

Stands in for the external ``simplellm`` package the labs import
(lab/tutorial_1b/primer/intro.py:1-18): ``LLama``, ``LLamaFirstStage``,
``LLamaStage``, ``LLamaLastStage`` and ``causalLLMLoss`` keep the call-site signatures used in
lab/tutorial_1b/PP/1F1B/intro_PP_1F1B.py:29-39.  Lab shape: dmodel 288, 6 heads (head dim 48),
6 layers, ctx 256, vocab 32 000.

Stage 0 is defined explicitly as embedding + its share of the transformer layers
(the reference only ever calls ``.embed`` on it); ``forward`` runs both.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F
from torch import nn


class RMSNorm(nn.Module):
    def __init__(self, dim: int, eps: float = 1e-5):
        super().__init__()
        self.eps = eps
        self.weight = nn.Parameter(torch.ones(dim))

    def forward(self, x):
        return x * torch.rsqrt(x.pow(2).mean(-1, keepdim=True) + self.eps) * self.weight


def rope_tables(head_dim: int, ctx: int, theta: float = 10000.0, device=None):
    inv = 1.0 / (theta ** (torch.arange(0, head_dim, 2, device=device).float() / head_dim))
    ang = torch.outer(torch.arange(ctx, device=device).float(), inv)
    return ang.cos(), ang.sin()


def apply_rope(x, cos, sin):
    # x: [B, H, T, D]; rotate pairs (even, odd)
    x1, x2 = x[..., 0::2], x[..., 1::2]
    T = x.shape[-2]
    c, s = cos[:T], sin[:T]
    return torch.stack((x1 * c - x2 * s, x1 * s + x2 * c), dim=-1).flatten(-2)


class Attention(nn.Module):
    def __init__(self, dmodel: int, num_heads: int, ctx_size: int):
        super().__init__()
        assert dmodel % num_heads == 0
        self.h, self.hd = num_heads, dmodel // num_heads
        self.wqkv = nn.Linear(dmodel, 3 * dmodel, bias=False)
        self.wo = nn.Linear(dmodel, dmodel, bias=False)
        cos, sin = rope_tables(self.hd, ctx_size)
        self.register_buffer("rope_cos", cos, persistent=False)
        self.register_buffer("rope_sin", sin, persistent=False)
        self.attn_fn = None        # ops.nn.use_kernels() installs the hand-written causal attention here
        self.qkv_fn = None         # ... and the fused QKV split + transpose + rotary embedding

    def forward(self, x):
        B, T, C = x.shape
        if self.qkv_fn is not None and x.is_cuda and x.dtype == torch.float32:
            q, k, v = self.qkv_fn(self.wqkv(x), self.rope_cos, self.rope_sin, self.h)
        else:
            q, k, v = self.wqkv(x).view(B, T, 3, self.h, self.hd).permute(2, 0, 3, 1, 4)
            q, k = apply_rope(q, self.rope_cos, self.rope_sin), apply_rope(k, self.rope_cos, self.rope_sin)
        o = self.attn_fn(q, k, v) if self.attn_fn is not None else F.scaled_dot_product_attention(q, k, v, is_causal=True)
        return self.wo(o.transpose(1, 2).reshape(B, T, C))


class GatedMLP(nn.Module):
    def __init__(self, dmodel: int, hidden: int | None = None):
        super().__init__()
        hidden = hidden or int(math.ceil(8 * dmodel / 3 / 32) * 32)
        self.w_gate = nn.Linear(dmodel, hidden, bias=False)
        self.w_up = nn.Linear(dmodel, hidden, bias=False)
        self.w_down = nn.Linear(hidden, dmodel, bias=False)
        self.act_fn = None         # ops.nn.use_kernels() installs the fused SwiGLU here ...
        self.gate_up_fn = None     # ... and the one-GEMM gate + up projection over the stacked weight [w_gate; w_up]

    def forward(self, x):
        if self.gate_up_fn is not None:
            return self.w_down(self.gate_up_fn(x, self.w_gate.weight, self.w_up.weight))
        if self.act_fn is not None:
            return self.w_down(self.act_fn(self.w_gate(x), self.w_up(x)))
        return self.w_down(F.silu(self.w_gate(x)) * self.w_up(x))


class Block(nn.Module):
    def __init__(self, dmodel: int, num_heads: int, ctx_size: int):
        super().__init__()
        self.norm1, self.attn = RMSNorm(dmodel), Attention(dmodel, num_heads, ctx_size)
        self.norm2, self.mlp = RMSNorm(dmodel), GatedMLP(dmodel)

    def forward(self, x):
        x = x + self.attn(self.norm1(x))
        return x + self.mlp(self.norm2(x))


class LLamaStage(nn.Module):
    """``n_layers`` transformer blocks: hidden -> hidden (intro_PP_1F1B.py:34-35,66)."""

    def __init__(self, dmodel: int, num_heads: int, device="cpu", n_layers: int = 2, ctx_size: int = 256):
        super().__init__()
        self.layers = nn.ModuleList(Block(dmodel, num_heads, ctx_size) for _ in range(n_layers))
        self.to(device)

    def forward(self, h):
        for blk in self.layers:
            h = blk(h)
        return h


class LLamaFirstStage(LLamaStage):
    """Embedding + first blocks (intro_PP_1F1B.py:29-30).  ``embed(tokens)`` alone is what the
    reference calls; ``forward(tokens)`` = blocks(embed(tokens))."""

    def __init__(self, vocab_size: int, dmodel: int, num_heads: int, device="cpu", n_layers: int = 2,
                 ctx_size: int = 256, padding_idx: int | None = None):
        super().__init__(dmodel, num_heads, "cpu", n_layers, ctx_size)
        self.tok = nn.Embedding(vocab_size, dmodel, padding_idx=padding_idx)
        self.to(device)

    def embed(self, tokens):
        return self.tok(tokens)

    def forward(self, tokens):
        return super().forward(self.embed(tokens))


class LLamaLastStage(LLamaStage):
    """Last blocks + final norm + LM head: hidden -> logits (intro_PP_1F1B.py:38-39,78)."""

    def __init__(self, vocab_size: int, dmodel: int, num_heads: int, device="cpu", n_layers: int = 2,
                 ctx_size: int = 256):
        super().__init__(dmodel, num_heads, "cpu", n_layers, ctx_size)
        self.norm = RMSNorm(dmodel)
        self.lm_head = nn.Linear(dmodel, vocab_size, bias=False)
        self.to(device)

    def forward(self, h):
        return self.lm_head(self.norm(super().forward(h)))


class CausalLLama:  # marker kept for call-site parity: LLama(CausalLLama, vocab, ...)
    pass


class LLama(nn.Module):
    """Full model: ``LLama(CausalLLama, vocab_size, dmodel, num_heads, device, n_layers, ctx_size,
    padding_idx)``; ``net(tokens)`` -> logits (primer/intro.py:17-18,28)."""

    def __init__(self, kind=CausalLLama, vocab_size: int = 32_000, dmodel: int = 288, num_heads: int = 6,
                 device="cpu", n_layers: int = 6, ctx_size: int = 256, padding_idx: int | None = None):
        super().__init__()
        self.tok = nn.Embedding(vocab_size, dmodel, padding_idx=padding_idx)
        self.layers = nn.ModuleList(Block(dmodel, num_heads, ctx_size) for _ in range(n_layers))
        self.norm = RMSNorm(dmodel)
        self.lm_head = nn.Linear(dmodel, vocab_size, bias=False)
        self.to(device)

    def forward(self, tokens):
        h = self.tok(tokens)
        for blk in self.layers:
            h = blk(h)
        return self.lm_head(self.norm(h))

    def split_stages(self, n_stages: int) -> list[nn.Module]:
        """Partition into pipeline stages that *share* this model's parameters (used to test
        PP-split == unsplit)."""
        L = len(self.layers)
        per = [L // n_stages + (1 if i < L % n_stages else 0) for i in range(n_stages)]
        stages, lo = [], 0
        for i, k in enumerate(per):
            stages.append(_SharedStage(self, lo, lo + k, first=(i == 0), last=(i == n_stages - 1)))
            lo += k
        return stages


class _SharedStage(nn.Module):
    def __init__(self, full: LLama, lo: int, hi: int, first: bool, last: bool):
        super().__init__()
        self.layers = nn.ModuleList(full.layers[lo:hi])
        self.tok = full.tok if first else None
        self.norm = full.norm if last else None
        self.lm_head = full.lm_head if last else None

    def embed(self, tokens):
        return self.tok(tokens)

    def forward(self, x):
        if self.tok is not None:
            x = self.tok(x)
        for blk in self.layers:
            x = blk(x)
        if self.lm_head is not None:
            x = self.lm_head(self.norm(x))
        return x


def causalLLMLoss(logits: torch.Tensor, target_tokens: torch.Tensor, vocab_size: int | None = None,
                  ignore_index: int = -100) -> torch.Tensor:
    """Shifted next-token cross-entropy (primer/intro.py:29)."""
    V = logits.shape[-1] if vocab_size is None else vocab_size
    tgt = target_tokens.to(logits.device)[:, 1:].reshape(-1)
    return F.cross_entropy(logits[:, :-1].reshape(-1, V), tgt, ignore_index=ignore_index)
