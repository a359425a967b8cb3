"""GPT-2 small (124 M) for the exchange stress config (BASELINE.json config 5).

12 layers, 12 heads, d_model 768, vocab 50257, context 1024, tied LM head.
Local step uses PyTorch ops (the product for this config is the
model exchange); the model is a next-token classifier over synthetic tokens.
Not part of the reference; required by BASELINE.json.
"""

from __future__ import annotations

import math
from typing import Optional

import torch
import torch.nn.functional as F
from torch import nn


class _Block(nn.Module):
    def __init__(self, d: int, heads: int) -> None:
        super().__init__()
        self.ln_1 = nn.LayerNorm(d)
        self.attn_qkv = nn.Linear(d, 3 * d)
        self.attn_proj = nn.Linear(d, d)
        self.ln_2 = nn.LayerNorm(d)
        self.mlp_fc = nn.Linear(d, 4 * d)
        self.mlp_proj = nn.Linear(4 * d, d)
        self.heads = heads

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        b, t, d = x.shape
        q, k, v = self.attn_qkv(self.ln_1(x)).split(d, dim=2)
        q = q.view(b, t, self.heads, d // self.heads).transpose(1, 2)
        k = k.view(b, t, self.heads, d // self.heads).transpose(1, 2)
        v = v.view(b, t, self.heads, d // self.heads).transpose(1, 2)
        a = F.scaled_dot_product_attention(q, k, v, is_causal=True)
        x = x + self.attn_proj(a.transpose(1, 2).reshape(b, t, d))
        return x + self.mlp_proj(F.gelu(self.mlp_fc(self.ln_2(x)), approximate="tanh"))


class GPT2(nn.Module):
    lr = 3e-4
    optimizer = "adam"
    fused_spec = "gpt2"
    is_language_model = True

    def __init__(
        self,
        vocab: int = 50257,
        ctx: int = 1024,
        d_model: int = 768,
        layers: int = 12,
        heads: int = 12,
        seed: Optional[int] = None,
    ) -> None:
        super().__init__()
        if seed is not None:
            torch.manual_seed(seed)
        self.wte = nn.Embedding(vocab, d_model)
        self.wpe = nn.Embedding(ctx, d_model)
        self.h = nn.ModuleList([_Block(d_model, heads) for _ in range(layers)])
        self.ln_f = nn.LayerNorm(d_model)
        self.vocab, self.ctx = vocab, ctx
        self.num_classes = vocab
        for m in self.modules():
            if isinstance(m, (nn.Linear, nn.Embedding)):
                nn.init.normal_(m.weight, std=0.02)
            if isinstance(m, nn.Linear) and m.bias is not None:
                nn.init.zeros_(m.bias)
        for blk in self.h:
            nn.init.normal_(blk.attn_proj.weight, std=0.02 / math.sqrt(2 * layers))
            nn.init.normal_(blk.mlp_proj.weight, std=0.02 / math.sqrt(2 * layers))

    def forward(self, idx: torch.Tensor) -> torch.Tensor:
        b, t = idx.shape
        pos = torch.arange(t, device=idx.device)
        x = self.wte(idx) + self.wpe(pos)[None]
        for blk in self.h:
            x = blk(x)
        x = self.ln_f(x)
        return x @ self.wte.weight.t()  # tied head: [b, t, vocab]


def gpt2_small(seed: Optional[int] = None) -> GPT2:
    return GPT2(seed=seed)


def gpt2_tiny(seed: Optional[int] = None) -> GPT2:
    """Same code path, toy size - used by CPU tests."""
    return GPT2(vocab=256, ctx=32, d_model=32, layers=2, heads=2, seed=seed)
