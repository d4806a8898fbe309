"""Stand-in for the third-party ``model`` module the reference imports.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  PARITY UNPINNED: the
reference does ``from model import GPT,GPTConfig,Block``
(partitions/gpt_model_parts.py:4) but ships no such file, lists it in no
requirements entry and pins no version.  The attribute names the reference
touches (``transformer.wte/wpe/h/ln_f``, ``lm_head``, ``config.block_size``;
gpt_model_parts.py:9-12,15,41-42) are those of the public nanoGPT ``model.py``,
so this file restates that model's published semantics, written from scratch:

  * GPTConfig(block_size=1024, vocab_size=50304, n_layer=12, n_head=12,
    n_embd=768, dropout=0.0, bias=True)
  * pre-LN blocks: x += attn(ln_1(x)); x += mlp(ln_2(x)); LayerNorm eps 1e-5
  * attn: c_attn Linear(C,3C) -> split q,k,v -> heads [B,H,T,C/H] ->
    causal softmax(q k^T / sqrt(C/H)) v -> merge heads -> c_proj Linear(C,C)
  * mlp: c_fc Linear(C,4C) -> exact (erf) GELU -> c_proj Linear(4C,C)
  * lm_head Linear(C,V,bias=False), weight tied to wte
  * init N(0,0.02); residual projections N(0, 0.02/sqrt(2*n_layer)); biases 0

It is cross-checked against transformers' GPT2LMHeadModel (activation "gelu")
in tests/test_oracle_gpt.py -- the best pin available offline.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import torch
import torch.nn as nn
import torch.nn.functional as F


@dataclass
class GPTConfig:
    block_size: int = 1024
    vocab_size: int = 50304
    n_layer: int = 12
    n_head: int = 12
    n_embd: int = 768
    dropout: float = 0.0
    bias: bool = True


class LayerNorm(nn.Module):
    def __init__(self, width: int, bias: bool):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(width))
        self.bias = nn.Parameter(torch.zeros(width)) if bias else None

    def forward(self, x):
        return F.layer_norm(x, self.weight.shape, self.weight, self.bias, 1e-5)


class CausalSelfAttention(nn.Module):
    def __init__(self, cfg: GPTConfig):
        super().__init__()
        assert cfg.n_embd % cfg.n_head == 0
        self.n_head, self.n_embd = cfg.n_head, cfg.n_embd
        self.c_attn = nn.Linear(cfg.n_embd, 3 * cfg.n_embd, bias=cfg.bias)
        self.c_proj = nn.Linear(cfg.n_embd, cfg.n_embd, bias=cfg.bias)

    def forward(self, x):
        b, t, c = x.shape
        hd = c // self.n_head
        q, k, v = self.c_attn(x).split(c, dim=2)
        q, k, v = (z.view(b, t, self.n_head, hd).transpose(1, 2) for z in (q, k, v))
        att = (q @ k.transpose(-2, -1)) * (1.0 / math.sqrt(hd))
        mask = torch.ones(t, t, dtype=torch.bool, device=x.device).tril()
        att = att.masked_fill(~mask, float("-inf")).softmax(dim=-1)
        y = (att @ v).transpose(1, 2).contiguous().view(b, t, c)
        return self.c_proj(y)


class MLP(nn.Module):
    def __init__(self, cfg: GPTConfig):
        super().__init__()
        self.c_fc = nn.Linear(cfg.n_embd, 4 * cfg.n_embd, bias=cfg.bias)
        self.gelu = nn.GELU()                      # exact erf form
        self.c_proj = nn.Linear(4 * cfg.n_embd, cfg.n_embd, bias=cfg.bias)

    def forward(self, x):
        return self.c_proj(self.gelu(self.c_fc(x)))


class Block(nn.Module):
    def __init__(self, cfg: GPTConfig):
        super().__init__()
        self.ln_1 = LayerNorm(cfg.n_embd, cfg.bias)
        self.attn = CausalSelfAttention(cfg)
        self.ln_2 = LayerNorm(cfg.n_embd, cfg.bias)
        self.mlp = MLP(cfg)

    def forward(self, x):
        x = x + self.attn(self.ln_1(x))
        return x + self.mlp(self.ln_2(x))


class GPT(nn.Module):
    def __init__(self, config: GPTConfig):
        super().__init__()
        self.config = config
        self.transformer = nn.ModuleDict(dict(
            wte=nn.Embedding(config.vocab_size, config.n_embd),
            wpe=nn.Embedding(config.block_size, config.n_embd),
            drop=nn.Dropout(config.dropout),
            h=nn.ModuleList(Block(config) for _ in range(config.n_layer)),
            ln_f=LayerNorm(config.n_embd, config.bias),
        ))
        self.lm_head = nn.Linear(config.n_embd, config.vocab_size, bias=False)
        self.transformer.wte.weight = self.lm_head.weight        # weight tying
        for name, mod in self.named_modules():
            if isinstance(mod, nn.Linear):
                std = 0.02 / math.sqrt(2 * config.n_layer) if name.endswith("c_proj") else 0.02
                nn.init.normal_(mod.weight, mean=0.0, std=std)
                if mod.bias is not None:
                    nn.init.zeros_(mod.bias)
            elif isinstance(mod, nn.Embedding):
                nn.init.normal_(mod.weight, mean=0.0, std=0.02)

    def forward(self, idx):
        """Full-sequence logits [B,T,V] (what the reference's three parts
        compose to: gpt_model_parts.py:13-22, 31-34, 44-50)."""
        b, t = idx.shape
        assert t <= self.config.block_size
        pos = torch.arange(t, dtype=torch.long, device=idx.device)
        x = self.transformer.drop(self.transformer.wte(idx) + self.transformer.wpe(pos))
        for blk in self.transformer.h:
            x = blk(x)
        return self.lm_head(self.transformer.ln_f(x))
