"""Functional CPU restatement of the reference's GPT pipeline parts.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).
Stage structure: PINNED against partitions/gpt_model_parts.py (golden vectors
made by running those very classes, oracle/make_golden.py).
Block arithmetic: PARITY UNPINNED -- see oracle/gpt_model.py for why and for
the pinned choice of semantics.

All functions take the FULL-model state dict with nanoGPT key names
(``transformer.wte.weight``, ``transformer.h.{i}.attn.c_attn.weight`` ...,
``lm_head.weight``), i.e. the ".pth" the reference's loader would read
(node.py:296).  Layer ranges are inclusive, like the reference constructors
(gpt_model_parts.py:7,12).
"""
from __future__ import annotations

import hashlib
import math

import torch
import torch.nn.functional as F

from .gpt_model import GPT, GPTConfig


def init_state_dict(cfg: GPTConfig, seed: int = 1234) -> dict[str, torch.Tensor]:
    torch.manual_seed(seed)
    sd = GPT(cfg).state_dict()
    return {k: v.detach().clone() for k, v in sd.items()}


def state_dict_digest(sd) -> str:
    h = hashlib.sha256()
    for k in sorted(sd):
        h.update(k.encode())
        h.update(sd[k].detach().cpu().contiguous().numpy().tobytes())
    return h.hexdigest()


def make_tokens(batch: int, seq: int, vocab: int, seed: int = 0) -> torch.Tensor:
    g = torch.Generator().manual_seed(seed)
    return torch.randint(0, vocab, (batch, seq), generator=g, dtype=torch.int64)


def _block(x, sd, i, n_head):
    """One transformer block (third-party ``model.Block``; semantics pinned in
    oracle/gpt_model.py)."""
    p = f"transformer.h.{i}."
    b, t, c = x.shape
    hd = c // n_head
    h = F.layer_norm(x, (c,), sd[p + "ln_1.weight"], sd.get(p + "ln_1.bias"), 1e-5)
    qkv = F.linear(h, sd[p + "attn.c_attn.weight"], sd.get(p + "attn.c_attn.bias"))
    q, k, v = (z.view(b, t, n_head, hd).transpose(1, 2) for z in qkv.split(c, dim=2))
    att = (q @ k.transpose(-2, -1)) / math.sqrt(hd)
    causal = torch.ones(t, t, dtype=torch.bool).tril()
    att = att.masked_fill(~causal, float("-inf")).softmax(-1)
    y = (att @ v).transpose(1, 2).reshape(b, t, c)
    x = x + F.linear(y, sd[p + "attn.c_proj.weight"], sd.get(p + "attn.c_proj.bias"))
    h = F.layer_norm(x, (c,), sd[p + "ln_2.weight"], sd.get(p + "ln_2.bias"), 1e-5)
    h = F.gelu(F.linear(h, sd[p + "mlp.c_fc.weight"], sd.get(p + "mlp.c_fc.bias")))
    return x + F.linear(h, sd[p + "mlp.c_proj.weight"], sd.get(p + "mlp.c_proj.bias"))


@torch.no_grad()
def part_first(idx, sd, start, end, n_head, block_size):
    """ModelPart0.forward, gpt_model_parts.py:13-22: wte[idx] + wpe[arange(T)],
    then blocks start..end inclusive.  idx [B,T] int64 -> [B,T,C]."""
    b, t = idx.shape
    assert t <= block_size, "Cannot forward, model block size is exhausted."   # :15
    x = sd["transformer.wte.weight"][idx] + sd["transformer.wpe.weight"][torch.arange(t)]
    for i in range(start, end + 1):
        x = _block(x, sd, i, n_head)
    return x


@torch.no_grad()
def part_middle(x, sd, start, end, n_head):
    """ModelPartIntermediate.forward, gpt_model_parts.py:31-34."""
    for i in range(start, end + 1):
        x = _block(x, sd, i, n_head)
    return x


@torch.no_grad()
def part_final(x, sd, start, end, n_head):
    """ModelPartFinal_GPT.forward, gpt_model_parts.py:44-50: blocks, ln_f,
    lm_head on ALL T positions -> logits [B,T,V]."""
    for i in range(start, end + 1):
        x = _block(x, sd, i, n_head)
    c = x.shape[-1]
    x = F.layer_norm(x, (c,), sd["transformer.ln_f.weight"], sd.get("transformer.ln_f.bias"), 1e-5)
    return F.linear(x, sd["lm_head.weight"])


def even_split(n_layer: int, n_parts: int) -> list[tuple[int, int]]:
    """Contiguous inclusive (start,end) ranges, earlier parts get the remainder."""
    base, rem = divmod(n_layer, n_parts)
    out, s = [], 0
    for p in range(n_parts):
        n = base + (1 if p < rem else 0)
        out.append((s, s + n - 1))
        s += n
    return out


@torch.no_grad()
def pipeline(idx, sd, cfg: GPTConfig, n_parts: int):
    """first -> middle* -> final, n_parts >= 1.  With one part, the single stage
    does embedding, all blocks and the head."""
    ranges = even_split(cfg.n_layer, n_parts)
    if n_parts == 1:
        x = part_first(idx, sd, 0, cfg.n_layer - 1, cfg.n_head, cfg.block_size)
        return part_final(x, sd, 1, 0, cfg.n_head)        # empty block range
    x = part_first(idx, sd, *ranges[0], cfg.n_head, cfg.block_size)
    for s, e in ranges[1:-1]:
        x = part_middle(x, sd, s, e, cfg.n_head)
    return part_final(x, sd, *ranges[-1], cfg.n_head)


@torch.no_grad()
def greedy_generate(idx, sd, cfg: GPTConfig, n_parts: int, max_new_tokens: int):
    """Greedy decoding through the stage chain, the way the public nanoGPT ``GPT.generate`` loops (crop the context
    to block_size, run the model, take the LAST position) with arg-max instead of sampling.  Also returns, per new
    token, the gap between the two largest logits so that tests can tell a real mismatch from a near-tie."""
    gaps = []
    for _ in range(max_new_tokens):
        ctx = idx if idx.shape[1] <= cfg.block_size else idx[:, -cfg.block_size:]
        logits = pipeline(ctx, sd, cfg, n_parts)[:, -1, :]
        top2 = logits.topk(2, -1).values
        gaps.append(top2[:, 0] - top2[:, 1])
        idx = torch.cat([idx, logits.argmax(-1, keepdim=True)], dim=1)
    return idx, torch.stack(gaps, dim=1)

