"""Host-side GPT parameter container (the engine's stand-in for the third-party
``model`` module the reference imports at partitions/gpt_model_parts.py:4).

Only what the reference's part classes touch is provided: ``GPTConfig`` with
``block_size`` (gpt_model_parts.py:15) and a ``GPT`` object exposing
``config``, ``transformer.wte/wpe/h/ln_f`` and ``lm_head`` with the nanoGPT
parameter names, i.e. the schema of the single full-model ``.pth``.  No CPU
arithmetic lives here -- ``forward`` chains the engine's CUDA parts.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import torch
import torch.nn as nn


@dataclass
class GPTConfig:
    block_size: int = 1024
    vocab_size: int = 50304
    n_layer: int = 12
    n_head: int = 12
    n_embd: int = 768
    dropout: float = 0.0
    bias: bool = True


GPT2_SMALL = dict(n_layer=12, n_head=12, n_embd=768)       # BASELINE.json config 3
GPT2_MEDIUM = dict(n_layer=24, n_head=16, n_embd=1024)     # BASELINE.json config 4


class _Params(nn.Module):
    """A bag of named parameters (never called)."""

    def __init__(self, **shapes):
        super().__init__()
        for name, shape in shapes.items():
            self.register_parameter(name, nn.Parameter(torch.empty(shape), requires_grad=False))


def _norm(c, bias):
    return _Params(weight=(c,), **({"bias": (c,)} if bias else {}))


def _linear(out_f, in_f, bias):
    return _Params(weight=(out_f, in_f), **({"bias": (out_f,)} if bias else {}))


class _Block(nn.Module):
    def __init__(self, cfg: GPTConfig):
        super().__init__()
        c = cfg.n_embd
        self.ln_1 = _norm(c, cfg.bias)
        self.attn = nn.ModuleDict(dict(c_attn=_linear(3 * c, c, cfg.bias), c_proj=_linear(c, c, cfg.bias)))
        self.ln_2 = _norm(c, cfg.bias)
        self.mlp = nn.ModuleDict(dict(c_fc=_linear(4 * c, c, cfg.bias), c_proj=_linear(c, 4 * c, cfg.bias)))


class GPT(nn.Module):
    def __init__(self, config: GPTConfig, seed: int | None = None):
        super().__init__()
        self.config = config
        c = config.n_embd
        self.transformer = nn.ModuleDict(dict(
            wte=_Params(weight=(config.vocab_size, c)),
            wpe=_Params(weight=(config.block_size, c)),
            h=nn.ModuleList(_Block(config) for _ in range(config.n_layer)),
            ln_f=_norm(c, config.bias),
        ))
        self.lm_head = _Params(weight=(config.vocab_size, c))
        self.transformer.wte.weight = self.lm_head.weight           # tied
        self.reset_parameters(seed)
        self._parts = None

    @torch.no_grad()
    def reset_parameters(self, seed: int | None = None):
        """N(0,0.02) matrices, N(0,0.02/sqrt(2L)) residual projections, unit LayerNorm
        gains, zero biases (the published nanoGPT scheme)."""
        g = torch.Generator().manual_seed(seed) if seed is not None else None
        for name, p in self.named_parameters():
            if name.endswith("bias"):
                p.zero_()
            elif ".ln_" in name or "ln_f" in name:
                p.fill_(1.0)
            else:
                std = 0.02 / math.sqrt(2 * self.config.n_layer) if name.endswith("c_proj.weight") else 0.02
                p.normal_(0.0, std, generator=g)

    def load_state_dict(self, *a, **k):
        self._parts = None
        return super().load_state_dict(*a, **k)

    def forward(self, idx, math: str = "fp32"):
        from .stages import gpt_parts_for
        dev = self.lm_head.weight.device
        if dev.type != "cuda":
            raise RuntimeError("GPT is not on a CUDA device; this engine has no CPU fallback")
        if self._parts is None or self._parts[0] != math:
            parts = [p.to(dev).eval() for p in gpt_parts_for(self, 1, math=math, max_tokens=idx.numel())]
            self._parts = (math, parts)
        x = idx
        for p in self._parts[1]:
            x = p(x)
        return x


def stage_cost_blocks(cfg: GPTConfig, seq_len: int) -> float:
    """LM-head cost expressed in transformer blocks (FLOP per token):
    2*C*V against 24*C^2 + 2*T*C (causal attention counted as executed)."""
    c = cfg.n_embd
    return (2.0 * c * cfg.vocab_size) / (24.0 * c * c + 2.0 * seq_len * c)


def balanced_ranges(cfg: GPTConfig, num_parts: int, seq_len: int = 1024) -> list[tuple[int, int]]:
    """Inclusive (start, end) layer ranges for ``num_parts`` stages that equalise FLOP per stage when
    the final stage also carries ``ln_f`` + ``lm_head`` (2.6x / 2.3x heavier than a 3-block stage for
    GPT-2 small / medium under an even split).  The reference constructors accept any
    inclusive range (gpt_model_parts.py:7,27,37); an empty final range is written (n_layer, n_layer-1)."""
    n = cfg.n_layer
    if num_parts <= 1:
        return [(0, n - 1)]
    head = stage_cost_blocks(cfg, seq_len)
    per_stage = (n + head) / num_parts
    last = min(n, max(0, round(per_stage - head)))
    rest = n - last
    base, rem = divmod(rest, num_parts - 1)
    out, s0 = [], 0
    for k in range(num_parts - 1):
        cnt = base + (1 if k < rem else 0)
        out.append((s0, s0 + cnt - 1))
        s0 += cnt
    out.append((s0, n - 1))
    return out


def from_hf_gpt2_state_dict(hf_sd: dict, n_layer: int | None = None) -> dict:
    """Hugging Face ``GPT2LMHeadModel`` checkpoint -> the nanoGPT-named full-model state dict this engine
    (and the reference's part classes) expect.  HF stores the four per-block projections as ``Conv1D``
    with weight [in, out]; ``nn.Linear`` wants [out, in], so those are transposed.  Attention-mask
    buffers (``attn.bias`` / ``attn.masked_bias``) are dropped; a missing ``lm_head.weight`` is tied to
    ``wte`` (GPT-2 ties them)."""
    out = {}
    for k, v in hf_sd.items():
        if k.endswith((".attn.bias", ".attn.masked_bias")):
            continue
        if k.endswith(("attn.c_attn.weight", "attn.c_proj.weight", "mlp.c_fc.weight", "mlp.c_proj.weight")):
            v = v.t()
        out[k] = v.contiguous()
    if "lm_head.weight" not in out and "transformer.wte.weight" in out:
        out["lm_head.weight"] = out["transformer.wte.weight"]
    return out
