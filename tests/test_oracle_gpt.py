"""GPT oracle: stage structure against the reference's part classes (golden), block
arithmetic against transformers' GPT-2 (the offline pin for the absent ``model``)."""
import os

import numpy as np
import pytest
import torch

from oracle import gpt_model, gpt_oracle


def _golden(golden_dir):
    g = np.load(os.path.join(golden_dir, "gpt_tiny.npz"))
    bs, v, nl, nh, c = (int(z) for z in g["cfg"])
    cfg = gpt_model.GPTConfig(block_size=bs, vocab_size=v, n_layer=nl, n_head=nh, n_embd=c)
    sd = {k[2:]: torch.from_numpy(g[k]) for k in g.files if k.startswith("w:")}
    return g, cfg, sd


def test_parts_match_reference_part_classes(golden_dir):
    torch.set_num_threads(1)
    g, cfg, sd = _golden(golden_dir)
    idx = torch.from_numpy(g["idx"])
    a = gpt_oracle.part_first(idx, sd, 0, 0, cfg.n_head, cfg.block_size)
    np.testing.assert_allclose(a.numpy(), g["stage0"], rtol=0, atol=2e-6)
    b = gpt_oracle.part_middle(torch.from_numpy(g["stage0"]), sd, 1, 1, cfg.n_head)
    np.testing.assert_allclose(b.numpy(), g["stage1"], rtol=0, atol=2e-6)
    c = gpt_oracle.part_final(torch.from_numpy(g["stage1"]), sd, 2, 2, cfg.n_head)
    np.testing.assert_allclose(c.numpy(), g["logits"], rtol=0, atol=5e-6)
    assert np.array_equal(c.argmax(-1).numpy(), g["logits"].argmax(-1))


@pytest.mark.parametrize("n_parts", [1, 2, 3])
def test_pipeline_equals_full_model(golden_dir, n_parts):
    g, cfg, sd = _golden(golden_dir)
    out = gpt_oracle.pipeline(torch.from_numpy(g["idx"]), sd, cfg, n_parts)
    np.testing.assert_allclose(out.numpy(), g["logits"], rtol=0, atol=1e-5)


def test_block_size_guard(golden_dir):
    _, cfg, sd = _golden(golden_dir)
    with pytest.raises(AssertionError):
        gpt_oracle.part_first(torch.zeros(1, cfg.block_size + 1, dtype=torch.int64), sd, 0, 0,
                              cfg.n_head, cfg.block_size)


def test_even_split():
    assert gpt_oracle.even_split(12, 4) == [(0, 2), (3, 5), (6, 8), (9, 11)]
    assert gpt_oracle.even_split(24, 8)[-1] == (21, 23)
    assert gpt_oracle.even_split(3, 2) == [(0, 1), (2, 2)]


def test_against_hf_gpt2(golden_dir):
    """Pin of the restated block arithmetic: HF GPT2LMHeadModel with the exact-erf
    GELU on the same weights (Conv1D stores [in,out]: transpose)."""
    tr = pytest.importorskip("transformers")
    g, cfg, sd = _golden(golden_dir)
    hf_cfg = tr.GPT2Config(vocab_size=cfg.vocab_size, n_positions=cfg.block_size, n_embd=cfg.n_embd,
                           n_layer=cfg.n_layer, n_head=cfg.n_head, activation_function="gelu",
                           resid_pdrop=0.0, embd_pdrop=0.0, attn_pdrop=0.0, layer_norm_epsilon=1e-5)
    hf = tr.GPT2LMHeadModel(hf_cfg).eval()
    new = {}
    for k, v in sd.items():
        if k == "lm_head.weight":
            new[k] = v
        elif k.endswith(("c_attn.weight", "c_proj.weight", "c_fc.weight")):
            new[k] = v.t().contiguous()
        else:
            new[k] = v
    missing, unexpected = hf.load_state_dict(new, strict=False)
    assert not unexpected
    assert all(("attn.bias" in m or "masked_bias" in m) for m in missing), missing
    with torch.no_grad():
        ref = hf(torch.from_numpy(g["idx"])).logits
    np.testing.assert_allclose(ref.numpy(), g["logits"], rtol=0, atol=2e-5)
    assert np.array_equal(ref.argmax(-1).numpy(), g["logits"].argmax(-1))
