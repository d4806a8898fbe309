"""Host-side behaviour of the stage operators (no GPU needed): the
load_state_dict(strict=False) contract of node.py:306 and GPT key mapping."""
import pytest
import torch


def test_cifar_parts_strict_false_reports_like_reference(engine):
    full = engine.NeuralNetwork()
    sd = full.state_dict()
    assert set(sd) == {"conv1.weight", "conv1.bias", "conv2.weight", "conv2.bias",
                       "fc1.weight", "fc1.bias", "fc2.weight", "fc2.bias"}
    p0 = engine.MODEL_PARTS_CLASSES[0](full)
    res = p0.load_state_dict(sd, strict=False)
    assert res.missing_keys == [] and sorted(res.unexpected_keys) == ["fc1.bias", "fc1.weight", "fc2.bias", "fc2.weight"]
    p1 = engine.MODEL_PARTS_CLASSES[1](full)
    res = p1.load_state_dict(sd, strict=False)
    assert res.missing_keys == [] and sorted(res.unexpected_keys) == ["conv1.bias", "conv1.weight", "conv2.bias", "conv2.weight"]
    with pytest.raises(RuntimeError, match="Unexpected key"):
        p1.load_state_dict(sd, strict=True)
    with pytest.raises(RuntimeError, match="size mismatch"):
        p1.load_state_dict({"fc2.bias": torch.zeros(11)}, strict=False)
    assert p1.load_state_dict({}, strict=False).missing_keys == list(p1.KEYS)
    assert p1.eval() is p1


def test_default_init_matches_reference_seeded_init(engine):
    from oracle import cifar_oracle
    torch.manual_seed(1234)
    sd = engine.NeuralNetwork().state_dict()
    ref = cifar_oracle.init_state_dict(1234)
    assert all(torch.equal(sd[k], ref[k]) for k in ref)


def test_gpt_parts_key_mapping(engine):
    cfg = engine.GPTConfig(block_size=64, vocab_size=256, n_layer=4, n_head=2, n_embd=128)
    g = engine.GPT(cfg, seed=3)
    sd = g.state_dict()
    assert "transformer.h.3.mlp.c_proj.bias" in sd and "lm_head.weight" in sd
    parts = engine.gpt_parts_for(g, 4)
    assert [type(p).__name__ for p in parts] == ["ModelPart0", "ModelPartIntermediate",
                                                 "ModelPartIntermediate", "ModelPartFinal_GPT"]
    assert [(p.start_layer, p.end_layer) for p in parts] == [(0, 0), (1, 1), (2, 2), (3, 3)]
    mid = parts[1]
    res = mid.load_state_dict(sd, strict=False)          # FULL-model names map by global index
    assert res.missing_keys == []
    assert "transformer.wte.weight" in res.unexpected_keys
    # names as the reference part would expose them (h.<local>.…)
    res = mid.load_state_dict({"h.0.ln_1.weight": torch.full((128,), 2.0)}, strict=False)
    assert res.unexpected_keys == []
    assert torch.equal(mid.state_dict()["transformer.h.1.ln_1.weight"], torch.full((128,), 2.0))
    # tied head: final stage accepts wte when no lm_head key is present
    fin = parts[-1]
    only_wte = {k: v for k, v in sd.items() if k != "lm_head.weight"}
    assert fin.load_state_dict(only_wte, strict=False).missing_keys == []
    with pytest.raises(IndexError):
        engine.ModelPartIntermediate(g, 2, 4)
    one = engine.gpt_parts_for(g, 1)
    assert [(p.start_layer, p.end_layer) for p in one] == [(0, 3), (4, 3)]


def test_balanced_ranges_offset_the_lm_head(engine):
    small = engine.GPTConfig(vocab_size=50257, **engine.GPT2_SMALL)
    medium = engine.GPTConfig(vocab_size=50257, **engine.GPT2_MEDIUM)
    r = engine.balanced_ranges(small, 4)
    assert r == [(0, 3), (4, 7), (8, 11), (12, 11)]          # head ~ 5 blocks: the final stage keeps none
    r = engine.balanced_ranges(medium, 8)
    assert r[0] == (0, 3) and r[-1] == (24, 23) and sum(e - s + 1 for s, e in r) == 24
    assert all(r[i][1] + 1 == r[i + 1][0] for i in range(7))      # contiguous, inclusive
    assert engine.balanced_ranges(small, 1) == [(0, 11)]
    tiny = engine.GPTConfig(block_size=64, vocab_size=256, n_layer=4, n_head=2, n_embd=128)
    r = engine.balanced_ranges(tiny, 2, seq_len=32)
    assert sum(e - s + 1 for s, e in r) == 4
    parts = engine.gpt_parts_for(engine.GPT(tiny, seed=0), 2, ranges=r)
    assert (parts[0].start_layer, parts[-1].end_layer) == (0, 3)


def test_hf_gpt2_checkpoint_import(engine):
    tr = pytest.importorskip("transformers")
    hf = tr.GPT2LMHeadModel(tr.GPT2Config(vocab_size=256, n_positions=64, n_embd=128, n_layer=2, n_head=2))
    sd = engine.from_hf_gpt2_state_dict(hf.state_dict())
    cfg = engine.GPTConfig(block_size=64, vocab_size=256, n_layer=2, n_head=2, n_embd=128)
    model = engine.GPT(cfg)
    missing, unexpected = model.load_state_dict(sd, strict=False)
    assert not missing and not unexpected
    assert sd["transformer.h.0.attn.c_attn.weight"].shape == (384, 128)          # [out, in] after the transpose
    assert torch.equal(sd["transformer.h.1.mlp.c_proj.weight"], hf.state_dict()["transformer.h.1.mlp.c_proj.weight"].t())
    assert torch.equal(sd["lm_head.weight"], sd["transformer.wte.weight"])


def test_input_image_fallback_and_preprocessing(engine, tmp_path):
    svc = engine.service
    x, real = svc.load_input_image(str(tmp_path / "missing.png"), seed=3)     # node.py:149-151 fallback
    assert not real and x.shape == (1, 3, 32, 32) and x.dtype == torch.float32
    from PIL import Image
    import numpy as np
    img = Image.fromarray((np.arange(64 * 48 * 3) % 256).astype(np.uint8).reshape(48, 64, 3), "RGB")
    p = tmp_path / "img.png"
    img.save(p)
    x, real = svc.load_input_image(str(p))
    assert real and x.shape == (1, 3, 32, 32) and float(x.min()) >= -1.0 and float(x.max()) <= 1.0
    pred, probs, real = svc.initiate_inference(svc.NodeServiceImpl(["n"], lambda a: np.eye(10, dtype=np.float32)[[7] * a.shape[0]]), None, seed=1)
    assert pred == 7 and probs.shape == (1, 10) and not real
