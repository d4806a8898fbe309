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
