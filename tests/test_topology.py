"""config.json resolver vs the reference's behaviour (node.py:222-277)."""
import json

import pytest

REF_CONFIG = {   # verbatim structure of the reference's config.json:2-18
    "nodes": [{"id": "node1", "address": "192.168.1.101:50051", "part_index": 0},
              {"id": "node2", "address": "192.168.1.120:50051", "part_index": 1}],
    "model_weights": "./cifar10_model.pth", "num_parts": 2, "return_to_node_id": "node1"}


def test_reference_example_resolves(engine, tmp_path):
    p = tmp_path / "config.json"
    p.write_text(json.dumps(REF_CONFIG))
    topo = engine.load_config(str(p))
    n1, n2 = topo.resolve("node1"), topo.resolve("node2")
    assert (n1.port, n1.part_index, n1.is_last, n1.next_address) == (50051, 0, False, "192.168.1.120:50051")
    assert (n2.is_last, n2.next_address, n2.first_address) == (True, None, "192.168.1.101:50051")
    assert n1.first_address is None and n1.model_weights == "./cifar10_model.pth"
    assert [n["id"] for n in topo.chain()] == ["node1", "node2"]


def test_errors_match_reference_conditions(engine, tmp_path):
    from importlib import import_module
    topo_mod = import_module("distributed-neural-networks_b200.topology")
    E = topo_mod.ConfigError
    with pytest.raises(E, match="not found at"):
        engine.load_config(str(tmp_path / "nope.json"))
    bad = tmp_path / "bad.json"
    bad.write_text("{not json")
    with pytest.raises(E, match="Invalid JSON"):
        engine.load_config(str(bad))
    t = topo_mod.parse_config(REF_CONFIG)
    with pytest.raises(E, match="not found in config"):
        t.resolve("node9")
    cfg = json.loads(json.dumps(REF_CONFIG)); del cfg["nodes"][0]["address"]
    with pytest.raises(E, match="missing required fields"):
        topo_mod.parse_config(cfg).resolve("node1")
    cfg = json.loads(json.dumps(REF_CONFIG)); cfg["nodes"][0]["address"] = "host:abc"
    with pytest.raises(E, match="Invalid format"):
        topo_mod.parse_config(cfg).resolve("node1")
    cfg = json.loads(json.dumps(REF_CONFIG)); cfg["nodes"].pop(1)
    with pytest.raises(E, match="next part index 1"):
        topo_mod.parse_config(cfg).resolve("node1")
    cfg = json.loads(json.dumps(REF_CONFIG)); cfg["num_parts"] = 0
    with pytest.raises(E, match="num_parts"):
        topo_mod.parse_config(cfg)


@pytest.mark.parametrize("n", [1, 4, 8])
def test_num_parts_other_than_two_is_accepted(engine, n):
    """Documented divergence: the reference exits unless num_parts == 2 (node.py:246)."""
    from importlib import import_module
    topo_mod = import_module("distributed-neural-networks_b200.topology")
    cfg = {"nodes": [{"id": f"n{k}", "address": f"127.0.0.1:{50051 + k}", "part_index": k} for k in range(n)],
           "model_weights": "w.pth", "num_parts": n, "return_to_node_id": "n0"}
    t = topo_mod.parse_config(cfg)
    chain = t.chain()
    assert len(chain) == n
    last = t.resolve(f"n{n - 1}")
    assert last.is_last and last.first_address == "127.0.0.1:50051"
    if n > 1:
        assert t.resolve("n0").next_address == "127.0.0.1:50052"
