"""config.json topology resolver (reference: node.py:222-277, schema config.json:2-18).

Same keys, same derivations, same error conditions -- raised as ``ConfigError``
instead of ``print(...); exit(1)``:

    nodes[].id / .address / .part_index, model_weights, num_parts, return_to_node_id

Differences, both deliberate (SURVEY 8b seam 2):
  * ``num_parts`` may be any value >= 1 (the reference rejects != 2 at node.py:246-248);
  * a node additionally maps to a GPU: ``part_index`` k runs on CUDA device
    ``nodes[k].get("device", k % visible_gpus)``.  ``address`` keeps its meaning for
    the front-door gRPC port.
"""
from __future__ import annotations

import json
from dataclasses import dataclass, field


class ConfigError(ValueError):
    pass


@dataclass
class NodeConfig:
    """What node.py keeps in module globals after start-up (node.py:17-24)."""
    node_id: str
    address: str
    port: int
    part_index: int
    num_parts: int
    is_last: bool
    next_address: str | None
    first_address: str | None
    model_weights: str
    device: int | None = None


@dataclass
class Topology:
    nodes: list[dict]
    model_weights: str
    num_parts: int
    return_to_node_id: str | None = None
    extra: dict = field(default_factory=dict)

    def chain(self) -> list[dict]:
        """Nodes ordered by part_index 0..num_parts-1 (next hop = part_index+1,
        node.py:262-264)."""
        out = []
        for k in range(self.num_parts):
            n = next((n for n in self.nodes if n.get("part_index") == k), None)
            if n is None:
                raise ConfigError(f"Could not find node config for next part index {k}")
            out.append(n)
        return out

    def resolve(self, node_id: str) -> NodeConfig:
        me = next((n for n in self.nodes if n.get("id") == node_id), None)        # node.py:234
        if not me:
            raise ConfigError(f"Node ID '{node_id}' not found in config file")
        address, part_index = me.get("address"), me.get("part_index")
        if None in (address, part_index, self.model_weights, self.num_parts):     # node.py:250
            raise ConfigError("Config file is missing required fields "
                              "(address, part_index, model_weights, num_parts)")
        try:
            port = int(str(address).split(":")[-1])                               # node.py:255
        except (ValueError, IndexError):
            raise ConfigError(f"Invalid format for MY_ADDRESS '{address}'. Expected IP:Port.")
        is_last = part_index == self.num_parts - 1                                # node.py:260
        next_address = first_address = None
        if not is_last:
            nxt = next((n for n in self.nodes if n.get("part_index") == part_index + 1), None)
            if not nxt:
                raise ConfigError(f"Could not find node config for next part index {part_index + 1}")
            next_address = nxt.get("address")
            if not next_address:
                raise ConfigError(f"Next node (index {part_index + 1}) is missing address in config")
        elif self.return_to_node_id:                                               # node.py:274-277
            ret = next((n for n in self.nodes if n.get("id") == self.return_to_node_id), None)
            if ret:
                first_address = ret.get("address")
        return NodeConfig(node_id, address, port, part_index, self.num_parts, is_last,
                          next_address, first_address, self.model_weights, me.get("device"))


def parse_config(cfg: dict) -> Topology:
    num_parts = cfg.get("num_parts")
    if num_parts is not None and (not isinstance(num_parts, int) or num_parts < 1):
        raise ConfigError(f"Configuration file specifies num_parts={num_parts}, expected an integer >= 1")
    known = {"nodes", "model_weights", "num_parts", "return_to_node_id"}
    return Topology(list(cfg.get("nodes", [])), cfg.get("model_weights"), num_parts,
                    cfg.get("return_to_node_id"), {k: v for k, v in cfg.items() if k not in known})


def load_config(path: str) -> Topology:
    try:
        with open(path, "r") as f:                                                 # node.py:223
            cfg = json.load(f)
    except FileNotFoundError:
        raise ConfigError(f"Config file not found at '{path}'")
    except json.JSONDecodeError as e:
        raise ConfigError(f"Invalid JSON in config file '{path}': {e}")
    return parse_config(cfg)
