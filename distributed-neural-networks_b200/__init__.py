"""b200pipe -- B200-native pipeline-partitioned inference engine.

Drop-in for the hot path of 123-code/Distributed-neural-networks: the stage
forward passes (cifar_model_parts.py, partitions/gpt_model_parts.py) and the
stage-to-stage activation handoff (node.py:35-105), behind the reference's own
surface: ``MODEL_PARTS_CLASSES``, the ``config.json`` topology, the single
full-model ``.pth`` and the ``node_service.proto`` client entry point.

The directory name carries a hyphen, so import it through the ``dnn_b200`` shim at
the repository root (``import dnn_b200 as engine``).
"""
from . import _lib
from ._lib import EngineError, launch_count
from .gpt_model import (GPT, GPT2_MEDIUM, GPT2_SMALL, GPTConfig, balanced_ranges, from_hf_gpt2_state_dict)
from .stages import (MODEL_PARTS_CLASSES, EngineStage, ModelPart0, ModelPart0_2Node,
                     ModelPart1_2Node, ModelPartFinal_GPT, ModelPartIntermediate,
                     NeuralNetwork, gpt_parts_for, greedy_generate)
from .topology import NodeConfig, Topology, load_config
from . import wire
from . import handoff, pipeline, service

__all__ = [
    "EngineError", "launch_count", "GPT", "GPTConfig", "GPT2_SMALL", "GPT2_MEDIUM", "balanced_ranges",
    "from_hf_gpt2_state_dict",
    "MODEL_PARTS_CLASSES", "EngineStage", "ModelPart0", "ModelPart0_2Node", "ModelPart1_2Node",
    "ModelPartFinal_GPT", "ModelPartIntermediate", "NeuralNetwork", "gpt_parts_for", "greedy_generate",
    "NodeConfig", "Topology", "load_config", "wire", "handoff", "pipeline", "service",
]
