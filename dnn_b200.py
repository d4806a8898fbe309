"""Import shim: the package directory is named ``distributed-neural-networks_b200``
(a hyphen is not a valid Python identifier), so ``import dnn_b200`` loads it via
importlib and aliases it."""
import importlib
import os
import sys

_root = os.path.dirname(os.path.abspath(__file__))
if _root not in sys.path:
    sys.path.insert(0, _root)
_pkg = importlib.import_module("distributed-neural-networks_b200")
sys.modules[__name__] = _pkg
