"""Stage operators: the engine's drop-ins for the reference's part classes.

The reference runtime does exactly this with a part (node.py:300-317, :52-53):

    cls  = MODEL_PARTS_CLASSES.get(part_index)
    part = cls(full_model).to(DEVICE)
    missing, unexpected = part.load_state_dict(full_state_dict, strict=False)
    part.eval()
    with torch.no_grad(): y = part(x)

Every class below supports that sequence with the same names, constructor
arguments (GPT parts take ``(original_model, start_layer, end_layer)``, inclusive,
partitions/gpt_model_parts.py:7,27,37) and output shape/dtype/layout, but the
forward pass runs hand-written sm_100a kernels through the C ABI
(include/b200pipe.h).  PyTorch is only the carrier for device buffers and streams.
There is no CPU fallback: a stage that is not on a CUDA device refuses to run.
"""
from __future__ import annotations

import ctypes as C
from collections import namedtuple
from typing import Mapping

import torch
import torch.nn as nn

from . import _lib

_IncompatibleKeys = namedtuple("IncompatibleKeys", ["missing_keys", "unexpected_keys"])


def _state_dict_of(model) -> dict[str, torch.Tensor]:
    if isinstance(model, Mapping):
        return dict(model)
    if hasattr(model, "state_dict"):
        return dict(model.state_dict())
    raise TypeError("original_model must be an nn.Module or a state dict")


class EngineStage(nn.Module):
    """Common machinery: weight table on the host until ``.to(cuda)``, then a
    ``b2p_stage`` on that GPU."""

    KIND: int = -1

    def __init__(self, math: str = "fp32"):
        super().__init__()
        if math not in _lib.MATH_NAMES:
            raise ValueError(f"math must be one of {sorted(_lib.MATH_NAMES)}")
        self.math = math
        self._handle: C.c_void_p | None = None
        self._device: torch.device = torch.device("cpu")
        self._host: dict[str, torch.Tensor] = {}     # full-model key -> fp32 host tensor
        self._finalized = False

    # ----- naming: full-model key <-> key as seen from inside the reference part ----
    def _full_keys(self) -> list[str]:
        raise NotImplementedError

    def _local_to_full(self, key: str) -> str | None:
        """Map a key as the reference *part* would name it to the full-model key."""
        return key if key in self._full_keys() else None

    def _desc(self, device_index: int) -> _lib.StageDesc:
        d = _lib.StageDesc()
        d.kind, d.math, d.device = self.KIND, _lib.MATH_NAMES[self.math], device_index
        return d

    # ----- nn.Module surface the reference runtime touches -------------------------------
    def to(self, device=None, *args, **kwargs):                      # node.py:305
        if device is None:
            return self
        dev = torch.device(device)
        if dev.type == "cpu":
            self._release()
            self._device = dev
            return self
        if dev.type != "cuda":
            raise RuntimeError(f"unsupported device {dev}: this engine runs on CUDA (sm_100a) only")
        if not torch.cuda.is_available():
            raise RuntimeError("no CUDA device is visible and this engine has no CPU fallback")
        index = dev.index if dev.index is not None else torch.cuda.current_device()
        if self._handle is not None and self._device.index == index:
            return self
        self._release()
        handle = C.c_void_p()
        desc = self._desc(index)
        _lib.check(_lib.lib().b2p_stage_create(C.byref(handle), C.byref(desc)))
        self._handle, self._device = handle, torch.device("cuda", index)
        for k, v in self._host.items():
            self._upload(k, v)
        return self

    def cuda(self, device=None):
        return self.to(torch.device("cuda", device) if isinstance(device, int) else (device or "cuda"))

    def cpu(self):
        return self.to("cpu")

    def load_state_dict(self, state_dict, strict: bool = True, assign: bool = False):   # node.py:306
        mine = self._full_keys()
        seen, unexpected = set(), []
        for key, value in state_dict.items():
            full = key if key in mine else self._local_to_full(key)
            if full is None:
                unexpected.append(key)
                continue
            t = value.detach().to("cpu", torch.float32).contiguous()
            want = self._shapes()[full]
            if tuple(t.shape) != tuple(want):
                raise RuntimeError(f"size mismatch for {key}: copying a param with shape "
                                   f"{tuple(t.shape)} from checkpoint, the shape in current model is {tuple(want)}.")
            self._host[full] = t
            seen.add(full)
            if self._handle is not None:
                self._upload(full, t)
        missing = [k for k in mine if k not in seen and self._required(k)]
        if strict and (missing or unexpected):
            raise RuntimeError(f"Error(s) in loading state_dict for {type(self).__name__}: "
                               f"Missing key(s): {missing}. Unexpected key(s): {unexpected}.")
        self._finalized = False
        return _IncompatibleKeys(missing, unexpected)

    def state_dict(self, *args, **kwargs):
        return {k: v.clone() for k, v in self._host.items()}

    def eval(self):                                                    # node.py:317
        if self._handle is not None:
            self._finalize()
        return super().eval()

    # ----- internals -----------------------------------------------------------------------
    def _shapes(self) -> dict[str, tuple]:
        raise NotImplementedError

    def _required(self, key: str) -> bool:
        return True

    def _upload(self, key: str, t: torch.Tensor):
        shape = (C.c_int64 * t.dim())(*t.shape)
        rc = _lib.check(_lib.lib().b2p_stage_load_weight(
            self._handle, key.encode(), C.c_void_p(t.data_ptr()), _lib.DT_F32, shape, t.dim()))
        if rc != 1:
            raise RuntimeError(f"engine did not accept weight {key}")
        self._finalized = False

    def _finalize(self):
        if not self._finalized:
            _lib.check(_lib.lib().b2p_stage_finalize(self._handle))
            self._finalized = True

    def _release(self):
        if self._handle is not None:
            _lib.lib().b2p_stage_destroy(self._handle)
            self._handle, self._finalized = None, False

    def __del__(self):
        try:
            self._release()
        except Exception:
            pass

    def _require_device(self):
        if self._handle is None:
            raise RuntimeError(f"{type(self).__name__} is not on a CUDA device (call .to('cuda')); "
                               "this engine has no CPU fallback")

    def _prepare_input(self, x: torch.Tensor, dtype: torch.dtype) -> torch.Tensor:
        if not isinstance(x, torch.Tensor):
            raise TypeError("stage input must be a torch.Tensor")
        if x.dtype != dtype:
            raise RuntimeError(f"{type(self).__name__} expects a {dtype} input, got {x.dtype}")
        return x.to(self._device, non_blocking=True).contiguous()

    def _run(self, x: torch.Tensor, out: torch.Tensor, batch: int, seq: int):
        self._finalize()
        stream = torch.cuda.current_stream(self._device).cuda_stream
        _lib.check(_lib.lib().b2p_stage_forward(
            self._handle, C.c_void_p(x.data_ptr()), C.c_void_p(out.data_ptr()), batch, seq, C.c_void_p(stream)))
        return out

    def reserve(self, batch: int, seq: int = 1):
        """Size the batch-dependent scratch for forwards of up to ``batch`` rows (b2p_stage_reserve): afterwards no
        forward of that size allocates, so the schedule can be captured in a CUDA graph.  Start-up call."""
        self._require_device()
        self._finalize()
        _lib.check(_lib.lib().b2p_stage_reserve(self._handle, batch, seq))

    def set_pdl(self, mask: int):
        """Programmatic-dependent-launch mask of THIS stage's kernels (b2p_stage_set_pdl); -1 = process default."""
        self._require_device()
        return _lib.lib().b2p_stage_set_pdl(self._handle, mask)

    def set_aux_stream(self, stream: int | None):
        """CIFAR part 1: run the tail kernel on a second stream (b2p_stage_set_aux_stream); ``join`` before reading."""
        self._require_device()
        _lib.check(_lib.lib().b2p_stage_set_aux_stream(self._handle, C.c_void_p(stream) if stream else None))

    def join(self, stream: int):
        _lib.check(_lib.lib().b2p_stage_join(self._handle, C.c_void_p(stream)))

    # raw-pointer entry for the pipeline scheduler (no tensor allocation, no checks)
    def forward_ptr(self, in_ptr: int, out_ptr: int, batch: int, seq: int, stream: int):
        _lib.check(_lib.lib().b2p_stage_forward(
            self._handle, C.c_void_p(in_ptr), C.c_void_p(out_ptr), batch, seq, C.c_void_p(stream)))

    def forward_ring(self, in_ptr: int | None, out_ptr: int | None, batch: int, seq_len: int, ring_in, ring_out,
                     seq: int, stream: int, rows_per_msg: int | None = None):
        """forward_ptr with the handoff handshake inside the stage's kernels (b2p_stage_forward_ring[_span]):
        ring_in (consumer side) replaces in_ptr, ring_out (producer side) replaces out_ptr.  rows_per_msg splits
        the batch into consecutive messages seq, seq+1, ... handled by one call (one launch where fused)."""
        _lib.check(_lib.lib().b2p_stage_forward_ring_span(
            self._handle, C.c_void_p(in_ptr), C.c_void_p(out_ptr), batch, seq_len,
            ring_in._h if ring_in is not None else None, ring_out._h if ring_out is not None else None,
            seq, batch if rows_per_msg is None else rows_per_msg, C.c_void_p(stream)))


# =============================================================================================
# CIFAR-10 parts (cifar_model_parts.py)
# =============================================================================================
_CIFAR_SHAPES = {
    "conv1.weight": (32, 3, 3, 3), "conv1.bias": (32,),
    "conv2.weight": (64, 32, 3, 3), "conv2.bias": (64,),
    "fc1.weight": (512, 4096), "fc1.bias": (512,),
    "fc2.weight": (10, 512), "fc2.bias": (10,),
}


class _CifarPart(EngineStage):
    KEYS: tuple = ()
    WHOLE_BATCH_OK = True       # no per-stage workspace bound: a pipeline on one GPU may run the batch in one call

    def check(self):
        """Synchronise and raise on deferred device-side range errors (b2p_stage_check): a part in a tensor-core math
        mode met a value (input, conv1 output, fc1 activation or weight) beyond the fp16 range its hi/lo split can carry."""
        self._require_device()
        stream = torch.cuda.current_stream(self._device).cuda_stream
        _lib.check(_lib.lib().b2p_stage_check(self._handle, C.c_void_p(stream)))

    def __init__(self, original_model=None, math: str = "fp32"):
        super().__init__(math)
        if original_model is not None:          # parts start out sharing the full model's weights
            sd = _state_dict_of(original_model)                      # cifar_model_parts.py:32-35
            self.load_state_dict({k: sd[k] for k in self.KEYS if k in sd}, strict=False)

    def _full_keys(self):
        return list(self.KEYS)

    def _shapes(self):
        return _CIFAR_SHAPES


class ModelPart0_2Node(_CifarPart):
    """conv1+ReLU+pool -> conv2+ReLU+pool -> flatten; [B,3,32,32] -> [B,4096]
    (cifar_model_parts.py:29-42)."""
    KIND = _lib.CIFAR_PART0
    KEYS = ("conv1.weight", "conv1.bias", "conv2.weight", "conv2.bias")

    def forward(self, x):
        self._require_device()
        x = self._prepare_input(x, torch.float32)
        if x.dim() != 4 or tuple(x.shape[1:]) != (3, 32, 32):
            raise RuntimeError(f"expected input [B,3,32,32], got {list(x.shape)}")
        out = torch.empty((x.shape[0], 4096), dtype=torch.float32, device=self._device)
        return self._run(x, out, x.shape[0], 0) if x.shape[0] else out


class ModelPart1_2Node(_CifarPart):
    """fc1+ReLU -> fc2 -> softmax(dim=1); [B,4096] -> [B,10] probabilities
    (cifar_model_parts.py:45-58)."""
    KIND = _lib.CIFAR_PART1
    KEYS = ("fc1.weight", "fc1.bias", "fc2.weight", "fc2.bias")

    def forward(self, x):
        self._require_device()
        x = self._prepare_input(x, torch.float32)
        if x.dim() != 2 or x.shape[1] != 4096:
            raise RuntimeError(f"mat1 and mat2 shapes cannot be multiplied ({list(x.shape)} and 4096x512)")
        out = torch.empty((x.shape[0], 10), dtype=torch.float32, device=self._device)
        return self._run(x, out, x.shape[0], 0) if x.shape[0] else out


class NeuralNetwork(nn.Module):
    """The un-split CIFAR model (cifar_model_parts.py:7-26).  It is the parameter
    schema of the ``.pth`` (node.py:298 builds one just for that) -- the nn layers
    here are parameter containers with PyTorch's default init, never executed.
    ``forward`` runs the two engine parts back to back on the model's GPU."""

    def __init__(self, math: str = "fp32"):
        super().__init__()
        self.conv1 = nn.Conv2d(3, 32, kernel_size=3, stride=1, padding=1)
        self.pool = nn.MaxPool2d(kernel_size=2, stride=2, padding=0)
        self.conv2 = nn.Conv2d(32, 64, kernel_size=3, stride=1, padding=1)
        self.fc1 = nn.Linear(64 * 8 * 8, 512)
        self.fc2 = nn.Linear(512, 10)
        self.relu = nn.ReLU()
        self.softmax = nn.Softmax(dim=1)
        self._math = math
        self._parts = None

    def load_state_dict(self, *a, **k):
        self._parts = None
        return super().load_state_dict(*a, **k)

    def forward(self, x):
        dev = self.conv1.weight.device
        if dev.type != "cuda":
            raise RuntimeError("NeuralNetwork is not on a CUDA device; this engine has no CPU fallback")
        if self._parts is None:
            self._parts = (ModelPart0_2Node(self, self._math).to(dev).eval(),
                           ModelPart1_2Node(self, self._math).to(dev).eval())
        return self._parts[1](self._parts[0](x))


# =============================================================================================
# GPT parts (partitions/gpt_model_parts.py)
# =============================================================================================
_BLOCK_PARAMS = {
    "ln_1.weight": lambda c: (c,), "ln_1.bias": lambda c: (c,),
    "attn.c_attn.weight": lambda c: (3 * c, c), "attn.c_attn.bias": lambda c: (3 * c,),
    "attn.c_proj.weight": lambda c: (c, c), "attn.c_proj.bias": lambda c: (c,),
    "ln_2.weight": lambda c: (c,), "ln_2.bias": lambda c: (c,),
    "mlp.c_fc.weight": lambda c: (4 * c, c), "mlp.c_fc.bias": lambda c: (4 * c,),
    "mlp.c_proj.weight": lambda c: (c, 4 * c), "mlp.c_proj.bias": lambda c: (c,),
}


class _GptPart(EngineStage):
    HAS_EMBED = False
    HAS_HEAD = False

    def __init__(self, original_model, start_layer: int, end_layer: int, math: str = "fp32",
                 max_tokens: int | None = None):
        super().__init__(math)
        cfg = original_model.config if hasattr(original_model, "config") else original_model[0]
        self.config = cfg                                           # gpt_model_parts.py:9,29,39
        self.start_layer, self.end_layer = int(start_layer), int(end_layer)
        if self.start_layer < 0 or self.end_layer >= cfg.n_layer or self.end_layer < self.start_layer - 1:
            raise IndexError(f"layer range [{start_layer},{end_layer}] outside 0..{cfg.n_layer - 1}")
        self.max_tokens = int(max_tokens) if max_tokens else cfg.block_size
        c = cfg.n_embd
        shapes = {}
        if self.HAS_EMBED:
            shapes["transformer.wte.weight"] = (cfg.vocab_size, c)
            shapes["transformer.wpe.weight"] = (cfg.block_size, c)
        for i in range(self.start_layer, self.end_layer + 1):
            for name, fn in _BLOCK_PARAMS.items():
                shapes[f"transformer.h.{i}.{name}"] = fn(c)
        if self.HAS_HEAD:
            shapes["transformer.ln_f.weight"] = (c,)
            shapes["transformer.ln_f.bias"] = (c,)
            shapes["lm_head.weight"] = (cfg.vocab_size, c)
        self._shape_table = shapes
        src = original_model if hasattr(original_model, "state_dict") else original_model[1]
        sd = _state_dict_of(src)
        self.load_state_dict({k: sd[k] for k in shapes if k in sd}, strict=False)

    # ---- greedy decoding with a key/value cache (include/b200pipe.h: b2p_stage_kv_reset / b2p_stage_forward_kv) ----
    def kv_reset(self, batch: int):
        """Size this stage's KV cache for ``batch`` sequences; call before a prefill (``forward_kv(x, 0)``)."""
        self._require_device()
        self._finalize()
        _lib.check(_lib.lib().b2p_stage_kv_reset(self._handle, batch))
        self._kv_batch = int(batch)

    def forward_kv(self, x, pos0: int, want_ids: bool = False):
        """The ``x.shape[1]`` NEW positions pos0.. of every sequence: pos0 == 0 is the prefill, afterwards one position
        per call, attending to the cache.  Returns what ``forward`` returns for those positions (``want_ids`` on the
        final part: greedy ids [B,T] instead of logits)."""
        self._require_device()
        first = self.KIND == _lib.GPT_FIRST
        x = self._prepare_input(x, torch.int64 if first else torch.float32)
        b, t = x.shape[0], x.shape[1]
        if pos0 > 0 and b * t > self.max_tokens:
            raise RuntimeError("forward_kv: the workspace cannot grow after the prefill (the cache would be lost)")
        self._ensure_capacity(b * t)
        self._finalize()
        stream = torch.cuda.current_stream(self._device).cuda_stream
        if want_ids:
            out = torch.empty((b, t), dtype=torch.int64, device=self._device)
            args = (None, C.c_void_p(out.data_ptr()))
        else:
            last = self.config.vocab_size if getattr(self, "HAS_HEAD", False) else self.config.n_embd
            out = torch.empty((b, t, last), dtype=torch.float32, device=self._device)
            args = (C.c_void_p(out.data_ptr()), None)
        _lib.check(_lib.lib().b2p_stage_forward_kv(self._handle, C.c_void_p(x.data_ptr()), args[0], args[1], b, t, pos0,
                                                   C.c_void_p(stream)))
        return out

    def forward_kv_at(self, x, pos_t: torch.Tensor, out: torch.Tensor, want_ids: bool = False):
        """One new position per sequence at the position stored in ``pos_t`` (int32 scalar tensor on this device), read
        by the kernels when they run (b2p_stage_forward_kv_at): no launch carries a step-dependent argument, so a decode
        step can live in a CUDA graph.  ``x``: [B,1] ids or [B,1,C] hidden states; ``out``: preallocated [B,1] int64
        (``want_ids``, final part) or [B,1,C] / [B,1,V] fp32."""
        self._require_device()
        self._finalize()
        assert pos_t.dtype == torch.int32 and pos_t.device == self._device and x.is_contiguous() and out.is_contiguous()
        stream = torch.cuda.current_stream(self._device).cuda_stream
        args = (None, C.c_void_p(out.data_ptr())) if want_ids else (C.c_void_p(out.data_ptr()), None)
        _lib.check(_lib.lib().b2p_stage_forward_kv_at(self._handle, C.c_void_p(x.data_ptr()), args[0], args[1], x.shape[0],
                                                      C.c_void_p(pos_t.data_ptr()), C.c_void_p(stream)))
        return out

    def _full_keys(self):
        return list(self._shape_table)

    def _shapes(self):
        return self._shape_table

    def _required(self, key):
        return not key.endswith(".bias")          # GPTConfig(bias=False) checkpoints carry none

    def _local_to_full(self, key):
        # names as seen from inside the reference part: "h.<local j>.*", "wte.weight",
        # "wpe.weight", "ln_f.*", "lm_head.weight"  (gpt_model_parts.py:10-12,41-42)
        if key.startswith("h."):
            _, j, rest = key.split(".", 2)
            full = f"transformer.h.{self.start_layer + int(j)}.{rest}"
        elif key.split(".")[0] in ("wte", "wpe", "ln_f"):
            full = "transformer." + key
        else:
            full = key
        return full if full in self._shape_table else None

    def _desc(self, device_index):
        d = super()._desc(device_index)
        cfg = self.config
        d.start_layer, d.end_layer = self.start_layer, self.end_layer
        d.n_head, d.n_embd, d.vocab_size = cfg.n_head, cfg.n_embd, cfg.vocab_size
        d.block_size, d.max_tokens = cfg.block_size, self.max_tokens
        return d

    def _ensure_capacity(self, tokens: int):
        """Workspaces are sized at stage creation; grow by re-creating the stage.  The re-created stage has no KV
        cache, so a cache sized by an earlier ``kv_reset`` is sized again (its contents are gone: growing is only
        legal before the prefill, which is the only forward_kv call that can carry more tokens than any before)."""
        if tokens > self.max_tokens:
            self.max_tokens = tokens
            if self._handle is not None:
                dev = self._device
                self._release()
                self.to(dev)
                kvb = getattr(self, "_kv_batch", 0)
                if kvb:
                    self.kv_reset(kvb)

    def _check_hidden(self, x):
        c = self.config.n_embd
        if x.dim() != 3 or x.shape[2] != c:
            raise RuntimeError(f"expected hidden states [B,T,{c}], got {list(x.shape)}")

    def check(self):
        """Synchronise and raise on deferred device-side input errors."""
        stream = torch.cuda.current_stream(self._device).cuda_stream
        rc = _lib.lib().b2p_stage_check(self._handle, C.c_void_p(stream))
        if rc < 0:
            raise IndexError(_lib.lib().b2p_last_error().decode())


class ModelPart0(_GptPart):
    """Token+position embedding, then blocks start..end; idx [B,T] int64 -> [B,T,C]
    (gpt_model_parts.py:6-22)."""
    KIND = _lib.GPT_FIRST
    HAS_EMBED = True

    def forward(self, idx):
        self._require_device()
        idx = self._prepare_input(idx, torch.int64)
        if idx.dim() != 2:
            raise ValueError(f"expected idx [B,T], got {list(idx.shape)}")
        b, t = idx.shape
        assert t <= self.config.block_size, "Cannot forward, model block size is exhausted."   # :15
        self._ensure_capacity(b * t)
        out = torch.empty((b, t, self.config.n_embd), dtype=torch.float32, device=self._device)
        self._run(idx, out, b, t)
        self.check()
        return out


class ModelPartIntermediate(_GptPart):
    """Blocks only; [B,T,C] -> [B,T,C] (gpt_model_parts.py:26-34)."""
    KIND = _lib.GPT_MIDDLE

    def forward(self, x):
        self._require_device()
        x = self._prepare_input(x, torch.float32)
        self._check_hidden(x)
        b, t, _ = x.shape
        self._ensure_capacity(b * t)
        return self._run(x, torch.empty_like(x), b, t)


class ModelPartFinal_GPT(_GptPart):
    """Blocks -> ln_f -> lm_head on all positions; [B,T,C] -> logits [B,T,V]
    (gpt_model_parts.py:36-50)."""
    KIND = _lib.GPT_FINAL
    HAS_HEAD = True

    def _local_to_full(self, key):
        full = super()._local_to_full(key)
        return full

    def load_state_dict(self, state_dict, strict=True, assign=False):
        # tied head: accept the embedding matrix when no lm_head.weight is given
        if "lm_head.weight" not in state_dict and "transformer.wte.weight" in state_dict:
            state_dict = dict(state_dict)
            state_dict["lm_head.weight"] = state_dict.pop("transformer.wte.weight")
        return super().load_state_dict(state_dict, strict, assign)

    def forward(self, x):
        self._require_device()
        x = self._prepare_input(x, torch.float32)
        self._check_hidden(x)
        b, t, _ = x.shape
        self._ensure_capacity(b * t)
        out = torch.empty((b, t, self.config.vocab_size), dtype=torch.float32, device=self._device)
        return self._run(x, out, b, t)

    def greedy_ids(self, x):
        """Row arg-max of the logits, int64 [B,T], without materialising [B,T,V]."""
        self._require_device()
        x = self._prepare_input(x, torch.float32)
        self._check_hidden(x)
        b, t, _ = x.shape
        self._ensure_capacity(b * t)
        self._finalize()
        ids = torch.empty((b, t), dtype=torch.int64, device=self._device)
        stream = torch.cuda.current_stream(self._device).cuda_stream
        _lib.check(_lib.lib().b2p_stage_forward_argmax(
            self._handle, C.c_void_p(x.data_ptr()), C.c_void_p(ids.data_ptr()), b, t, C.c_void_p(stream)))
        return ids


@torch.no_grad()
def greedy_generate(parts, idx, max_new_tokens: int, kv_cache: bool = True, graph: bool = False):
    """Greedy decoding through a chain of GPT parts: the arg-max of the LAST position is appended, the
    head never materialises [B,T,V].  The reference only ever does one full-sequence forward
    (gpt_model_parts.py:44-50).  With ``kv_cache`` the prompt is run once (prefill) and every further token costs one
    position per stage (keys / values of earlier positions are kept per block, csrc/kv_decode.cuh); positions are
    absolute, so once the context would exceed block_size the loop falls back to what the public nanoGPT ``generate``
    does: re-run the chain on the context cropped to block_size.  ``graph`` (parts on one GPU): the decode step -- every
    stage, the feedback of the new token and the advance of the position, which lives in device memory -- is captured
    once in a CUDA graph and replayed per token (a step is ~110 launches for 12 blocks; the graph frees the host thread,
    the step time itself is set by the kernels' fixed latencies).  idx: int64 [B,T0]; returns [B,T0+new]."""
    block_size = parts[0].config.block_size
    idx = idx.to(parts[0]._device)

    def chain(x, pos0=None):
        for k, p in enumerate(parts):
            x = x.to(p._device)
            lastp = k == len(parts) - 1
            if pos0 is None:
                x = p.greedy_ids(x) if lastp else p(x.contiguous())
            else:
                x = p.forward_kv(x.contiguous(), pos0, want_ids=lastp)
        return x[:, -1:].to(idx.device)

    done = 0
    if kv_cache and idx.shape[1] < block_size and max_new_tokens > 0:
        for p in parts:
            p._ensure_capacity(idx.shape[0] * idx.shape[1])   # grow BEFORE the cache exists (re-creating a stage drops it)
            p.kv_reset(idx.shape[0])
        nxt = chain(idx, 0)                                   # prefill
        idx = torch.cat([idx, nxt], dim=1)
        done = 1
        n_graph = min(max_new_tokens - done, block_size - idx.shape[1] + 1)
        if graph and n_graph > 1 and len({p._device for p in parts}) == 1:
            idx = torch.cat([idx, _decode_steps_graph(parts, nxt, idx.shape[1] - 1, n_graph)], dim=1)
            done += n_graph
        while done < max_new_tokens and idx.shape[1] <= block_size:
            nxt = chain(idx[:, -1:], idx.shape[1] - 1)        # the token just appended sits at position len-1
            idx = torch.cat([idx, nxt], dim=1)
            done += 1
    for _ in range(max_new_tokens - done):
        x = idx if idx.shape[1] <= block_size else idx[:, -block_size:]
        idx = torch.cat([idx, chain(x)], dim=1)
    return idx


def _decode_steps_graph(parts, tok, pos: int, n: int):
    """n decode steps as n replays of ONE CUDA graph: tok [B,1] is the token at position ``pos`` (already appended to the
    context); returns the n following tokens [B,n]."""
    dev = parts[0]._device
    b = tok.shape[0]
    cfg = parts[0].config
    tok = tok.to(dev).contiguous().clone()
    pos_t = torch.tensor([pos], dtype=torch.int32, device=dev)
    hid = [torch.empty((b, 1, cfg.n_embd), dtype=torch.float32, device=dev) for _ in parts[:-1]]
    ids = torch.empty((b, 1), dtype=torch.int64, device=dev)
    gen = torch.zeros((b, n), dtype=torch.int64, device=dev)

    def stages():
        x = tok
        for k, p in enumerate(parts):
            lastp = k == len(parts) - 1
            x = p.forward_kv_at(x, pos_t, ids if lastp else hid[k], want_ids=lastp)

    def advance():
        gen.index_copy_(1, (pos_t - pos).to(torch.int64), ids)      # column = steps done so far, computed on the device
        tok.copy_(ids)
        pos_t.add_(1)

    stages()                                                  # warm-up outside the capture (idempotent: same token, same position)
    torch.cuda.synchronize(dev)
    side = torch.cuda.Stream(device=dev)
    side.wait_stream(torch.cuda.current_stream(dev))
    g = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        with torch.cuda.graph(g, stream=side):
            stages()
            advance()
        for _ in range(n):
            g.replay()
    torch.cuda.current_stream(dev).wait_stream(side)
    return gen


# node.py:29-32
MODEL_PARTS_CLASSES = {
    0: ModelPart0_2Node,
    1: ModelPart1_2Node,
}


def gpt_parts_for(model, num_parts: int, math: str = "fp32", max_tokens: int | None = None,
                  ranges: list[tuple[int, int]] | None = None):
    """Builds the reference's first/intermediate/final chain over contiguous layer
    ranges (even split unless ``ranges`` is given; uneven splits are API-legal,
    gpt_model_parts.py:7).  With one part the single GPU runs first+final back to
    back (first gets every block, final gets none)."""
    n = model.config.n_layer
    if ranges is None:
        if num_parts == 1:
            ranges = [(0, n - 1)]
        else:
            base, rem = divmod(n, num_parts)
            ranges, s = [], 0
            for p in range(num_parts):
                cnt = base + (1 if p < rem else 0)
                ranges.append((s, s + cnt - 1))
                s += cnt
    kw = dict(math=math, max_tokens=max_tokens)
    if num_parts == 1:
        return [ModelPart0(model, *ranges[0], **kw), ModelPartFinal_GPT(model, n, n - 1, **kw)]
    parts = [ModelPart0(model, *ranges[0], **kw)]
    parts += [ModelPartIntermediate(model, s, e, **kw) for s, e in ranges[1:-1]]
    parts.append(ModelPartFinal_GPT(model, *ranges[-1], **kw))
    return parts
