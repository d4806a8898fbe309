// stage.cuh -- the stage object behind the C ABI: weight table, packed weights,
// workspaces.  One b2p_stage corresponds to one instance of a reference part class
// (MODEL_PARTS_CLASSES value, node.py:29-32 / partitions/gpt_model_parts.py).
#pragma once
#include <string>
#include <vector>
#include "common.cuh"

struct b2p_weight {
  std::string name;                 // key in the FULL-model state dict
  std::vector<int64_t> shape;
  bool required = true;
  bool loaded = false;
  float* dev = nullptr;             // fp32 copy in the original layout
  void* packed = nullptr;           // math-mode copy: bf16 (BF16) or tf32 hi plane (TF32X3)
  float* packed_lo = nullptr;       // TF32X3: lo plane
  void* packed16 = nullptr;         // CIFAR fc1.weight, TF32X3: fp16 hi plane | fp16 lo plane (x 2^11), back to back (fc_stage_tc.cuh)
  size_t elems() const {
    size_t n = 1;
    for (auto d : shape) n *= (size_t)d;
    return n;
  }
};

struct b2p_gpt_block {
  int ln1_w, ln1_b, attn_w, attn_b, proj_w, proj_b, ln2_w, ln2_b, fc_w, fc_b, fc2_w, fc2_b;
};

struct b2p_stage {
  b2p_stage_desc d{};
  std::vector<b2p_weight> w;
  bool finalized = false;
  int pdl_mask = -1;                // b2p_stage_set_pdl: -1 = the process-wide default (b2p_set_pdl)

  // ---- CIFAR part 0 ----
  float* w1t = nullptr;             // [27][32]
  float w1_host[27 * 32 + 32] = {};  // host copy of w1t and conv1.bias: kernel-parameter operand of part0_tc
  float* w2t = nullptr;             // [288][64]
  unsigned char* w2img = nullptr;   // tensor-core modes: smem byte image of conv2's [W_hi ; W_lo] tiles, one per tap (split fp16)
  unsigned char* a1img = nullptr;   // tensor-core modes: conv1 as pool-window GEMM, A tiles hi | lo (cifar_tc2.cuh)
  unsigned char* w2tap8 = nullptr;  // ... and its tap 8 as a shared-memory tile image
  float in_limit = 65504.f;         // |input| from which on a conv1 output may leave the fp16 range (range guard)
  uint32_t* w2tm = nullptr;         // tensor-core modes: tensor-memory image of conv2's [W_hi ; W_lo], taps 0..7
  long long* trace = nullptr;       // optional kernel timeline buffer (b2p_stage_trace)
  // ---- CIFAR part 1 ----
  float* fc_ws = nullptr;           // split-K partials [splits][M][512]
  size_t fc_ws_bytes = 0;
  // optional second stream for the tail kernel (b2p_stage_set_aux_stream): fc1 of step s+1 then overlaps the tail of step s
  cudaStream_t aux = nullptr;
  float* fc_ws2 = nullptr;          // second partial-sum buffer: steps alternate, so fc1(s+1) never writes what tail(s) reads
  size_t fc_ws2_bytes = 0;
  cudaEvent_t ev_fc1 = nullptr, ev_tail[2] = {nullptr, nullptr};
  bool ev_tail_recorded[2] = {false, false};
  uint64_t fc_step = 0;

  // ---- GPT ----
  std::vector<b2p_gpt_block> blocks;
  int wte = -1, wpe = -1, lnf_w = -1, lnf_b = -1, head = -1;
  float *ws_ln = nullptr, *ws_qkv = nullptr, *ws_att = nullptr, *ws_h = nullptr;
  float *ws_x[2] = {nullptr, nullptr};
  float* ws_split = nullptr;        // TF32X3: hi | lo planes of the current GEMM's A operand
  size_t ws_split_elems = 0;        // elements per plane
  float* amax_val = nullptr;        // forward_argmax: LM-head epilogue partials [M][slots]
  int* amax_idx = nullptr;
  size_t amax_elems = 0;
  float* ws_logits = nullptr;       // forward_argmax chunk buffer (fp32 CUDA-core mode)
  size_t ws_logits_rows = 0;
  int* bad_flag = nullptr;          // device flag: out-of-range token id seen
  float* kv_k = nullptr;            // KV cache (kv_decode.cuh): [block][B][H][block_size][64] keys ...
  float* kv_v = nullptr;            // ... and values, allocated by b2p_stage_kv_reset
  int64_t kv_batch = 0;             // sequences the cache is sized for

  int find(const std::string& n) const {
    for (size_t i = 0; i < w.size(); ++i)
      if (w[i].name == n) return (int)i;
    return -1;
  }
  int add(const std::string& n, std::vector<int64_t> shape, bool required = true) {
    b2p_weight x;
    x.name = n;
    x.shape = std::move(shape);
    x.required = required;
    w.push_back(std::move(x));
    return (int)w.size() - 1;
  }
  float* dev(int i) const { return i >= 0 && w[i].loaded ? w[i].dev : nullptr; }
};
