// b200pipe.cu -- C ABI of libb200pipe.so (declared in include/b200pipe.h).
//
// Replaces, behind the reference's stage-operator seam, what node.py does with a
// part object: construct (node.py:300-305), load_state_dict(strict=False) (:306),
// eval (:317) and the per-request forward (:52-53, :159-160).
#include <algorithm>
#include <cstring>
#include <new>

#include "attn_tc.cuh"
#include "cifar_simt.cuh"
#include "cifar_tc.cuh"
#include "cifar_tc2.cuh"
#include "common.cuh"
#include "gpt_simt.cuh"
#include "stage.cuh"
#include "tc_gemm2.cuh"
#include "fc_stage_tc.cuh"
#include "kv_decode.cuh"

using namespace b2p;

namespace {

__global__ void transpose2d_f32(const float* __restrict__ in, float* __restrict__ out, int R, int C) {
  // out[c][r] = in[r][c]; tiny one-off weight repack, not on the hot path
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < R * C) {
    int r = i / C, c = i % C;
    out[(size_t)c * R + r] = in[i];
  }
}

template <typename T>
int dev_alloc(T** p, size_t n_elems) {
  B2P_CUDA(cudaMalloc((void**)p, std::max<size_t>(n_elems, 1) * sizeof(T)));
  return 0;
}

// Scratch that depends on the batch (split-K partials, hi/lo planes, arg-max partials) is sized by
// b2p_stage_finalize (CIFAR part 1: 1024 rows; GPT: max_tokens) and b2p_stage_reserve.  A forward that needs more than
// that grows it on the spot -- cudaFree + cudaMalloc, which synchronise the device -- and refuses to do so while its
// stream is being captured (the allocation would invalidate the capture).
template <typename T>
int ensure_scratch(T** p, size_t* have, size_t need, size_t elem_bytes, cudaStream_t s, bool check_capture, const char* what) {
  if (need <= *have) return 0;
  if (check_capture) {
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(s, &cs) == cudaSuccess && cs != cudaStreamCaptureStatusNone)
      return fail(B2P_ESTATE, "forward needs a larger %s buffer than reserved and its stream is being captured: call "
                              "b2p_stage_reserve for the largest batch before starting the capture", what);
    cudaGetLastError();
  }
  if (*p) B2P_CUDA(cudaFree(*p));
  *p = nullptr;
  *have = 0;
  B2P_CUDA(cudaMalloc((void**)p, need * elem_bytes));
  *have = need;
  return 0;
}

bool is_cifar(int kind) { return kind == B2P_CIFAR_PART0 || kind == B2P_CIFAR_PART1; }
bool is_gpt(int kind) { return kind == B2P_GPT_FIRST || kind == B2P_GPT_MIDDLE || kind == B2P_GPT_FINAL; }

int declare_weights(b2p_stage* st) {
  const b2p_stage_desc& d = st->d;
  switch (d.kind) {
    case B2P_CIFAR_PART0:                                   // cifar_model_parts.py:32-35
      st->add("conv1.weight", {32, 3, 3, 3});
      st->add("conv1.bias", {32});
      st->add("conv2.weight", {64, 32, 3, 3});
      st->add("conv2.bias", {64});
      return 0;
    case B2P_CIFAR_PART1:                                   // cifar_model_parts.py:48-51
      st->add("fc1.weight", {512, 4096});
      st->add("fc1.bias", {512});
      st->add("fc2.weight", {10, 512});
      st->add("fc2.bias", {10});
      return 0;
    default:
      break;
  }
  const int64_t C = d.n_embd;
  if (d.kind == B2P_GPT_FIRST) {                            // gpt_model_parts.py:10-11
    st->wte = st->add("transformer.wte.weight", {d.vocab_size, C});
    st->wpe = st->add("transformer.wpe.weight", {d.block_size, C});
  }
  for (int i = d.start_layer; i <= d.end_layer; ++i) {      // gpt_model_parts.py:12,30,40
    std::string p = "transformer.h." + std::to_string(i) + ".";
    b2p_gpt_block b;
    b.ln1_w = st->add(p + "ln_1.weight", {C});
    b.ln1_b = st->add(p + "ln_1.bias", {C}, false);
    b.attn_w = st->add(p + "attn.c_attn.weight", {3 * C, C});
    b.attn_b = st->add(p + "attn.c_attn.bias", {3 * C}, false);
    b.proj_w = st->add(p + "attn.c_proj.weight", {C, C});
    b.proj_b = st->add(p + "attn.c_proj.bias", {C}, false);
    b.ln2_w = st->add(p + "ln_2.weight", {C});
    b.ln2_b = st->add(p + "ln_2.bias", {C}, false);
    b.fc_w = st->add(p + "mlp.c_fc.weight", {4 * C, C});
    b.fc_b = st->add(p + "mlp.c_fc.bias", {4 * C}, false);
    b.fc2_w = st->add(p + "mlp.c_proj.weight", {C, 4 * C});
    b.fc2_b = st->add(p + "mlp.c_proj.bias", {C}, false);
    st->blocks.push_back(b);
  }
  if (d.kind == B2P_GPT_FINAL) {                            // gpt_model_parts.py:41-42
    st->lnf_w = st->add("transformer.ln_f.weight", {C});
    st->lnf_b = st->add("transformer.ln_f.bias", {C}, false);
    st->head = st->add("lm_head.weight", {d.vocab_size, C});
  }
  return 0;
}

int alloc_gpt_workspace(b2p_stage* st) {
  B2P_CUDA(cudaFuncSetAttribute(attention_causal_f32<float>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)sizeof(AttnSmem)));
  B2P_CUDA(cudaFuncSetAttribute(attention_causal_f32<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)sizeof(AttnSmem)));
  const size_t M = (size_t)st->d.max_tokens, C = (size_t)st->d.n_embd;
  B2P_TRY(dev_alloc(&st->ws_ln, M * C));
  B2P_TRY(dev_alloc(&st->ws_qkv, M * 3 * C));
  B2P_TRY(dev_alloc(&st->ws_att, M * C));
  B2P_TRY(dev_alloc(&st->ws_h, M * 4 * C));
  B2P_TRY(dev_alloc(&st->ws_x[0], M * C));
  B2P_TRY(dev_alloc(&st->ws_x[1], M * C));
  B2P_TRY(dev_alloc(&st->bad_flag, 1));
  B2P_CUDA(cudaMemset(st->bad_flag, 0, sizeof(int)));
  return 0;
}

// ---- forward implementations -------------------------------------------------------

const RingSpan kNoSpan{};

bool part0_v1() {
  static const bool on = [] { const char* e = getenv("B2P_PART0_V1"); return e && e[0] == '1'; }();
  return on;
}

int cifar_part0_forward(b2p_stage* st, const float* in, float* out, int64_t batch, cudaStream_t s,
                        RingSpan span = kNoSpan) {
  const int dev = st->d.device;
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int grid = (int)std::min<int64_t>(batch, sms);
  if (st->d.math == B2P_MATH_FP32)
    B2P_LAUNCH(part0_fused_f32, grid, 256, sizeof(Part0Smem), s, in, st->w1t, st->dev(1), st->w2t,
               st->dev(3), out, (int)batch);
  else if (!part0_v1())   // both convolutions on the tensor cores (split-fp16, fp32-grade) in every tensor-core math mode
  {
    // ring mode with bit 2 of the stage's PDL mask: consecutive steps use different slots and the input is not the
    // preceding kernel's output, so the kernel need not wait for that kernel (b200pipe.h, b2p_stage_set_pdl)
    const int no_wait = (span.data && pdl_call_mask() >= 0 && (pdl_call_mask() & 4) && pdl_enabled(2) &&
                         2 * ceil_div(batch, span.rows_per_msg) <= span.slots) ? 1 : 0;   // (this launch and the next never share a slot)
    B2P_LAUNCH_PDL_BIT(2, cifar_tc2::part0_tc2, grid, cifar_tc2::kThreads, cifar_tc2::SMEM_BYTES, s, in, st->a1img, st->w2tm,
                       st->w2tap8, st->dev(1), st->dev(3), out, (int)batch, st->bad_flag, st->in_limit, st->trace, span, no_wait);
  }
  else   // round 1's kernel (conv1 on the CUDA cores), kept for A/B measurements: B2P_PART0_V1=1
  {
    cifar_tc::Conv1Const c1;
    static_assert(sizeof(c1) == sizeof(st->w1_host), "host copy layout");
    memcpy(&c1, st->w1_host, sizeof(c1));
    B2P_LAUNCH_PDL_BIT(2, cifar_tc::part0_tc, grid, cifar_tc::kThreads, cifar_tc::SMEM_BYTES, s, in, c1,
                       st->w2img, st->dev(3), out, (int)batch, st->trace, span);
  }
  return 0;
}

// Fixed split count: a row's summation order must not depend on the batch it travels
// in, so that micro-batching (8 x 128 vs 1 x 1024) is bit-exact.
constexpr int FC1_SPLITS = 8;
int fc1_splits(int64_t) { return FC1_SPLITS; }

// Tensor-core modes use more, shorter K slices: the 128x512 fc1 output is only 4 tiles,
// so the K dimension is what spreads the 8.4 MB weight read over the SMs.
constexpr int FC1_SPLITS_TC = 16;

int stage_linear(b2p_stage* st, int act, const void* A, int widx, int bidx, const float* res, void* out,
                 bool out_bf16, int M, int N, int K, cudaStream_t s, int splits = 1, int64_t split_stride = 0,
                 float* amax_val = nullptr, int* amax_idx = nullptr, int amax_slots = 0, int amax_bn = 128);

// B2P_FC_FUSED=0 keeps the three-kernel path (split, split-K linear, tail) for A/B measurements
bool fc_fused_enabled() {
  static const bool on = [] { const char* e = getenv("B2P_FC_FUSED"); return !(e && e[0] == '0'); }();
  return on;
}

int cifar_part1_forward(b2p_stage* st, const float* in, float* out, int64_t batch, cudaStream_t s,
                        RingSpan span = kNoSpan) {
  const bool tcm = st->d.math != B2P_MATH_FP32;
  if (st->d.math == B2P_MATH_TF32X3 && fc_fused_enabled()) {
    // fc1 with the activation split inside the kernel (fc_stage_tc.cuh), 32 K slices, then the shared tail
    B2P_TRY(ensure_scratch(&st->fc_ws, &st->fc_ws_bytes, fcs::ws_bytes(batch), 1, s, true, "fc1 split-K"));
    // Two-stream mode (b2p_stage_set_aux_stream): the tail runs on the stage's second stream behind an event, the next
    // step's fc1 does not wait for it (the partial sums alternate between two buffers), so on a consumer GPU the
    // 9 us tail of step s hides behind fc1 of step s+1, which is itself waiting for the producer's messages.
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    const bool two_streams = st->aux && cudaStreamIsCapturing(s, &cs) == cudaSuccess && cs == cudaStreamCaptureStatusNone;
    float* ws = st->fc_ws;
    cudaStream_t tail_stream = s;
    const int buf = (int)(st->fc_step & 1);
    if (two_streams) {
      B2P_TRY(ensure_scratch(&st->fc_ws2, &st->fc_ws2_bytes, fcs::ws_bytes(batch), 1, s, false, "fc1 split-K"));
      if (buf) ws = st->fc_ws2;
      if (st->ev_tail_recorded[buf]) B2P_CUDA(cudaStreamWaitEvent(s, st->ev_tail[buf], 0));   // tail(s-2) has read this buffer
      tail_stream = st->aux;
    }
    const b2p_weight& w = st->w[0];
    static const bool tf32_fc1 = [] { const char* e = getenv("B2P_FC1_TF32"); return e && e[0] == '1'; }();   // A/B: round 1's 3xTF32 kernel
    int parts = 0, per_item = 0;
    if (!tf32_fc1 && w.packed16) {
      const __half* w16 = static_cast<const __half*>(w.packed16);
      per_item = fcs::launch_fc1_f16x3_splitk(in, w16, w16 + w.elems(), ws, batch, st->bad_flag, s, span);
      if (per_item < 0) return per_item;
    } else {
      parts = fcs::launch_fc1_x3_splitk(in, static_cast<const float*>(w.packed), w.packed_lo, ws, batch, s, span);
      if (parts < 0) return parts;
    }
    if (two_streams) {
      B2P_CUDA(cudaEventRecord(st->ev_fc1, s));
      B2P_CUDA(cudaStreamWaitEvent(st->aux, st->ev_fc1, 0));
    }
    const int tail_grid = (int)batch + (span.data ? 1 : 0);
    if (per_item == 1)
      B2P_LAUNCH_PDL((fc_tail_f32<fcs::NSLICES, 4>), tail_grid, 128, 0, tail_stream, ws, (int)batch, st->dev(1), st->dev(2),
                     st->dev(3), out, span);
    else if (per_item == 4)
      B2P_LAUNCH_PDL((fc_tail_f32<fcs::NSLICES / 4, 1>), tail_grid, 128, 0, tail_stream, ws, (int)batch, st->dev(1), st->dev(2),
                     st->dev(3), out, span);
    else if (parts == fcs::NSLICES)
      B2P_LAUNCH_PDL((fc_tail_f32<fcs::NSLICES, 2>), tail_grid, 128, 0, tail_stream, ws, (int)batch, st->dev(1), st->dev(2),
                     st->dev(3), out, span);
    else
      B2P_LAUNCH_PDL((fc_tail_f32<fcs::NSLICES / 2, 1>), tail_grid, 128, 0, tail_stream, ws, (int)batch, st->dev(1), st->dev(2),
                     st->dev(3), out, span);
    if (two_streams) {
      B2P_CUDA(cudaEventRecord(st->ev_tail[buf], st->aux));
      st->ev_tail_recorded[buf] = true;
      ++st->fc_step;
    }
    return 0;
  }
  const int splits = tcm ? FC1_SPLITS_TC : fc1_splits(batch);
  B2P_TRY(ensure_scratch(&st->fc_ws, &st->fc_ws_bytes, (size_t)splits * batch * FC1_N * sizeof(float), 1, s, true,
                         "fc1 split-K"));
  if (tcm) {
    B2P_TRY(stage_linear(st, EPI_NONE, in, 0, -1, nullptr, st->fc_ws, false, (int)batch, FC1_N, FC1_K, s, splits,
                         (int64_t)batch * FC1_N));
  } else {
    dim3 grid(FC1_N / 64, ceil_div(batch, 64), splits);
    B2P_LAUNCH(fc1_splitk_f32, grid, 256, 0, s, in, st->dev(0), st->fc_ws, (int)batch, FC1_K / splits);
  }
  if (tcm)
    B2P_LAUNCH_PDL(fc_tail_f32<FC1_SPLITS_TC>, (int)batch, 128, 0, s, st->fc_ws, (int)batch, st->dev(1), st->dev(2),
                   st->dev(3), out, kNoSpan);
  else
    B2P_LAUNCH_PDL(fc_tail_f32<FC1_SPLITS>, (int)batch, 128, 0, s, st->fc_ws, (int)batch, st->dev(1), st->dev(2),
                   st->dev(3), out, kNoSpan);
  return 0;
}

int linear(int epi, const float* A, const float* W, const float* bias, const float* res, float* out,
           int M, int N, int K, cudaStream_t s) {
  dim3 grid(ceil_div(N, 64), ceil_div(M, 64));
  if (K % 16) return fail(B2P_EINVAL, "linear: K=%d must be a multiple of 16", K);
  switch (epi) {
    case EPI_NONE: B2P_LAUNCH(linear_f32<EPI_NONE>, grid, 256, 0, s, A, W, bias, res, out, M, N, K); break;
    case EPI_RELU: B2P_LAUNCH(linear_f32<EPI_RELU>, grid, 256, 0, s, A, W, bias, res, out, M, N, K); break;
    case EPI_GELU: B2P_LAUNCH(linear_f32<EPI_GELU>, grid, 256, 0, s, A, W, bias, res, out, M, N, K); break;
    default: return fail(B2P_EINVAL, "linear: bad epilogue %d", epi);
  }
  return 0;
}

int elementwise_grid(size_t n) { return (int)std::min<size_t>((n + 255) / 256, 148 * 8); }

// hi | lo planes of an fp32 activation for the 3xTF32 contraction (grows on demand)
int split_activation(b2p_stage* st, const float* A, size_t n, cudaStream_t s) {
  B2P_TRY(ensure_scratch(&st->ws_split, &st->ws_split_elems, n, 2 * sizeof(float), s, true, "hi/lo plane"));
  B2P_LAUNCH_PDL(split_tf32, elementwise_grid(n), 256, 0, s, A, st->ws_split, st->ws_split + st->ws_split_elems, n);
  return 0;
}

// One nn.Linear call site in the stage's math mode.  A is fp32 except in BF16 mode,
// where the producer wrote it as bf16; `out` is fp32 unless out_bf16.
int stage_linear(b2p_stage* st, int act, const void* A, int widx, int bidx, const float* res, void* out,
                 bool out_bf16, int M, int N, int K, cudaStream_t s, int splits, int64_t split_stride,
                 float* amax_val, int* amax_idx, int amax_slots, int amax_bn) {
  const float* bias = st->dev(bidx);
  const b2p_weight& w = st->w[widx];
  tc::EpiParams ep{bias, res, out, N, N, out_bf16 ? 1 : 0, act, split_stride, amax_val, amax_idx, amax_slots};
  int rc = 0;
  switch (st->d.math) {
    case B2P_MATH_FP32:
      return linear(act, static_cast<const float*>(A), w.dev, bias, res, static_cast<float*>(out), M, N, K, s);
    case B2P_MATH_TF32:
      rc = tc::linear_dispatch(1, A, nullptr, K, w.dev, nullptr, K, ep, M, N, K, splits, amax_val ? amax_bn : 0, s);
      break;
    case B2P_MATH_TF32X3:
      B2P_TRY(split_activation(st, static_cast<const float*>(A), (size_t)M * K, s));
      rc = tc::linear_dispatch(2, st->ws_split, st->ws_split + st->ws_split_elems, K, w.packed, w.packed_lo, K,
                                  ep, M, N, K, splits, 0, s);
      break;
    case B2P_MATH_BF16:
      rc = tc::linear_dispatch(0, A, nullptr, K, w.packed, nullptr, K, ep, M, N, K, splits, amax_val ? amax_bn : 0, s);
      break;
    default:
      return fail(B2P_EINVAL, "unknown math mode %d", st->d.math);
  }
  return rc < 0 ? rc : 0;
}

template <typename OutT, int NVEC>
int layernorm_launch(const float* x, const float* g, const float* b, OutT* y, int M, int C, cudaStream_t s) {
  B2P_LAUNCH_PDL((layernorm_rows<OutT, NVEC>), ceil_div((int64_t)M * 32, 256), 256, 0, s, x, g, b, y, M, C);
  return 0;
}

template <typename OutT>
int layernorm_typed(const float* x, const float* g, const float* b, OutT* y, int M, int C, cudaStream_t s) {
  switch (C / 128) {                       // n_embd is a multiple of 128, at most 2048 (checked at stage creation)
    case 1: return layernorm_launch<OutT, 1>(x, g, b, y, M, C, s);
    case 2: return layernorm_launch<OutT, 2>(x, g, b, y, M, C, s);
    case 3: return layernorm_launch<OutT, 3>(x, g, b, y, M, C, s);
    case 4: return layernorm_launch<OutT, 4>(x, g, b, y, M, C, s);
    case 5: return layernorm_launch<OutT, 5>(x, g, b, y, M, C, s);
    case 6: return layernorm_launch<OutT, 6>(x, g, b, y, M, C, s);
    case 7: return layernorm_launch<OutT, 7>(x, g, b, y, M, C, s);
    case 8: return layernorm_launch<OutT, 8>(x, g, b, y, M, C, s);
    case 9: return layernorm_launch<OutT, 9>(x, g, b, y, M, C, s);
    case 10: return layernorm_launch<OutT, 10>(x, g, b, y, M, C, s);
    case 11: return layernorm_launch<OutT, 11>(x, g, b, y, M, C, s);
    case 12: return layernorm_launch<OutT, 12>(x, g, b, y, M, C, s);
    case 13: return layernorm_launch<OutT, 13>(x, g, b, y, M, C, s);
    case 14: return layernorm_launch<OutT, 14>(x, g, b, y, M, C, s);
    case 15: return layernorm_launch<OutT, 15>(x, g, b, y, M, C, s);
    case 16: return layernorm_launch<OutT, 16>(x, g, b, y, M, C, s);
    default: return fail(B2P_EINVAL, "layernorm: unsupported width %d", C);
  }
}

int layernorm(b2p_stage* st, const float* x, int gi, int bi, void* y, int M, int C, cudaStream_t s) {
  if (st->d.math == B2P_MATH_BF16)
    return layernorm_typed<__nv_bfloat16>(x, st->dev(gi), st->dev(bi), static_cast<__nv_bfloat16*>(y), M, C, s);
  return layernorm_typed<float>(x, st->dev(gi), st->dev(bi), static_cast<float*>(y), M, C, s);
}

// Runs this stage's blocks.  `x` is the stage input (never written); the result of
// the last block lands in `final_out` (caller-provided) -- intermediates ping-pong
// between the two workspace activations.  The residual stream stays fp32 in every
// math mode; in BF16 mode the GEMM A operands (LN output, attention output, GELU
// output) are written as bf16 by their producers.
// pos0 < 0: plain forward.  pos0 >= 0: the rows are positions pos0 .. pos0+T-1 of their sequences; their keys and values
// go into the stage's KV cache, and a single new position (T == 1, pos0 > 0) attends to the cache instead of recomputing.
int gpt_blocks(b2p_stage* st, const float* x, float* final_out, int B, int T, cudaStream_t s, int pos0 = -1,
               const int* pos_dev = nullptr) {      // pos_dev: one new position per sequence, its index in device memory (pos0 > 0 then only says "decode")
  const int C = st->d.n_embd, H = st->d.n_head, M = B * T;
  const bool bf = st->d.math == B2P_MATH_BF16;
  const float* cur = x;
  const size_t nb = st->blocks.size();
  for (size_t i = 0; i < nb; ++i) {
    const b2p_gpt_block& k = st->blocks[i];
    float* x1 = st->ws_x[0];
    float* x2 = (i + 1 == nb) ? final_out : st->ws_x[1];
    B2P_TRY(layernorm(st, cur, k.ln1_w, k.ln1_b, st->ws_ln, M, C, s));
    B2P_TRY(stage_linear(st, EPI_NONE, st->ws_ln, k.attn_w, k.attn_b, nullptr, st->ws_qkv, bf, M, 3 * C, C, s));
    const bool decode = pos0 > 0;
    if (pos0 >= 0) {
      const int Tmax = st->d.block_size;
      const size_t per_block = (size_t)st->kv_batch * H * Tmax * 64;
      float* kc = st->kv_k + i * per_block;
      float* vc = st->kv_v + i * per_block;
      const int64_t n = (int64_t)M * H * 16;
      if (bf)
        B2P_LAUNCH_PDL(kv::kv_append<__nv_bfloat16>, ceil_div(n, 256), 256, 0, s, reinterpret_cast<const __nv_bfloat16*>(st->ws_qkv),
                       kc, vc, B, T, C, H, Tmax, pos0, pos_dev);
      else
        B2P_LAUNCH_PDL(kv::kv_append<float>, ceil_div(n, 256), 256, 0, s, st->ws_qkv, kc, vc, B, T, C, H, Tmax, pos0, pos_dev);
      if (decode) {
        const size_t smem = ((size_t)Tmax + 64 + kv::DEC_THREADS) * sizeof(float);
        dim3 dgrid(H, B);
        if (bf)
          B2P_LAUNCH_PDL(kv::attention_decode<__nv_bfloat16>, dgrid, kv::DEC_THREADS, smem, s, reinterpret_cast<const __nv_bfloat16*>(st->ws_qkv),
                         kc, vc, reinterpret_cast<__nv_bfloat16*>(st->ws_att), C, H, Tmax, pos0, pos_dev);
        else
          B2P_LAUNCH_PDL(kv::attention_decode<float>, dgrid, kv::DEC_THREADS, smem, s, st->ws_qkv, kc, vc, st->ws_att, C, H, Tmax, pos0, pos_dev);
      }
    }
    if (decode) {
      // (attention done above, from the cache)
    } else if (bf) {
      B2P_TRY(attn::launch_attention_tc(reinterpret_cast<const __nv_bfloat16*>(st->ws_qkv),
                                        reinterpret_cast<__nv_bfloat16*>(st->ws_att), B, T, C, H, s));
    } else {
      dim3 agrid(ceil_div(T, 64), H, B);
      B2P_LAUNCH(attention_causal_f32<float>, agrid, 256, sizeof(AttnSmem), s, st->ws_qkv, st->ws_att, T, C);
    }
    B2P_TRY(stage_linear(st, EPI_NONE, st->ws_att, k.proj_w, k.proj_b, cur, x1, false, M, C, C, s));
    B2P_TRY(layernorm(st, x1, k.ln2_w, k.ln2_b, st->ws_ln, M, C, s));
    B2P_TRY(stage_linear(st, EPI_GELU, st->ws_ln, k.fc_w, k.fc_b, nullptr, st->ws_h, bf, M, 4 * C, C, s));
    B2P_TRY(stage_linear(st, EPI_NONE, st->ws_h, k.fc2_w, k.fc2_b, x1, x2, false, M, C, 4 * C, s));
    cur = x2;
  }
  if (nb == 0 && final_out != x)
    B2P_CUDA(cudaMemcpyAsync(final_out, x, (size_t)M * C * sizeof(float), cudaMemcpyDeviceToDevice, s));
  return 0;
}

int check_gpt_geometry(b2p_stage* st, int64_t batch, int64_t seq) {
  if (batch <= 0 || seq <= 0) return fail(B2P_EINVAL, "batch and seq must be positive");
  if (seq > st->d.block_size)                                 // gpt_model_parts.py:15
    return fail(B2P_EINVAL, "Cannot forward, model block size is exhausted. (T=%lld > block_size=%d)",
                (long long)seq, st->d.block_size);
  if (batch * seq > st->d.max_tokens)
    return fail(B2P_ELIMIT, "batch*seq=%lld exceeds the stage's max_tokens=%d", (long long)(batch * seq),
                st->d.max_tokens);
  return 0;
}

int reserve_amax(b2p_stage* st, size_t need, cudaStream_t s, bool check_capture) {
  if (need <= st->amax_elems) return 0;
  size_t have_v = st->amax_elems, have_i = st->amax_elems;
  st->amax_elems = 0;
  B2P_TRY(ensure_scratch(&st->amax_val, &have_v, need, sizeof(float), s, check_capture, "arg-max partial"));
  B2P_TRY(ensure_scratch(&st->amax_idx, &have_i, need, sizeof(int), s, false, "arg-max partial"));
  st->amax_elems = need;
  return 0;
}

// Everything a forward of `rows` rows (CIFAR: batch; GPT: batch*seq tokens) may need beyond the stage's fixed workspaces.
int reserve_rows(b2p_stage* st, int64_t rows) {
  const size_t M = (size_t)rows;
  switch (st->d.kind) {
    case B2P_CIFAR_PART0:
      return 0;                                              // no scratch at all: everything lives in shared / tensor memory
    case B2P_CIFAR_PART1: {
      size_t need = (size_t)(st->d.math == B2P_MATH_FP32 ? FC1_SPLITS : FC1_SPLITS_TC) * M * FC1_N * sizeof(float);
      if (st->d.math == B2P_MATH_TF32X3) {
        need = std::max(need, fcs::ws_bytes((int64_t)M));
        B2P_TRY(ensure_scratch(&st->ws_split, &st->ws_split_elems, M * FC1_K, 2 * sizeof(float), nullptr, false, ""));   // B2P_FC_FUSED=0 path
      }
      return ensure_scratch(&st->fc_ws, &st->fc_ws_bytes, need, 1, nullptr, false, "");
    }
    default: {
      const size_t C = (size_t)st->d.n_embd;
      if (st->d.math == B2P_MATH_TF32X3)
        B2P_TRY(ensure_scratch(&st->ws_split, &st->ws_split_elems, M * 4 * C, 2 * sizeof(float), nullptr, false, ""));
      if (st->d.kind == B2P_GPT_FINAL) {
        if (st->d.math != B2P_MATH_FP32)
          B2P_TRY(reserve_amax(st, M * 2 * (size_t)ceil_div(st->d.vocab_size, 128), nullptr, false));
        else
          B2P_TRY(ensure_scratch(&st->ws_logits, &st->ws_logits_rows, std::min<size_t>(M, 2048),
                                 (size_t)st->d.vocab_size * sizeof(float), nullptr, false, ""));
      }
      return 0;
    }
  }
}

int gpt_forward(b2p_stage* st, const void* in, void* out, int64_t batch, int64_t seq, cudaStream_t s,
                int64_t* ids_out, int pos0 = -1, const int* pos_dev = nullptr) {
  B2P_TRY(check_gpt_geometry(st, batch, seq));
  if (pos0 >= 0) {
    if (!st->blocks.empty() && (!st->kv_k || batch > st->kv_batch))
      return fail(B2P_ESTATE, "forward_kv: call b2p_stage_kv_reset for at least %lld sequences first", (long long)batch);
    if (pos0 > 0 && seq != 1) return fail(B2P_EINVAL, "forward_kv: after the prefill (pos0 = 0) positions come one at a time");
    if (pos0 + seq > st->d.block_size)
      return fail(B2P_EINVAL, "forward_kv: position %lld is beyond block_size=%d (the cache holds absolute positions)",
                  (long long)(pos0 + seq - 1), st->d.block_size);
    if (st->d.block_size > 8192) return fail(B2P_ELIMIT, "forward_kv: block_size above 8192 does not fit the score buffer");
  }
  const int B = (int)batch, T = (int)seq, M = B * T, C = st->d.n_embd, V = st->d.vocab_size;
  const float* x = static_cast<const float*>(in);
  if (st->d.kind == B2P_GPT_FIRST) {
    // The embedding is block 0's input: it goes to ws_x[1], which block 0 overwrites
    // (with its own output) only after reading it for the last time.
    float* emb = st->blocks.empty() ? static_cast<float*>(out) : st->ws_x[1];
    B2P_LAUNCH(embed_f32, ceil_div((int64_t)M * 32, 256), 256, 0, s, static_cast<const int64_t*>(in),
               st->dev(st->wte), st->dev(st->wpe), emb, M, T, C, V, st->bad_flag, pos0 > 0 ? pos0 : 0, pos_dev, st->d.block_size);
    if (st->blocks.empty()) return 0;
    return gpt_blocks(st, emb, static_cast<float*>(out), B, T, s, pos0, pos_dev);
  }
  if (st->d.kind == B2P_GPT_MIDDLE) return gpt_blocks(st, x, static_cast<float*>(out), B, T, s, pos0, pos_dev);

  // final: blocks -> ln_f -> lm_head                          gpt_model_parts.py:44-50
  const float* hx = x;
  if (!st->blocks.empty()) {
    B2P_TRY(gpt_blocks(st, x, st->ws_att, B, T, s, pos0, pos_dev));    // ws_att is free after the last block's proj
    hx = st->ws_att;
  }
  B2P_TRY(layernorm(st, hx, st->lnf_w, st->lnf_b, st->ws_ln, M, C, s));
  if (!ids_out) return stage_linear(st, EPI_NONE, st->ws_ln, st->head, -1, nullptr, out, false, M, V, C, s);
  if (st->d.math != B2P_MATH_FP32) {
    // tensor-core modes: arg-max fused into the LM-head epilogue, logits never materialised
    // tile width of the head: 256 wide (half the shared-memory traffic per MMA, N = 256 MMAs at the tensor pipe's rate)
    // whenever such tiles still fill the GPU, else 128; two partial results (column halves) per tile and row
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, st->d.device);
    const int head_bn = (st->d.math != B2P_MATH_TF32X3 && (int64_t)ceil_div(M, 128) * ceil_div(V, 256) >= sms) ? 256 : 128;
    const int slots = 2 * ceil_div(V, head_bn);
    B2P_TRY(reserve_amax(st, (size_t)M * slots, s, true));
    B2P_TRY(stage_linear(st, EPI_NONE, st->ws_ln, st->head, -1, nullptr, st->amax_val /*unused*/, false, M, V, C, s, 1, 0,
                         st->amax_val, st->amax_idx, slots, head_bn));      // (the CTA-pair kernel measured the same: 2.75 vs 2.77 M tokens/s)
    B2P_LAUNCH_PDL(argmax_partials, ceil_div((int64_t)M * 32, 256), 256, 0, s, st->amax_val, st->amax_idx, ids_out, M, slots);
    return 0;
  }
  // fp32 CUDA-core mode: logits in row chunks through a bounded scratch buffer, then row arg-max
  const size_t chunk = std::min<size_t>((size_t)M, 2048);
  B2P_TRY(ensure_scratch(&st->ws_logits, &st->ws_logits_rows, chunk, (size_t)V * sizeof(float), s, true, "logit chunk"));
  for (size_t m0 = 0; m0 < (size_t)M; m0 += chunk) {
    const int rows = (int)std::min(chunk, (size_t)M - m0);
    const char* a_rows = reinterpret_cast<const char*>(st->ws_ln) +
                         m0 * C * (st->d.math == B2P_MATH_BF16 ? sizeof(__nv_bfloat16) : sizeof(float));
    B2P_TRY(stage_linear(st, EPI_NONE, a_rows, st->head, -1, nullptr, st->ws_logits, false, rows, V, C, s));
    B2P_LAUNCH(argmax_rows, rows, 256, 0, s, st->ws_logits, ids_out + m0, V, (int64_t)V);
  }
  return 0;
}

}  // namespace

// =====================================================================================
extern "C" {

int b2p_abi_version(void) { return B2P_ABI_VERSION; }

const char* b2p_last_error(void) { return err_slot().c_str(); }

int b2p_device_count(int* count) {
  if (!count) return fail(B2P_EINVAL, "count is null");
  *count = 0;
  B2P_CUDA(cudaGetDeviceCount(count));
  return 0;
}

int64_t b2p_launch_count(void) { return launch_counter().load(); }

int b2p_set_pdl(int mask) {
  const int prev = pdl_switch().exchange(mask & 3);
  return prev;
}

int b2p_stage_create(b2p_stage** out, const b2p_stage_desc* desc) {
  if (!out || !desc) return fail(B2P_EINVAL, "null argument");
  *out = nullptr;
  if (!is_cifar(desc->kind) && !is_gpt(desc->kind)) return fail(B2P_EINVAL, "unknown stage kind %d", desc->kind);
  if (desc->math < B2P_MATH_FP32 || desc->math > B2P_MATH_BF16) return fail(B2P_EINVAL, "unknown math mode %d", desc->math);
  if (is_cifar(desc->kind) && desc->math == B2P_MATH_BF16)
    return fail(B2P_EINVAL, "the CIFAR parts compute in fp32, tf32x3 or tf32 (the reference is fp32); bf16 is a GPT mode");
  if (is_gpt(desc->kind)) {
    if (desc->n_embd <= 0 || desc->n_embd % 128 || desc->n_embd > 2048)
      return fail(B2P_EINVAL, "n_embd=%d must be a positive multiple of 128, at most 2048", desc->n_embd);
    if (desc->n_head <= 0 || desc->n_embd != desc->n_head * 64)
      return fail(B2P_EINVAL, "head dim must be 64 (n_embd=%d, n_head=%d)", desc->n_embd, desc->n_head);
    if (desc->start_layer < 0 || desc->end_layer < desc->start_layer - 1)
      return fail(B2P_EINVAL, "bad layer range [%d,%d]", desc->start_layer, desc->end_layer);
    if (desc->block_size <= 0 || desc->max_tokens <= 0) return fail(B2P_EINVAL, "block_size and max_tokens must be positive");
    if (desc->kind != B2P_GPT_MIDDLE && desc->vocab_size <= 0) return fail(B2P_EINVAL, "vocab_size must be positive");
  }
  int ndev = 0;
  B2P_CUDA(cudaGetDeviceCount(&ndev));
  if (desc->device < 0 || desc->device >= ndev)
    return fail(B2P_ECUDA, "device %d not present (%d CUDA devices visible); there is no CPU fallback", desc->device, ndev);
  b2p_stage* st = new (std::nothrow) b2p_stage();
  if (!st) return fail(B2P_EINVAL, "out of host memory");
  st->d = *desc;
  declare_weights(st);
  DeviceGuard g(desc->device);
  for (auto& w : st->w) {
    int rc = dev_alloc(&w.dev, w.elems());
    if (rc < 0) { b2p_stage_destroy(st); return rc; }
  }
  if (is_gpt(desc->kind)) {
    int rc = alloc_gpt_workspace(st);
    if (rc < 0) { b2p_stage_destroy(st); return rc; }
  }
  if (desc->kind == B2P_CIFAR_PART1) {
    int rc = dev_alloc(&st->bad_flag, 1);                       // "a value left the split-fp16 range" (fc_stage_tc.cuh)
    if (rc == 0 && cudaMemset(st->bad_flag, 0, sizeof(int)) != cudaSuccess) rc = fail(B2P_ECUDA, "cudaMemset(range flag)");
    if (rc < 0) { b2p_stage_destroy(st); return rc; }
  }
  if (desc->kind == B2P_CIFAR_PART0) {
    int rc = dev_alloc(&st->bad_flag, 1);                       // "a value left the split-fp16 range" (cifar_tc2.cuh)
    if (rc == 0 && cudaMemset(st->bad_flag, 0, sizeof(int)) != cudaSuccess) rc = fail(B2P_ECUDA, "cudaMemset(range flag)");
    if (rc < 0) { b2p_stage_destroy(st); return rc; }
    // per-device function attribute: set with the stage's device current
    cudaError_t e = cudaFuncSetAttribute(part0_fused_f32, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)sizeof(Part0Smem));
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(cifar_tc::part0_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, cifar_tc::SMEM_BYTES);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(cifar_tc2::part0_tc2, cudaFuncAttributeMaxDynamicSharedMemorySize, cifar_tc2::SMEM_BYTES);
    if (e != cudaSuccess) {
      b2p_stage_destroy(st);
      return fail(B2P_ECUDA, "cudaFuncSetAttribute(part0 kernels): %s", cudaGetErrorString(e));
    }
  }
  *out = st;
  return 0;
}

int b2p_stage_destroy(b2p_stage* st) {
  if (!st) return 0;
  DeviceGuard g(st->d.device);
  for (auto& w : st->w) { cudaFree(w.dev); cudaFree(w.packed); cudaFree(w.packed_lo); cudaFree(w.packed16); }
  cudaFree(st->ws_split); cudaFree(st->amax_val); cudaFree(st->amax_idx);
  cudaFree(st->kv_k); cudaFree(st->kv_v); cudaFree(st->w1t); cudaFree(st->w2t); cudaFree(st->w2img); cudaFree(st->a1img); cudaFree(st->w2tm); cudaFree(st->w2tap8); cudaFree(st->trace); cudaFree(st->fc_ws); cudaFree(st->fc_ws2);
  if (st->ev_fc1) { cudaEventDestroy(st->ev_fc1); cudaEventDestroy(st->ev_tail[0]); cudaEventDestroy(st->ev_tail[1]); }
  cudaFree(st->ws_ln); cudaFree(st->ws_qkv); cudaFree(st->ws_att); cudaFree(st->ws_h);
  cudaFree(st->ws_x[0]); cudaFree(st->ws_x[1]); cudaFree(st->ws_logits); cudaFree(st->bad_flag);
  delete st;
  return 0;
}

int b2p_stage_load_weight(b2p_stage* st, const char* name, const void* host_data, int32_t dtype,
                          const int64_t* shape, int32_t ndim) {
  if (!st || !name || !host_data || (ndim > 0 && !shape)) return fail(B2P_EINVAL, "null argument");
  std::string n(name);
  int i = st->find(n);
  // tied head: a final stage accepts the embedding matrix under its wte name when the
  // checkpoint carries no separate lm_head.weight (same tensor upstream).
  if (i < 0 && st->d.kind == B2P_GPT_FINAL && n == "transformer.wte.weight" && !st->w[st->head].loaded) i = st->head;
  if (i < 0) return 0;                                      // "unexpected key": not ours
  b2p_weight& w = st->w[i];
  if (dtype != B2P_DT_F32) return fail(B2P_EINVAL, "%s: only float32 checkpoints are supported (dtype code %d)", name, dtype);
  bool same = (size_t)ndim == w.shape.size();
  for (int k = 0; same && k < ndim; ++k) same = shape[k] == w.shape[k];
  if (!same) {
    std::string got, want;
    for (int k = 0; k < ndim; ++k) got += (k ? "," : "") + std::to_string(shape[k]);
    for (size_t k = 0; k < w.shape.size(); ++k) want += (k ? "," : "") + std::to_string(w.shape[k]);
    return fail(B2P_EINVAL, "size mismatch for %s: checkpoint has [%s], stage expects [%s]", name, got.c_str(), want.c_str());
  }
  DeviceGuard g(st->d.device);
  B2P_CUDA(cudaMemcpy(w.dev, host_data, w.elems() * sizeof(float), cudaMemcpyHostToDevice));
  w.loaded = true;
  st->finalized = false;
  return 1;
}

int b2p_stage_missing(b2p_stage* st, char* names_buf, size_t buf_len) {
  if (!st) return fail(B2P_EINVAL, "null stage");
  int n = 0;
  std::string all;
  for (auto& w : st->w)
    if (w.required && !w.loaded) { ++n; all += (all.empty() ? "" : "\n") + w.name; }
  if (names_buf && buf_len) {
    strncpy(names_buf, all.c_str(), buf_len - 1);
    names_buf[buf_len - 1] = 0;
  }
  return n;
}

int b2p_stage_finalize(b2p_stage* st) {
  if (!st) return fail(B2P_EINVAL, "null stage");
  for (auto& w : st->w)
    if (w.required && !w.loaded) return fail(B2P_EMISSING, "weight %s was never loaded", w.name.c_str());
  DeviceGuard g(st->d.device);
  if (st->d.kind == B2P_CIFAR_PART0) {
    if (!st->w1t) B2P_TRY(dev_alloc(&st->w1t, 27 * 32));
    if (!st->w2t) B2P_TRY(dev_alloc(&st->w2t, 288 * 64));
    B2P_LAUNCH(transpose2d_f32, ceil_div(32 * 27, 256), 256, 0, 0, st->dev(0), st->w1t, 32, 27);
    B2P_LAUNCH(transpose2d_f32, ceil_div(64 * 288, 256), 256, 0, 0, st->dev(2), st->w2t, 64, 288);
    if (st->d.math != B2P_MATH_FP32) {
      if (!st->w2img) B2P_CUDA(cudaMalloc((void**)&st->w2img, cifar_tc::W2_BYTES));
      B2P_LAUNCH(cifar_tc::pack_conv2_f16x2, ceil_div(9 * 128 * 4, 256), 256, 0, 0, st->dev(2), st->w2img);
      if (!st->a1img) B2P_CUDA(cudaMalloc((void**)&st->a1img, 2 * cifar_tc2::PLANE_BYTES));
      if (!st->w2tm) B2P_CUDA(cudaMalloc((void**)&st->w2tm, 128 * cifar_tc2::W2T_WORDS * sizeof(uint32_t)));
      B2P_CUDA(cudaMemset(st->a1img, 0, 2 * cifar_tc2::PLANE_BYTES));
      B2P_LAUNCH(cifar_tc2::pack_conv1_poolwin, ceil_div(128 * 64, 256), 256, 0, 0, st->dev(0), st->a1img);
      B2P_LAUNCH(cifar_tc2::pack_conv2_tmem, ceil_div(128 * cifar_tc2::W2T_WORDS, 256), 256, 0, 0, st->dev(2), st->w2tm);
      if (!st->w2tap8) B2P_CUDA(cudaMalloc((void**)&st->w2tap8, cifar_tc::W2_TAP_BYTES));
      B2P_LAUNCH(cifar_tc2::pack_conv2_tap8, 2, 256, 0, 0, st->dev(2), st->w2tap8);
      // host copy of the transposed conv1 weights + bias: they travel as a kernel parameter (cifar_tc::Conv1Const)
      B2P_CUDA(cudaMemcpy(st->w1_host, st->w1t, 27 * 32 * sizeof(float), cudaMemcpyDeviceToHost));
      B2P_CUDA(cudaMemcpy(st->w1_host + 27 * 32, st->dev(1), 32 * sizeof(float), cudaMemcpyDeviceToHost));
      // input magnitude from which on a conv1 output can reach the fp16 limit of the split scheme (cifar_tc2.cuh)
      float wsum = 0.f, bmax = 0.f;
      for (int oc = 0; oc < 32; ++oc) {
        float a = 0.f;
        for (int k = 0; k < 27; ++k) a += fabsf(st->w1_host[k * 32 + oc]);
        wsum = std::max(wsum, a);
        bmax = std::max(bmax, fabsf(st->w1_host[27 * 32 + oc]));
      }
      st->in_limit = (wsum > 0.f && bmax < 65504.f) ? std::min(65504.f, (65504.f - bmax) / wsum) : (bmax < 65504.f ? 65504.f : 0.f);
    }
  }
  // math-mode copies of the matrices that feed the tensor cores
  for (auto& w : st->w) {
    if (!w.loaded || w.shape.size() != 2) continue;
    const bool gemm_operand = is_gpt(st->d.kind) ? (w.name.find("wpe") == std::string::npos &&
                                                   (st->d.kind != B2P_GPT_FIRST || w.name.find("wte") == std::string::npos))
                                                 : w.name == "fc1.weight";
    if (!gemm_operand) continue;
    const size_t n = w.elems();
    if (st->d.math == B2P_MATH_BF16) {
      if (!w.packed) B2P_CUDA(cudaMalloc(&w.packed, n * sizeof(__nv_bfloat16)));
      B2P_LAUNCH(f32_to_bf16, elementwise_grid(n), 256, 0, 0, w.dev, static_cast<__nv_bfloat16*>(w.packed), n);
    } else if (st->d.math == B2P_MATH_TF32X3) {
      if (!w.packed) B2P_CUDA(cudaMalloc(&w.packed, n * sizeof(float)));
      if (!w.packed_lo) B2P_CUDA(cudaMalloc((void**)&w.packed_lo, n * sizeof(float)));
      B2P_LAUNCH(split_tf32, elementwise_grid(n), 256, 0, 0, w.dev, static_cast<float*>(w.packed), w.packed_lo, n);
      if (st->d.kind == B2P_CIFAR_PART1) {       // fc1 runs in split fp16 (fc_stage_tc.cuh): hi plane | lo plane
        if (!w.packed16) B2P_CUDA(cudaMalloc(&w.packed16, 2 * n * sizeof(__half)));
        B2P_LAUNCH(fcs::split_f16_planes, elementwise_grid(n), 256, 0, 0, w.dev, static_cast<__half*>(w.packed16),
                   static_cast<__half*>(w.packed16) + n, n, st->bad_flag);
      }
    }
  }
  // batch-dependent scratch, so that no forward up to these sizes ever allocates (b2p_stage_reserve raises them)
  B2P_TRY(reserve_rows(st, is_cifar(st->d.kind) ? 1024 : st->d.max_tokens));
  B2P_CUDA(cudaDeviceSynchronize());
  st->finalized = true;
  return 0;
}

int b2p_stage_reserve(b2p_stage* st, int64_t batch, int64_t seq) {
  if (!st) return fail(B2P_EINVAL, "null stage");
  if (!st->finalized) return fail(B2P_ESTATE, "stage_reserve called before stage_finalize");
  if (batch <= 0) return fail(B2P_EINVAL, "batch must be positive (got %lld)", (long long)batch);
  int64_t rows = batch;
  if (is_gpt(st->d.kind)) {
    if (seq <= 0) return fail(B2P_EINVAL, "seq must be positive for a GPT stage");
    rows = batch * seq;
    if (rows > st->d.max_tokens)
      return fail(B2P_ELIMIT, "batch*seq=%lld exceeds the stage's max_tokens=%d (fixed at creation)", (long long)rows,
                  st->d.max_tokens);
  }
  if (rows > (1LL << 30)) return fail(B2P_ELIMIT, "batch too large");
  DeviceGuard g(st->d.device);
  return reserve_rows(st, rows);
}

int b2p_stage_set_aux_stream(b2p_stage* st, void* stream) {
  if (!st) return fail(B2P_EINVAL, "null stage");
  if (st->d.kind != B2P_CIFAR_PART1) return fail(B2P_EINVAL, "set_aux_stream: only the CIFAR part-1 stage has a tail kernel to move");
  DeviceGuard g(st->d.device);
  if (stream && !st->ev_fc1) {
    B2P_CUDA(cudaEventCreateWithFlags(&st->ev_fc1, cudaEventDisableTiming));
    B2P_CUDA(cudaEventCreateWithFlags(&st->ev_tail[0], cudaEventDisableTiming));
    B2P_CUDA(cudaEventCreateWithFlags(&st->ev_tail[1], cudaEventDisableTiming));
  }
  st->aux = static_cast<cudaStream_t>(stream);
  return 0;
}

int b2p_stage_join(b2p_stage* st, void* stream) {
  if (!st) return fail(B2P_EINVAL, "null stage");
  DeviceGuard g(st->d.device);
  for (int b = 0; b < 2; ++b)
    if (st->ev_tail_recorded[b]) B2P_CUDA(cudaStreamWaitEvent(static_cast<cudaStream_t>(stream), st->ev_tail[b], 0));
  return 0;
}

int b2p_stage_set_pdl(b2p_stage* st, int mask) {
  if (!st) return fail(B2P_EINVAL, "null stage");
  const int prev = st->pdl_mask;
  st->pdl_mask = mask < 0 ? -1 : (mask & 7);
  return prev < 0 ? pdl_switch().load() : prev;
}

int b2p_stage_out_elems(b2p_stage* st, int64_t batch, int64_t seq, int64_t* elems) {
  if (!st || !elems) return fail(B2P_EINVAL, "null argument");
  switch (st->d.kind) {
    case B2P_CIFAR_PART0: *elems = batch * 4096; break;
    case B2P_CIFAR_PART1: *elems = batch * 10; break;
    case B2P_GPT_FINAL: *elems = batch * seq * st->d.vocab_size; break;
    default: *elems = batch * seq * st->d.n_embd; break;
  }
  return 0;
}

int b2p_stage_forward(b2p_stage* st, const void* in_dev, void* out_dev, int64_t batch, int64_t seq, void* stream) {
  if (!st || !in_dev || !out_dev) return fail(B2P_EINVAL, "null argument");
  if (!st->finalized) return fail(B2P_ESTATE, "stage_forward called before stage_finalize");
  if (batch <= 0) return fail(B2P_EINVAL, "batch must be positive (got %lld)", (long long)batch);
  if (batch > (1LL << 30)) return fail(B2P_ELIMIT, "batch too large");
  DeviceGuard g(st->d.device);
  PdlScope pdl(st->pdl_mask);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  switch (st->d.kind) {
    case B2P_CIFAR_PART0:
      return cifar_part0_forward(st, static_cast<const float*>(in_dev), static_cast<float*>(out_dev), batch, s);
    case B2P_CIFAR_PART1:
      return cifar_part1_forward(st, static_cast<const float*>(in_dev), static_cast<float*>(out_dev), batch, s);
    default:
      return gpt_forward(st, in_dev, out_dev, batch, seq, s, nullptr);
  }
}

int b2p_split_tf32(const float* x_dev, float* hi_dev, float* lo_dev, int64_t n, void* stream) {
  if (!x_dev || !hi_dev || !lo_dev || n < 0) return fail(B2P_EINVAL, "bad argument");
  if (n == 0) return 0;
  B2P_LAUNCH(split_tf32, elementwise_grid((size_t)n), 256, 0, static_cast<cudaStream_t>(stream), x_dev, hi_dev, lo_dev,
             (size_t)n);
  return 0;
}

int b2p_linear(int operand_kind, const void* a_dev, const void* a_lo_dev, int64_t lda, const void* w_dev,
               const void* w_lo_dev, int64_t ldw, const float* bias_dev, const float* res_dev, int64_t ld_res,
               void* out_dev, int64_t ldo, int out_bf16, int act, int64_t M, int64_t N, int64_t K, void* stream) {
  if (!a_dev || !w_dev || !out_dev) return fail(B2P_EINVAL, "null argument");
  if (M > INT32_MAX || N > INT32_MAX || K > INT32_MAX) return fail(B2P_ELIMIT, "dimension too large");
  if (act < 0 || act > 2) return fail(B2P_EINVAL, "unknown activation %d", act);
  tc::EpiParams ep{bias_dev, res_dev, out_dev, ldo, ld_res, out_bf16, act, 0, nullptr, nullptr, 0};
  const int tile_n = (operand_kind >> 8) & 0xfff;       // test hook: force the tile width (0 = pick)
  int rc = tc::linear_dispatch(operand_kind & 0xff, a_dev, a_lo_dev, lda, w_dev, w_lo_dev, ldw, ep, (int)M, (int)N,
                                  (int)K, 1, tile_n, static_cast<cudaStream_t>(stream));
  return rc < 0 ? rc : 0;
}

int b2p_stage_trace(b2p_stage* st, long long* host_out, int n) {
  // developer aid: clock64 timeline of CTA 0 of the CIFAR part-0 tensor-core kernel
  if (!st || !host_out || n < 256) return fail(B2P_EINVAL, "trace buffer must hold at least 256 entries");
  DeviceGuard g(st->d.device);
  if (!st->trace) {
    B2P_CUDA(cudaMalloc((void**)&st->trace, 1024 * sizeof(long long)));
    B2P_CUDA(cudaMemset(st->trace, 0, 1024 * sizeof(long long)));
    memset(host_out, 0, 256 * sizeof(long long));
    return 0;
  }
  B2P_CUDA(cudaDeviceSynchronize());
  B2P_CUDA(cudaMemcpy(host_out, st->trace, (size_t)std::min(n, 1024) * sizeof(long long), cudaMemcpyDeviceToHost));
  return 0;
}

int b2p_debug_attn_trace(long long* host_out, int n) {
  if (!host_out) return fail(B2P_EINVAL, "null argument");
  return attn::read_attn_trace(host_out, n);
}

int b2p_debug_fc_trace(long long* host_out, int n) {
  if (!host_out) return fail(B2P_EINVAL, "null argument");
  return fcs::read_fc_trace(host_out, n);
}

int b2p_stage_check(b2p_stage* st, void* stream) {
  if (!st) return fail(B2P_EINVAL, "null stage");
  if (!st->bad_flag) return 0;
  DeviceGuard g(st->d.device);
  int flag = 0;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  B2P_CUDA(cudaMemcpyAsync(&flag, st->bad_flag, sizeof(int), cudaMemcpyDeviceToHost, s));
  B2P_CUDA(cudaStreamSynchronize(s));
  if (flag && st->d.kind == B2P_CIFAR_PART0) {
    B2P_CUDA(cudaMemsetAsync(st->bad_flag, 0, sizeof(int), s));
    return fail(B2P_ELIMIT, "an input or conv1 output reached the fp16 range limit (|v| >= 65504) of the tensor-core math "
                            "modes: the affected samples are not finite where the reference's are; use math=fp32 for such data");
  }
  if (flag && st->d.kind == B2P_CIFAR_PART1) {
    B2P_CUDA(cudaMemsetAsync(st->bad_flag, 0, sizeof(int), s));
    return fail(B2P_ELIMIT, "an activation or fc1 weight reached the fp16 range limit (|v| >= 65504) of the tensor-core math "
                            "mode: the affected rows are not finite where the reference's are; use math=fp32 for such data");
  }
  if (flag) {
    B2P_CUDA(cudaMemsetAsync(st->bad_flag, 0, sizeof(int), s));
    return fail(B2P_EINVAL, "index out of range in self: a token id was outside [0, vocab_size=%d)", st->d.vocab_size);
  }
  return 0;
}

int b2p_stage_kv_reset(b2p_stage* st, int64_t batch) {
  if (!st) return fail(B2P_EINVAL, "null stage");
  if (st->d.kind < B2P_GPT_FIRST) return fail(B2P_EINVAL, "kv_reset: only GPT stages keep a key/value cache");
  if (!st->finalized) return fail(B2P_ESTATE, "kv_reset called before stage_finalize");
  if (batch <= 0 || batch > 65535) return fail(B2P_EINVAL, "kv_reset: batch must be in 1..65535");
  DeviceGuard g(st->d.device);
  if (batch > st->kv_batch) {
    if (st->kv_k) B2P_CUDA(cudaFree(st->kv_k));
    if (st->kv_v) B2P_CUDA(cudaFree(st->kv_v));
    st->kv_k = st->kv_v = nullptr;
    st->kv_batch = 0;
    const size_t n = std::max<size_t>(1, st->blocks.size()) * (size_t)batch * st->d.n_head * st->d.block_size * 64;
    B2P_CUDA(cudaMalloc((void**)&st->kv_k, n * sizeof(float)));
    B2P_CUDA(cudaMalloc((void**)&st->kv_v, n * sizeof(float)));
    st->kv_batch = batch;
  }
  return 0;      // nothing to clear: a prefill at position 0 overwrites what the decode steps will read
}

int b2p_stage_forward_kv(b2p_stage* st, const void* in_dev, void* out_dev, int64_t* ids_dev, int64_t batch, int64_t seq,
                         int64_t pos0, void* stream) {
  if (!st || !in_dev) return fail(B2P_EINVAL, "null argument");
  if (st->d.kind < B2P_GPT_FIRST) return fail(B2P_EINVAL, "forward_kv: only GPT stages keep a key/value cache");
  if (!st->finalized) return fail(B2P_ESTATE, "stage_forward called before stage_finalize");
  if (pos0 < 0) return fail(B2P_EINVAL, "forward_kv: pos0 must be >= 0");
  if (st->d.kind == B2P_GPT_FINAL ? (!out_dev && !ids_dev) : !out_dev) return fail(B2P_EINVAL, "forward_kv: no output");
  if (ids_dev && st->d.kind != B2P_GPT_FINAL) return fail(B2P_EINVAL, "forward_kv: greedy ids come from the final stage");
  DeviceGuard g(st->d.device);
  PdlScope pdl(st->pdl_mask);
  return gpt_forward(st, in_dev, out_dev, batch, seq, static_cast<cudaStream_t>(stream), ids_dev, (int)pos0);
}

int b2p_stage_forward_kv_at(b2p_stage* st, const void* in_dev, void* out_dev, int64_t* ids_dev, int64_t batch,
                            const int32_t* pos_dev, void* stream) {
  if (!st || !in_dev || !pos_dev) return fail(B2P_EINVAL, "null argument");
  if (st->d.kind < B2P_GPT_FIRST) return fail(B2P_EINVAL, "forward_kv_at: only GPT stages keep a key/value cache");
  if (!st->finalized) return fail(B2P_ESTATE, "stage_forward called before stage_finalize");
  if (st->d.kind == B2P_GPT_FINAL ? (!out_dev && !ids_dev) : !out_dev) return fail(B2P_EINVAL, "forward_kv_at: no output");
  if (ids_dev && st->d.kind != B2P_GPT_FINAL) return fail(B2P_EINVAL, "forward_kv_at: greedy ids come from the final stage");
  DeviceGuard g(st->d.device);
  PdlScope pdl(st->pdl_mask);
  // pos0 = 1 only selects the decode path; the kernels read the position from pos_dev
  return gpt_forward(st, in_dev, out_dev, batch, 1, static_cast<cudaStream_t>(stream), ids_dev, 1, reinterpret_cast<const int*>(pos_dev));
}

int b2p_stage_forward_argmax(b2p_stage* st, const void* in_dev, int64_t* ids_dev, int64_t batch, int64_t seq,
                             void* stream) {
  if (!st || !in_dev || !ids_dev) return fail(B2P_EINVAL, "null argument");
  if (st->d.kind != B2P_GPT_FINAL) return fail(B2P_EINVAL, "forward_argmax is only defined for the GPT final stage");
  if (!st->finalized) return fail(B2P_ESTATE, "stage_forward_argmax called before stage_finalize");
  DeviceGuard g(st->d.device);
  PdlScope pdl(st->pdl_mask);
  return gpt_forward(st, in_dev, nullptr, batch, seq, static_cast<cudaStream_t>(stream), ids_dev);
}

}  // extern "C"

#include "handoff.cuh"

// =====================================================================================
// Stage forward with the handoff-ring handshake folded into the stage's own kernels.
// Same meaning as  recv -> forward -> release  (ring_in)  and  acquire -> forward -> publish
// (ring_out); for the CIFAR tensor-core kernels the four one-thread kernels disappear
// (ring_hooks.cuh), every other stage composes the existing calls.
// =====================================================================================
namespace {
int64_t stage_out_bytes(b2p_stage* st, int64_t batch, int64_t seq) {
  switch (st->d.kind) {
    case B2P_CIFAR_PART0: return batch * 4096 * 4;
    case B2P_CIFAR_PART1: return batch * 10 * 4;
    case B2P_GPT_FINAL: return batch * seq * (int64_t)st->d.vocab_size * 4;
    default: return batch * seq * (int64_t)st->d.n_embd * 4;
  }
}
}  // namespace

namespace {
b2p::RingSpan span_of(b2p_ring* r, uint64_t seq, int64_t rows_per_msg) {
  b2p::RingSpan sp{};
  sp.count = r->local_count;
  sp.published = r->flags;
  sp.released = r->flags + r->slots;
  sp.data = r->data;
  sp.slot_bytes = r->slot_bytes;
  sp.slots = r->slots;
  sp.first_slot = (int)(seq % r->slots);
  sp.rows_per_msg = (int)rows_per_msg;
  sp.tickets = r->ticket;
  sp.acq_seen = r->local_count + 2 * r->slots;
  return sp;
}
int64_t stage_in_bytes(b2p_stage* st, int64_t batch, int64_t seq) {
  switch (st->d.kind) {
    case B2P_CIFAR_PART0: return batch * 3 * 32 * 32 * 4;
    case B2P_CIFAR_PART1: return batch * 4096 * 4;
    case B2P_GPT_FIRST: return batch * seq * 8;                       // int64 token ids
    default: return batch * seq * (int64_t)st->d.n_embd * 4;
  }
}
}  // namespace

extern "C" int b2p_stage_forward_ring_span(b2p_stage* st, const void* in_dev, void* out_dev, int64_t batch, int64_t seq_len,
                                           b2p_ring* ring_in, b2p_ring* ring_out, uint64_t seq, int64_t rows_per_msg,
                                           void* stream) {
  using namespace b2p;
  if (!st) return fail(B2P_EINVAL, "null stage");
  if (!ring_in && !in_dev) return fail(B2P_EINVAL, "forward_ring: no input (neither a ring nor a device pointer)");
  if (!ring_out && !out_dev) return fail(B2P_EINVAL, "forward_ring: no output (neither a ring nor a device pointer)");
  if (!st->finalized) return fail(B2P_ESTATE, "stage_forward called before stage_finalize");
  if (batch <= 0) return fail(B2P_EINVAL, "batch must be positive (got %lld)", (long long)batch);
  if (rows_per_msg <= 0) return fail(B2P_EINVAL, "forward_ring: rows_per_msg must be positive");
  if (ring_in && !ring_in->owner) return fail(B2P_ESTATE, "forward_ring: ring_in must be the consumer side of a ring");
  if (ring_out && ring_out->owner) return fail(B2P_ESTATE, "forward_ring: ring_out must be the producer side of a ring");
  const int64_t msgs = (batch + rows_per_msg - 1) / rows_per_msg;
  const int64_t msg_rows = std::min(batch, rows_per_msg);
  if (ring_out && (uint64_t)stage_out_bytes(st, msg_rows, seq_len) > ring_out->slot_bytes)
    return fail(B2P_ELIMIT, "forward_ring: output of %lld bytes exceeds the slot size %zu",
                (long long)stage_out_bytes(st, msg_rows, seq_len), ring_out->slot_bytes);
  if ((ring_in && msgs > ring_in->slots) || (ring_out && msgs > ring_out->slots))
    return fail(B2P_ELIMIT, "forward_ring: %lld messages in one call, the ring has fewer slots", (long long)msgs);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const bool hooked_out = ring_out && st->d.kind == B2P_CIFAR_PART0 && st->d.math != B2P_MATH_FP32 && !ring_in;
  const bool hooked_in = ring_in && st->d.kind == B2P_CIFAR_PART1 && st->d.math == B2P_MATH_TF32X3 && fc_fused_enabled() &&
                         !ring_out && (rows_per_msg == fcs::BM || (msgs == 1 && batch <= fcs::BM)) &&
                         ring_in->slot_bytes == (size_t)fcs::BM * fcs::K_IN * 4;
  PdlScope pdl(st->pdl_mask);
  if (hooked_out) {
    DeviceGuard g(st->d.device);
    return cifar_part0_forward(st, static_cast<const float*>(in_dev), static_cast<float*>(out_dev), batch, s,
                               span_of(ring_out, seq, rows_per_msg));
  }
  if (hooked_in) {
    DeviceGuard g(st->d.device);
    return cifar_part1_forward(st, nullptr, static_cast<float*>(out_dev), batch, s, span_of(ring_in, seq, fcs::BM));
  }
  // every other stage: one message at a time through the one-thread handshake kernels
  for (int64_t m = 0; m < msgs; ++m) {
    const int64_t rows = std::min(rows_per_msg, batch - m * rows_per_msg);
    void* slot = nullptr;
    const void* in = nullptr;
    void* out = nullptr;
    if (ring_in) {
      B2P_TRY(b2p_ring_recv(ring_in, seq + m, &slot, stream));
      in = slot;
    } else {
      in = static_cast<const unsigned char*>(in_dev) + m * stage_in_bytes(st, rows_per_msg, seq_len);
    }
    if (ring_out) {
      B2P_TRY(b2p_ring_acquire(ring_out, seq + m, &slot, stream));
      out = slot;
    } else {
      out = static_cast<unsigned char*>(out_dev) + m * stage_out_bytes(st, rows_per_msg, seq_len);
    }
    B2P_TRY(b2p_stage_forward(st, in, out, rows, seq_len, stream));
    if (ring_in) B2P_TRY(b2p_ring_release(ring_in, seq + m, stream));
    if (ring_out) B2P_TRY(b2p_ring_publish(ring_out, seq + m, stream));
  }
  return 0;
}

extern "C" int b2p_stage_forward_ring(b2p_stage* st, const void* in_dev, void* out_dev, int64_t batch, int64_t seq_len,
                                      b2p_ring* ring_in, b2p_ring* ring_out, uint64_t seq, void* stream) {
  return b2p_stage_forward_ring_span(st, in_dev, out_dev, batch, seq_len, ring_in, ring_out, seq, batch > 0 ? batch : 1, stream);
}

// =====================================================================================
// Allocation-free variant of the CIFAR part-1 forward (fc1 split-K + tail).
//
// b2p_stage_forward() keeps the fc1 split-K partial sums in a buffer it obtains with
// cudaMalloc the first time a batch of a given size shows up.  Some hosts cannot let a
// library do that: a serving thread with a latency budget must never run into the
// device-wide synchronisation cudaMalloc/cudaFree imply, and a process whose device
// memory was handed out in full to its own pool at start-up has nothing left for a
// library to allocate behind its back (the call would fail, or break the host's
// accounting).  For them the partial sums live in the fixed array below, which is part of
// the library image: every GPU that loads libb200pipe.so holds these 4 MiB until the
// process exits, whether or not this entry point is ever called.
//
// Sharing rules.  There is ONE array per GPU and every call on that GPU uses it, keeping
// whatever the previous call left there.  Nothing in it has to start at zero: a call with
// batch M first overwrites every partial sum of rows 0..M-1 (all 8 K-slices, all 512
// columns) and only then reads exactly those, so stale rows are never observed.  Two
// calls whose kernels can overlap on the GPU (different streams, no ordering between
// them) race on the array and corrupt each other's result: order them on one stream or
// with events.
// =====================================================================================
static __device__ __align__(256) float g_noalloc_fc_scratch[FC1_SPLITS * B2P_NOALLOC_MAX_BATCH * FC1_N];

namespace {
// In host code the array's name is only a stand-in for the symbol, not a device address.
// The address valid on a given GPU is obtained once, with that GPU current, and cached.
constexpr int kMaxNoallocDevices = 64;
float* g_noalloc_ptr[kMaxNoallocDevices] = {};
std::atomic<bool> g_noalloc_ready[kMaxNoallocDevices];

int noalloc_scratch_for_current_device(float** out) {
  int dev = -1;
  B2P_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxNoallocDevices) return fail(B2P_ELIMIT, "device ordinal %d not supported", dev);
  if (!g_noalloc_ready[dev].load(std::memory_order_acquire)) {
    float* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noalloc_fc_scratch);
    if (e != cudaSuccess || p == nullptr)
      return fail(B2P_ECUDA, "cudaGetSymbolAddress(g_noalloc_fc_scratch) on device %d failed: %s", dev,
                  cudaGetErrorString(e));
    g_noalloc_ptr[dev] = p;     // racing first calls compute the same address
    g_noalloc_ready[dev].store(true, std::memory_order_release);
  }
  *out = g_noalloc_ptr[dev];
  return 0;
}
}  // namespace

extern "C" int b2p_cifar_part1_forward_noalloc(const float* x_dev, const float* fc1_weight_dev,
                                               const float* fc1_bias_dev, const float* fc2_weight_dev,
                                               const float* fc2_bias_dev, float* out_dev, int64_t batch,
                                               void* stream) {
  if (!x_dev || !fc1_weight_dev || !fc1_bias_dev || !fc2_weight_dev || !fc2_bias_dev || !out_dev)
    return fail(B2P_EINVAL, "null argument");
  if (batch <= 0) return fail(B2P_EINVAL, "batch must be positive (got %lld)", (long long)batch);
  if ((reinterpret_cast<uintptr_t>(x_dev) | reinterpret_cast<uintptr_t>(fc1_weight_dev) |
       reinterpret_cast<uintptr_t>(fc1_bias_dev) | reinterpret_cast<uintptr_t>(fc2_weight_dev)) & 15)
    return fail(B2P_EINVAL, "x, fc1_weight, fc1_bias and fc2_weight must be 16-byte aligned (they are read with 128-bit loads)");
  if (batch > B2P_NOALLOC_MAX_BATCH)
    return fail(B2P_ELIMIT,
                "batch %lld exceeds the fixed scratch of the allocation-free entry point (at most %d rows per "
                "call); split the batch or use b2p_stage_forward",
                (long long)batch, B2P_NOALLOC_MAX_BATCH);
  float* scratch = nullptr;
  B2P_TRY(noalloc_scratch_for_current_device(&scratch));   // first call on a GPU: one-time set-up
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  dim3 grid(FC1_N / 64, ceil_div(batch, 64), FC1_SPLITS);
  B2P_LAUNCH(fc1_splitk_f32, grid, 256, 0, s, x_dev, fc1_weight_dev, scratch, (int)batch, FC1_K / FC1_SPLITS);
  B2P_LAUNCH(fc_tail_f32<FC1_SPLITS>, (int)batch, 128, 0, s, scratch, (int)batch, fc1_bias_dev, fc2_weight_dev,
             fc2_bias_dev, out_dev, b2p::RingSpan{});
  return 0;
}
