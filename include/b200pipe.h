/* b200pipe.h -- C ABI of the B200 pipeline-stage engine (libb200pipe.so).
 *
 * The reference (123-code/Distributed-neural-networks) is pure Python and has no
 * FFI of its own; its seams are Python-level.  This ABI
 * sits *under* seam 1, the stage-operator seam: everything the reference does
 * with a stage object --
 *     part = MODEL_PARTS_CLASSES[part_index](full_model).to(DEVICE)   node.py:300-305
 *     part.load_state_dict(full_state_dict, strict=False)             node.py:306
 *     part.eval()                                                     node.py:317
 *     with torch.no_grad(): y = part(x)                               node.py:52-53,159-160
 * -- maps onto stage_create / stage_load_weight / stage_finalize / stage_forward
 * below, and the activation hop  output_np.tobytes() -> TensorRequest ->
 * stub.SendTensor -> np.frombuffer().copy()  (node.py:73-89, 170-181) maps onto
 * the b2p_ring_* peer-handoff calls.
 *
 * Conventions
 *   - plain C types only; no torch, no C++ types, no exceptions cross the ABI;
 *   - every call returns 0 on success or a negative B2P_E* code; the message is
 *     available from b2p_last_error() (thread-local, valid until the thread's
 *     next failing call);
 *   - the caller owns all input/output DEVICE buffers; the library owns weights,
 *     workspaces and handoff rings;
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream);
 *     calls only enqueue work, they never synchronise unless stated.  The forward calls
 *     allocate nothing as long as the batch stays within what b2p_stage_finalize /
 *     b2p_stage_reserve sized (see b2p_stage_reserve); a larger batch grows the scratch
 *     on the spot, which synchronises the device and is refused (B2P_ESTATE) while the
 *     stream is being captured;
 *   - there is NO CPU fallback: without a CUDA device every compute call fails
 *     with B2P_ECUDA.
 */
#ifndef B200PIPE_H_
#define B200PIPE_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2P_ABI_VERSION 1

#if defined(__GNUC__)
#define B2P_API __attribute__((visibility("default")))
#else
#define B2P_API
#endif

/* error codes */
#define B2P_OK        0
#define B2P_EINVAL   -1   /* bad argument (shape, kind, null pointer ...)        */
#define B2P_ECUDA    -2   /* a CUDA runtime/driver call failed                   */
#define B2P_ESTATE   -3   /* call out of order (e.g. forward before finalize)    */
#define B2P_EMISSING -4   /* finalize: a required weight was never loaded        */
#define B2P_ELIMIT   -5   /* size beyond a documented maximum                    */
#define B2P_ETIMEOUT -6   /* a device-side ring wait gave up (b2p_ring_check)    */

/* stage kinds: one per reference part class */
#define B2P_CIFAR_PART0   0   /* ModelPart0_2Node   cifar_model_parts.py:29-42   */
#define B2P_CIFAR_PART1   1   /* ModelPart1_2Node   cifar_model_parts.py:45-58   */
#define B2P_GPT_FIRST     2   /* ModelPart0          partitions/gpt_model_parts.py:6-22  */
#define B2P_GPT_MIDDLE    3   /* ModelPartIntermediate          ...:26-34        */
#define B2P_GPT_FINAL     4   /* ModelPartFinal_GPT             ...:36-50        */

/* arithmetic modes (what the contractions compute in) */
#define B2P_MATH_FP32     0   /* CUDA-core fp32 FMA: reference-grade parity mode           */
#define B2P_MATH_TF32X3   1   /* tcgen05 kind::tf32, 3-term split ~ fp32: tensor parity    */
#define B2P_MATH_TF32     2   /* tcgen05 kind::tf32, single pass                           */
#define B2P_MATH_BF16     3   /* tcgen05 kind::f16 (bf16 in, fp32 accumulate): throughput  */

/* element types of host weight tensors handed to stage_load_weight */
#define B2P_DT_F32        0
#define B2P_DT_BF16       1
#define B2P_DT_I64        2

typedef struct b2p_stage b2p_stage;   /* opaque */
typedef struct b2p_ring  b2p_ring;    /* opaque */

/* Stage geometry.  CIFAR parts ignore every field except `math`.
 * GPT parts: local blocks are the reference's inclusive [start_layer,end_layer]
 * (gpt_model_parts.py:7,12) -- n_local = end-start+1 may be 0. */
typedef struct b2p_stage_desc {
  int32_t kind;          /* B2P_CIFAR_* / B2P_GPT_*                               */
  int32_t math;          /* B2P_MATH_*                                            */
  int32_t device;        /* CUDA device ordinal that will own the weights         */
  int32_t start_layer;   /* GPT: first global block index of this stage           */
  int32_t end_layer;     /* GPT: last global block index (inclusive)              */
  int32_t n_head;        /* GPT                                                   */
  int32_t n_embd;        /* GPT: C, multiple of 128 (at most 2048), head dim 64   */
  int32_t vocab_size;    /* GPT first/final                                       */
  int32_t block_size;    /* GPT: max T (gpt_model_parts.py:15 assert)             */
  int32_t max_tokens;    /* GPT: largest B*T one forward() will carry (workspace) */
  int32_t reserved[6];
} b2p_stage_desc;

/* ---- library ---------------------------------------------------------------- */
B2P_API int         b2p_abi_version(void);
B2P_API const char* b2p_last_error(void);
B2P_API int         b2p_device_count(int* count);
/* number of kernels this library has launched in this process (bench's gpu_launches) */
B2P_API int64_t     b2p_launch_count(void);

/* Programmatic dependent launch (overlap each kernel's prologue with its predecessor's tail).
 * mask bit 0: the light kernel chains (GEMMs, LayerNorm, attention, fc tail); bit 1: the CIFAR
 * part-0 kernel, whose waiting CTAs each hold a whole SM -- clear it when two stages share one
 * GPU on two streams.  Default 3.  Returns the previous mask.
 * b2p_set_pdl sets the process-wide DEFAULT; b2p_stage_set_pdl sets it for one stage (mask < 0:
 * back to the default) and is what a scheduler should use -- two pipelines, or two threads, in
 * one process then never see each other's setting.
 * b2p_stage_set_pdl only, bit 2 (value 4), CIFAR part 0 writing into a handoff ring
 * (b2p_stage_forward_ring_span): the kernel does not wait for the kernel that precedes it in the
 * stream at all.  Consecutive steps write different ring slots, so step k+1's first samples are
 * computed while step k's last messages are still being published (5-8 us of fences per step
 * otherwise stand between two steps: 18 -> 26 M samples/s on two GPUs).  The caller promises that
 * the stage's INPUT is not written by that preceding kernel (a resident buffer, or a copy ordered
 * by an event -- not a kernel of the same stream); the library applies the bit only when the
 * launch uses at most half of the ring's slots (this step and the next never share a slot). */
B2P_API int b2p_set_pdl(int mask);

/* ---- stage lifecycle (seam 1; node.py:300-317) ------------------------------- */
B2P_API int b2p_stage_create(b2p_stage** out, const b2p_stage_desc* desc);
B2P_API int b2p_stage_destroy(b2p_stage* st);

/* Upload one named tensor of the FULL-model state dict from host memory.
 * Names are the reference's .pth keys: "conv1.weight", "fc2.bias", ... for CIFAR
 * (cifar_model_parts.py:10-14); nanoGPT names "transformer.wte.weight",
 * "transformer.h.<global i>.attn.c_attn.weight", "lm_head.weight" ... for GPT.
 * Returns 1 if the stage consumed the tensor, 0 if the name is not one of this
 * stage's weights (the strict=False "unexpected key" case, node.py:306), <0 on
 * error (e.g. shape mismatch).  Synchronous w.r.t. the host buffer. */
B2P_API int b2p_stage_load_weight(b2p_stage* st, const char* name, const void* host_data,
                          int32_t dtype, const int64_t* shape, int32_t ndim);

/* Number of weights still missing, and their names (strict=False "missing keys"). */
B2P_API int b2p_stage_missing(b2p_stage* st, char* names_buf, size_t buf_len);

/* Pack weights for the selected math mode (bf16 copies, tf32 hi/lo planes,
 * transposes).  Must be called once after loading, before forward. */
B2P_API int b2p_stage_finalize(b2p_stage* st);

/* Size the stage's batch-dependent scratch (fc1 split-K partial sums, hi/lo planes of the
 * 3xTF32 GEMM operands, arg-max partials of the fused LM head) for forwards of up to `batch`
 * rows (GPT: batch*seq tokens, at most max_tokens).  finalize already does this for 1024 rows
 * (CIFAR) / max_tokens (GPT); call reserve when larger batches will come, at start-up, NOT on
 * the serving path and not inside a stream capture: it may cudaFree/cudaMalloc (device-wide
 * synchronisation).  After it, no forward of that size or smaller allocates. */
B2P_API int b2p_stage_reserve(b2p_stage* st, int64_t batch, int64_t seq);
B2P_API int b2p_stage_set_pdl(b2p_stage* st, int mask);

/* CIFAR part 1 only, no reference counterpart (the reference has one forward per request, node.py:52-53): give the
 * stage a second stream for its tail kernel (bias + ReLU + fc2 + softmax over fc1's partial sums).  forward then
 * launches fc1 on the caller's stream and the tail on `aux_stream` behind an event, alternating between two
 * partial-sum buffers, so fc1 of the next forward does not wait for this forward's tail: on the consumer GPU of a
 * 2-stage pipeline, where fc1 spends its time waiting for the producer's messages, the tail of step s hides behind
 * fc1 of step s+1.  The OUTPUT is then complete on `aux_stream`, not on the caller's stream: call b2p_stage_join
 * before anything on another stream reads it.  Ignored while the caller's stream is being captured.  NULL turns it off. */
B2P_API int b2p_stage_set_aux_stream(b2p_stage* st, void* aux_stream);
/* Make `stream` wait for every tail kernel this stage has enqueued on its aux stream so far. */
B2P_API int b2p_stage_join(b2p_stage* st, void* stream);

/* Output element count for a given input geometry (host helper). */
B2P_API int b2p_stage_out_elems(b2p_stage* st, int64_t batch, int64_t seq, int64_t* elems);

/* ---- the hot path: one stage forward (node.py:52-53) -------------------------
 *   CIFAR part0 : in  fp32 [batch,3,32,32] NCHW   -> out fp32 [batch,4096]
 *   CIFAR part1 : in  fp32 [batch,4096]            -> out fp32 [batch,10] (probabilities)
 *   GPT first   : in  int64 [batch,seq]            -> out fp32 [batch,seq,C]
 *   GPT middle  : in  fp32 [batch,seq,C]           -> out fp32 [batch,seq,C]
 *   GPT final   : in  fp32 [batch,seq,C]           -> out fp32 [batch,seq,V] logits
 * `seq` is ignored by the CIFAR kinds.  in/out are device pointers on the
 * stage's device; `in` is not modified; in and out must not overlap. */
B2P_API int b2p_stage_forward(b2p_stage* st, const void* in_dev, void* out_dev,
                      int64_t batch, int64_t seq, void* stream);

/* Synchronising health check: waits for `stream`, then reports (and clears) deferred
 * device-side input errors of earlier forwards:
 *   - a token id outside [0,vocab_size) seen by a GPT first stage (an IndexError in the
 *     reference's nn.Embedding, gpt_model_parts.py:17): B2P_EINVAL;
 *   - CIFAR part 0 in a tensor-core math mode: those modes carry inputs and conv1 outputs
 *     as fp16 hi + lo parts, so a value with |v| >= 65504 comes out non-finite where the
 *     reference (cifar_model_parts.py:37-42, fp32) stays finite.  The guard is on the
 *     input: B2P_ELIMIT when some |x| >= (65504 - max|conv1.bias|) / max_oc sum|conv1.weight[oc]|,
 *     the magnitude from which on a conv1 output CAN reach the limit (conservative: it never
 *     misses an overflow, it may flag a batch whose outputs happened to stay below).
 *     math = B2P_MATH_FP32 has the reference's range.  (Un-normalised 0..255 images with
 *     ordinary weights are far inside the range; NaN inputs propagate to NaN outputs as in
 *     torch and are not flagged; +-inf inputs give NaN.)
 *   - CIFAR part 1 in B2P_MATH_TF32X3: fc1 (cifar_model_parts.py:55) runs in the same split-fp16
 *     scheme, so an activation or fc1 weight with |v| >= 65504 is flagged: B2P_ELIMIT (the
 *     reference, fp32, stays finite there; math = B2P_MATH_FP32 has its range).
 * Returns 0 when clean. */
B2P_API int b2p_stage_check(b2p_stage* st, void* stream);

/* Developer aid (no reference counterpart): the first call arms a per-stage trace buffer,
 * later calls synchronise and copy out min(n, 1024) stamps of the CIFAR part-0 tensor-core
 * kernel: clock64 of CTA 0 as [role 0..3][sample 0..7][event 0..7] (entry 192 = end of set-up),
 * from entry 256 on the %globaltimer entry / release / exit time of every CTA (3 entries each),
 * 710..712 CTA 0's clock64 at entry / release / exit, 719.. a ring of the last 16 launches'
 * release and exit times, 760.. who published which ring message and what it cost
 * (scripts/trace_part0.py, scripts/probe_prod2.py decode it).  n >= 256. */
B2P_API int b2p_stage_trace(b2p_stage* st, long long* host_out, int n);

/* Developer aid (no reference counterpart): clock64 timeline of CTA (0,0,0) of the last tensor-core
 * attention launch on the current device; 193 entries [role 0..2][key tile 0..7][event 0..7] + start. */
B2P_API int b2p_debug_attn_trace(long long* host_out, int n);

/* Developer aid (no reference counterpart): with B2P_FC_TRACE=1 in the environment, clock64 timeline of the
 * 128 CTAs of row tile 0 of the last fused CIFAR part-1 launch; 2048 entries [hidden tile 0..3][K slice 0..31][event 0..15]. */
B2P_API int b2p_debug_fc_trace(long long* host_out, int n);

/* GPT final only: greedy token ids (row arg-max of the logits) without
 * materialising [B,T,V]; out int64 [batch,seq].  (north_star: "greedy token ids") */
B2P_API int b2p_stage_forward_argmax(b2p_stage* st, const void* in_dev, int64_t* ids_dev,
                             int64_t batch, int64_t seq, void* stream);

/* Greedy decoding with a key/value cache (no counterpart in the reference, whose parts only ever run one
 * full-sequence forward, gpt_model_parts.py:13-50).  kv_reset sizes the stage's cache for `batch` sequences of
 * block_size positions.  forward_kv processes `seq` NEW positions pos0 .. pos0+seq-1 of every sequence: pos0 == 0 is the
 * prefill (ordinary causal attention over the prompt, keys and values are stored); pos0 > 0 takes one position at a
 * time (seq == 1) and attends to the cache.  Inputs and outputs as b2p_stage_forward ([B,seq] ids or [B,seq,C] hidden
 * states in, hidden states or logits out); on the final stage ids_dev != NULL returns greedy ids [B,seq] instead of
 * logits.  Positions are absolute (learned position embeddings): pos0 + seq <= block_size, beyond that a caller has to
 * re-run the cropped context.  B2P_ESTATE without kv_reset, B2P_EINVAL for positions out of order or out of range. */
B2P_API int b2p_stage_kv_reset(b2p_stage* st, int64_t batch);
B2P_API int b2p_stage_forward_kv(b2p_stage* st, const void* in_dev, void* out_dev, int64_t* ids_dev, int64_t batch,
                                 int64_t seq, int64_t pos0, void* stream);
/* The decode step with the position in DEVICE memory (int32, the same for every sequence): one new position per
 * sequence at *pos_dev, read by the kernels when they run.  No launch then carries a step-dependent argument, so a whole
 * decode step through every stage (about 110 kernels for 12 blocks) can be captured ONCE in a CUDA graph and replayed
 * per token -- one launch per token from a host thread that has better things to do; the caller advances *pos_dev (and
 * feeds the new token back) inside the same graph.  (The step itself is bound by the fixed latency of its small kernels,
 * not by their launches: 1.03 ms per position either way for GPT-2 small, B = 8.)  Requires the prefill through b2p_stage_forward_kv(pos0 = 0).  A position at or beyond
 * block_size stores nothing and attends to the last cached position (the caller bounds the number of replays).
 * No counterpart in the reference (gpt_model_parts.py:13-50 runs one full-sequence forward). */
B2P_API int b2p_stage_forward_kv_at(b2p_stage* st, const void* in_dev, void* out_dev, int64_t* ids_dev, int64_t batch,
                            const int32_t* pos_dev, void* stream);

/* ---- dense contraction building block (tcgen05 tensor cores) ------------------------
 *   out[m,n] = act( sum_k A[m,k] * W[n,k] + bias[n] ) + res[m,n]
 * i.e. every nn.Linear call site on the path (cifar_model_parts.py:13-14,55-56; the
 * c_attn / c_proj / c_fc / lm_head linears of the block the reference imports at
 * partitions/gpt_model_parts.py:4).  A [M,K] and W [N,K] are row-major with K contiguous
 * (nn.Linear's own [out,in] weight layout), row pitches lda/ldw in elements.
 *   operand_kind 0: A and W are bf16;  1: A and W are fp32, multiplied as tf32;
 *                2: "3xTF32": a_dev/w_dev are the tf32-exact hi planes and a_lo_dev/w_lo_dev
 *                   the lo planes made by b2p_split_tf32 (lo*hi + hi*lo + hi*hi, fp32-grade).
 *   a_lo_dev / w_lo_dev are ignored (may be NULL) for kinds 0 and 1.
 *   bias [N] fp32 or NULL; res [M,ld_res] fp32 or NULL (added after the activation);
 *   act 0 none, 1 ReLU, 2 exact-erf GELU; out [M,ldo] fp32 (out_bf16=0) or bf16 (=1).
 * Accumulation is fp32 in tensor memory.  Pointers 16-byte aligned, lda/ldw multiples
 * of 16 bytes.  Used by the stages in the tensor-core math modes; exported because it
 * is also the unit the parity tests exercise directly. */
B2P_API int b2p_linear(int operand_kind, const void* a_dev, const void* a_lo_dev, int64_t lda,
                       const void* w_dev, const void* w_lo_dev, int64_t ldw, const float* bias_dev,
                       const float* res_dev, int64_t ld_res, void* out_dev, int64_t ldo, int out_bf16,
                       int act, int64_t M, int64_t N, int64_t K, void* stream);
/* x = hi + lo, both exactly representable in tf32 (|x-hi-lo| <= 2^-22 |x|); n elements. */
B2P_API int b2p_split_tf32(const float* x_dev, float* hi_dev, float* lo_dev, int64_t n, void* stream);

/* ---- allocation-free variant (no counterpart in the reference) ----------------
 * Same arithmetic as b2p_stage_forward on a B2P_CIFAR_PART1 stage (fc1+ReLU, fc2,
 * softmax; fp32), for hosts that cannot let a library allocate device memory: a
 * serving thread that must never meet the device-wide synchronisation cudaMalloc and
 * cudaFree imply, or a process that gave all device memory to its own pool at
 * start-up.  The call allocates nothing and frees nothing.  The caller passes every
 * buffer, weights included, as device pointers on the CURRENT device:
 *   x [batch,4096], fc1_weight [512,4096], fc1_bias [512], fc2_weight [10,512],
 *   fc2_bias [10]  ->  out [batch,10] (probabilities).
 * The fc1 split-K partial sums go to a fixed 4 MiB array inside the library image
 * (held by every GPU that loads the library, until exit), so
 *   - batch is limited to B2P_NOALLOC_MAX_BATCH rows; larger returns B2P_ELIMIT;
 *   - all calls on a GPU share that array: calls that may overlap on the GPU
 *     (unordered streams) race and must be ordered by the caller; stale contents
 *     from an earlier call are harmless (every value read is first rewritten);
 *   - the first call on each GPU looks the array's device address up once
 *     (cudaGetSymbolAddress); make that first call before starting any stream
 *     capture on that GPU;
 *   - "allocates nothing" holds from a GPU's SECOND call on.  With CUDA's default lazy
 *     module loading the very first use of this library on a GPU (this call, or any
 *     other) loads the library image there, which is when the driver sets the 4 MiB
 *     array aside -- an implicit allocation plus a synchronisation.  So make one call
 *     per GPU at start-up, off the latency-critical thread and BEFORE the host's own
 *     pool claims the remaining memory (or run with CUDA_MODULE_LOADING=EAGER);
 *     otherwise that first call can fail with B2P_ECUDA for lack of memory;
 *   - the cached address does not survive cudaDeviceReset(): a process that resets a
 *     device must not call this entry point on it again. */
#define B2P_NOALLOC_MAX_BATCH 256
B2P_API int b2p_cifar_part1_forward_noalloc(const float* x_dev, const float* fc1_weight_dev,
                                            const float* fc1_bias_dev, const float* fc2_weight_dev,
                                            const float* fc2_bias_dev, float* out_dev,
                                            int64_t batch, void* stream);

/* ---- activation handoff (replaces node.py:73-89 / :170-181) ------------------
 * A ring is a set of `slots` fixed-size receive buffers living on the CONSUMER's
 * GPU plus two 64-bit sequence flags per slot (full / free).  The consumer
 * process creates it, exports an opaque handle (CUDA IPC), and the producer
 * process opens it; inside one process ring_open_local() skips IPC.  Data moves
 * with one device-to-device copy over NVLink (or rides on the producing kernel's own
 * stores); signalling is done on the device, by one-thread kernels on the caller's
 * stream or inside the stage kernels, so neither side ever blocks the host.
 *
 * Device-side waits give up after a time-out (default 20 s) instead of hanging: the wait
 * records what it was waiting for, every later ring wait on that GPU returns at once, the
 * enqueued kernels drain with meaningless results, and b2p_ring_check reports
 * B2P_ETIMEOUT (and clears the record).  Nothing traps, so the context, the stages and
 * the rings stay usable.  b2p_ring_set_timeout(device, 0) waits for ever -- for a serving
 * process whose front door may sit idle between requests. */
#define B2P_RING_HANDLE_BYTES 256
B2P_API int b2p_ring_set_timeout(int device, double seconds);   /* per GPU; 0 = no time-out */
B2P_API int b2p_ring_check(int device);                         /* synchronous read of the GPU's time-out record */
B2P_API int b2p_ring_create(b2p_ring** out, int device, int slots, size_t slot_bytes);
B2P_API int b2p_ring_export(b2p_ring* ring, void* handle_out /* B2P_RING_HANDLE_BYTES */);
B2P_API int b2p_ring_open(b2p_ring** out, const void* handle, int producer_device);
B2P_API int b2p_ring_open_local(b2p_ring** out, b2p_ring* consumer_side, int producer_device);
B2P_API int b2p_ring_destroy(b2p_ring* ring);
/* How b2p_ring_send moves the payload: 0 = cudaMemcpyAsync (copy engine),
 * 1 = SM copy kernel with 128-bit peer stores (default). */
B2P_API int b2p_ring_set_mode(b2p_ring* ring, int mode);
/* Zero-copy variant for a producer whose kernel can write its output anywhere: acquire
 * waits (on the stream) until slot (seq % slots) is free and returns its address -- mapped
 * memory of the consumer's GPU; run the producing stage with out = that address, then
 * publish(seq).  The transfer then rides on the stage kernel's own stores over NVLink. */
B2P_API int b2p_ring_acquire(b2p_ring* ring, uint64_t seq, void** slot_dev, void* stream);
B2P_API int b2p_ring_publish(b2p_ring* ring, uint64_t seq, void* stream);
/* producer: wait until slot (seq % slots) is free, copy `bytes` from src_dev into
 * it and publish sequence number seq (seq counts from 0, strictly increasing). */
B2P_API int b2p_ring_send(b2p_ring* ring, const void* src_dev, size_t bytes, uint64_t seq, void* stream);
/* consumer: make `stream` wait until message seq has landed; returns its device address. */
B2P_API int b2p_ring_recv(b2p_ring* ring, uint64_t seq, void** slot_dev, void* stream);
/* consumer: mark message seq consumed (enqueue after the last kernel that reads the slot). */
B2P_API int b2p_ring_release(b2p_ring* ring, uint64_t seq, void* stream);

/* One node's whole SendTensor body (node.py:45-89: receive the tensor, run the part, forward the result) as ONE
 * call: the stage's forward with the ring handshake folded into its own kernels.
 *   ring_in  != NULL (consumer side): the input is message `seq` of ring_in (in_dev ignored); same meaning as
 *             b2p_ring_recv -> forward -> b2p_ring_release.
 *   ring_out != NULL (producer side): the output is stored straight into slot `seq` of ring_out (out_dev ignored);
 *             same meaning as b2p_ring_acquire -> forward -> b2p_ring_publish.
 * The CIFAR tensor-core kernels do the waiting and the flag stores themselves (csrc/ring_hooks.cuh), so the four
 * one-thread kernels and their launch latencies disappear; every other stage composes the calls above.  Messages of
 * one slot must be used in order, as with the separate calls.  Errors: B2P_EINVAL (missing input/output),
 * B2P_ESTATE (stage not finalized, ring used from the wrong side), B2P_ELIMIT (output larger than a slot). */
B2P_API int b2p_stage_forward_ring(b2p_stage* st, const void* in_dev, void* out_dev, int64_t batch, int64_t seq_len,
                                   b2p_ring* ring_in, b2p_ring* ring_out, uint64_t seq, void* stream);

/* A whole step in one call: rows [i*rows_per_msg, (i+1)*rows_per_msg) of the batch travel as message seq+i (at most
 * `slots` messages per call).  For the CIFAR tensor-core kernels this is ONE launch per stage per step: the producing
 * kernel publishes each message as soon as its last row is stored, the consuming kernel waits message by message, so
 * the two GPUs overlap at microbatch granularity (the micro-batched pipelining of the north star) without paying a
 * launch per microbatch.  The fused consumer needs rows_per_msg == 128 and 2 MiB slots; anything else, and every other
 * stage, runs message by message through the calls above with the same results. */
B2P_API int b2p_stage_forward_ring_span(b2p_stage* st, const void* in_dev, void* out_dev, int64_t batch, int64_t seq_len,
                                        b2p_ring* ring_in, b2p_ring* ring_out, uint64_t seq, int64_t rows_per_msg,
                                        void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200PIPE_H_ */
