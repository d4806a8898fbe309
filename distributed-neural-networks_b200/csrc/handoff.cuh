// handoff.cuh -- stage-to-stage activation handoff over NVLink peer memory.
//
// Replaces the reference's hop  output_np.tobytes() -> TensorRequest -> stub.SendTensor ->
// np.frombuffer(...).copy()  (node.py:73-89, 170-181, 45-48): no host staging, no
// serialisation, no per-request channel.  A ring is `slots` receive buffers on the
// CONSUMER's GPU plus two 64-bit counters per slot (published / released).  The producer
// maps the ring (CUDA IPC across processes, peer access inside one process) and either
//   - copies its output into the slot (b2p_ring_send: copy engine or SM copy kernel), or
//   - lets the producing stage kernel store its output STRAIGHT into the mapped slot
//     (b2p_ring_acquire -> stage forward with out = slot -> b2p_ring_publish), so the
//     transfer rides on the kernel's own epilogue stores: zero extra copies.
// (The CIFAR tensor-core kernels do the same handshake themselves: ring_hooks.cuh, b2p_stage_forward_ring_span.)
// Here all waiting is done on the device by one-thread kernels (acquire loads with back-off and a
// time-out that traps instead of hanging), so the host never blocks and the schedule of a
// whole batch can be enqueued ahead -- or captured once in a CUDA graph and replayed: the message
// counters live in device memory, so no launch carries a step-dependent argument.  `seq` in the
// C ABI only selects the slot (seq % slots); messages of one slot must be used in order.  The consumer's waits are on local memory; the
// producer's back-pressure wait polls the consumer's counter over NVLink.
#pragma once
#include <cstring>

#include "common.cuh"
#include "ring_hooks.cuh"

struct b2p_ring {
  int consumer_dev = -1, producer_dev = -1;
  int slots = 0;
  size_t slot_bytes = 0;
  unsigned char* data = nullptr;      // slots * slot_bytes
  unsigned long long* flags = nullptr;  // [slots] published count, then [slots] released count
  bool owner = false, ipc = false;
  int mode = 1;                       // 0: cudaMemcpyAsync (copy engine), 1: SM copy kernel
  void* base = nullptr;               // allocation base (data and flags live in one allocation)
  // Message counters per slot, in the memory of the GPU that uses them.  Keeping the sequence state
  // on the device (instead of passing targets as kernel arguments) makes every wait / publish kernel
  // launch identical from step to step, so a whole step can be captured in a CUDA graph and replayed.
  unsigned long long* local_count = nullptr;   // producer side: published per slot; consumer side: consumed per slot
  int* ticket = nullptr;              // producer side, [slots]: rows a publishing stage kernel has stored into each slot (ring_hooks.cuh)
};

namespace b2p {

struct RingWire {                      // what travels in the opaque 256-byte handle
  cudaIpcMemHandle_t mem;
  int32_t consumer_dev, slots;
  uint64_t slot_bytes;
  uint64_t magic;
};
static_assert(sizeof(RingWire) <= B2P_RING_HANDLE_BYTES, "handle too small");
constexpr uint64_t kRingMagic = 0x62327072696e6731ull;   // "b2pring1"

// All four are one-thread, stream-ordered kernels whose arguments never change between steps.
// producer, before writing a slot: the consumer must have released every earlier message of that slot
__global__ void ring_acquire_k(const unsigned long long* count, const unsigned long long* released_flag) {
  const unsigned long long k = *count;               // messages already published into this slot
  if (k > 0) spin_until_ge(released_flag, k);
}
// producer, after writing: bump the local count and publish it
__global__ void ring_publish_k(unsigned long long* count, unsigned long long* published_flag) {
  const unsigned long long k = *count + 1;
  *count = k;
  __threadfence_system();                            // everything earlier in the stream is visible system-wide
  st_release_sys(published_flag, k);
}
// consumer, before reading a slot
__global__ void ring_recv_k(const unsigned long long* count, const unsigned long long* published_flag) {
  spin_until_ge(published_flag, *count + 1);
}
// consumer, after the last kernel that reads the slot
__global__ void ring_release_k(unsigned long long* count, unsigned long long* released_flag) {
  const unsigned long long k = *count + 1;
  *count = k;
  __threadfence_system();
  st_release_sys(released_flag, k);
}
// SM-driven copy with 128-bit peer stores (dst may be mapped memory of another GPU)
__global__ void __launch_bounds__(256)
ring_copy(uint4* __restrict__ dst, const uint4* __restrict__ src, size_t n16, unsigned char* dst_tail,
          const unsigned char* src_tail, int tail) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (; i + 3 * stride < n16; i += 4 * stride) {           // 4 independent 16-byte loads in flight
    uint4 a = src[i], b = src[i + stride], c = src[i + 2 * stride], d = src[i + 3 * stride];
    dst[i] = a; dst[i + stride] = b; dst[i + 2 * stride] = c; dst[i + 3 * stride] = d;
  }
  for (; i < n16; i += stride) dst[i] = src[i];
  if (blockIdx.x == 0 && (int)threadIdx.x < tail) dst_tail[threadIdx.x] = src_tail[threadIdx.x];
}

// the GPU whose stream runs this side's handshake kernels: the consumer's for the owning side, else the producer's
inline int ring_side_device(const b2p_ring* r) { return r->owner ? r->consumer_dev : r->producer_dev; }

inline unsigned long long* published(b2p_ring* r, uint64_t seq) { return r->flags + (seq % r->slots); }
inline unsigned long long* released(b2p_ring* r, uint64_t seq) { return r->flags + r->slots + (seq % r->slots); }

inline int ring_copy_into(b2p_ring* r, void* dst, const void* src, size_t bytes, cudaStream_t s) {
  if (r->mode == 0 || ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15)) {
    B2P_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, s));
    return 0;
  }
  const size_t n16 = bytes / 16;
  const int tail = (int)(bytes % 16);
  const int grid = (int)std::max<size_t>(1, std::min<size_t>((n16 + 1023) / 1024, 148 * 4));
  B2P_LAUNCH(ring_copy, grid, 256, 0, s, static_cast<uint4*>(dst), static_cast<const uint4*>(src), n16,
             static_cast<unsigned char*>(dst) + n16 * 16, static_cast<const unsigned char*>(src) + n16 * 16, tail);
  return 0;
}

}  // namespace b2p

extern "C" {

int b2p_ring_create(b2p_ring** out, int device, int slots, size_t slot_bytes) {
  using namespace b2p;
  if (!out) return fail(B2P_EINVAL, "null argument");
  *out = nullptr;
  if (slots < 1 || slots > 1024 || slot_bytes == 0) return fail(B2P_EINVAL, "ring: need 1..1024 slots of non-zero size");
  int ndev = 0;
  B2P_CUDA(cudaGetDeviceCount(&ndev));
  if (device < 0 || device >= ndev) return fail(B2P_ECUDA, "ring: device %d not present", device);
  DeviceGuard g(device);
  const size_t slot_pad = (slot_bytes + 255) & ~size_t(255);
  const size_t flag_bytes = ((size_t)2 * slots * sizeof(unsigned long long) + 255) & ~size_t(255);
  void* base = nullptr;
  B2P_CUDA(cudaMalloc(&base, flag_bytes + slot_pad * slots));
  cudaError_t ce0 = cudaMemset(base, 0, flag_bytes);
  if (ce0 == cudaSuccess) ce0 = cudaDeviceSynchronize();
  if (ce0 != cudaSuccess) {
    cudaFree(base);
    return fail(B2P_ECUDA, "ring: clearing the flags failed: %s", cudaGetErrorString(ce0));
  }
  b2p_ring* r = new b2p_ring();
  r->consumer_dev = device;
  r->slots = slots;
  r->slot_bytes = slot_pad;
  r->base = base;
  r->flags = static_cast<unsigned long long*>(base);
  r->data = static_cast<unsigned char*>(base) + flag_bytes;
  r->owner = true;
  cudaError_t ce = cudaMalloc((void**)&r->local_count, (3 * slots + 4) * sizeof(unsigned long long));
  if (ce == cudaSuccess) ce = cudaMemset(r->local_count, 0, (3 * slots + 4) * sizeof(unsigned long long));
  r->ticket = reinterpret_cast<int*>(r->local_count + slots);
  if (ce == cudaSuccess) ce = cudaDeviceSynchronize();
  if (ce != cudaSuccess) {
    cudaFree(r->local_count);
    cudaFree(base);
    delete r;
    return fail(B2P_ECUDA, "ring: counter allocation failed: %s", cudaGetErrorString(ce));
  }
  *out = r;
  return 0;
}

int b2p_ring_export(b2p_ring* ring, void* handle_out) {
  using namespace b2p;
  if (!ring || !handle_out) return fail(B2P_EINVAL, "null argument");
  if (!ring->owner) return fail(B2P_ESTATE, "ring: only the consumer side can export");
  RingWire w;
  memset(&w, 0, sizeof(w));
  DeviceGuard g(ring->consumer_dev);
  B2P_CUDA(cudaIpcGetMemHandle(&w.mem, ring->base));
  w.consumer_dev = ring->consumer_dev;
  w.slots = ring->slots;
  w.slot_bytes = ring->slot_bytes;
  w.magic = kRingMagic;
  memset(handle_out, 0, B2P_RING_HANDLE_BYTES);
  memcpy(handle_out, &w, sizeof(w));
  return 0;
}

int b2p_ring_open(b2p_ring** out, const void* handle, int producer_device) {
  using namespace b2p;
  if (!out || !handle) return fail(B2P_EINVAL, "null argument");
  *out = nullptr;
  RingWire w;
  memcpy(&w, handle, sizeof(w));
  if (w.magic != kRingMagic) return fail(B2P_EINVAL, "ring: not a ring handle");
  DeviceGuard g(producer_device);
  void* base = nullptr;
  B2P_CUDA(cudaIpcOpenMemHandle(&base, w.mem, cudaIpcMemLazyEnablePeerAccess));
  b2p_ring* r = new b2p_ring();
  r->consumer_dev = w.consumer_dev;
  r->producer_dev = producer_device;
  r->slots = w.slots;
  r->slot_bytes = (size_t)w.slot_bytes;
  const size_t flag_bytes = ((size_t)2 * r->slots * sizeof(unsigned long long) + 255) & ~size_t(255);
  r->base = base;
  r->flags = static_cast<unsigned long long*>(base);
  r->data = static_cast<unsigned char*>(base) + flag_bytes;
  r->ipc = true;
  cudaError_t ce = cudaMalloc((void**)&r->local_count, (3 * r->slots + 4) * sizeof(unsigned long long));
  if (ce == cudaSuccess) ce = cudaMemset(r->local_count, 0, (3 * r->slots + 4) * sizeof(unsigned long long));
  if (ce == cudaSuccess) ce = cudaDeviceSynchronize();
  if (ce != cudaSuccess) {
    cudaFree(r->local_count);
    cudaIpcCloseMemHandle(base);
    delete r;
    return fail(B2P_ECUDA, "ring: counter allocation failed: %s", cudaGetErrorString(ce));
  }
  r->ticket = reinterpret_cast<int*>(r->local_count + r->slots);
  *out = r;
  return 0;
}

int b2p_ring_open_local(b2p_ring** out, b2p_ring* consumer_side, int producer_device) {
  using namespace b2p;
  if (!out || !consumer_side) return fail(B2P_EINVAL, "null argument");
  *out = nullptr;
  if (producer_device != consumer_side->consumer_dev) {
    DeviceGuard g(producer_device);
    int can = 0;
    B2P_CUDA(cudaDeviceCanAccessPeer(&can, producer_device, consumer_side->consumer_dev));
    if (!can) return fail(B2P_ECUDA, "ring: GPU %d cannot access GPU %d as a peer", producer_device, consumer_side->consumer_dev);
    cudaError_t e = cudaDeviceEnablePeerAccess(consumer_side->consumer_dev, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
    else if (e != cudaSuccess) return fail(B2P_ECUDA, "cudaDeviceEnablePeerAccess: %s", cudaGetErrorString(e));
  }
  b2p_ring* r = new b2p_ring(*consumer_side);
  r->owner = false;
  r->ipc = false;
  r->producer_dev = producer_device;
  r->local_count = nullptr;
  r->ticket = nullptr;
  {
    DeviceGuard g(producer_device);
    cudaError_t ce = cudaMalloc((void**)&r->local_count, (3 * r->slots + 4) * sizeof(unsigned long long));
    if (ce == cudaSuccess) ce = cudaMemset(r->local_count, 0, (3 * r->slots + 4) * sizeof(unsigned long long));
    if (ce == cudaSuccess) ce = cudaDeviceSynchronize();
    if (ce != cudaSuccess) {
      cudaFree(r->local_count);
      delete r;
      return fail(B2P_ECUDA, "ring: counter allocation failed: %s", cudaGetErrorString(ce));
    }
    r->ticket = reinterpret_cast<int*>(r->local_count + r->slots);
  }
  *out = r;
  return 0;
}

int b2p_ring_destroy(b2p_ring* ring) {
  if (!ring) return 0;
  if (ring->owner) {
    b2p::DeviceGuard g(ring->consumer_dev);
    cudaFree(ring->base);
    cudaFree(ring->local_count);
  } else {
    b2p::DeviceGuard g(ring->producer_dev);
    cudaFree(ring->local_count);
    if (ring->ipc) cudaIpcCloseMemHandle(ring->base);
  }
  delete ring;
  return 0;
}

int b2p_ring_set_timeout(int device, double seconds) {
  using namespace b2p;
  if (!(seconds >= 0.0) || seconds > 1e6) return fail(B2P_EINVAL, "ring: time-out must be 0 (wait for ever) .. 1e6 seconds");
  DeviceGuard g(device);
  if (!g.ok) return fail(B2P_ECUDA, "ring: device %d not present", device);
  const unsigned long long ns = (unsigned long long)(seconds * 1e9);
  B2P_CUDA(cudaMemcpyToSymbol(g_wait_limit_ns, &ns, sizeof(ns)));
  return 0;
}

int b2p_ring_check(int device) {
  using namespace b2p;
  DeviceGuard g(device);
  if (!g.ok) return fail(B2P_ECUDA, "ring: device %d not present", device);
  unsigned long long rec[4] = {0, 0, 0, 0};
  B2P_CUDA(cudaMemcpyFromSymbol(rec, g_ring_error, sizeof(rec)));
  if (rec[0] == 0) return 0;
  const unsigned long long zero[4] = {0, 0, 0, 0};
  B2P_CUDA(cudaMemcpyToSymbol(g_ring_error, zero, sizeof(zero)));
  return fail(B2P_ETIMEOUT,
              "ring: a device-side wait on GPU %d gave up (%s flag at %#llx: waited for %llu, last saw %llu); results of "
              "the steps enqueued since are invalid -- the peer stage stalled for longer than the ring time-out",
              device, rec[0] == 2 ? "remote `released`" : "local `published`/`released`", rec[1], rec[2], rec[3]);
}

int b2p_ring_set_mode(b2p_ring* ring, int mode) {
  if (!ring || mode < 0 || mode > 1) return b2p::fail(B2P_EINVAL, "ring: mode must be 0 (copy engine) or 1 (SM copy kernel)");
  ring->mode = mode;
  return 0;
}

int b2p_ring_acquire(b2p_ring* ring, uint64_t seq, void** slot_dev, void* stream) {
  using namespace b2p;
  if (!ring || !slot_dev) return fail(B2P_EINVAL, "null argument");
  DeviceGuard g(ring_side_device(ring));
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  B2P_LAUNCH(ring_acquire_k, 1, 1, 0, s, ring->local_count + (seq % ring->slots), released(ring, seq));
  *slot_dev = ring->data + (seq % ring->slots) * ring->slot_bytes;
  return 0;
}

int b2p_ring_publish(b2p_ring* ring, uint64_t seq, void* stream) {
  using namespace b2p;
  if (!ring) return fail(B2P_EINVAL, "null argument");
  DeviceGuard g(ring_side_device(ring));
  B2P_LAUNCH(ring_publish_k, 1, 1, 0, static_cast<cudaStream_t>(stream), ring->local_count + (seq % ring->slots),
             published(ring, seq));
  return 0;
}

int b2p_ring_send(b2p_ring* ring, const void* src_dev, size_t bytes, uint64_t seq, void* stream) {
  using namespace b2p;
  if (!ring || !src_dev) return fail(B2P_EINVAL, "null argument");
  if (bytes > ring->slot_bytes) return fail(B2P_ELIMIT, "ring: message of %zu bytes exceeds the slot size %zu", bytes, ring->slot_bytes);
  void* slot = nullptr;
  DeviceGuard g(ring_side_device(ring));
  B2P_TRY(b2p_ring_acquire(ring, seq, &slot, stream));
  B2P_TRY(ring_copy_into(ring, slot, src_dev, bytes, static_cast<cudaStream_t>(stream)));
  return b2p_ring_publish(ring, seq, stream);
}

int b2p_ring_recv(b2p_ring* ring, uint64_t seq, void** slot_dev, void* stream) {
  using namespace b2p;
  if (!ring || !slot_dev) return fail(B2P_EINVAL, "null argument");
  DeviceGuard g(ring_side_device(ring));
  B2P_LAUNCH(ring_recv_k, 1, 1, 0, static_cast<cudaStream_t>(stream), ring->local_count + (seq % ring->slots),
             published(ring, seq));
  *slot_dev = ring->data + (seq % ring->slots) * ring->slot_bytes;
  return 0;
}

int b2p_ring_release(b2p_ring* ring, uint64_t seq, void* stream) {
  using namespace b2p;
  if (!ring) return fail(B2P_EINVAL, "null argument");
  DeviceGuard g(ring_side_device(ring));
  B2P_LAUNCH(ring_release_k, 1, 1, 0, static_cast<cudaStream_t>(stream), ring->local_count + (seq % ring->slots),
             released(ring, seq));
  return 0;
}

}  // extern "C"
