// ring_hooks.cuh -- one side of a handoff-ring handshake, small enough to live INSIDE a stage kernel.
//
// handoff.cuh does the slot handshake with one-thread kernels (wait / publish / recv / release), which keeps the
// compute kernels unaware of rings but costs a serialised launch (~2.5 us) each: five kernels per 128-sample
// microbatch on the FC rank of the 2-GPU CIFAR pipeline, more time than the arithmetic.  A RingHook hands the same
// two pointers to the compute kernel instead:
//   consumer: the thread that loads the slot waits for `published`; the first kernel that runs after the last
//             reader stores `released`;
//   producer: the storing warps wait for `released` while the tensor core is still busy, and the last CTA to
//             finish (arrival ticket) stores `published` after a system-scope fence.
// Counters stay in device memory exactly as in handoff.cuh, so launches still carry no step-dependent argument and
// a captured CUDA graph replays verbatim.  (Replaces node.py:73-89 / :45-48, like handoff.cuh.)
#pragma once
#include <cstdint>

#include "common.cuh"

namespace b2p {

struct RingHook {                  // all null: no ring on this side
  unsigned long long* count;       // this GPU's message counter of the slot (local device memory)
  unsigned long long* flag;        // the slot's published / released counter (consumer GPU's memory)
};

__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_count(const unsigned long long* p) {      // written by an earlier kernel
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");     // (never from a stale L1 line)
  return v;
}
__device__ __forceinline__ uint64_t hook_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ unsigned long long ld_relaxed_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
// A wait that has run out of time (g_wait_limit_ns, common.cuh) records itself -- first one wins -- and returns; once
// a record exists every later ring wait on this GPU returns immediately.  No trap: a trap is a sticky context error
// that takes every stage, ring and graph of the process with it, and the peer rank soon after.
__device__ __forceinline__ unsigned long long ld_volatile_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ bool ring_gave_up() { return ld_volatile_u64(&g_ring_error[0]) != 0ull; }
__device__ __noinline__ void ring_wait_timed_out(const unsigned long long* flag, unsigned long long target,
                                                 unsigned long long seen, int kind) {
  if (atomicCAS(&g_ring_error[0], 0ull, (unsigned long long)(kind + 1)) == 0ull) {
    g_ring_error[1] = (unsigned long long)(uintptr_t)flag;
    g_ring_error[2] = target;
    g_ring_error[3] = seen;
    __threadfence_system();
  }
}
// Measured on 2 x B200: a system-scope fence costs ~2.5 us when peer traffic is outstanding, so the hooks use as few
// as the memory model allows and keep them off the critical path where they can.
// Flag in LOCAL memory (the consumer waiting for `published`): acquire loads.
__device__ __forceinline__ void spin_until_ge(const unsigned long long* flag, unsigned long long target) {
  if (ld_acquire_sys(flag) >= target) return;
  const uint64_t t0 = hook_ns();
  const unsigned long long limit = ld_volatile_u64(&g_wait_limit_ns);
  unsigned ns = 32;
  unsigned long long v;
  while ((v = ld_acquire_sys(flag)) < target) {     // one thread; gives up after the limit instead of hanging
    __nanosleep(ns);
    if (ns < 1024) ns *= 2;
    if (ring_gave_up()) return;
    if (limit && hook_ns() - t0 > limit) { ring_wait_timed_out(flag, target, v, 0); return; }
  }
}
// Flag on ANOTHER GPU (the producer waiting for `released`): every poll is an NVLink round trip, so poll relaxed
// and pay for the acquire once, after the value has been seen.
__device__ __forceinline__ void spin_until_ge_remote(const unsigned long long* flag, unsigned long long target) {
  unsigned long long v;
  if ((v = ld_relaxed_sys(flag)) < target) {
    const uint64_t t0 = hook_ns();
    const unsigned long long limit = ld_volatile_u64(&g_wait_limit_ns);
    unsigned ns = 64;
    while ((v = ld_relaxed_sys(flag)) < target) {
      __nanosleep(ns);
      if (ns < 1024) ns *= 2;
      if (ring_gave_up()) break;
      if (limit && hook_ns() - t0 > limit) { ring_wait_timed_out(flag, target, v, 1); break; }
    }
  }
  asm volatile("fence.acq_rel.sys;" ::: "memory");
}

// A run of consecutive messages of one ring handled by ONE launch: rows [i*rows_per_msg, (i+1)*rows_per_msg) of the
// launch's batch travel as message i, in slot (first_slot + i) % slots.  The producing kernel publishes each message
// when its last row has been stored, the consuming kernel waits per message: the stages overlap at microbatch
// granularity although each of them is a single launch per step.  data == nullptr: no ring.
struct RingSpan {
  unsigned long long* count;       // [slots] this GPU's message counters (local)
  unsigned long long* published;   // [slots] on the consumer GPU
  unsigned long long* released;    // [slots] on the consumer GPU
  unsigned char* data;             // slot 0 (consumer GPU's memory)
  unsigned long long slot_bytes;
  int slots, first_slot, rows_per_msg;
  int* tickets;                    // [slots] producer side: rows stored so far into the slot (local)
  unsigned long long* acq_seen;    // [slots] producer side: highest `released` value some CTA has observed (local)
};
__device__ __forceinline__ int span_slot(const RingSpan& r, int msg) { return (r.first_slot + msg) % r.slots; }

// CTA-scope hand-offs between warp roles of one kernel (shared-memory words)
__device__ __forceinline__ int ld_acquire_cta(const int* p) {
  int v;
  asm volatile("ld.acquire.cta.shared.s32 %0, [%1];" : "=r"(v) : "r"((unsigned)__cvta_generic_to_shared(p)) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_cta(int* p, int v) {
  asm volatile("st.release.cta.shared.s32 [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(p)), "r"(v) : "memory");
}
__device__ __forceinline__ void red_release_cta_add(int* p, int v) {
  asm volatile("red.release.cta.shared.add.s32 [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(p)), "r"(v) : "memory");
}

__device__ __forceinline__ unsigned long long ld_acquire_gpu(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_gpu(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// consumer, before reading the slot (one thread)
__device__ __forceinline__ void hook_wait_published(const RingHook& h) {
  if (h.flag) spin_until_ge(h.flag, ld_count(h.count) + 1);
}
// consumer, after the last read of the slot (one thread of one CTA)
__device__ __forceinline__ void hook_release(const RingHook& h) {
  if (h.flag) {
    const unsigned long long k = ld_count(h.count) + 1;
    *h.count = k;
    st_relaxed_sys(h.flag, k);       // nothing to publish but "my reads are over", and they are: their kernel has completed
  }
}
// producer, before the first store into the slot (any number of threads may call it; it only reads)
__device__ __forceinline__ void hook_wait_released(const RingHook& h) {
  if (h.flag) {
    const unsigned long long k = ld_count(h.count);      // messages already published into this slot
    if (k > 0) spin_until_ge_remote(h.flag, k);
  }
}
// producer, one thread of the last CTA, after it has observed every other CTA's arrival (each behind a gpu-scope
// fence): the release store's system-scope fence is cumulative over all of them
__device__ __forceinline__ void hook_publish(const RingHook& h) {
  if (h.flag) {
    const unsigned long long k = ld_count(h.count) + 1;
    *h.count = k;
    st_release_sys(h.flag, k);
  }
}

}  // namespace b2p
