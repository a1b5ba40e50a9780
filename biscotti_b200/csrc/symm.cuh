// Cross-GPU signalling primitives for kernels that read and write peer memory over NVLink.
//
// Every rank owns a flag array flags[slot][world] of 64-bit epochs in its symmetric buffer.  A producer
// publishes "my part of <slot> for block <epoch> is complete" by storing the epoch into its column of
// every peer's array (st.release.sys after a system-scope fence); a consumer polls only its own local
// array with ld.acquire.sys.  Polls are bounded in time; a timeout is reported to the host (g_symm_error),
// never a GPU that spins forever and never a trap.
//
// Replaces the reference's blocking RPC fan-outs and WaitGroups (DistSys/main.go:1666-1711,
// 1403-1444) for the data that now lives on the GPUs.
#pragma once
#include <stdint.h>
#include <stdio.h>

namespace bsc {

constexpr int SYMM_MAX_WORLD = 8;

enum SymmSlot { SLOT_PLANES = 0, SLOT_GRAM = 1, SLOT_SHARES = 2, SLOT_WEIGHTS = 3, SLOT_POINTS = 4, SLOT_COUNT = 8 };

struct PeerTable {
  void* ptr[SYMM_MAX_WORLD];
};

__device__ __forceinline__ void st_release_sys_u64(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

// one thread: publish `epoch` for `slot` to every rank (including self)
__device__ __forceinline__ void symm_signal_all(const PeerTable& flags, int world, int my_rank, int slot,
                                                unsigned long long epoch) {
  __threadfence_system();
  for (int q = 0; q < world; q++)
    st_release_sys_u64(reinterpret_cast<unsigned long long*>(flags.ptr[q]) + (size_t)slot * SYMM_MAX_WORLD + my_rank, epoch);
}

__device__ __forceinline__ unsigned long long symm_now_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// Cross-GPU waits are bounded by TIME (a slow peer - first-iteration warm-up, a long CIFAR step, a host pause -
// must stall the job, not kill it) and end in a reported error, never in __trap(): a trap would destroy the
// context of a rank whose memory its peers have mapped.  On a timeout the waiter records a code in g_symm_error
// (the host reads it with bsc_symm_error() when it collects the block) and carries on with whatever data it has.
constexpr unsigned long long SYMM_TIMEOUT_NS = 20ull * 1000 * 1000 * 1000;
static __device__ int g_symm_error;   // 0 = none, else 1 + slot of the first wait that timed out (one copy per .cu
                                       // file that waits; bsc_symm_error() in fused.cu folds them)

// one thread: wait until every rank has published >= epoch for `slot` in my local flag array
__device__ __forceinline__ bool symm_wait_all(const unsigned long long* my_flags, int world, int slot,
                                              unsigned long long epoch) {
  unsigned long long t0 = 0;
  for (int q = 0; q < world; q++) {
    const unsigned long long* f = my_flags + (size_t)slot * SYMM_MAX_WORLD + q;
    unsigned spins = 0;
    while (ld_acquire_sys_u64(f) < epoch) {
      __nanosleep(32);
      if ((++spins & 1023u) == 0) {
        const unsigned long long now = symm_now_ns();
        if (t0 == 0) t0 = now;
        if (now - t0 > SYMM_TIMEOUT_NS) {
          atomicCAS(&g_symm_error, 0, 1 + slot);
          return false;
        }
      }
    }
  }
  return true;
}

// same-GPU hand-off of "the remote data is here": one elected thread of the grid polls the NVLink-visible flags
// and republishes the epoch in a local word; every other CTA polls that word through L2 only
__device__ __forceinline__ void symm_local_publish(unsigned long long* word, unsigned long long epoch) {
  asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(word), "l"(epoch) : "memory");
}
__device__ __forceinline__ bool symm_local_wait(const unsigned long long* word, unsigned long long epoch) {
  unsigned long long t0 = 0;
  unsigned spins = 0;
  for (;;) {
    unsigned long long v;
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(word) : "memory");
    if (v >= epoch) return true;
    __nanosleep(32);
    if ((++spins & 1023u) == 0) {
      const unsigned long long now = symm_now_ns();
      if (t0 == 0) t0 = now;
      if (now - t0 > SYMM_TIMEOUT_NS) {
        atomicCAS(&g_symm_error, 0, 100);
        return false;
      }
    }
  }
}

}  // namespace bsc
