// Cross-GPU signalling primitives for kernels that read and write peer memory over NVLink.
//
// Every rank owns a flag array flags[slot][world] of 64-bit epochs in its symmetric buffer.  A producer
// publishes "my part of <slot> for block <epoch> is complete" by storing the epoch into its column of
// every peer's array (st.release.sys after a system-scope fence); a consumer polls only its own local
// array with ld.acquire.sys.  Polls are bounded: a protocol bug must surface as a failed launch, never
// as a GPU that spins forever.
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

// one thread: wait until every rank has published >= epoch for `slot` in my local flag array
__device__ __forceinline__ void symm_wait_all(const unsigned long long* my_flags, int world, int slot,
                                              unsigned long long epoch) {
  for (int q = 0; q < world; q++) {
    const unsigned long long* f = my_flags + (size_t)slot * SYMM_MAX_WORLD + q;
    unsigned long long spins = 0;
    while (ld_acquire_sys_u64(f) < epoch) {
      __nanosleep(64);
      if (++spins > (1ull << 24)) {   // ~seconds
        printf("[biscotti_b200] symm_wait_all: rank %d never published slot %d epoch %llu\n", q, slot, epoch);
        __trap();
      }
    }
  }
}

}  // namespace bsc
