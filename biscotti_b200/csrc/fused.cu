// Fused compute + collective kernels over NVLink peer memory (the two communication-bound paths).
//
// 1. krum_gram_fused_kernel — the reference's workers->verifiers fan-out (every noised update sent to
//    every verifier, DistSys/main.go:1666-1775, krum.go:227-365) fused with the Gram GEMM of
//    get_krum_scores (ML/Pytorch/client_obj.py:132-143).  Instead of all-gathering the n x d updates,
//    the K (= d) dimension is sharded over the GPUs: GPU r computes the partial Gram over its slice of
//    d for ALL peers, pulling the other GPUs' rows of that slice tile by tile with TMA straight from
//    their memory into the tcgen05 pipeline (1/8 of the all-gather bytes cross NVLink), and pushes its
//    128x128 fp32 partial into every GPU's reduction buffer with peer stores.  krum_select_fused_kernel
//    then reduces the partials in fixed order (bit-identical on every GPU), forms distances, sorts,
//    scores and selects.
// 2. secagg_fused_kernel — the workers->miners share scatter, the miners' fold and the leader's gather
//    (main.go:1847-1952, 2157-2235; kyber.go:244-287, 809-857; honest.go:391-502) as one
//    reduce-scatter + reconstruct + SGD step + all-gather: the GPU that owns a chunk range pulls every
//    GPU's partial share sums with peer loads, adds them, recovers the polynomial exactly, verifies the
//    redundant shares, de-quantises, steps the weights and stores the new slice into every GPU's copy.
//
// Cross-GPU ordering uses the epoch flags of symm.cuh; all waits are bounded.
#include "common.cuh"
#include <mutex>
#include <unordered_map>

#include "symm.cuh"
#include "tc_ptx.cuh"
#include "secagg_math.cuh"
#include <cuda_bf16.h>
#include <string.h>

namespace bsc {

constexpr int FG_BM = 128;
constexpr int FG_BK = 64;
constexpr int FG_STAGES = 3;
constexpr int FG_TILE_BYTES = FG_BM * FG_BK * 2;
constexpr int FG_STAGE_BYTES = 4 * FG_TILE_BYTES;
constexpr int FG_THREADS = 192;
constexpr int FG_TMEM_COLS = 128;
constexpr size_t FG_SMEM = (size_t)FG_STAGES * FG_STAGE_BYTES + 1024 + 256;

struct GramFusedParams {
  CUtensorMap hi[SYMM_MAX_WORLD];   // rank q's bf16 hi plane [Sr][ldh], box (64 x Sr), SWIZZLE_128B
  CUtensorMap lo[SYMM_MAX_WORLD];
  PeerTable gpart;                  // rank q's reduction buffer [world][ldg][ldg] fp32 (one partial per source rank)
  PeerTable flags;                  // rank q's flag array
  float* lpart;                     // local split-K scratch [ksplit][npairs][128][128] fp32
  unsigned int* done_counter;       // local, monotonic: [0] = CTAs finished, [1 + pair] = splits of a pair written
  unsigned int seq;                 // 1-based launch sequence number of this kernel on this rank
  unsigned long long epoch;
  int world, rank, Sr, ntile, ldg;
  int kb_lo, kb_cnt, kb_per, ksplit;
};

__global__ void __launch_bounds__(FG_THREADS, 1) krum_gram_fused_kernel(const __grid_constant__ GramFusedParams P) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t tiles = (raw + 1023u) & ~1023u;
  const uint32_t bars = tiles + FG_STAGES * FG_STAGE_BYTES;
  const uint32_t bar_full = bars, bar_empty = bars + 8 * FG_STAGES, bar_tmem = bars + 16 * FG_STAGES;
  const uint32_t tmem_slot = bar_tmem + 8;
  uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(smem_raw + (tmem_slot - raw));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int ti = 0, tj = 0;
  {
    int p = blockIdx.x;
    for (ti = 0; ti < P.ntile; ti++) {
      const int rowlen = P.ntile - ti;
      if (p < rowlen) {
        tj = ti + p;
        break;
      }
      p -= rowlen;
    }
  }
  const bool diag = (ti == tj);
  const int split = blockIdx.y;
  // The planes of this block were written by the previous kernel on this stream, so they are complete when
  // this kernel starts: publish them to every rank before anything else (no separate signal launch).
  if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) symm_signal_all(P.flags, P.world, P.rank, SLOT_PLANES, P.epoch);
  const int kb0 = P.kb_lo + split * P.kb_per;
  int nkb = min(P.kb_lo + P.kb_cnt, kb0 + P.kb_per) - kb0;
  if (nkb < 0) nkb = 0;

  if (threadIdx.x == 0) {
    for (int s = 0; s < FG_STAGES; s++) {
      mbar_init(bar_full + 8 * s, 1);
      mbar_init(bar_empty + 8 * s, 1);
    }
    mbar_init(bar_tmem, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(FG_TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;
  // Sr <= 128: a 128-row tile is assembled from 128/Sr owners (one box of Sr rows each).  Sr a multiple of 128 (more
  // than 128 peers per rank): a tile is one box of 128 rows out of one owner's plane, tpo tiles per owner.
  const int sub = P.Sr >= FG_BM ? 1 : FG_BM / P.Sr;
  const int tpo = P.Sr >= FG_BM ? P.Sr / FG_BM : 1;

  if (warp == 0) {
    if (lane == 0 && nkb > 0) {
      // every rank must have finished writing this block's planes before its rows are pulled
      symm_wait_all(reinterpret_cast<const unsigned long long*>(P.flags.ptr[P.rank]), P.world, SLOT_PLANES, P.epoch);
      asm volatile("fence.proxy.async;" ::: "memory");
      for (int kb = 0; kb < nkb; kb++) {
        const int s = kb % FG_STAGES;
        const uint32_t ph = (uint32_t)(kb / FG_STAGES) & 1u;
        mbar_wait(bar_empty + 8 * s, ph ^ 1u);
        const uint32_t st = tiles + s * FG_STAGE_BYTES;
        mbar_expect_tx(bar_full + 8 * s, diag ? 2 * FG_TILE_BYTES : 4 * FG_TILE_BYTES);
        const int kx = (kb0 + kb) * FG_BK;
        for (int j = 0; j < sub; j++) {
          const int owner = tpo > 1 ? ti / tpo : ti * sub + j;
          const int y = tpo > 1 ? (ti % tpo) * FG_BM : 0;
          const uint32_t off = (uint32_t)j * P.Sr * 128u;   // Sr % 8 == 0 keeps every box on a 1024-B swizzle atom
          tma_load_2d(st + off, &P.hi[owner], kx, y, bar_full + 8 * s);
          tma_load_2d(st + FG_TILE_BYTES + off, &P.lo[owner], kx, y, bar_full + 8 * s);
        }
        if (!diag) {
          for (int j = 0; j < sub; j++) {
            const int owner = tpo > 1 ? tj / tpo : tj * sub + j;
            const int y = tpo > 1 ? (tj % tpo) * FG_BM : 0;
            const uint32_t off = (uint32_t)j * P.Sr * 128u;
            tma_load_2d(st + 2 * FG_TILE_BYTES + off, &P.hi[owner], kx, y, bar_full + 8 * s);
            tma_load_2d(st + 3 * FG_TILE_BYTES + off, &P.lo[owner], kx, y, bar_full + 8 * s);
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && nkb > 0) {
      const uint32_t idesc = umma_idesc_bf16_128x128();
      for (int kb = 0; kb < nkb; kb++) {
        const int s = kb % FG_STAGES;
        const uint32_t ph = (uint32_t)(kb / FG_STAGES) & 1u;
        mbar_wait(bar_full + 8 * s, ph);
        tc_fence_after();
        const uint32_t st = tiles + s * FG_STAGE_BYTES;
        const uint64_t a_hi = umma_desc_sw128(st), a_lo = umma_desc_sw128(st + FG_TILE_BYTES);
        const uint64_t b_hi = diag ? a_hi : umma_desc_sw128(st + 2 * FG_TILE_BYTES);
        const uint64_t b_lo = diag ? a_lo : umma_desc_sw128(st + 3 * FG_TILE_BYTES);
#pragma unroll
        for (int k = 0; k < FG_BK / 16; k++) {
          const uint64_t off = (uint64_t)((k * 16 * 2) >> 4);
          umma_bf16(tmem_base, a_hi + off, b_hi + off, idesc, (kb > 0 || k > 0) ? 1u : 0u);
          umma_bf16(tmem_base, a_hi + off, b_lo + off, idesc, 1u);
          umma_bf16(tmem_base, a_lo + off, b_hi + off, idesc, 1u);
        }
        umma_commit(bar_empty + 8 * s);
      }
      umma_commit(bar_tmem);
    }
  } else {
    // epilogue, stage 1: TMEM -> registers -> this rank's local split-K scratch
    const int q4 = warp & 3;
    if (nkb > 0) {
      mbar_wait(bar_tmem, 0);
      tc_fence_after();
    }
    float* L = P.lpart + ((size_t)split * gridDim.x + blockIdx.x) * (FG_BM * FG_BM);
    const int lrow = q4 * 32 + lane;
#pragma unroll 1
    for (int cc = 0; cc < FG_BM / 32; cc++) {
      uint32_t v[32];
      if (nkb > 0) {
        tmem_ld32(tmem_base + ((uint32_t)(q4 * 32) << 16) + (uint32_t)(cc * 32), v);
      } else {
#pragma unroll
        for (int c = 0; c < 32; c++) v[c] = 0u;
      }
      float4* dst = reinterpret_cast<float4*>(L + (size_t)lrow * FG_BM + cc * 32);
#pragma unroll
      for (int c = 0; c < 8; c++)
        dst[c] = make_float4(__uint_as_float(v[4 * c]), __uint_as_float(v[4 * c + 1]), __uint_as_float(v[4 * c + 2]),
                             __uint_as_float(v[4 * c + 3]));
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(FG_TMEM_COLS) : "memory");
  }
  // epilogue, stage 2: all CTAs of a tile pair meet at a counter barrier (every CTA of this grid is resident:
  // grid <= #SMs at one CTA per SM), then each reduces its 1/ksplit slice of the pair's tile over the split-K
  // partials in split order (L2-resident) and stores the slice, and its mirror, into every rank's reduction
  // buffer over NVLink.  One reduced tile per (source rank, pair) crosses the links, and the reduction
  // itself is spread over all of the pair's CTAs.  Counters only grow (target = launch sequence * count).
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(P.done_counter + 1 + blockIdx.x, 1u);
    const unsigned int target = P.seq * gridDim.y;
    // bounded by time like the cross-GPU waits (symm.cuh): on a timeout the error word is set (the host reads it
    // with bsc_symm_error() when it collects the block) and the CTA carries on - no trap, the context survives
    unsigned int spins = 0;
    unsigned long long t0 = 0;
    for (;;) {
      unsigned int v;
      asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(P.done_counter + 1 + blockIdx.x) : "memory");
      if (v >= target) break;
      __nanosleep(32);
      if ((++spins & 1023u) == 0) {
        const unsigned long long now = symm_now_ns();
        if (t0 == 0) t0 = now;
        if (now - t0 > SYMM_TIMEOUT_NS) {
          atomicCAS(&g_symm_error, 0, 200);
          break;
        }
      }
    }
  }
  __syncthreads();
  {
    const float* Lp = P.lpart + (size_t)blockIdx.x * (FG_BM * FG_BM);
    const size_t split_stride = (size_t)gridDim.x * (FG_BM * FG_BM);
    const size_t part = (size_t)P.rank * P.ldg * P.ldg;
    const int chunk = (FG_BM * FG_BM + gridDim.y - 1) / gridDim.y;
    const int e0 = split * chunk, e1 = min(FG_BM * FG_BM, e0 + chunk);
    for (int e = e0 + threadIdx.x; e < e1; e += blockDim.x) {
      float acc = 0.f;
      for (unsigned int sp = 0; sp < gridDim.y; sp++) acc += __ldcg(Lp + sp * split_stride + e);
      const int r = ti * FG_BM + e / FG_BM, c = tj * FG_BM + e % FG_BM;
      for (int q = 0; q < P.world; q++) {
        float* G = reinterpret_cast<float*>(P.gpart.ptr[q]) + part;
        G[(size_t)r * P.ldg + c] = acc;
        if (!diag) G[(size_t)c * P.ldg + r] = acc;
      }
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    const unsigned int total = gridDim.x * gridDim.y;
    if (atomicAdd(P.done_counter, 1u) + 1u == P.seq * total) symm_signal_all(P.flags, P.world, P.rank, SLOT_GRAM, P.epoch);
  }
}

// Tail of the fused Krum path.  Identical on every rank: waits until all ranks' partials have landed
// in the local reduction buffer, reduces them in (rank, split) order, takes squared norms from the
// diagonal of the reduced Gram (no norm exchange needed), and runs the Multi-Krum scoring/selection
// over the participating slots rows[0..n).
__global__ void __launch_bounds__(256) krum_select_fused_kernel(const float* __restrict__ Gpart, int nsplit, int ldg,
                                                                const int* __restrict__ rows, int n, int clip,
                                                                const unsigned long long* __restrict__ my_flags,
                                                                int world, unsigned long long epoch,
                                                                float* __restrict__ scores,
                                                                unsigned char* __restrict__ accept, int np2) {
  extern __shared__ float srow[];   // [nwarp][np2] rows, [n] scores, [n] diag
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarp = blockDim.x >> 5;
  float* myrow = srow + (size_t)warp * np2;
  float* ssc = srow + (size_t)nwarp * np2;
  float* sdiag = ssc + n;
  __shared__ int srows[1024];
  if (tid == 0) symm_wait_all(my_flags, world, SLOT_GRAM, epoch);
  __syncthreads();
  for (int i = tid; i < n; i += blockDim.x) {
    const int r = rows[i];
    srows[i] = r;
    float g = 0.f;
    for (int s = 0; s < nsplit; s++) g += Gpart[((size_t)s * ldg + r) * ldg + r];
    sdiag[i] = g;
  }
  __syncthreads();
  const int group = n - clip;
  for (int i = warp; i < n; i += nwarp) {
    const float ni = sdiag[i];
    const int ri = srows[i];
    for (int j = lane; j < np2; j += 32) {
      float v = INFINITY;
      if (j < n) {
        const int rj = srows[j];
        float g = 0.f;
        for (int s = 0; s < nsplit; s++) g += Gpart[((size_t)s * ldg + ri) * ldg + rj];
        v = ni + sdiag[j] - 2.f * g;
      }
      myrow[j] = v;
    }
    __syncwarp();
    for (int k = 2; k <= np2; k <<= 1) {
      for (int j = k >> 1; j > 0; j >>= 1) {
        for (int t = lane; t < np2; t += 32) {
          const int ixj = t ^ j;
          if (ixj > t) {
            const float a = myrow[t], b = myrow[ixj];
            const bool up = ((t & k) == 0);
            if ((a > b) == up) {
              myrow[t] = b;
              myrow[ixj] = a;
            }
          }
        }
        __syncwarp();
      }
    }
    float s = 0.f;
    for (int j = 1 + lane; j < group - 1; j += 32) s += myrow[j];
    s = warp_sum(s);
    if (lane == 0) {
      ssc[i] = s;
      scores[i] = s;
    }
    __syncwarp();
  }
  __syncthreads();
  for (int i = tid; i < n; i += blockDim.x) {
    const float si = ssc[i];
    int rank = 0;
    for (int j = 0; j < n; j++) {
      const float sj = ssc[j];
      rank += (sj < si) || (sj == si && j < i);
    }
    accept[i] = rank < (n - clip) ? 1 : 0;
  }
}

// ---- secure aggregation -------------------------------------------------------------------------------
// Reduce-scatter + exact reconstruction + SGD step + all-gather in one kernel.
//   * A CTA of 256 threads owns SF_CH consecutive chunks of this rank's range.  Their S aggregated shares are
//     contiguous in every rank's ypart ([chunk][S] int64), so the reduce-scatter is `world` coalesced reads of one
//     contiguous SF_CH*S*8-byte run per peer, summed straight into shared memory.
//   * Recovery is the constant-matrix form (secagg_math.cuh: RecoverMat): one thread per (chunk, coefficient), so
//     consecutive threads produce consecutive weights and the all-gather is `world` coalesced stores of one
//     contiguous run of doubles (and floats) per peer.
//   * One elected thread of the grid polls the NVLink-visible flags and republishes the epoch in a local word; the
//     other CTAs poll that word through L2.  The last CTA to finish publishes SLOT_WEIGHTS and waits for the other
//     ranks' publications, so when the kernel retires every rank's new slice is in local memory (no wait launch).
//   * gate: device-resident contributor count (block_plan_kernel); below gate_min the block will be empty and the
//     weights stay as they are - the flag protocol still runs so that the ranks stay in step.
constexpr int SF_CH = 16;
struct SecaggFusedParams {
  PeerTable ypart;   // rank q's partial share sums [nchunk][S] int64 over its local contributors
  PeerTable w64;     // rank q's fp64 master weights [d]
  PeerTable w32;     // rank q's fp32 working copy [d]
  PeerTable flags;
  unsigned int* done_counter;
  unsigned long long* local_ready;
  int* bad;          // local int[4]: {inexact, mismatched shares, first timed-out wait (0 = none), unused}
  const int* gate;
  unsigned long long epoch;
  double scale;
  int world, rank, S, d, c0, nc, gate_min, use_mat;
  unsigned char avail[MAX_SHARES];
  RecoverMat R;      // R.xs is always valid; M/D only when use_mat
};

__global__ void __launch_bounds__(256) secagg_fused_kernel(const __grid_constant__ SecaggFusedParams P) {
  __shared__ long long sy[SF_CH * MAX_SHARES];
  __shared__ long long sc[SF_CH * POLY_SIZE];
  __shared__ RecoverMat sR;   // matrix rows are indexed by the thread's coefficient: shared memory, not the constant bank
  const int tid = threadIdx.x;
  if (tid < (int)(sizeof(RecoverMat) / 4)) reinterpret_cast<int*>(&sR)[tid] = reinterpret_cast<const int*>(&P.R)[tid];
  // this rank's partial share sums were produced by the previous kernel on this stream: publish, then wait
  if (tid == 0) {
    if (blockIdx.x == 0) {
      symm_signal_all(P.flags, P.world, P.rank, SLOT_SHARES, P.epoch);
      if (!symm_wait_all(reinterpret_cast<const unsigned long long*>(P.flags.ptr[P.rank]), P.world, SLOT_SHARES, P.epoch))
        atomicCAS(&P.bad[2], 0, 1 + SLOT_SHARES);   // reported with the share-check counters when the block is collected
      symm_local_publish(P.local_ready, P.epoch);
    } else {
      if (!symm_local_wait(P.local_ready, P.epoch)) atomicCAS(&P.bad[2], 0, 100);
    }
  }
  __syncthreads();
  const bool open = !(P.gate && *P.gate < P.gate_min);
  const int ch0 = blockIdx.x * SF_CH;
  const int nch = min(SF_CH, P.nc - ch0);
  if (open && nch > 0) {
    const size_t base = (size_t)(P.c0 + ch0) * P.S;
    for (int e = tid; e < nch * P.S; e += blockDim.x) {
      uint64_t acc = 0;   // int64 adds commute: the sum is the same on every rank whatever the order
      for (int q = 0; q < P.world; q++) acc += (uint64_t) __ldcv(reinterpret_cast<const long long*>(P.ypart.ptr[q]) + base + e);
      sy[(e / P.S) * MAX_SHARES + (e % P.S)] = (long long)acc;
    }
    __syncthreads();
    int inexact = 0;
    const double* wl = reinterpret_cast<const double*>(P.w64.ptr[P.rank]);
    if (P.use_mat) {
      for (int e = tid; e < nch * POLY_SIZE; e += blockDim.x) {
        const int lc = e / POLY_SIZE, j = e % POLY_SIZE;
        int64_t y[POLY_SIZE];
#pragma unroll
        for (int i = 0; i < POLY_SIZE; i++) y[i] = sy[lc * MAX_SHARES + sR.xs[i]];
        int64_t coef;
        inexact += !recover_coef_mat(sR, y, j, coef);
        sc[e] = coef;
      }
    } else {
      for (int lc = tid; lc < nch; lc += blockDim.x) {   // abscissae without a safe matrix: divided differences
        int lx[POLY_SIZE];
        int64_t ly[POLY_SIZE], coef[POLY_SIZE];
#pragma unroll
        for (int j = 0; j < POLY_SIZE; j++) {
          lx[j] = P.R.xs[j];
          ly[j] = sy[lc * MAX_SHARES + lx[j]];
        }
        inexact += !shamir_recover(lx, ly, coef);
        for (int j = 0; j < POLY_SIZE; j++) sc[lc * POLY_SIZE + j] = coef[j];
      }
    }
    __syncthreads();
    // all-gather: consecutive threads own consecutive coordinates; the owner stores them into every rank's copy
    for (int e = tid; e < nch * POLY_SIZE; e += blockDim.x) {
      const int k = (P.c0 + ch0) * POLY_SIZE + e;
      if (k < P.d) {
        const double nw = wl[k] + (double)sc[e] / P.scale;
        const float nf = (float)nw;
        for (int q = 0; q < P.world; q++) {
          reinterpret_cast<double*>(P.w64.ptr[q])[k] = nw;
          reinterpret_cast<float*>(P.w32.ptr[q])[k] = nf;
        }
      }
    }
    int mism = 0;
    for (int e = tid; e < nch * P.S; e += blockDim.x) {
      const int lc = e / P.S, sidx = e % P.S;
      if (P.avail[sidx]) mism += (shamir_eval((const int64_t*)(sc + lc * POLY_SIZE), sidx) != sy[lc * MAX_SHARES + sidx]);
    }
    if (inexact) atomicAdd(&P.bad[0], inexact);
    if (mism) atomicAdd(&P.bad[1], mism);
  }
  __syncthreads();
  if (tid == 0) {
    __threadfence_system();
    const unsigned int old = atomicAdd(P.done_counter, 1u);
    if (old == gridDim.x - 1) {
      *P.done_counter = 0;
      symm_signal_all(P.flags, P.world, P.rank, SLOT_WEIGHTS, P.epoch);
      if (!symm_wait_all(reinterpret_cast<const unsigned long long*>(P.flags.ptr[P.rank]), P.world, SLOT_WEIGHTS, P.epoch))
        atomicCAS(&P.bad[2], 0, 1 + SLOT_WEIGHTS);
    }
  }
}

__global__ void symm_signal_kernel(PeerTable flags, int world, int rank, int slot, unsigned long long epoch) {
  if (threadIdx.x == 0 && blockIdx.x == 0) symm_signal_all(flags, world, rank, slot, epoch);
}
__global__ void symm_wait_kernel(const unsigned long long* my_flags, int world, int slot, unsigned long long epoch) {
  if (threadIdx.x == 0 && blockIdx.x == 0) symm_wait_all(my_flags, world, slot, epoch);
}

// All-gather of a small per-rank block (the Jacobian full commitments that go into the block, ~1 KB per rank):
// one CTA stores this rank's block into every rank's buffer, publishes the slot's flag and waits for the others'
// flags, so that when the kernel retires the whole gathered array is in local memory.  Replaces the last NCCL
// collective of the fused path (ncclAllGather of the commitment bytes, once per block).
__global__ void __launch_bounds__(256) symm_allgather_rows_kernel(PeerTable dst, PeerTable flags,
                                                                  const unsigned long long* my_flags, int world, int rank,
                                                                  const uint4* __restrict__ src, int n16, size_t dst_off16,
                                                                  int slot, unsigned long long epoch) {
  for (int q = 0; q < world; q++) {
    uint4* d = reinterpret_cast<uint4*>(dst.ptr[q]) + dst_off16;
    for (int i = threadIdx.x; i < n16; i += blockDim.x) d[i] = src[i];
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    symm_signal_all(flags, world, rank, slot, epoch);
    symm_wait_all(my_flags, world, slot, epoch);
  }
}

int make_plane_map(CUtensorMap* map, const void* base, int rows, int ld, int box_rows);   // gram_tc.cu

}  // namespace bsc

using namespace bsc;

extern "C" {

// ---- symmetric memory plumbing (cudaMalloc + CUDA IPC; handles travel over torch.distributed) ----------
int bsc_symm_alloc(size_t bytes, void** out) {
  BSC_CHECK(cudaMalloc(out, bytes));
  BSC_CHECK(cudaMemset(*out, 0, bytes));
  return 0;
}
int bsc_symm_free(void* p) {
  BSC_CHECK(cudaFree(p));
  return 0;
}
int bsc_ipc_get_handle(void* p, unsigned char out[64]) {
  cudaIpcMemHandle_t h;
  BSC_CHECK(cudaIpcGetMemHandle(&h, p));
  static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
  memcpy(out, &h, 64);
  return 0;
}
int bsc_ipc_open_handle(const unsigned char in[64], void** out) {
  cudaIpcMemHandle_t h;
  memcpy(&h, in, 64);
  BSC_CHECK(cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}
int bsc_ipc_close_handle(void* p) {
  BSC_CHECK(cudaIpcCloseMemHandle(p));
  return 0;
}
int bsc_symm_signal(void* const* flag_ptrs, int world, int rank, int slot, unsigned long long epoch, void* stream) {
  PeerTable t;
  for (int q = 0; q < SYMM_MAX_WORLD; q++) t.ptr[q] = q < world ? flag_ptrs[q] : nullptr;
  symm_signal_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(t, world, rank, slot, epoch);
  return finish_launch("symm_signal");
}
int bsc_symm_wait(const void* my_flags, int world, int slot, unsigned long long epoch, void* stream) {
  symm_wait_kernel<<<1, 32, 0, (cudaStream_t)stream>>>((const unsigned long long*)my_flags, world, slot, epoch);
  return finish_launch("symm_wait");
}
// dst_ptrs[q]: rank q's gather buffer [world][bytes_per_rank] (peer-mapped); src: this rank's block (16-byte aligned,
// bytes_per_rank a multiple of 16).  On return of the kernel every rank's block for `epoch` is in the local buffer.
int bsc_symm_allgather_rows(void* const* dst_ptrs, void* const* flag_ptrs, const void* my_flags, int world, int rank,
                            const void* src, long long bytes_per_rank, int slot, unsigned long long epoch, void* stream) {
  if (world < 1 || world > SYMM_MAX_WORLD || bytes_per_rank % 16 != 0 || ((uintptr_t)src & 15)) return -1;
  PeerTable d, f;
  for (int q = 0; q < SYMM_MAX_WORLD; q++) {
    d.ptr[q] = q < world ? dst_ptrs[q] : nullptr;
    f.ptr[q] = q < world ? flag_ptrs[q] : nullptr;
  }
  const int n16 = (int)(bytes_per_rank / 16);
  symm_allgather_rows_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(d, f, (const unsigned long long*)my_flags, world, rank,
                                                                  (const uint4*)src, n16, (size_t)rank * n16, slot, epoch);
  return finish_launch("symm_allgather_rows");
}

// Fused all-gather + Gram.  plane_hi/lo[q]: rank q's [Sr][ldh] bf16 planes (peer-mapped), gpart[q]: rank q's
// [world][ldg][ldg] fp32 reduction buffer (slot r receives rank r's reduced partial), flags[q]: rank q's flag
// array, lpart: local scratch [ksplit][npairs][128*128] fp32, done_counter: local uint32[1 + npairs] (zeroed once;
// the counters only grow), seq: 1-based count of launches of this kernel on this rank.
// world*Sr must be a multiple of 128; Sr a multiple of 8 that divides 128, or a multiple of 128 (more than 128 peers
// per rank).
int bsc_krum_gram_fused(void* const* plane_hi, void* const* plane_lo, void* const* gpart, void* const* flags, int world,
                        int rank, int Sr, int ldh, int ksplit, int ldg, float* lpart, unsigned int* done_counter,
                        unsigned int seq, unsigned long long epoch, void* stream) {
  if (world < 1 || world > SYMM_MAX_WORLD || Sr % 8 != 0 || (FG_BM % Sr != 0 && Sr % FG_BM != 0) ||
      (world * Sr) % FG_BM != 0 || ldh % FG_BK != 0)
    return -1;
  const int box_rows = Sr < FG_BM ? Sr : FG_BM;
  GramFusedParams P;
  memset(&P, 0, sizeof(P));
  for (int q = 0; q < world; q++) {
    int rc = make_plane_map(&P.hi[q], plane_hi[q], Sr, ldh, box_rows);
    if (rc) return rc;
    rc = make_plane_map(&P.lo[q], plane_lo[q], Sr, ldh, box_rows);
    if (rc) return rc;
    P.gpart.ptr[q] = gpart[q];
    P.flags.ptr[q] = flags[q];
  }
  P.lpart = lpart;
  P.done_counter = done_counter;
  P.seq = seq;
  P.epoch = epoch;
  P.world = world;
  P.rank = rank;
  P.Sr = Sr;
  P.ntile = world * Sr / FG_BM;
  P.ldg = ldg;
  if (ldg != P.ntile * FG_BM) return -1;
  const int kb_total = ldh / FG_BK;
  const int per_rank = (kb_total + world - 1) / world;     // this rank's slice of the K dimension
  P.kb_lo = min(kb_total, rank * per_rank);
  P.kb_cnt = min(kb_total, P.kb_lo + per_rank) - P.kb_lo;
  P.ksplit = ksplit;
  P.kb_per = (max(P.kb_cnt, 1) + ksplit - 1) / ksplit;
  static bool configured = false;
  if (!configured) {
    BSC_CHECK(cudaFuncSetAttribute(krum_gram_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FG_SMEM));
    configured = true;
  }
  dim3 grid(P.ntile * (P.ntile + 1) / 2, ksplit);
  // The in-kernel split barrier needs every CTA of the grid resident at once: the limit is what THIS device can hold
  // of this kernel (fewer SMs under MIG / green contexts, another SKU), not a constant.
  static int max_resident[64] = {0};
  int dev = 0;
  BSC_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return -2;
  if (!max_resident[dev]) {
    int sms = 0, per_sm = 0;
    BSC_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    BSC_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, krum_gram_fused_kernel, FG_THREADS, FG_SMEM));
    max_resident[dev] = sms * per_sm;
  }
  if ((int)(grid.x * grid.y) > max_resident[dev]) return -5;
  // The counters only grow and their targets are seq * (grid extent): a launch with another grid shape on the same
  // counter buffer would wait for a count that never comes.  Allowed only together with seq == 1, i.e. after the
  // caller has zeroed the counters and restarted its launch count.
  {
    static std::mutex mu;
    static std::unordered_map<const void*, unsigned long long> shapes;
    const unsigned long long shape = ((unsigned long long)grid.x << 32) | grid.y;
    std::lock_guard<std::mutex> lk(mu);
    auto it = shapes.find(done_counter);
    if (it == shapes.end() || seq == 1) shapes[done_counter] = shape;
    else if (it->second != shape) return -6;
  }
  krum_gram_fused_kernel<<<grid, FG_THREADS, FG_SMEM, (cudaStream_t)stream>>>(P);
  return finish_launch("krum_gram_fused");
}

int bsc_krum_select_fused(const float* gpart_local, int nsplit, int ldg, const int* rows, int n, int clip,
                          const void* my_flags, int world, unsigned long long epoch, float* scores,
                          unsigned char* accept, void* stream) {
  if (n > 1024 || n < 3) return -1;
  int np2 = 32;
  while (np2 < n) np2 <<= 1;
  const size_t smem = ((size_t)8 * np2 + 2 * n) * sizeof(float);
  krum_select_fused_kernel<<<1, 256, smem, (cudaStream_t)stream>>>(gpart_local, nsplit, ldg, rows, n, clip,
                                                                    (const unsigned long long*)my_flags, world, epoch,
                                                                    scores, accept, np2);
  return finish_launch("krum_select_fused");
}

// Fused reduce-scatter + reconstruct + SGD step + all-gather for this rank's chunk range [c0, c0+nc).
//   rm          : host RecoverMat (secagg_math.cuh) for the abscissae in use; use_mat = 0 keeps only rm->xs and
//                 recovers by divided differences
//   local_ready : local uint64 word (zeroed once), done_counter: local uint32 (zeroed once)
//   gate        : device int, may be null: the step is applied only when *gate >= gate_min
// When the kernel retires, every rank's new weight slice has landed in this rank's w64 / w32.
int bsc_secagg_fused(void* const* ypart, void* const* w64, void* const* w32, void* const* flags, int world, int rank,
                     int S, int d, int c0, int nc, const void* rm, int use_mat, const unsigned char* avail_host, double scale,
                     unsigned int* done_counter, unsigned long long* local_ready, int* bad, const int* gate, int gate_min,
                     unsigned long long epoch, void* stream) {
  if (world < 1 || world > SYMM_MAX_WORLD || S > MAX_SHARES) return -1;
  SecaggFusedParams P;
  memset(&P, 0, sizeof(P));
  for (int q = 0; q < world; q++) {
    P.ypart.ptr[q] = ypart[q];
    P.w64.ptr[q] = w64[q];
    P.w32.ptr[q] = w32[q];
    P.flags.ptr[q] = flags[q];
  }
  P.done_counter = done_counter;
  P.local_ready = local_ready;
  P.bad = bad;
  P.gate = gate;
  P.gate_min = gate_min;
  P.epoch = epoch;
  P.scale = scale;
  P.world = world;
  P.rank = rank;
  P.S = S;
  P.d = d;
  P.c0 = c0;
  P.nc = nc;
  P.use_mat = use_mat;
  memcpy(&P.R, rm, sizeof(P.R));
  for (int s = 0; s < S; s++) P.avail[s] = avail_host ? avail_host[s] : 1;
  const int blocks = max(1, (nc + SF_CH - 1) / SF_CH);
  secagg_fused_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(P);
  return finish_launch("secagg_fused");
}

// first cross-GPU wait that timed out since the last call (0 = none); clears the record
int bsc_symm_error() {
  int v = 0, zero = 0;
  if (cudaMemcpyFromSymbol(&v, g_symm_error, sizeof(int)) != cudaSuccess) return -1;
  if (v) cudaMemcpyToSymbol(g_symm_error, &zero, sizeof(int));
  return v;
}

}  // extern "C"
