// Krum Gram matrix on the 5th-generation tensor cores: tcgen05.mma with TMEM accumulators, operands
// staged by TMA (cp.async.bulk.tensor) into 128B-swizzled shared memory, warp-specialised
// producer / MMA-issuer / epilogue roles, split-K over CTAs with a deterministic fixed-order
// reduction in krum_select_kernel.
//
// What it computes (ML/Pytorch/client_obj.py:132-143, the np.dot(X, X.T) inside get_krum_scores):
//   G = U U^T for the noised updates U [n, d].  U arrives as two bf16 planes (hi = bf16(u),
//   lo = bf16(u - hi), written by noise_mix_kernel) and the product is formed as
//   hi*hi^T + hi*lo^T + lo*hi^T with fp32 accumulation in TMEM ("bf16x3"), which keeps ~16 mantissa
//   bits per factor: Krum ranks peers by small differences of ||a||^2 + ||b||^2 - 2 a.b, plain bf16
//   would not survive the cancellation.
//
// Tiling: 128x128 output tile per CTA (UMMA M=128, N=128, K=16), BLOCK_K = 64 bf16 = one 128-byte
// swizzle atom, 3-stage mbarrier ring.  Only tiles with ti <= tj are computed; the epilogue also
// writes the mirrored tile.  Diagonal tiles reuse the A operand tiles as B (half the traffic — the
// whole problem for n <= 128, i.e. the 100-peer headline configuration).
#include "common.cuh"
#include "tc_ptx.cuh"
#include <cuda.h>
#include <cuda_bf16.h>

namespace bsc {

constexpr int GT_BM = 128;
constexpr int GT_BK = 64;
constexpr int GT_STAGES = 3;
constexpr int GT_TILE_BYTES = GT_BM * GT_BK * 2;  // 16 KiB: 128 rows x 128 B
constexpr int GT_STAGE_BYTES = 4 * GT_TILE_BYTES;
constexpr int GT_THREADS = 192;                   // warp0 TMA, warp1 MMA+TMEM, warps 2-5 epilogue
constexpr int GT_TMEM_COLS = 128;
constexpr size_t GT_SMEM = (size_t)GT_STAGES * GT_STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;

// mbarrier / TMA / UMMA descriptor / tcgen05 helpers: tc_ptx.cuh (shared with fused.cu and kernels_ml.cu)

//   tm_hi / tm_lo : 2-D tensor maps over the bf16 planes, dims (ldh, rows), box (64, 128), SWIZZLE_128B
//   Gpart         : [gridDim.y][ldg][ldg] fp32, ldg = ntile*128
//   kb_total      : number of 64-wide k-blocks ; kb_per : k-blocks per split (gridDim.y * kb_per >= kb_total)
__global__ void __launch_bounds__(GT_THREADS, 1)
    gram_tc_kernel(const __grid_constant__ CUtensorMap tm_hi, const __grid_constant__ CUtensorMap tm_lo,
                   float* __restrict__ Gpart, int ldg, int ntile, int kb_total, int kb_per) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t tiles = (raw + 1023u) & ~1023u;              // SWIZZLE_128B needs 1024-B alignment
  const uint32_t bars = tiles + GT_STAGES * GT_STAGE_BYTES;   // full[3], empty[3], tmem_full, tmem slot
  const uint32_t bar_full = bars, bar_empty = bars + 8 * GT_STAGES, bar_tmem = bars + 16 * GT_STAGES;
  const uint32_t tmem_slot = bar_tmem + 8;
  uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(smem_raw + (tmem_slot - raw));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // decode the upper-triangular tile pair
  int ti = 0, tj = 0;
  {
    int p = blockIdx.x;
    for (ti = 0; ti < ntile; ti++) {
      const int rowlen = ntile - ti;
      if (p < rowlen) {
        tj = ti + p;
        break;
      }
      p -= rowlen;
    }
  }
  const bool diag = (ti == tj);
  const int split = blockIdx.y;
  const int kb0 = split * kb_per;
  const int nkb = min(kb_total, kb0 + kb_per) - kb0;   // host guarantees nkb >= 1

  if (threadIdx.x == 0) {
    for (int s = 0; s < GT_STAGES; s++) {
      mbar_init(bar_full + 8 * s, 1);
      mbar_init(bar_empty + 8 * s, 1);
    }
    mbar_init(bar_tmem, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(GT_TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      for (int kb = 0; kb < nkb; kb++) {
        const int s = kb % GT_STAGES;
        const uint32_t ph = (uint32_t)(kb / GT_STAGES) & 1u;
        mbar_wait(bar_empty + 8 * s, ph ^ 1u);
        const uint32_t st = tiles + s * GT_STAGE_BYTES;
        mbar_expect_tx(bar_full + 8 * s, diag ? 2 * GT_TILE_BYTES : 4 * GT_TILE_BYTES);
        const int kx = (kb0 + kb) * GT_BK;
        tma_load_2d(st, &tm_hi, kx, ti * GT_BM, bar_full + 8 * s);
        tma_load_2d(st + GT_TILE_BYTES, &tm_lo, kx, ti * GT_BM, bar_full + 8 * s);
        if (!diag) {
          tma_load_2d(st + 2 * GT_TILE_BYTES, &tm_hi, kx, tj * GT_BM, bar_full + 8 * s);
          tma_load_2d(st + 3 * GT_TILE_BYTES, &tm_lo, kx, tj * GT_BM, bar_full + 8 * s);
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (one thread) =====
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_bf16_128x128();
      for (int kb = 0; kb < nkb; kb++) {
        const int s = kb % GT_STAGES;
        const uint32_t ph = (uint32_t)(kb / GT_STAGES) & 1u;
        mbar_wait(bar_full + 8 * s, ph);
        tc_fence_after();
        const uint32_t st = tiles + s * GT_STAGE_BYTES;
        const uint64_t a_hi = umma_desc_sw128(st), a_lo = umma_desc_sw128(st + GT_TILE_BYTES);
        const uint64_t b_hi = diag ? a_hi : umma_desc_sw128(st + 2 * GT_TILE_BYTES);
        const uint64_t b_lo = diag ? a_lo : umma_desc_sw128(st + 3 * GT_TILE_BYTES);
#pragma unroll
        for (int k = 0; k < GT_BK / 16; k++) {
          const uint64_t off = (uint64_t)((k * 16 * 2) >> 4);   // advance 32 B inside the swizzle atom
          umma_bf16(tmem_base, a_hi + off, b_hi + off, idesc, (kb > 0 || k > 0) ? 1u : 0u);
          umma_bf16(tmem_base, a_hi + off, b_lo + off, idesc, 1u);
          umma_bf16(tmem_base, a_lo + off, b_hi + off, idesc, 1u);
        }
        umma_commit(bar_empty + 8 * s);   // frees the smem stage once these MMAs have read it
      }
      umma_commit(bar_tmem);              // accumulator complete
    }
  } else {
    // ===== epilogue: TMEM -> registers -> global (direct + mirrored tile) =====
    const int q = warp & 3;               // TMEM lane quarter this warp may touch
    mbar_wait(bar_tmem, 0);
    tc_fence_after();
    float* G = Gpart + (size_t)split * ldg * ldg;
    const int row = ti * GT_BM + q * 32 + lane;
#pragma unroll 1
    for (int cc = 0; cc < GT_BM / 32; cc++) {
      uint32_t v[32];
      tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(cc * 32), v);
      const int col0 = tj * GT_BM + cc * 32;
      float4* dst = reinterpret_cast<float4*>(G + (size_t)row * ldg + col0);
#pragma unroll
      for (int c = 0; c < 8; c++)
        dst[c] = make_float4(__uint_as_float(v[4 * c]), __uint_as_float(v[4 * c + 1]), __uint_as_float(v[4 * c + 2]),
                             __uint_as_float(v[4 * c + 3]));
      if (!diag) {
#pragma unroll
        for (int c = 0; c < 32; c++) G[(size_t)(col0 + c) * ldg + row] = __uint_as_float(v[c]);   // coalesced over lanes
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(GT_TMEM_COLS) : "memory");
  }
}

// ---------------------------------------------------------------------------------------------------
// Batched model evaluation on the tensor cores (K14 / K15: test error, attack rate, RONI):
//   logits[n][m][c] = X[n] . W_m[c] + b_m[c]   for N samples and M candidate weight vectors,
//   wrong[m] = #{n : argmax_c logits[n][m][c] != y[n]}
// (ML/Pytorch/client.py:136-172, client_obj.py:100-112).  It is an A*B^T GEMM with A = samples [N, F] and
// B = candidate class rows [M*16, F] (classes padded to 16 per candidate so that a 128-column tile holds exactly 8
// candidates), both K-major bf16x3 planes like the Gram kernel, same TMA / mbarrier / tcgen05 pipeline.  The
// epilogue never writes the logits: each TMEM lane is one sample, it adds the bias, takes the arg-max of each
// 16-column group and counts disagreements with the label; one integer atomicAdd per (warp, candidate).
constexpr int EV_CPAD = 16;   // class slots per candidate
__global__ void __launch_bounds__(GT_THREADS, 1)
    eval_tc_kernel(const __grid_constant__ CUtensorMap ta_hi, const __grid_constant__ CUtensorMap ta_lo,
                   const __grid_constant__ CUtensorMap tb_hi, const __grid_constant__ CUtensorMap tb_lo,
                   const float* __restrict__ bias, const int* __restrict__ y, int* __restrict__ wrong, int N, int M,
                   int kb_total) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t tiles = (raw + 1023u) & ~1023u;
  const uint32_t bars = tiles + GT_STAGES * GT_STAGE_BYTES;
  const uint32_t bar_full = bars, bar_empty = bars + 8 * GT_STAGES, bar_tmem = bars + 16 * GT_STAGES;
  const uint32_t tmem_slot = bar_tmem + 8;
  uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(smem_raw + (tmem_slot - raw));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tm = blockIdx.x, tn = blockIdx.y;   // sample tile, candidate tile (8 candidates)

  if (threadIdx.x == 0) {
    for (int s = 0; s < GT_STAGES; s++) {
      mbar_init(bar_full + 8 * s, 1);
      mbar_init(bar_empty + 8 * s, 1);
    }
    mbar_init(bar_tmem, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(GT_TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  if (warp == 0) {
    if (lane == 0) {
      for (int kb = 0; kb < kb_total; kb++) {
        const int s = kb % GT_STAGES;
        const uint32_t ph = (uint32_t)(kb / GT_STAGES) & 1u;
        mbar_wait(bar_empty + 8 * s, ph ^ 1u);
        const uint32_t st = tiles + s * GT_STAGE_BYTES;
        mbar_expect_tx(bar_full + 8 * s, 4 * GT_TILE_BYTES);
        const int kx = kb * GT_BK;
        tma_load_2d(st, &ta_hi, kx, tm * GT_BM, bar_full + 8 * s);
        tma_load_2d(st + GT_TILE_BYTES, &ta_lo, kx, tm * GT_BM, bar_full + 8 * s);
        tma_load_2d(st + 2 * GT_TILE_BYTES, &tb_hi, kx, tn * GT_BM, bar_full + 8 * s);
        tma_load_2d(st + 3 * GT_TILE_BYTES, &tb_lo, kx, tn * GT_BM, bar_full + 8 * s);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_bf16_128x128();
      for (int kb = 0; kb < kb_total; kb++) {
        const int s = kb % GT_STAGES;
        const uint32_t ph = (uint32_t)(kb / GT_STAGES) & 1u;
        mbar_wait(bar_full + 8 * s, ph);
        tc_fence_after();
        const uint32_t st = tiles + s * GT_STAGE_BYTES;
        const uint64_t a_hi = umma_desc_sw128(st), a_lo = umma_desc_sw128(st + GT_TILE_BYTES);
        const uint64_t b_hi = umma_desc_sw128(st + 2 * GT_TILE_BYTES), b_lo = umma_desc_sw128(st + 3 * GT_TILE_BYTES);
#pragma unroll
        for (int k = 0; k < GT_BK / 16; k++) {
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
    const int q = warp & 3;
    mbar_wait(bar_tmem, 0);
    tc_fence_after();
    const int row = tm * GT_BM + q * 32 + lane;
    const int label = row < N ? y[row] : -1;
#pragma unroll 1
    for (int cc = 0; cc < GT_BM / 32; cc++) {
      uint32_t v[32];
      tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(cc * 32), v);
#pragma unroll
      for (int h = 0; h < 2; h++) {                       // two candidates per 32-column chunk
        const int m = tn * (GT_BM / EV_CPAD) + cc * 2 + h;
        const float* b = bias + (size_t)m * EV_CPAD;
        float best = -INFINITY;
        int arg = 0;
        if (m < M) {
#pragma unroll
          for (int c = 0; c < EV_CPAD; c++) {
            const float l = __uint_as_float(v[h * EV_CPAD + c]) + b[c];   // padded classes carry a -1e30 bias
            if (l > best) {
              best = l;
              arg = c;
            }
          }
        }
        const unsigned bad = __ballot_sync(0xffffffffu, m < M && row < N && arg != label);
        if (lane == 0 && bad) atomicAdd(&wrong[m], __popc(bad));
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(GT_TMEM_COLS) : "memory");
  }
}

// fp32 rows -> bf16x3 planes, zero padded to [rows_pad][ldk].  Row r = (o, i) with o = r / inner, i = r % inner reads
// src + o*ld_outer + i*ld_inner (+ base + i*ld_inner when base != null) and lands in plane row o*inner_pad + i.
// Used for the sample matrix (inner = 1) and for candidate weights (inner = C classes, inner_pad = 16), for which it
// also emits the bias vector (value at column `cols` of the class row block: flat layout W[C][F] then b[C]).
__global__ void __launch_bounds__(256) split_planes_kernel(const float* __restrict__ src, const float* __restrict__ base,
                                                           int outer, int inner, int inner_pad, long long ld_outer,
                                                           int ld_inner, int cols, __nv_bfloat16* __restrict__ hi,
                                                           __nv_bfloat16* __restrict__ lo, int ldk, int rows_pad,
                                                           float* __restrict__ bias_out, int bias_off) {
  const int r = blockIdx.x;               // plane row
  if (r >= rows_pad) return;
  const int o = r / inner_pad, i = r % inner_pad;
  const bool live = o < outer && i < inner;
  const float* s = src + (size_t)o * ld_outer + (size_t)i * ld_inner;
  const float* b = base ? base + (size_t)i * ld_inner : nullptr;
  for (int k = threadIdx.x; k < ldk; k += blockDim.x) {
    float x = 0.f;
    if (live && k < cols) x = s[k] + (b ? b[k] : 0.f);
    const __nv_bfloat16 h = __float2bfloat16(x);
    hi[(size_t)r * ldk + k] = h;
    lo[(size_t)r * ldk + k] = __float2bfloat16(x - __bfloat162float(h));
  }
  if (bias_out && threadIdx.x == 0) {
    float bv = -1e30f;
    if (live) bv = src[(size_t)o * ld_outer + bias_off + i] + (base ? base[bias_off + i] : 0.f);
    bias_out[r] = o < outer ? bv : 0.f;
  }
}

typedef CUresult (*PFN_tmapEncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                        const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                        CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_tmapEncodeTiled get_encode() {
  static PFN_tmapEncodeTiled fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess)
      return nullptr;
    fn = (PFN_tmapEncodeTiled)p;
  }
  return fn;
}

// bf16 [rows][ld] row-major plane -> tensor map with a (64 x box_rows) box, 128-byte swizzle, zero OOB fill
int make_plane_map(CUtensorMap* map, const void* base, int rows, int ld, int box_rows) {
  PFN_tmapEncodeTiled enc = get_encode();
  if (!enc) return -2;
  cuuint64_t dims[2] = {(cuuint64_t)ld, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {(cuuint32_t)GT_BK, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : -3;
}

}  // namespace bsc

using namespace bsc;

extern "C" {

// G partials of hi/lo planes [rows][ldh] (ldh % 64 == 0, 16-byte aligned base), first n rows valid.
// Gpart: [ksplit][ldg][ldg] with ldg = ceil(n/128)*128; every split must own >= 1 k-block.
int bsc_gram_tc(const void* hi, const void* lo, int n, int rows, int ldh, int ksplit, float* Gpart, int ldg,
                void* stream) {
  if (ldh % GT_BK != 0 || n < 1 || ksplit < 1) return -1;
  const int ntile = (n + GT_BM - 1) / GT_BM;
  if (ldg != ntile * GT_BM) return -1;
  const int kb_total = ldh / GT_BK;
  const int kb_per = (kb_total + ksplit - 1) / ksplit;
  if ((ksplit - 1) * kb_per >= kb_total) return -4;   // an empty split would leave its partial unwritten
  CUtensorMap mh, ml;
  int rc = make_plane_map(&mh, hi, rows, ldh, GT_BM);
  if (rc) return rc;
  rc = make_plane_map(&ml, lo, rows, ldh, GT_BM);
  if (rc) return rc;
  static bool configured = false;
  if (!configured) {
    BSC_CHECK(cudaFuncSetAttribute(gram_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GT_SMEM));
    configured = true;
  }
  dim3 grid(ntile * (ntile + 1) / 2, ksplit);
  gram_tc_kernel<<<grid, GT_THREADS, GT_SMEM, (cudaStream_t)stream>>>(mh, ml, Gpart, ldg, ntile, kb_total, kb_per);
  return finish_launch("gram_tc");
}

// fp32 rows -> zero-padded bf16x3 planes [rows_pad][ldk] (+ bias vector); see split_planes_kernel.
int bsc_split_planes(const float* src, const float* base, int outer, int inner, int inner_pad, long long ld_outer,
                     int ld_inner, int cols, void* hi, void* lo, int ldk, int rows_pad, float* bias_out, int bias_off,
                     void* stream) {
  if (ldk % GT_BK != 0 || cols > ldk || inner > inner_pad || rows_pad < 1) return -1;
  split_planes_kernel<<<rows_pad, 256, 0, (cudaStream_t)stream>>>(src, base, outer, inner, inner_pad, ld_outer, ld_inner,
                                                                  cols, (__nv_bfloat16*)hi, (__nv_bfloat16*)lo, ldk,
                                                                  rows_pad, bias_out, bias_off);
  return finish_launch("split_planes");
}

// wrong[m] += #{n < N : argmax_c (A[n] . B[m*16 + c] + bias[m*16 + c]) != y[n]} for m < M.  A planes: [a_rows_pad][ldk]
// (a_rows_pad % 128 == 0, >= N), B planes: [b_rows_pad][ldk] (b_rows_pad % 128 == 0, >= 16*M).  Caller zero-fills wrong.
int bsc_eval_tc(const void* a_hi, const void* a_lo, int a_rows_pad, const void* b_hi, const void* b_lo, int b_rows_pad,
                int ldk, const float* bias, const int* y, int* wrong, int N, int M, void* stream) {
  if (ldk % GT_BK != 0 || a_rows_pad % GT_BM != 0 || b_rows_pad % GT_BM != 0 || a_rows_pad < N ||
      b_rows_pad < EV_CPAD * M || N < 1 || M < 1)
    return -1;
  CUtensorMap ma_hi, ma_lo, mb_hi, mb_lo;
  int rc = make_plane_map(&ma_hi, a_hi, a_rows_pad, ldk, GT_BM);
  if (!rc) rc = make_plane_map(&ma_lo, a_lo, a_rows_pad, ldk, GT_BM);
  if (!rc) rc = make_plane_map(&mb_hi, b_hi, b_rows_pad, ldk, GT_BM);
  if (!rc) rc = make_plane_map(&mb_lo, b_lo, b_rows_pad, ldk, GT_BM);
  if (rc) return rc;
  static bool configured = false;
  if (!configured) {
    BSC_CHECK(cudaFuncSetAttribute(eval_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GT_SMEM));
    configured = true;
  }
  dim3 grid((N + GT_BM - 1) / GT_BM, (M + GT_BM / EV_CPAD - 1) / (GT_BM / EV_CPAD));
  eval_tc_kernel<<<grid, GT_THREADS, GT_SMEM, (cudaStream_t)stream>>>(ma_hi, ma_lo, mb_hi, mb_lo, bias, y, wrong, N, M,
                                                                      ldk / GT_BK);
  return finish_launch("eval_tc");
}

}  // extern "C"
