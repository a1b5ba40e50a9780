// Batched-over-peers local training steps, DP-noise mixing, evaluation, and the SIMT Krum kernels.
//
// Reference behaviour reproduced (see each kernel): ML/Pytorch/client.py:38-65 (one minibatch,
// CrossEntropyLoss, clip_grad_norm 100, flatten in named_parameters order, update = -grad),
// ML/code/logistic_model.py:92-140 (L2-regularised logistic step, delta = -alpha*g + own noise),
// ML/Pytorch/client_obj.py:97-98 + DistSys/main.go:1530-1537, 1653-1656 (noiser noise, averaged),
// ML/Pytorch/client_obj.py:114-143 (Multi-Krum scores and selection),
// ML/Pytorch/client.py:136-172 (argmax error / attack rate).
//
// The reference runs these once per peer process on the CPU; here one CTA handles one simulated
// peer and the whole population of a GPU is one launch.
#include "common.cuh"
#include "secagg_math.cuh"
#include "symm.cuh"
#include "tc_ptx.cuh"
#include <cuda_bf16.h>
#include <mutex>

namespace bsc {

unsigned long long g_launch_count = 0;
static bool g_softmax_simt = false;

// ------------------------------------------------------------------------------------------------
// K1: softmax regression, one CTA per peer.
//   w      : [C*F + C] fp32  (weight row-major [C][F], then bias[C]) — the layout of
//            SoftmaxModel.named_parameters() (ML/Pytorch/softmax_model.py)
//   X      : [P][Ns][F] fp32 shard (or [P][B][F] minibatch when idx == nullptr, Ns == B)
//   y      : [P][Ns] int32 labels
//   idx    : [P][B] int32 rows of the shard that form this block's minibatch (may be null)
//   U      : [P][ldu] fp32 out, U = -clip(grad)
//   loss   : [P] fp32 out (mean CE of the minibatch)
// dynamic smem: B*F (minibatch) + B*C (logits/dlogits) + 64 floats
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) local_softmax_step_kernel(
    const float* __restrict__ w, const float* __restrict__ X, const int* __restrict__ y,
    const int* __restrict__ idx, float* __restrict__ U, float* __restrict__ loss, int Ns, int B, int F,
    int C, int ldu, float clip) {
  extern __shared__ float smem[];
  float* sX = smem;              // [B][F]
  float* sL = sX + B * F;        // [B][C]
  float* red = sL + B * C;       // 40
  __shared__ int sy[64];
  const int p = blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarp = blockDim.x >> 5;

  // stage the minibatch (vectorised when F % 4 == 0)
  for (int b = 0; b < B; b++) {
    const int row = idx ? idx[p * B + b] : b;
    const float* src = X + ((size_t)p * Ns + row) * F;
    if ((F & 3) == 0) {
      const float4* s4 = reinterpret_cast<const float4*>(src);
      float4* d4 = reinterpret_cast<float4*>(sX + b * F);
      for (int k = tid; k < (F >> 2); k += blockDim.x) d4[k] = s4[k];
    } else {
      for (int k = tid; k < F; k += blockDim.x) sX[b * F + k] = src[k];
    }
    if (tid == 0) sy[b] = y[(size_t)p * Ns + row];
  }
  __syncthreads();

  // logits[b][c] = X[b] . W[c] + bias[c]
  for (int pair = warp; pair < B * C; pair += nwarp) {
    const int b = pair / C, c = pair % C;
    const float* wr = w + (size_t)c * F;
    float acc = 0.f;
    for (int k = lane; k < F; k += 32) acc = fmaf(sX[b * F + k], wr[k], acc);
    acc = warp_sum(acc);
    if (lane == 0) sL[b * C + c] = acc + w[(size_t)C * F + c];
  }
  __syncthreads();

  // softmax + dlogits = (softmax - onehot)/B ; mean CE loss
  float myloss = 0.f;
  if (tid < B) {
    float m = -INFINITY;
    for (int c = 0; c < C; c++) m = fmaxf(m, sL[tid * C + c]);
    float s = 0.f;
    for (int c = 0; c < C; c++) s += expf(sL[tid * C + c] - m);
    const float lse = m + logf(s);
    const int lab = sy[tid];
    myloss = (lse - sL[tid * C + lab]) / (float)B;
    for (int c = 0; c < C; c++) {
      float pr = expf(sL[tid * C + c] - lse);
      sL[tid * C + c] = (pr - (c == lab ? 1.f : 0.f)) / (float)B;
    }
  }
  float tl = block_sum(myloss, red);
  if (tid == 0 && loss) loss[p] = tl;

  // pass 1: squared global norm of (gW, gb)
  float nrm = 0.f;
  for (int k = tid; k < F; k += blockDim.x) {
    for (int c = 0; c < C; c++) {
      float g = 0.f;
      for (int b = 0; b < B; b++) g = fmaf(sL[b * C + c], sX[b * F + k], g);
      nrm = fmaf(g, g, nrm);
    }
  }
  if (tid < C) {
    float g = 0.f;
    for (int b = 0; b < B; b++) g += sL[b * C + tid];
    nrm = fmaf(g, g, nrm);
  }
  const float total = sqrtf(block_sum(nrm, red));
  // torch.nn.utils.clip_grad_norm: coef = max_norm / (total + 1e-6), applied when < 1
  float coef = clip / (total + 1e-6f);
  if (coef > 1.f) coef = 1.f;

  // pass 2: write U = -coef * grad
  float* out = U + (size_t)p * ldu;
  for (int k = tid; k < F; k += blockDim.x) {
    for (int c = 0; c < C; c++) {
      float g = 0.f;
      for (int b = 0; b < B; b++) g = fmaf(sL[b * C + c], sX[b * F + k], g);
      out[(size_t)c * F + k] = -coef * g;
    }
  }
  if (tid < C) {
    float g = 0.f;
    for (int b = 0; b < B; b++) g += sL[b * C + tid];
    out[(size_t)C * F + tid] = -coef * g;
  }
  for (int k = C * F + C + tid; k < ldu; k += blockDim.x) out[k] = 0.f;  // zero the padding
}

// ------------------------------------------------------------------------------------------------
// K1 on the tensor cores.  One CTA takes PPC peers (PPC * B <= 32 minibatch rows) and runs three phases:
//   1. all threads gather the rows (fp32, shard-resident) and the weight matrix, split each value into bf16 hi + lo
//      and store both planes straight into shared memory in the K-major 128-byte-swizzled layout tcgen05.mma reads
//      (the whole K extent stays resident: 32 rows x 832 columns per plane);
//   2. one thread issues the [128 x 16] += A[128 x 16] * W^T[16 x 16] UMMAs for every 16-column K step, three per
//      step (hi*hi, hi*lo, lo*hi - the Gram kernel's fp32-grade split), accumulating the logits in TMEM.  The A tile
//      is declared 128 rows tall; rows past PPC*B alias the next K blocks' bytes and produce TMEM lanes nobody reads;
//   3. warp 0 pulls its 32 TMEM lanes (one minibatch row each), adds the bias and turns the logits into
//      (softmax - onehot) / B and the row's loss; then every thread accumulates its columns of the PPC gradients in
//      registers (exact fp32, rows re-read from L2), the global-norm clip factor is reduced per peer, and U is written.
// Same inputs / outputs as local_softmax_step_kernel; C is fixed at 10 classes, F <= 1024 and F % 4 == 0.
// ------------------------------------------------------------------------------------------------
constexpr int LS_C = 10;            // classes
constexpr int LS_NP = 16;           // padded N of the MMA
constexpr int LS_ROWS = 32;         // minibatch rows per CTA (one TMEM lane quarter)
constexpr int LS_THREADS = 512;
constexpr int LS_KI = 2;            // columns per thread in the gradient phase (F <= LS_KI * LS_THREADS)
constexpr int LS_A_KB_BYTES = LS_ROWS * 128;   // one 64-column K block of the A plane
constexpr int LS_B_KB_BYTES = LS_NP * 128;
constexpr int LS_A_TAIL = (128 - LS_ROWS) * 128;   // what the 128-row descriptor of the last K block reaches past its rows

// umma_idesc_bf16_128xN, tmem_ld16, split4, sw128_off: tc_ptx.cuh

template <int PPC>
__global__ void __launch_bounds__(LS_THREADS, 1) local_softmax_tc_kernel(
    const float* __restrict__ w, const float* __restrict__ X, const int* __restrict__ y,
    const int* __restrict__ idx, float* __restrict__ U, float* __restrict__ loss, int P, int Ns, int B, int F,
    int ldu, float clip, int nkb) {
  extern __shared__ uint8_t ls_smem[];
  __shared__ const float* sRow[LS_ROWS];
  __shared__ int sLab[LS_ROWS];
  __shared__ float sDL[LS_ROWS][LS_NP];
  __shared__ float sLoss[LS_ROWS];
  __shared__ float sRed[PPC][LS_THREADS / 32];
  __shared__ float sCoef[PPC];
  __shared__ __align__(8) unsigned long long sBar;
  __shared__ uint32_t sTmem;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int p0 = blockIdx.x * PPC;
  const int rows = PPC * B;                       // <= LS_ROWS
  const uint32_t raw = smem_u32(ls_smem);
  const uint32_t base = (raw + 1023u) & ~1023u;
  uint8_t* gen = ls_smem + (base - raw);
  const uint32_t a_plane = (uint32_t)nkb * LS_A_KB_BYTES + LS_A_TAIL, b_plane = (uint32_t)nkb * LS_B_KB_BYTES;
  const uint32_t oAhi = 0, oAlo = a_plane, oBhi = 2 * a_plane, oBlo = 2 * a_plane + b_plane;
  const uint32_t bar = smem_u32(&sBar);

  if (tid < LS_ROWS) {
    const int p = p0 + tid / B, b = tid % B;
    const bool ok = tid < rows && p < P;
    const int row = ok ? (idx ? idx[p * B + b] : b) : 0;
    sRow[tid] = ok ? X + ((size_t)p * Ns + row) * F : nullptr;
    sLab[tid] = ok ? y[(size_t)p * Ns + row] : 0;
  }
  if (tid == 0) {
    mbar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&sTmem)), "r"(32)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sTmem;

  // ---- phase 1: operands -> swizzled bf16 planes -----------------------------------------------------------
  const int QP = nkb * 16;                        // float4 slots per padded row
  const int F4 = F >> 2;
  {
    // only the rows that exist are staged (rows past `rows` of the 128-row tile are never read back from TMEM);
    // their K padding must be zero, the loads of a batch are all in flight before the first conversion
    constexpr int BATCH = 7;
    const int items = rows * QP;
    for (int it0 = tid; it0 < items; it0 += LS_THREADS * BATCH) {
      float4 v[BATCH];
      int rr[BATCH], qq[BATCH];
#pragma unroll
      for (int u = 0; u < BATCH; u++) {
        const int it = it0 + u * LS_THREADS;
        rr[u] = it / QP;
        qq[u] = it - rr[u] * QP;
        v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (it < items) {
          const float* src = sRow[rr[u]];
          if (src && qq[u] < F4) v[u] = __ldg(reinterpret_cast<const float4*>(src) + qq[u]);
        }
      }
#pragma unroll
      for (int u = 0; u < BATCH; u++) {
        if (it0 + u * LS_THREADS < items) {
          uint2 hi, lo;
          split4(v[u], hi, lo);
          const uint32_t off = sw128_off(rr[u], qq[u], LS_A_KB_BYTES);
          *reinterpret_cast<uint2*>(gen + oAhi + off) = hi;
          *reinterpret_cast<uint2*>(gen + oAlo + off) = lo;
        }
      }
    }
    const int witems = LS_NP * QP;
    for (int it = tid; it < witems; it += LS_THREADS) {
      const int r = it / QP, q = it - r * QP;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (r < LS_C && q < F4) v = __ldg(reinterpret_cast<const float4*>(w + (size_t)r * F) + q);
      uint2 hi, lo;
      split4(v, hi, lo);
      const uint32_t off = sw128_off(r, q, LS_B_KB_BYTES);
      *reinterpret_cast<uint2*>(gen + oBhi + off) = hi;
      *reinterpret_cast<uint2*>(gen + oBlo + off) = lo;
    }
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy stores -> visible to the MMA's reads
  __syncthreads();

  // ---- phase 2: logits on the tensor cores ------------------------------------------------------------------
  if (tid == 0) {
    tc_fence_after();
    const uint32_t idesc = umma_idesc_bf16_128xN(LS_NP);
    for (int kb = 0; kb < nkb; kb++) {
      const uint64_t a_hi = umma_desc_sw128(base + oAhi + kb * LS_A_KB_BYTES), a_lo = umma_desc_sw128(base + oAlo + kb * LS_A_KB_BYTES);
      const uint64_t b_hi = umma_desc_sw128(base + oBhi + kb * LS_B_KB_BYTES), b_lo = umma_desc_sw128(base + oBlo + kb * LS_B_KB_BYTES);
#pragma unroll
      for (int k = 0; k < 4; k++) {
        const uint64_t off = (uint64_t)((k * 16 * 2) >> 4);
        umma_bf16(tmem, a_hi + off, b_hi + off, idesc, (kb > 0 || k > 0) ? 1u : 0u);
        umma_bf16(tmem, a_hi + off, b_lo + off, idesc, 1u);
        umma_bf16(tmem, a_lo + off, b_hi + off, idesc, 1u);
      }
    }
    umma_commit(bar);
  }
  // ---- phase 3a: softmax / loss per row (warp 0 owns TMEM lanes 0..31) ---------------------------------------
  if (warp == 0) {
    mbar_wait(bar, 0);
    tc_fence_after();
    uint32_t v[16];
    tmem_ld16(tmem, v);
    float l[LS_C];
    float m = -INFINITY;
#pragma unroll
    for (int c = 0; c < LS_C; c++) {
      l[c] = __uint_as_float(v[c]) + __ldg(w + (size_t)LS_C * F + c);
      m = fmaxf(m, l[c]);
    }
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < LS_C; c++) s += expf(l[c] - m);
    const float lse = m + logf(s);
    const bool ok = sRow[lane] != nullptr;
    const int lab = sLab[lane];
    float myloss = 0.f;
#pragma unroll
    for (int c = 0; c < LS_C; c++) {
      const float pr = expf(l[c] - lse);
      sDL[lane][c] = ok ? (pr - (c == lab ? 1.f : 0.f)) / (float)B : 0.f;
      if (c == lab) myloss = (lse - l[c]) / (float)B;
    }
    sLoss[lane] = ok ? myloss : 0.f;
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(32) : "memory");
  }
  if (tid < PPC && p0 + tid < P && loss) {
    float t = 0.f;
    for (int b = 0; b < B; b++) t += sLoss[tid * B + b];
    loss[p0 + tid] = t;
  }

  // ---- phase 3b: gradients in registers, clip, write ----------------------------------------------------------
  float g[PPC][LS_KI][LS_C];
  float nrm[PPC];
#pragma unroll
  for (int j = 0; j < PPC; j++) {
    nrm[j] = 0.f;
#pragma unroll
    for (int i = 0; i < LS_KI; i++) {
      const int k = tid + i * LS_THREADS;
#pragma unroll
      for (int c = 0; c < LS_C; c++) g[j][i][c] = 0.f;
    }
    if (p0 + j < P) {
      // the rows were read a moment ago (phase 1) and sit in L2; all loads of a 10-row chunk are issued before the
      // first FMA needs one, so a peer costs one L2 round trip per chunk instead of one per row
      for (int b0 = 0; b0 < B; b0 += 10) {
        float x[LS_KI][10];
#pragma unroll
        for (int i = 0; i < LS_KI; i++) {
          const int k = tid + i * LS_THREADS;
#pragma unroll
          for (int u = 0; u < 10; u++) x[i][u] = (k < F && b0 + u < B) ? __ldg(sRow[j * B + b0 + u] + k) : 0.f;
        }
#pragma unroll
        for (int u = 0; u < 10; u++) {
          if (b0 + u < B) {
#pragma unroll
            for (int c = 0; c < LS_C; c++) {
              const float dl = sDL[j * B + b0 + u][c];
#pragma unroll
              for (int i = 0; i < LS_KI; i++) g[j][i][c] = fmaf(dl, x[i][u], g[j][i][c]);
            }
          }
        }
      }
#pragma unroll
      for (int i = 0; i < LS_KI; i++)
#pragma unroll
        for (int c = 0; c < LS_C; c++) nrm[j] = fmaf(g[j][i][c], g[j][i][c], nrm[j]);
    }
  }
  float gb[PPC];                                   // bias gradient: thread c < C of each peer
#pragma unroll
  for (int j = 0; j < PPC; j++) {
    gb[j] = 0.f;
    if (tid < LS_C) {
      for (int b = 0; b < B; b++) gb[j] += sDL[j * B + b][tid];
      nrm[j] = fmaf(gb[j], gb[j], nrm[j]);
    }
    const float t = warp_sum(nrm[j]);
    if (lane == 0) sRed[j][warp] = t;
  }
  __syncthreads();
  if (tid < PPC) {
    float t = 0.f;
    for (int wi = 0; wi < LS_THREADS / 32; wi++) t += sRed[tid][wi];
    // torch.nn.utils.clip_grad_norm: coef = max_norm / (total + 1e-6), applied when < 1
    const float coef = clip / (sqrtf(t) + 1e-6f);
    sCoef[tid] = coef > 1.f ? 1.f : coef;
  }
  __syncthreads();
#pragma unroll
  for (int j = 0; j < PPC; j++) {
    if (p0 + j >= P) break;
    const float coef = sCoef[j];
    float* out = U + (size_t)(p0 + j) * ldu;
#pragma unroll
    for (int i = 0; i < LS_KI; i++) {
      const int k = tid + i * LS_THREADS;
      if (k < F) {
#pragma unroll
        for (int c = 0; c < LS_C; c++) out[(size_t)c * F + k] = -coef * g[j][i][c];
      }
    }
    if (tid < LS_C) out[(size_t)LS_C * F + tid] = -coef * gb[j];
    for (int k = LS_C * F + LS_C + tid; k < ldu; k += LS_THREADS) out[k] = 0.f;
  }
}

// ------------------------------------------------------------------------------------------------
// K1b: logistic regression (creditcard path), labels in {-1,+1}.
//   g = (1/B) X^T(-y / (1+exp(y Xw))) + lambda w ;  U = -alpha g + own_noise
//   (ML/code/logistic_model.py:92-140; own noise = (-alpha/B) * sigma * sum_B N(0,1), :79-83,108-109)
// One warp per peer.
// ------------------------------------------------------------------------------------------------
__global__ void local_logistic_step_kernel(const float* __restrict__ w, const float* __restrict__ X,
                                           const float* __restrict__ y, const int* __restrict__ idx,
                                           float* __restrict__ U, int P, int Ns, int B, int F, int ldu,
                                           float alpha, float lambda, float own_noise_scale,
                                           unsigned long long seed, const int* __restrict__ peer_ids,
                                           int slot) {
  const int gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (gw >= P) return;
  const int p = gw;
  // res[b] = -y / (1 + exp(y * x.w))
  float g_acc[4] = {0.f, 0.f, 0.f, 0.f};  // supports F <= 128
  for (int b = 0; b < B; b++) {
    const int row = idx ? idx[p * B + b] : b;
    const float* xr = X + ((size_t)p * Ns + row) * F;
    float dot = 0.f;
    for (int k = lane; k < F; k += 32) dot = fmaf(xr[k], w[k], dot);
    dot = warp_sum(dot);
    const float yy = y[(size_t)p * Ns + row];
    const float res = -yy / (1.f + expf(yy * dot));
    for (int k = lane, j = 0; k < F; k += 32, j++) g_acc[j] = fmaf(xr[k], res, g_acc[j]);
  }
  const int gid = peer_ids ? peer_ids[p] : p;
  for (int k = lane, j = 0; k < F; k += 32, j++) {
    float g = g_acc[j] / (float)B + lambda * w[k];
    float nz = 0.f;
    if (own_noise_scale != 0.f) nz = own_noise_scale * gauss1(seed ^ 0x6f776e6eull, gid, slot, k);
    U[(size_t)p * ldu + k] = -alpha * g + nz;
  }
  for (int k = F + lane; k < ldu; k += 32) U[(size_t)p * ldu + k] = 0.f;
}

// ------------------------------------------------------------------------------------------------
// K3/K17: noiser noise, regenerated from counters and averaged (DistSys/main.go:1592-1660), fused
// with everything the Krum Gram kernel needs: the noised update in fp32, its split into two bf16
// planes (hi = bf16(u), lo = bf16(u - hi); hi*hi + hi*lo + lo*hi reproduces the fp32 product to
// ~2^-16 relative) and the fp32 squared row norm.  One CTA per peer.
//   noisers : [P][nn] global ids of each peer's noisers (-1 = none)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) noise_mix_kernel(
    const float* __restrict__ U, const int* __restrict__ noisers, int nn, unsigned long long seed, int slot,
    float noise_scale, float* __restrict__ Un, __nv_bfloat16* __restrict__ hi, __nv_bfloat16* __restrict__ lo,
    float* __restrict__ norms, int d, int ldu, int ldh, const int* __restrict__ dst_row) {
  __shared__ float red[40];
  __shared__ int snz[64];
  const int p = blockIdx.x, tid = threadIdx.x;
  // dst_row (optional): plane / norm row of peer p, so that the Krum candidates land compacted and in Krum order
  // (no gather afterwards); a negative entry means the peer is not a candidate and needs no planes
  const int pr = dst_row ? dst_row[p] : p;
  if (pr < 0) {
    hi = nullptr;
    norms = nullptr;
    if (!Un) return;
  }
  int cnt = 0;
  if (tid == 0) {
    for (int q = 0; q < nn; q++) {
      int id = noisers ? noisers[p * nn + q] : -1;
      if (id >= 0) snz[cnt++] = id;
    }
    snz[63] = cnt;
  }
  __syncthreads();
  cnt = snz[63];
  const float inv = cnt > 0 ? noise_scale / (float)cnt : 0.f;
  float nrm = 0.f;
  const int d4 = (d + 3) >> 2;
  for (int cb = tid; cb < d4; cb += blockDim.x) {
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    for (int q = 0; q < cnt; q++) {
      float g[4];
      gauss4(seed, (uint32_t)snz[q], (uint32_t)slot, (uint32_t)cb, g);
#pragma unroll
      for (int j = 0; j < 4; j++) acc[j] += g[j];
    }
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int k = 4 * cb + j;
      if (k < d) {
        const float v = U[(size_t)p * ldu + k] + inv * acc[j];
        if (Un) Un[(size_t)p * ldu + k] = v;
        const __nv_bfloat16 h = __float2bfloat16_rn(v);
        const float r = v - __bfloat162float(h);
        if (hi) {
          hi[(size_t)pr * ldh + k] = h;
          lo[(size_t)pr * ldh + k] = __float2bfloat16_rn(r);
        }
        nrm = fmaf(v, v, nrm);
      }
    }
  }
  // zero padding columns so the Gram kernel can run over ldh without masks
  for (int k = d + tid; k < ldh; k += blockDim.x) {
    if (hi) {
      hi[(size_t)pr * ldh + k] = __float2bfloat16_rn(0.f);
      lo[(size_t)pr * ldh + k] = __float2bfloat16_rn(0.f);
    }
  }
  for (int k = d + tid; k < ldu; k += blockDim.x)
    if (Un) Un[(size_t)p * ldu + k] = 0.f;
  const float t = block_sum(nrm, red);
  if (tid == 0 && norms) norms[pr] = t;
}

// Minibatch draw: B distinct rows out of a peer's Ns, uniformly at random (logistic_model.py:121-125; the torch
// DataLoader shuffle of client.py).  One warp per peer: every row gets a Philox key from (seed, iteration, peer,
// row) and the B smallest keys win, so the draw depends only on those four numbers — any rank computes the same
// rows for a peer, and no generator state lives anywhere.  Ns <= 32 * SMB_MAXK.
constexpr int SMB_MAXK = 32;
__global__ void __launch_bounds__(128) sample_minibatch_kernel(unsigned long long seed, int it, const int* __restrict__ peer_ids,
                                                               int P, int Ns, int B, int* __restrict__ idx) {
  const int w = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (w >= P) return;
  const uint32_t peer = peer_ids ? (uint32_t)peer_ids[w] : (uint32_t)w;
  uint32_t key[SMB_MAXK];
#pragma unroll
  for (int k = 0; k < SMB_MAXK; k++) {
    const int r = lane + 32 * k;
    key[k] = 0xffffffffu;
    if (r < Ns) {
      const Philox4 ph = philox4x32_10((uint32_t)r, (uint32_t)it, peer, 0x6d696e69u, (uint32_t)seed, (uint32_t)(seed >> 32));
      key[k] = ph.x >> 1;   // 31 bits: 0xffffffff stays reserved for "taken / out of range"
    }
  }
  for (int b = 0; b < B; b++) {
    unsigned long long best = ~0ull;
#pragma unroll
    for (int k = 0; k < SMB_MAXK; k++) {
      const unsigned long long c = ((unsigned long long)key[k] << 32) | (unsigned)(lane + 32 * k);
      best = c < best ? c : best;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const unsigned long long other = __shfl_xor_sync(0xffffffffu, best, o);
      best = other < best ? other : best;
    }
    const int row = (int)(best & 0xffffffffu);
    if (lane == (row & 31)) {
#pragma unroll
      for (int k = 0; k < SMB_MAXK; k++)
        if (k == (row >> 5)) key[k] = 0xffffffffu;
    }
    if (lane == 0) idx[(size_t)w * B + b] = row;
  }
}

// raw noise table row (RequestNoise parity + tests): out[k] = scale * N(0,1)(peer, slot, k)
__global__ void noise_row_kernel(float* __restrict__ out, int d, unsigned long long seed, int peer, int slot,
                                 float scale) {
  const int cb = blockIdx.x * blockDim.x + threadIdx.x;
  if (4 * cb >= d) return;
  float g[4];
  gauss4(seed, peer, slot, cb, g);
  for (int j = 0; j < 4; j++)
    if (4 * cb + j < d) out[4 * cb + j] = scale * g[j];
}

// ------------------------------------------------------------------------------------------------
// K5 (SIMT baseline / oracle on device): G = A A^T in fp32, 32x32 tiles, K streamed.
// The product path is krum_gram_tc (tcgen05); this kernel is the un-fused baseline of SURVEY 8.11
// and the on-device cross-check.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) gram_simt_kernel(const float* __restrict__ A, float* __restrict__ G,
                                                        int n, int d, int lda, int ksplit) {
  __shared__ float sa[32][33], sb[32][33];
  const int ti = blockIdx.x, tj = blockIdx.y, ks = blockIdx.z;
  if (tj < ti) return;  // symmetric: upper tiles only
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
  const int kchunk = ((d + ksplit - 1) / ksplit + 31) & ~31;
  const int k0 = ks * kchunk, k1 = min(d, k0 + kchunk);
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  for (int kb = k0; kb < k1; kb += 32) {
    for (int r = ty; r < 32; r += 8) {
      const int ri = ti * 32 + r, rj = tj * 32 + r, k = kb + tx;
      sa[r][tx] = (ri < n && k < k1) ? A[(size_t)ri * lda + k] : 0.f;
      sb[r][tx] = (rj < n && k < k1) ? A[(size_t)rj * lda + k] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < 32; kk++) {
      const float bv = sb[tx][kk];
#pragma unroll
      for (int q = 0; q < 4; q++) acc[q] = fmaf(sa[ty + 8 * q][kk], bv, acc[q]);
    }
    __syncthreads();
  }
  for (int q = 0; q < 4; q++) {
    const int i = ti * 32 + ty + 8 * q, j = tj * 32 + tx;
    if (i < n && j < n) {
      float* dst = G + (size_t)ks * n * n;
      dst[(size_t)i * n + j] = acc[q];
      if (ti != tj) dst[(size_t)j * n + i] = acc[q];
    }
  }
}

// ------------------------------------------------------------------------------------------------
// K5 tail: fixed-order reduction of split-K partial Grams, distances, per-row sort, Krum scores and
// the Multi-Krum accept set.  Single CTA (n <= 1024); one warp sorts one row at a time (bitonic in
// shared memory).  Deterministic: partials are summed in index order and score ties break by id.
//   Gpart  : [nsplit][ldg][ldg] fp32 partial Gram matrices (ldg >= n)
//   scores : [n] out ; accept : [n] uint8 out ; D : optional [n][n] distances out
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) krum_select_kernel(const float* __restrict__ Gpart, int nsplit, int ldg,
                                                          const float* __restrict__ norms, int n, int clip,
                                                          float* __restrict__ scores,
                                                          unsigned char* __restrict__ accept,
                                                          float* __restrict__ Dout, int np2) {
  extern __shared__ float srow[];  // [nwarp][np2] + [n] scores
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarp = blockDim.x >> 5;
  float* myrow = srow + (size_t)warp * np2;
  float* ssc = srow + (size_t)nwarp * np2;
  const int group = n - clip;
  for (int i = warp; i < n; i += nwarp) {
    const float ni = norms[i];
    for (int j = lane; j < np2; j += 32) {
      float v = INFINITY;
      if (j < n) {
        float g = 0.f;
        for (int s = 0; s < nsplit; s++) g += Gpart[((size_t)s * ldg + i) * ldg + j];
        v = ni + norms[j] - 2.f * g;
        if (Dout) Dout[(size_t)i * n + j] = v;
      }
      myrow[j] = v;
    }
    __syncwarp();
    // bitonic sort ascending over np2 elements
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
    // score = sum(sorted[1 : group-1])
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

// Multi-CTA form of the Krum tail: one warp per row, 8 rows per CTA (scores), then a single small CTA
// ranks the scores (accept).  Same arithmetic and the same fixed reduction order as krum_select_kernel,
// so both produce identical bits; this one removes the single-CTA serialisation over rows.
//   rows (optional): participant i lives at row/column rows[i] of the partial Grams
//   norms == nullptr: squared norms are taken from the reduced Gram's diagonal
__global__ void __launch_bounds__(256) krum_scores_kernel(const float* __restrict__ Gpart, int nsplit, int ldg,
                                                          const float* __restrict__ norms, const int* __restrict__ rows,
                                                          int n, int clip, float* __restrict__ scores,
                                                          float* __restrict__ Dout, int np2,
                                                          const unsigned long long* __restrict__ wait_flags, int world,
                                                          unsigned long long epoch) {
  extern __shared__ float srow[];  // [nwarp][np2] + [n] norms
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarp = blockDim.x >> 5;
  if (wait_flags) {   // fused multi-GPU path: every rank's partial Gram must have landed in Gpart first
    if (tid == 0) symm_wait_all(wait_flags, world, SLOT_GRAM, epoch);
    __syncthreads();
  }
  float* myrow = srow + (size_t)warp * np2;
  float* snorm = srow + (size_t)nwarp * np2;
  for (int i = tid; i < n; i += blockDim.x) {
    if (norms) {
      snorm[i] = norms[i];
    } else {
      const int r = rows ? rows[i] : i;
      float g = 0.f;
      for (int s = 0; s < nsplit; s++) g += Gpart[((size_t)s * ldg + r) * ldg + r];
      snorm[i] = g;
    }
  }
  __syncthreads();
  const int i = blockIdx.x * nwarp + warp;
  if (i >= n) return;
  const int group = n - clip;
  const int ri = rows ? rows[i] : i;
  const float ni = snorm[i];
  for (int j = lane; j < np2; j += 32) {
    float v = INFINITY;
    if (j < n) {
      const int rj = rows ? rows[j] : j;
      float g = 0.f;
      for (int s = 0; s < nsplit; s++) g += Gpart[((size_t)s * ldg + ri) * ldg + rj];
      v = ni + snorm[j] - 2.f * g;
      if (Dout) Dout[(size_t)i * n + j] = v;
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
  if (lane == 0) scores[i] = s;
}
__global__ void __launch_bounds__(256) krum_accept_kernel(const float* __restrict__ scores, int n, int clip,
                                                          unsigned char* __restrict__ accept) {
  extern __shared__ float ssc[];
  for (int i = threadIdx.x; i < n; i += blockDim.x) ssc[i] = scores[i];
  __syncthreads();
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const float si = ssc[i];
    int rank = 0;
    for (int j = 0; j < n; j++) {
      const float sj = ssc[j];
      rank += (sj < si) || (sj == si && j < i);
    }
    accept[i] = rank < (n - clip) ? 1 : 0;
  }
}

// ------------------------------------------------------------------------------------------------
// K14/K15: argmax classification error of linear models over a held-out set, batched over
// candidate weight vectors (RONI = error(w + delta_i) - error(w), client_obj.py:100-112).
//   Wc : [M][ldw] candidate weights (softmax layout) ; X : [N][F]; y : [N]
//   wrong : [M] int32 out (count of misclassified rows; caller zero-fills)
// grid (ceil(N/8), M), 8 warps per CTA, one row per warp.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) eval_softmax_kernel(const float* __restrict__ Wc, const float* __restrict__ base,
                                                           const float* __restrict__ X, const int* __restrict__ y,
                                                           int* __restrict__ wrong, int N, int F, int C, int ldw) {
  const int m = blockIdx.y;
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= N) return;
  const float* w = Wc + (size_t)m * ldw;
  float best = -INFINITY;
  int arg = 0;
  for (int c = 0; c < C; c++) {
    float acc = 0.f;
    for (int k = lane; k < F; k += 32) {
      float wv = w[(size_t)c * F + k];
      if (base) wv += base[(size_t)c * F + k];
      acc = fmaf(X[(size_t)row * F + k], wv, acc);
    }
    acc = warp_sum(acc);
    float bias = w[(size_t)C * F + c];
    if (base) bias += base[(size_t)C * F + c];
    acc += bias;
    if (acc > best) {
      best = acc;
      arg = c;
    }
  }
  if (lane == 0 && arg != y[row]) atomicAdd(&wrong[m], 1);
}

// logistic (sign classifier) error: ML/code/logistic_model_test.py:13-24
__global__ void eval_logistic_kernel(const float* __restrict__ Wc, const float* __restrict__ base,
                                     const float* __restrict__ X, const float* __restrict__ y,
                                     int* __restrict__ wrong, int N, int F, int ldw) {
  const int m = blockIdx.y;
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= N) return;
  float acc = 0.f;
  for (int k = lane; k < F; k += 32) {
    float wv = Wc[(size_t)m * ldw + k];
    if (base) wv += base[k];
    acc = fmaf(X[(size_t)row * F + k], wv, acc);
  }
  acc = warp_sum(acc);
  const float pred = acc >= 0.f ? 1.f : -1.f;
  if (lane == 0 && pred != y[row]) atomicAdd(&wrong[m], 1);
}

}  // namespace bsc

using namespace bsc;

extern "C" {

// test hook: 1 = keep K1 on the SIMT kernel (the tensor-core kernel is compared against it and against fp32 PyTorch)
void bsc_softmax_force_simt(int on) { g_softmax_simt = on != 0; }
unsigned long long bsc_launch_count() { return g_launch_count; }
void bsc_launch_count_reset() { g_launch_count = 0; }

int bsc_local_softmax_step(const float* w, const float* X, const int* y, const int* idx, float* U, float* loss,
                           int P, int Ns, int B, int F, int C, int ldu, float clip, void* stream) {
  if (B > 64 || C > 32) return -1;
  if (C == LS_C && B <= LS_ROWS && (F & 3) == 0 && F <= LS_KI * LS_THREADS && !g_softmax_simt) {
    // tensor-core path: PPC peers per CTA, all of K resident in shared memory
    // as many CTAs as there are SMs to hold them: one peer per CTA up to 148 peers, then two, then three
    const int ppc_max = B <= 10 ? 3 : (B <= 16 ? 2 : 1);
    int ppc = (P + 147) / 148;
    if (ppc > ppc_max) ppc = ppc_max;
    if (ppc < 1) ppc = 1;
    const int nkb = (F + 63) / 64;
    const size_t smem_tc = 1024 + 2 * ((size_t)nkb * LS_A_KB_BYTES + LS_A_TAIL) + 2 * (size_t)nkb * LS_B_KB_BYTES;
    if (smem_tc <= 220 * 1024) {
      const int grid = (P + ppc - 1) / ppc;
      static bool configured_tc = false;
      if (!configured_tc) {
        BSC_CHECK(cudaFuncSetAttribute(local_softmax_tc_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
        BSC_CHECK(cudaFuncSetAttribute(local_softmax_tc_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
        BSC_CHECK(cudaFuncSetAttribute(local_softmax_tc_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
        configured_tc = true;
      }
      cudaStream_t st = (cudaStream_t)stream;
      if (ppc == 3) local_softmax_tc_kernel<3><<<grid, LS_THREADS, smem_tc, st>>>(w, X, y, idx, U, loss, P, Ns, B, F, ldu, clip, nkb);
      else if (ppc == 2) local_softmax_tc_kernel<2><<<grid, LS_THREADS, smem_tc, st>>>(w, X, y, idx, U, loss, P, Ns, B, F, ldu, clip, nkb);
      else local_softmax_tc_kernel<1><<<grid, LS_THREADS, smem_tc, st>>>(w, X, y, idx, U, loss, P, Ns, B, F, ldu, clip, nkb);
      return finish_launch("local_softmax_tc");
    }
  }
  const size_t smem = ((size_t)B * F + (size_t)B * C + 64) * sizeof(float);
  static size_t configured = 0;
  if (smem > 48 * 1024 && smem > configured) {
    BSC_CHECK(cudaFuncSetAttribute(local_softmax_step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured = smem;
  }
  local_softmax_step_kernel<<<P, 256, smem, (cudaStream_t)stream>>>(w, X, y, idx, U, loss, Ns, B, F, C, ldu, clip);
  return finish_launch("local_softmax_step");
}

int bsc_local_logistic_step(const float* w, const float* X, const float* y, const int* idx, float* U, int P, int Ns,
                            int B, int F, int ldu, float alpha, float lambda, float own_noise_scale,
                            unsigned long long seed, const int* peer_ids, int slot, void* stream) {
  if (F > 128) return -1;
  const int threads = 128, wpb = threads / 32;
  local_logistic_step_kernel<<<(P + wpb - 1) / wpb, threads, 0, (cudaStream_t)stream>>>(
      w, X, y, idx, U, P, Ns, B, F, ldu, alpha, lambda, own_noise_scale, seed, peer_ids, slot);
  return finish_launch("local_logistic_step");
}

int bsc_noise_mix(const float* U, const int* noisers, int nn, unsigned long long seed, int slot, float noise_scale,
                  float* Un, void* hi, void* lo, float* norms, int P, int d, int ldu, int ldh, void* stream) {
  if (nn > 62) return -1;
  noise_mix_kernel<<<P, 256, 0, (cudaStream_t)stream>>>(U, noisers, nn, seed, slot, noise_scale, Un,
                                                         (__nv_bfloat16*)hi, (__nv_bfloat16*)lo, norms, d, ldu, ldh, nullptr);
  return finish_launch("noise_mix");
}
// Same, with the bf16 planes and norms of peer p written to row dst_row[p] (skipped when negative): the Krum
// candidates come out compacted in Krum order and the Gram kernel reads them without a gather.
int bsc_noise_mix_rows(const float* U, const int* noisers, int nn, unsigned long long seed, int slot, float noise_scale,
                       float* Un, void* hi, void* lo, float* norms, int P, int d, int ldu, int ldh, const int* dst_row,
                       void* stream) {
  if (nn > 62) return -1;
  noise_mix_kernel<<<P, 256, 0, (cudaStream_t)stream>>>(U, noisers, nn, seed, slot, noise_scale, Un,
                                                         (__nv_bfloat16*)hi, (__nv_bfloat16*)lo, norms, d, ldu, ldh, dst_row);
  return finish_launch("noise_mix");
}

// idx[p][0..B) = B distinct rows in [0, Ns) for peer peer_ids[p] (or p) at iteration `it`; Ns <= 1024, B <= Ns.
int bsc_sample_minibatch(unsigned long long seed, int it, const int* peer_ids, int P, int Ns, int B, int* idx, void* stream) {
  if (Ns > 32 * SMB_MAXK || B > Ns || B < 1 || P < 1) return -1;
  sample_minibatch_kernel<<<(P + 3) / 4, 128, 0, (cudaStream_t)stream>>>(seed, it, peer_ids, P, Ns, B, idx);
  return finish_launch("sample_minibatch");
}

int bsc_noise_row(float* out, int d, unsigned long long seed, int peer, int slot, float scale, void* stream) {
  const int n4 = (d + 3) / 4;
  noise_row_kernel<<<(n4 + 255) / 256, 256, 0, (cudaStream_t)stream>>>(out, d, seed, peer, slot, scale);
  return finish_launch("noise_row");
}

// host restatement of the same generator (CPU path + oracle)
void bsc_noise_row_host(float* out, int d, unsigned long long seed, int peer, int slot, float scale) {
  for (int cb = 0; 4 * cb < d; cb++) {
    float g[4];
    gauss4(seed, peer, slot, cb, g);
    for (int j = 0; j < 4 && 4 * cb + j < d; j++) out[4 * cb + j] = scale * g[j];
  }
}

int bsc_gram_simt(const float* A, float* Gpart, int n, int d, int lda, int ksplit, void* stream) {
  const int nt = (n + 31) / 32;
  dim3 grid(nt, nt, ksplit);
  gram_simt_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(A, Gpart, n, d, lda, ksplit);
  return finish_launch("gram_simt");
}

int bsc_krum_select(const float* Gpart, int nsplit, int ldg, const float* norms, int n, int clip, float* scores,
                    unsigned char* accept, float* Dout, void* stream) {
  if (n > 1024 || n < 3) return -1;
  int np2 = 32;
  while (np2 < n) np2 <<= 1;
  const size_t smem = ((size_t)8 * np2 + n) * sizeof(float);
  krum_select_kernel<<<1, 256, smem, (cudaStream_t)stream>>>(Gpart, nsplit, ldg, norms, n, clip, scores, accept, Dout,
                                                              np2);
  return finish_launch("krum_select");
}

// ---- allocation-free Multi-Krum ----------------------------------------------------------------------
// Split-K partial Gram matrices normally live in a buffer the caller allocates (bsc_gram_simt ->
// bsc_krum_select).  Some callers cannot allocate device memory at the point where they need a Krum
// verdict: code running inside a region where cudaMalloc is forbidden or would serialise the device
// (a verifier loop that must not touch the allocator between blocks, a host that embeds this library
// under a fixed memory budget handed out at start-up, a plug-in called from a framework that owns the
// caching allocator).  For them the partials go into the fixed array below, which belongs to the
// library image: every GPU that loads libbiscotti_b200.so holds these 512 KiB until the process exits.
#define BSC_KRUM_STATIC_MAX_PEERS 128
#define BSC_KRUM_STATIC_MAX_SPLITS 8
static __device__ __align__(256) float g_krum_static_scratch[BSC_KRUM_STATIC_MAX_SPLITS * BSC_KRUM_STATIC_MAX_PEERS *
                                                             BSC_KRUM_STATIC_MAX_PEERS];
static float* g_krum_static_ptr[64] = {nullptr};   // device address of the array, per device ordinal
static std::mutex g_krum_static_mu;

// split count used by bsc_krum_scores_static for a given dimension (so a caller can reproduce the same
// reduction order with the allocating entry points)
int bsc_krum_static_splits(int d) {
  int ks = d / 512;
  if (ks < 1) ks = 1;
  if (ks > BSC_KRUM_STATIC_MAX_SPLITS) ks = BSC_KRUM_STATIC_MAX_SPLITS;
  return ks;
}

// No counterpart in the reference (its Krum runs in NumPy on the host, client_obj.py:114-143); this is
// the allocation-free variant of bsc_gram_simt + bsc_krum_select for callers that cannot call
// cudaMalloc.  Inputs: A [n][lda] fp32 updates (first d columns used), norms [n] squared row norms.
// Outputs: scores [n], accept [n] (1 = in the Multi-Krum accept set).  Limits: 3 <= n <= 128 peers, any
// d >= 1; larger n is rejected with -1 and a message on stderr.  Nothing is allocated or freed, here or
// in any library it calls.  The first call on each device looks up the device address of the scratch
// array (cudaGetSymbolAddress, which may load the CUDA module), so make that first call before starting
// any stream capture.  All calls on a device share the one scratch array: a call rewrites every word it
// later reads (nothing has to start at zero, and contents left by the previous call are never
// consumed), but two calls whose kernels overlap on the GPU — different streams, or different host
// threads — race on it; serialise them on one stream.
int bsc_krum_scores_static(const float* A, const float* norms, int n, int d, int lda, int clip, float* scores,
                           unsigned char* accept, void* stream) {
  if (n > BSC_KRUM_STATIC_MAX_PEERS || n < 3 || d < 1 || clip < 0 || clip >= n) {
    fprintf(stderr,
            "[biscotti_b200] bsc_krum_scores_static: n=%d d=%d clip=%d outside the static scratch limits "
            "(3 <= n <= %d peers, d >= 1, 0 <= clip < n); use bsc_gram_simt + bsc_krum_select with your own buffer\n",
            n, d, clip, BSC_KRUM_STATIC_MAX_PEERS);
    return -1;
  }
  if (lda < d) {   // a row stride shorter than the row would make the Gram kernel read past the matrix
    fprintf(stderr, "[biscotti_b200] bsc_krum_scores_static: lda=%d is smaller than d=%d\n", lda, d);
    return -1;
  }
  int dev = 0;
  BSC_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return -2;
  float* scratch = nullptr;
  {
    std::lock_guard<std::mutex> lk(g_krum_static_mu);
    if (!g_krum_static_ptr[dev]) {
      // the symbol's name is only a host-side handle; kernels need its address on the current device
      float* p = nullptr;
      BSC_CHECK(cudaGetSymbolAddress((void**)&p, g_krum_static_scratch));
      g_krum_static_ptr[dev] = p;
    }
    scratch = g_krum_static_ptr[dev];
  }
  const int ksplit = bsc_krum_static_splits(d);
  const int nt = (n + 31) / 32;
  dim3 grid(nt, nt, ksplit);
  // partials are stored [ksplit][n][n]; every (i, j) of every split is written by this launch
  gram_simt_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(A, scratch, n, d, lda, ksplit);
  int rc = finish_launch("krum_scores_static/gram");
  if (rc) return rc;
  int np2 = 32;
  while (np2 < n) np2 <<= 1;
  const size_t smem = ((size_t)8 * np2 + n) * sizeof(float);
  krum_select_kernel<<<1, 256, smem, (cudaStream_t)stream>>>(scratch, ksplit, n, norms, n, clip, scores, accept,
                                                              nullptr, np2);
  return finish_launch("krum_scores_static/select");
}

// Multi-CTA Krum tail (scores over a grid of row-groups, then the ranking).  rows / norms may be null
// (identity mapping / norms from the Gram diagonal).  Bit-identical to bsc_krum_select.
int bsc_krum_select_mc(const float* Gpart, int nsplit, int ldg, const float* norms, const int* rows, int n, int clip,
                       float* scores, unsigned char* accept, float* Dout, void* stream) {
  if (n > 1024 || n < 3) return -1;
  int np2 = 32;
  while (np2 < n) np2 <<= 1;
  const size_t smem = ((size_t)8 * np2 + n) * sizeof(float);
  krum_scores_kernel<<<(n + 7) / 8, 256, smem, (cudaStream_t)stream>>>(Gpart, nsplit, ldg, norms, rows, n, clip, scores,
                                                                        Dout, np2, nullptr, 0, 0ull);
  krum_accept_kernel<<<1, 256, (size_t)n * sizeof(float), (cudaStream_t)stream>>>(scores, n, clip, accept);
  return finish_launch("krum_select_mc", 2);
}
// Same tail for the fused multi-GPU path: the scores kernel first waits (bounded) until every rank has
// published SLOT_GRAM for `epoch` in my_flags, i.e. all partial Grams have landed in Gpart.
int bsc_krum_select_mc_wait(const float* Gpart, int nsplit, int ldg, const int* rows, int n, int clip,
                            const void* my_flags, int world, unsigned long long epoch, float* scores,
                            unsigned char* accept, void* stream) {
  if (n > 1024 || n < 3) return -1;
  int np2 = 32;
  while (np2 < n) np2 <<= 1;
  const size_t smem = ((size_t)8 * np2 + n) * sizeof(float);
  krum_scores_kernel<<<(n + 7) / 8, 256, smem, (cudaStream_t)stream>>>(Gpart, nsplit, ldg, nullptr, rows, n, clip,
                                                                        scores, nullptr, np2,
                                                                        (const unsigned long long*)my_flags, world,
                                                                        epoch);
  krum_accept_kernel<<<1, 256, (size_t)n * sizeof(float), (cudaStream_t)stream>>>(scores, n, clip, accept);
  return finish_launch("krum_select_mc_wait", 2);
}

int bsc_eval_softmax(const float* Wc, const float* base, const float* X, const int* y, int* wrong, int M, int N, int F,
                     int C, int ldw, void* stream) {
  dim3 grid((N + 7) / 8, M);
  eval_softmax_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(Wc, base, X, y, wrong, N, F, C, ldw);
  return finish_launch("eval_softmax");
}

int bsc_eval_logistic(const float* Wc, const float* base, const float* X, const float* y, int* wrong, int M, int N,
                      int F, int ldw, void* stream) {
  dim3 grid((N + 7) / 8, M);
  eval_logistic_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(Wc, base, X, y, wrong, N, F, ldw);
  return finish_launch("eval_logistic");
}

}  // extern "C"
