// Secure aggregation on the device: fixed-point quantisation, exact-integer Shamir share evaluation,
// homomorphic share aggregation, reconstruction and the global SGD step.
//
// Reference path replaced: DistSys/kyber.go:698-743 (quantise + chunk by 10), :579-609 (share
// evaluation, x = s-10), :244-287 (miners add shares at equal x), :809-857 + DistSys/honest.go:442-502
// (pool shares, solve, round, /10^4) and honest.go:391-440 (w_{t+1} = w_t + sum of deltas).
//
// Layout: the d coordinates are padded to nchunk*10; chunk c owns coordinates [10c, 10c+10).  Chunks
// are the unit of d-axis sharding across GPUs (each GPU owns a contiguous chunk range and plays all
// miners for it), see kernels_fused.cu for the multi-GPU form.
#include "common.cuh"
#include "secagg_math.cuh"
#include <string.h>

namespace bsc {

// Q[p][k] = int64(U[p][k] * scale), zero in the padding
__global__ void quantize_kernel(const float* __restrict__ U, long long* __restrict__ Q, int P, int d, int ldu,
                                int dpad, double scale) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)P * dpad) return;
  const int p = (int)(i / dpad), k = (int)(i % dpad);
  Q[i] = k < d ? quantize_fixed(U[(size_t)p * ldu + k], scale) : 0;
}

// Per-peer shares (what a worker sends to the miners, DistSys/main.go:1847-1952):
//   Y[p][c][s] = poly_{p,c}(s - 10)
__global__ void shares_eval_kernel(const long long* __restrict__ Q, long long* __restrict__ Y, int P, int nchunk,
                                   int dpad, int S) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)P * nchunk) return;
  const int p = (int)(i / nchunk), c = (int)(i % nchunk);
  int64_t a[POLY_SIZE];
#pragma unroll
  for (int j = 0; j < POLY_SIZE; j++) a[j] = Q[(size_t)p * dpad + (size_t)c * POLY_SIZE + j];
  for (int s = 0; s < S; s++) Y[((size_t)p * nchunk + c) * S + s] = shamir_eval(a, s);
}

// Miner-side homomorphic aggregation fused with share generation: for the chunk range
// [c0, c0+nc) sum the shares of every accepted peer.  Thread per (chunk, share).
//   sel : [nsel] indices of contributing peers (rows of U) in fixed (ascending id) order
//   Ysum: [nc][S]
//   nsel_dev : optional device count (<= nsel) written by block_plan_kernel: the contributor set is decided
//              on the device and never visits the host before the block's tail
__global__ void share_sum_kernel(const float* __restrict__ U, const int* __restrict__ sel, int nsel, int d, int ldu,
                                 int c0, int nc, int S, double scale, long long* __restrict__ Ysum,
                                 const int* __restrict__ nsel_dev) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nc * S) return;
  if (nsel_dev) nsel = min(nsel, *nsel_dev);
  const int c = c0 + i / S, s = i % S;
  uint64_t acc = 0;
  for (int q = 0; q < nsel; q++) {
    const float* row = U + (size_t)sel[q] * ldu;
    int64_t a[POLY_SIZE];
#pragma unroll
    for (int j = 0; j < POLY_SIZE; j++) {
      const int k = c * POLY_SIZE + j;
      a[j] = k < d ? quantize_fixed(row[k], scale) : 0;
    }
    acc += (uint64_t)shamir_eval(a, s);
  }
  Ysum[(size_t)(c - c0) * S + s] = (long long)acc;
}

// Reconstruct sum-of-updates for a chunk range from the first POLY_SIZE available shares, verify the
// remaining shares against the recovered polynomial, de-quantise and take the SGD step.
//   xs     : [POLY_SIZE] share indices used for the solve ; S shares total per chunk in Ysum
//   avail  : [S] 1 if that share is present (missing miners => 0)
//   w64    : [d] fp64 master weights (in/out) ; w32 : [d] fp32 working copy (out, may be null)
//   dsum   : [d] fp64 sum of deltas (out, may be null)
//   bad    : int[2] -> {inexact divisions, shares inconsistent with the recovered polynomial}
__global__ void reconstruct_step_kernel(const long long* __restrict__ Ysum, const int* __restrict__ xs,
                                        const unsigned char* __restrict__ avail, int S, int c0, int nc, int d,
                                        double scale, double* __restrict__ w64, float* __restrict__ w32,
                                        double* __restrict__ dsum, int* __restrict__ bad,
                                        const int* __restrict__ gate, int gate_min) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nc) return;
  // too few contributors (main.go:2079): the block will be empty and the weights must stay as they are
  if (gate && *gate < gate_min) return;
  const int c = c0 + i;
  int lx[POLY_SIZE];
  int64_t ly[POLY_SIZE];
#pragma unroll
  for (int j = 0; j < POLY_SIZE; j++) {
    lx[j] = xs[j];
    ly[j] = Ysum[(size_t)i * S + lx[j]];
  }
  int64_t coef[POLY_SIZE];
  const bool exact = shamir_recover(lx, ly, coef);
  if (!exact && bad) atomicAdd(&bad[0], 1);
  if (bad) {
    int mism = 0;
    for (int s = 0; s < S; s++)
      if (avail == nullptr || avail[s]) mism += (shamir_eval(coef, s) != Ysum[(size_t)i * S + s]);
    if (mism) atomicAdd(&bad[1], mism);
  }
#pragma unroll
  for (int j = 0; j < POLY_SIZE; j++) {
    const int k = c * POLY_SIZE + j;
    if (k < d) {
      const double delta = (double)coef[j] / scale;  // updateIntToFloat, kyber.go:745-757
      const double nw = w64[k] + delta;
      w64[k] = nw;
      if (w32) w32[k] = (float)nw;
      if (dsum) dsum[k] = delta;
    }
  }
}

// Same contract as reconstruct_step_kernel, recovery by the constant inverse matrix (secagg_math.cuh: RecoverMat).
// A CTA of 256 threads takes RM_CH chunks: (1) stage their S aggregated shares in shared memory, (2) one thread
// per (chunk, coefficient) forms the coefficient, de-quantises and steps the weight, (3) one thread per (chunk,
// share) re-evaluates the recovered polynomial and checks the redundant shares.
constexpr int RM_CH = 16;
__global__ void __launch_bounds__(256) reconstruct_mat_kernel(const long long* __restrict__ Ysum, const RecoverMat R,
                                                              const unsigned char* __restrict__ avail, int S, int c0,
                                                              int nc, int d, double scale, double* __restrict__ w64,
                                                              float* __restrict__ w32, double* __restrict__ dsum,
                                                              int* __restrict__ bad, const int* __restrict__ gate,
                                                              int gate_min) {
  __shared__ long long sy[RM_CH * MAX_SHARES];
  __shared__ long long sc[RM_CH * POLY_SIZE];
  __shared__ RecoverMat sR;   // the matrix rows are indexed by the thread's coefficient: shared memory, not the constant bank
  if (gate && *gate < gate_min) return;
  if (threadIdx.x < (int)(sizeof(RecoverMat) / 4)) reinterpret_cast<int*>(&sR)[threadIdx.x] = reinterpret_cast<const int*>(&R)[threadIdx.x];
  const int ch0 = blockIdx.x * RM_CH;
  const int nch = min(RM_CH, nc - ch0);
  for (int e = threadIdx.x; e < nch * S; e += blockDim.x) sy[(e / S) * MAX_SHARES + (e % S)] = Ysum[(size_t)ch0 * S + e];
  __syncthreads();
  int inexact = 0;
  for (int e = threadIdx.x; e < nch * POLY_SIZE; e += blockDim.x) {
    const int lc = e / POLY_SIZE, j = e % POLY_SIZE;
    int64_t y[POLY_SIZE];
#pragma unroll
    for (int i = 0; i < POLY_SIZE; i++) y[i] = sy[lc * MAX_SHARES + sR.xs[i]];
    int64_t coef;
    inexact += !recover_coef_mat(sR, y, j, coef);
    sc[e] = coef;
    const int k = (c0 + ch0 + lc) * POLY_SIZE + j;
    if (k < d) {
      const double delta = (double)coef / scale;
      const double nw = w64[k] + delta;
      w64[k] = nw;
      if (w32) w32[k] = (float)nw;
      if (dsum) dsum[k] = delta;
    }
  }
  __syncthreads();
  int mism = 0;
  if (bad) {
    for (int e = threadIdx.x; e < nch * S; e += blockDim.x) {
      const int lc = e / S, s = e % S;
      if (avail == nullptr || avail[s]) mism += (shamir_eval((const int64_t*)(sc + lc * POLY_SIZE), s) != sy[lc * MAX_SHARES + s]);
    }
    if (inexact) atomicAdd(&bad[0], inexact);
    if (mism) atomicAdd(&bad[1], mism);
  }
}

// Column sums of the selected rows of the quantised updates: out[k] = sum_{q < nsel} Q[sel[q]][k].  The miners' homomorphic
// aggregate of the contributors' commitments / witnesses (kyber.go:244-287) is the commitment / witness of this summed
// polynomial (same group elements: sum_p sum_j a_pj T_j = sum_j (sum_p a_pj) T_j), so it costs one more MSM row instead of
// a point addition per contributor and group.
__global__ void sum_rows_i64_kernel(const long long* __restrict__ Q, int ldq, const int* __restrict__ sel, int nsel,
                                    const int* __restrict__ nsel_dev, int ncols, long long* __restrict__ out) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= ncols) return;
  if (nsel_dev) nsel = min(nsel, *nsel_dev);
  long long acc = 0;
  for (int q = 0; q < nsel; q++) acc += Q[(size_t)sel[q] * ldq + k];
  out[k] = acc;
}

// Plain aggregation (-sa=false, DistSys/honest.go:346-388; also the FedSys server,
// FedSys/honest.go:311-337): w += sum of selected deltas, summed in fixed order in fp64.
__global__ void plain_sum_step_kernel(const float* __restrict__ U, const int* __restrict__ sel, int nsel, int d,
                                      int ldu, double* __restrict__ w64, float* __restrict__ w32,
                                      double* __restrict__ dsum, const int* __restrict__ nsel_dev,
                                      const int* __restrict__ gate, int gate_min) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= d) return;
  if (gate && *gate < gate_min) return;
  if (nsel_dev) nsel = min(nsel, *nsel_dev);
  double acc = 0.0;
  for (int q = 0; q < nsel; q++) acc += (double)U[(size_t)sel[q] * ldu + k];
  const double nw = w64[k] + acc;
  w64[k] = nw;
  if (w32) w32[k] = (float)nw;
  if (dsum) dsum[k] = acc;
}

}  // namespace bsc

using namespace bsc;

extern "C" {

int bsc_quantize(const float* U, long long* Q, int P, int d, int ldu, int dpad, double scale, void* stream) {
  const size_t n = (size_t)P * dpad;
  quantize_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(U, Q, P, d, ldu, dpad, scale);
  return finish_launch("quantize");
}

int bsc_shares_eval(const long long* Q, long long* Y, int P, int nchunk, int dpad, int S, void* stream) {
  const size_t n = (size_t)P * nchunk;
  shares_eval_kernel<<<(unsigned)((n + 127) / 128), 128, 0, (cudaStream_t)stream>>>(Q, Y, P, nchunk, dpad, S);
  return finish_launch("shares_eval");
}

// nsel_dev (may be null): device-resident contributor count, see share_sum_kernel
int bsc_share_sum_dc(const float* U, const int* sel, int nsel, const int* nsel_dev, int d, int ldu, int c0, int nc, int S,
                     double scale, long long* Ysum, void* stream) {
  if (S > MAX_SHARES) return -1;
  const int n = nc * S;
  if (n <= 0) return 0;
  share_sum_kernel<<<(n + 127) / 128, 128, 0, (cudaStream_t)stream>>>(U, sel, nsel, d, ldu, c0, nc, S, scale, Ysum, nsel_dev);
  return finish_launch("share_sum");
}
int bsc_share_sum(const float* U, const int* sel, int nsel, int d, int ldu, int c0, int nc, int S, double scale,
                  long long* Ysum, void* stream) {
  return bsc_share_sum_dc(U, sel, nsel, nullptr, d, ldu, c0, nc, S, scale, Ysum, stream);
}

// gate (may be null): the step is applied only if *gate >= gate_min (device-resident contributor count)
int bsc_reconstruct_step_gated(const long long* Ysum, const int* xs, const unsigned char* avail, int S, int c0, int nc, int d,
                               double scale, double* w64, float* w32, double* dsum, int* bad, const int* gate,
                               int gate_min, void* stream) {
  if (nc <= 0) return 0;
  reconstruct_step_kernel<<<(nc + 127) / 128, 128, 0, (cudaStream_t)stream>>>(Ysum, xs, avail, S, c0, nc, d, scale,
                                                                               w64, w32, dsum, bad, gate, gate_min);
  return finish_launch("reconstruct_step");
}
int bsc_reconstruct_step(const long long* Ysum, const int* xs, const unsigned char* avail, int S, int c0, int nc, int d,
                         double scale, double* w64, float* w32, double* dsum, int* bad, void* stream) {
  return bsc_reconstruct_step_gated(Ysum, xs, avail, S, c0, nc, d, scale, w64, w32, dsum, bad, nullptr, 0, stream);
}

// Matrix form of bsc_reconstruct_step_gated; `rm` points at a host RecoverMat (see secagg_math.cuh).
int bsc_reconstruct_mat(const long long* Ysum, const void* rm, const unsigned char* avail, int S, int c0, int nc, int d,
                        double scale, double* w64, float* w32, double* dsum, int* bad, const int* gate, int gate_min,
                        void* stream) {
  if (nc <= 0) return 0;
  if (S > MAX_SHARES) return -1;
  RecoverMat R;
  memcpy(&R, rm, sizeof(R));
  reconstruct_mat_kernel<<<(nc + RM_CH - 1) / RM_CH, 256, 0, (cudaStream_t)stream>>>(Ysum, R, avail, S, c0, nc, d, scale,
                                                                                      w64, w32, dsum, bad, gate, gate_min);
  return finish_launch("reconstruct_mat");
}
// host restatement: coefficients of nc chunks through the matrix; returns the number of inexact coefficients
int bsc_recover_mat_host(const long long* Ysum, const void* rm, int S, int nc, long long* coef_out) {
  RecoverMat R;
  memcpy(&R, rm, sizeof(R));
  int badc = 0;
  for (int i = 0; i < nc; i++) {
    int64_t y[POLY_SIZE];
    for (int k = 0; k < POLY_SIZE; k++) y[k] = Ysum[(size_t)i * S + R.xs[k]];
    for (int j = 0; j < POLY_SIZE; j++) {
      int64_t c;
      badc += !recover_coef_mat(R, y, j, c);
      coef_out[(size_t)i * POLY_SIZE + j] = c;
    }
  }
  return badc;
}
int bsc_recover_mat_sizeof() { return (int)sizeof(RecoverMat); }

int bsc_sum_rows_i64_dc(const long long* Q, int ldq, const int* sel, int nsel, const int* nsel_dev, int ncols,
                        long long* out, void* stream) {
  if (ncols <= 0) return 0;
  sum_rows_i64_kernel<<<(ncols + 255) / 256, 256, 0, (cudaStream_t)stream>>>(Q, ldq, sel, nsel, nsel_dev, ncols, out);
  return finish_launch("sum_rows_i64");
}

int bsc_plain_sum_step_dc(const float* U, const int* sel, int nsel, const int* nsel_dev, int d, int ldu, double* w64,
                          float* w32, double* dsum, const int* gate, int gate_min, void* stream) {
  plain_sum_step_kernel<<<(d + 255) / 256, 256, 0, (cudaStream_t)stream>>>(U, sel, nsel, d, ldu, w64, w32, dsum, nsel_dev,
                                                                           gate, gate_min);
  return finish_launch("plain_sum_step");
}
int bsc_plain_sum_step(const float* U, const int* sel, int nsel, int d, int ldu, double* w64, float* w32, double* dsum,
                       void* stream) {
  return bsc_plain_sum_step_dc(U, sel, nsel, nullptr, d, ldu, w64, w32, dsum, nullptr, 0, stream);
}

// ---- host restatements (CPU plumbing path and oracles; same arithmetic, same header) -------------
void bsc_quantize_host(const float* U, long long* Q, int P, int d, int ldu, int dpad, double scale) {
  for (int p = 0; p < P; p++)
    for (int k = 0; k < dpad; k++) Q[(size_t)p * dpad + k] = k < d ? quantize_fixed(U[(size_t)p * ldu + k], scale) : 0;
}
void bsc_shares_eval_host(const long long* Q, long long* Y, int P, int nchunk, int dpad, int S) {
  for (int p = 0; p < P; p++)
    for (int c = 0; c < nchunk; c++) {
      int64_t a[POLY_SIZE];
      for (int j = 0; j < POLY_SIZE; j++) a[j] = Q[(size_t)p * dpad + (size_t)c * POLY_SIZE + j];
      for (int s = 0; s < S; s++) Y[((size_t)p * nchunk + c) * S + s] = shamir_eval(a, s);
    }
}
// returns number of chunks whose recovery was inexact or inconsistent
int bsc_shamir_recover_host(const long long* Ysum, const int* xs, const unsigned char* avail, int S, int nc,
                            long long* coef_out) {
  int badc = 0;
  for (int i = 0; i < nc; i++) {
    int64_t ly[POLY_SIZE], coef[POLY_SIZE];
    for (int j = 0; j < POLY_SIZE; j++) ly[j] = Ysum[(size_t)i * S + xs[j]];
    bool ok = shamir_recover(xs, ly, coef);
    for (int s = 0; s < S; s++)
      if (!avail || avail[s]) ok = ok && (shamir_eval(coef, s) == Ysum[(size_t)i * S + s]);
    badc += !ok;
    for (int j = 0; j < POLY_SIZE; j++) coef_out[(size_t)i * POLY_SIZE + j] = coef[j];
  }
  return badc;
}

}  // extern "C"
