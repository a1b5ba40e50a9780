// BN256 G1 multi-scalar multiplication for polynomial commitments and KZG witnesses, batched over
// peers x chunks x shares.
//
// Reference work replaced (all double-and-add per term on the CPU, amd64 asm field):
//   createCommitment            DistSys/kyber.go:533-562   sum_i coeff_i * PKG1_i
//   fillPolynomialMap           DistSys/kyber.go:172-202   per-chunk commitment with SRS slice [10c,10c+10)
//   createShareAndWitness       DistSys/kyber.go:579-646   witness = commit(quotient (p(X)-p(x))/(X-x))
//   aggregateSecret             DistSys/kyber.go:244-287   G1 adds of commitments / witnesses
//
// B200 design: witnesses are rewritten as 9-term MSMs with the *update coefficients themselves* as
// scalars over a precomputed per-(chunk, share) base table
//     W_{c,s} = sum_{j=1..9} a_j * T_{c,s,j},   T_{c,s,j} = sum_{k<j} x_s^{j-1-k} * PKG1[10c+k]
// (identical group element to committing to the quotient polynomial), so every scalar is a small
// signed fixed-point integer (|a| <~ 2^20) instead of a ~2^50 quotient coefficient.  One thread
// computes one output with an interleaved (Straus) doubling chain; consecutive threads share the
// bases of a group so base loads are warp-broadcast.  Results are normalised with Montgomery's
// batched-inversion trick.
#include <stdlib.h>
#include "common.cuh"
#include <mutex>
#include "g1.cuh"

namespace bsc {

constexpr int MSM_TMAX = 16;

//   Q      : [rows][ldq] int64 scalars ; sel[p] (optional) maps output peer p to a row of Q
//   sidx[g]: offset of group g's first scalar within a row ; bidx[g]: index of its first base
//   og[g]  : output offset of group g (optional, default g) ; out[p*ostride_p + og[g]]
__global__ void __launch_bounds__(128) msm_groups_kernel(const long long* __restrict__ Q, int ldq,
                                                         const int* __restrict__ sel, int P,
                                                         const int* __restrict__ sidx, const int* __restrict__ bidx,
                                                         const int* __restrict__ og, int G, int T,
                                                         const G1Affine* __restrict__ bases, G1Jac* __restrict__ out,
                                                         long long ostride_p) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const bool live = i < (long long)G * P;
  const int g = live ? (int)(i / P) : 0, p = live ? (int)(i % P) : 0;
  const long long* row = Q + (size_t)(sel ? sel[p] : p) * ldq + sidx[g];
  int64_t s[MSM_TMAX];
  uint64_t maxmag = 0;
  for (int t = 0; t < T; t++) {
    s[t] = live ? row[t] : 0;
    const uint64_t m = s[t] < 0 ? (uint64_t)0 - (uint64_t)s[t] : (uint64_t)s[t];
    maxmag |= m;
  }
  int nbits = 64 - __clzll((long long)maxmag);
  if (maxmag == 0) nbits = 0;
  // keep the warp in lock-step on the shared doubling chain
  nbits = __reduce_max_sync(0xffffffffu, nbits);
  const G1Jac acc = g1_straus_i64<MSM_TMAX>(bases + bidx[g], s, T, nbits);
  if (live) out[(size_t)p * ostride_p + (og ? og[g] : g)] = acc;
}

// ---- fixed-base window tables ----------------------------------------------------------------------
// Every base of the commitment / witness MSMs is fixed for the life of a run (SRS points and the
// witness base table), so the doubling chain can be paid once: table[b][i][k-1] = k * 2^(w*i) * B_b for
// k = 1..2^(w-1).  A scalar is recoded into NW signed w-bit digits and each term costs NW mixed
// additions and no doublings.  With w = 7, NW = 3 (|scalar| < 2^20, the clipped fixed-point range)
// a 9-term witness drops from ~20 doublings + ~90 additions to 27 additions.
// Two-width tables.  Clipped fixed-point updates are small (|u * 10^4| < 2^12 for ~99 % of the coordinates of a softmax
// step, a handful of bits more for the rest), so the first window should be as wide as the memory budget allows - one
// mixed addition per term - and the windows above it only have to cover the rare large values and can be narrow.
// `nw` carries both numbers: low byte = window count, next byte = width of windows 1.. (0 = same width as window 0).
HD int tw_count(int nw) { return nw & 0xff; }
HD int tw_w1(int w, int nw) { const int w1 = (nw >> 8) & 0xff; return w1 ? w1 : w; }
HD int tw_half(int w, int nw, int i) { return 1 << ((i == 0 ? w : tw_w1(w, nw)) - 1); }
HD long long tw_offset(int w, int nw, int i) { return i == 0 ? 0 : (long long)tw_half(w, nw, 0) + (long long)(i - 1) * tw_half(w, nw, 1); }
HD long long tw_entries(int w, int nw) { return tw_offset(w, nw, tw_count(nw)); }
HD int tw_shift(int w, int nw, int i) { return i == 0 ? 0 : w + (i - 1) * tw_w1(w, nw); }
HD int64_t table_max_mag(int w, int nw) { return (int64_t)1 << (tw_shift(w, nw, tw_count(nw)) - 1); }

// one thread per (base, window): emits the 2^(w-1) multiples as Jacobian points (normalised afterwards)
__global__ void __launch_bounds__(128) msm_table_build_kernel(const G1Affine* __restrict__ bases, long long nb, int w,
                                                              int nw, G1Jac* __restrict__ out) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int cnt = tw_count(nw);
  if (t >= nb * cnt) return;
  const long long b = t / cnt;
  const int i = (int)(t % cnt);
  const int H = tw_half(w, nw, i);
  G1Jac J = g1j_from_affine(bases[b]);
  for (int k = 0; k < tw_shift(w, nw, i); k++) J = g1j_dbl(J);
  G1Jac acc = J;
  G1Jac* dst = out + (size_t)b * tw_entries(w, nw) + tw_offset(w, nw, i);
  dst[0] = acc;
  for (int k = 1; k < H; k++) {
    acc = g1j_add(acc, J);
    dst[k] = acc;
  }
}

template <int TMAX>
HDN inline G1Jac g1_table_msm(const G1Affine* __restrict__ table, long long first_base, const int64_t* s, int T, int w,
                              int nw) {
  const int cnt = tw_count(nw);
  const long long per_base = tw_entries(w, nw);
  const uint64_t lim = (uint64_t)table_max_mag(w, nw);
  bool overflow = false;
  for (int t = 0; t < T; t++) {
    const uint64_t m = s[t] < 0 ? (uint64_t)0 - (uint64_t)s[t] : (uint64_t)s[t];
    overflow = overflow || (m >= lim);
  }
  if (overflow) {
    // rare: a scalar outside the table range -> interleaved double-and-add over the k=1 entries
    G1Affine b[TMAX];
    uint64_t maxmag = 0;
    for (int t = 0; t < T; t++) {
      b[t] = table[(size_t)(first_base + t) * per_base];
      maxmag |= s[t] < 0 ? (uint64_t)0 - (uint64_t)s[t] : (uint64_t)s[t];
    }
    int nbits = 0;
    while (nbits < 64 && (maxmag >> nbits)) nbits++;
    return g1_straus_i64<TMAX>(b, s, T, nbits);
  }
  G1Jac acc = g1j_inf();
  for (int t = 0; t < T; t++) {
    const bool neg = s[t] < 0;
    uint64_t m = neg ? (uint64_t)0 - (uint64_t)s[t] : (uint64_t)s[t];
    const G1Affine* tb = table + (size_t)(first_base + t) * per_base;
    for (int i = 0; i < cnt && m != 0; i++) {
      const int wi = i == 0 ? w : tw_w1(w, nw);
      const uint64_t H = (uint64_t)1 << (wi - 1);
      uint64_t dgt = m & (((uint64_t)1 << wi) - 1);
      m >>= wi;
      bool dneg = neg;
      if (dgt > H) {
        dgt = ((uint64_t)1 << wi) - dgt;
        dneg = !neg;
        m += 1;
      }
      if (dgt) {
        G1Affine pt = tb[tw_offset(w, nw, i) + (long long)(dgt - 1)];
        if (dneg) pt.y = fe_neg(pt.y);
        acc = g1j_add_affine(acc, pt);
      }
    }
  }
  return acc;
}

// same indexing contract as msm_groups_kernel, bases replaced by their window tables
template <int MINB>
__global__ void __launch_bounds__(128, MINB) msm_table_kernel(const long long* __restrict__ Q, int ldq,
                                                        const int* __restrict__ sel, int P,
                                                        const int* __restrict__ sidx, const int* __restrict__ bidx,
                                                        const int* __restrict__ og, int G, int T,
                                                        const G1Affine* __restrict__ table, int w, int nw,
                                                        G1Jac* __restrict__ out, long long ostride_p,
                                                        const int* __restrict__ pcount) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  // Row count decided on the device (block_plan_kernel); P is only its upper bound (the grid is sized for it).  The live
  // count, not P, is the stride of the (group, row) mapping: warps stay full of live rows and whole CTAs retire at once,
  // instead of every warp carrying P - count idle lanes.
  if (pcount) P = min(P, *pcount);
  if (i >= (long long)G * P) return;
  // Groups vary fastest along a warp.  The 21 witnesses of one (peer, chunk) share their nine scalars, and consecutive
  // chunks of one peer share the class row whose gradient is (or is not) negligible - so the lanes of a warp agree on
  // which digits are zero and a zero digit really skips its addition, instead of idling next to 31 other peers' digits.
  const int p = (int)(i / G), g = (int)(i % G);
  const long long* row = Q + (size_t)(sel ? sel[p] : p) * ldq + sidx[g];
  // scalars are read in place (twice: range check, then digits) rather than staged in registers
  out[(size_t)p * ostride_p + (og ? og[g] : g)] = g1_table_msm<MSM_TMAX>(table, bidx[g], (const int64_t*)row, T, w, nw);
}

// Fixed-base comb for full-width scalars: out[i] = k_i * B with B's window table (w bits, nw windows,
// nw*w >= 257 so the final carry has a window).  Used for the Schnorr commitments T = v*G of the
// verifiers' approvals (kyber.go:873-896): 33 mixed additions instead of ~380 point operations.
// The approvals sit on the path to the block's tail with only ~200 scalars to hide latency behind, so one scalar
// is spread over COMB_LANES threads: every lane recodes the scalar (cheap integer work), adds the windows
// lane, lane + COMB_LANES, ... and the partial sums meet in a shared-memory tree - ceil(33/8) = 5 dependent mixed
// additions + 3 Jacobian additions instead of 33.
constexpr int COMB_LANES = 8;
__global__ void __launch_bounds__(128) g1_comb_mul_kernel(const U256* __restrict__ k, int N,
                                                          const G1Affine* __restrict__ table, int w, int nw,
                                                          G1Jac* __restrict__ out) {
  __shared__ G1Jac sh[128];
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int i = t / COMB_LANES, lane = t % COMB_LANES;
  const int H = 1 << (w - 1);
  const uint32_t mask = (1u << w) - 1u;
  G1Jac acc = g1j_inf();
  if (i < N) {
    const U256 s = k[i];
    // pass 1: the signed digits (the recoding carry is sequential, but it is a few integer operations per window)
    short dg[48];
    uint32_t carry = 0;
    for (int win = 0; win < nw; win++) {
      const int bit = win * w;
      uint32_t d = carry;
      if (bit < 256) {
        const int limb = bit >> 5, shft = bit & 31;
        uint64_t two = s.v[limb];
        if (limb + 1 < 8) two |= (uint64_t)s.v[limb + 1] << 32;
        d += (uint32_t)(two >> shft) & mask;
      }
      carry = 0;
      int sd = (int)d;
      if (d > (uint32_t)H) {
        sd = (int)d - (1 << w);
        carry = 1;
      }
      dg[win] = (short)sd;
    }
    // pass 2: every lane walks ITS windows (lane, lane + 8, ...), so the lanes of a warp add at the same time - a branch
    // on "is this window mine" inside one loop over all windows would serialise the additions again
    for (int win = lane; win < nw; win += COMB_LANES) {
      const int sd = dg[win];
      if (sd) {
        G1Affine pt = table[(size_t)win * H + ((sd < 0 ? -sd : sd) - 1)];
        if (sd < 0) pt.y = fe_neg(pt.y);
        acc = g1j_add_affine(acc, pt);
      }
    }
  }
  sh[threadIdx.x] = acc;
  __syncthreads();
  for (int step = COMB_LANES >> 1; step > 0; step >>= 1) {
    if (lane < step) sh[threadIdx.x] = g1j_add(sh[threadIdx.x], sh[threadIdx.x + step]);
    __syncthreads();
  }
  if (i < N && lane == 0) out[i] = sh[threadIdx.x];
}

// k_i * B for full 256-bit scalars (SRS generation: PKG1[i] = tau^i * G, DistSys/publicKey.go:28-66)
__global__ void g1_fixed_base_mul_kernel(const U256* __restrict__ k, int N, G1Affine base, G1Jac* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  out[i] = g1_mul_u256(base, k[i]);
}

// Jacobian -> affine (Montgomery form), NORM_BATCH points per thread share one field inversion.  Large inputs (table
// builds, millions of points) use batches of 8: an inversion costs ~35 multiplications, so sharing it cuts the work by 3x.
// Small inputs (a block's 100 commitments, its 210 approval points) sit on the block's critical path and are latency-
// bound: one point per thread (NORM_BATCH = 1) makes the kernel a single inversion deep instead of 8 + 1 + 8 * 6
// dependent multiplications on 13 threads.
template <int NORM_BATCH>
__global__ void __launch_bounds__(128) g1_normalize_kernel(const G1Jac* __restrict__ in, G1Affine* __restrict__ out,
                                                           long long N) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long base = t * NORM_BATCH;
  if (base >= N) return;
  Fp z[NORM_BATCH], pre[NORM_BATCH];
  const int cnt = (int)min((long long)NORM_BATCH, N - base);
  Fp run = fe_one<BnP>();
  for (int i = 0; i < NORM_BATCH; i++) {
    z[i] = (i < cnt) ? in[base + i].z : fe_one<BnP>();
    if (fe_is_zero(z[i])) z[i] = fe_one<BnP>();  // infinity: handled by the flag below
    run = fe_mul(run, z[i]);
    pre[i] = run;
  }
  Fp inv = fe_inv(run);
  for (int i = NORM_BATCH - 1; i >= 0; i--) {
    const Fp zi = (i > 0) ? fe_mul(inv, pre[i - 1]) : inv;
    inv = fe_mul(inv, z[i]);
    if (i < cnt) {
      const G1Jac pt = in[base + i];
      out[base + i] = g1j_to_affine_with_inv(pt, zi);
    }
  }
}

__global__ void g1_affine_to_bytes_kernel(const G1Affine* __restrict__ in, unsigned char* __restrict__ out, long long N) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  g1a_to_bytes(out + 64 * i, in[i]);
}
__global__ void g1_bytes_to_affine_kernel(const unsigned char* __restrict__ in, G1Affine* __restrict__ out, long long N) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  out[i] = g1a_from_bytes(in + 64 * i);
}

// Miner aggregation (kyber.go:244-287): out[g] = sum_{q<nsel} pts[sel[q]*G + g], affine inputs.
__global__ void __launch_bounds__(128) g1_sum_rows_kernel(const G1Affine* __restrict__ pts, const int* __restrict__ sel,
                                                          int nsel, long long G, G1Jac* __restrict__ out,
                                                          const int* __restrict__ nsel_dev) {
  const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= G) return;
  if (nsel_dev) nsel = min(nsel, *nsel_dev);
  G1Jac acc = g1j_inf();
  for (int q = 0; q < nsel; q++) acc = g1j_add_affine(acc, pts[(size_t)sel[q] * G + g]);
  out[g] = acc;
}

// Full-vector commitment = sum of the chunk commitments: one CTA per row,
// strided partial sums then a shared-memory Jacobian tree.  out[p] = sum_{g<G} pts[p*G + g].
__global__ void __launch_bounds__(128) g1_sum_cols_kernel(const G1Affine* __restrict__ pts, int G,
                                                          G1Jac* __restrict__ out) {
  __shared__ G1Jac sh[128];
  const int p = blockIdx.x, tid = threadIdx.x;
  G1Jac acc = g1j_inf();
  for (int g = tid; g < G; g += blockDim.x) acc = g1j_add_affine(acc, pts[(size_t)p * G + g]);
  sh[tid] = acc;
  __syncthreads();
  for (int s = blockDim.x >> 1; s > 0; s >>= 1) {
    if (tid < s) sh[tid] = g1j_add(sh[tid], sh[tid + s]);
    __syncthreads();
  }
  if (tid == 0) out[p] = sh[0];
}

// Jacobian-input variants: points stay in Jacobian form between kernels (no field inversion on the
// block's critical path); only what enters the ledger is normalised.  sel_rows: rows are strided by
// `row_stride` points; the peer dimension is split over gridDim.y partial sums (fixed order) to shorten
// the dependent addition chain, and the partials are added by the y == 0 thread's owner afterwards.
__global__ void __launch_bounds__(128) g1_sum_rows_jac_kernel(const G1Jac* __restrict__ pts, const int* __restrict__ sel,
                                                              int nsel, long long G, G1Jac* __restrict__ out,
                                                              const int* __restrict__ nsel_dev) {
  const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= G) return;
  if (nsel_dev) nsel = min(nsel, *nsel_dev);
  G1Jac acc = g1j_inf();
  for (int q = 0; q < nsel; q++) acc = g1j_add(acc, pts[(size_t)sel[q] * G + g]);
  out[g] = acc;
}
// Column sum in two stages: grid (P, S) CTAs each tree-reduce a slice of <= 128 points of one row (one point per thread,
// no serial prefix), then one warp per row adds the S partial sums.  The single-stage version (one CTA per row, six
// serial additions per thread before the tree) occupied 100 of 148 SMs with four warps each and sat on the path
// commitments -> approvals -> read-back of every block.
__global__ void __launch_bounds__(128) g1_sum_cols_jac_part_kernel(const G1Jac* __restrict__ pts, int G, int per,
                                                                   G1Jac* __restrict__ part) {
  __shared__ G1Jac sh[128];
  const int p = blockIdx.x, sl = blockIdx.y, tid = threadIdx.x;
  const int g0 = sl * per, g1 = min(G, g0 + per);
  G1Jac acc = g1j_inf();
  for (int g = g0 + tid; g < g1; g += blockDim.x) acc = g1j_add(acc, pts[(size_t)p * G + g]);
  sh[tid] = acc;
  __syncthreads();
  for (int s = blockDim.x >> 1; s > 0; s >>= 1) {
    if (tid < s) sh[tid] = g1j_add(sh[tid], sh[tid + s]);
    __syncthreads();
  }
  if (tid == 0) part[(size_t)p * gridDim.y + sl] = sh[0];
}
__global__ void __launch_bounds__(32) g1_sum_cols_jac_final_kernel(const G1Jac* __restrict__ part, int S,
                                                                   G1Jac* __restrict__ out) {
  __shared__ G1Jac sh[32];
  const int p = blockIdx.x, tid = threadIdx.x;
  G1Jac acc = g1j_inf();
  for (int q = tid; q < S; q += 32) acc = g1j_add(acc, part[(size_t)p * S + q]);
  sh[tid] = acc;
  __syncwarp();
  for (int s = 16; s > 0; s >>= 1) {
    if (tid < s && tid + s < S) sh[tid] = g1j_add(sh[tid], sh[tid + s]);
    __syncwarp();
  }
  if (tid == 0) out[p] = sh[0];
}

}  // namespace bsc

using namespace bsc;

extern "C" {

int bsc_msm_groups(const long long* Q, int ldq, const int* sel, int P, const int* sidx, const int* bidx, const int* og,
                   int G, int T, const void* bases_affine, void* out_jac, long long ostride_p, void* stream) {
  if (T > MSM_TMAX) return -1;
  const long long n = (long long)G * P;
  if (n == 0) return 0;
  msm_groups_kernel<<<(unsigned)((n + 127) / 128), 128, 0, (cudaStream_t)stream>>>(
      Q, ldq, sel, P, sidx, bidx, og, G, T, (const G1Affine*)bases_affine, (G1Jac*)out_jac, ostride_p);
  return finish_launch("msm_groups");
}

int bsc_msm_table_build(const void* bases_affine, long long nb, int w, int nw, void* out_jac, void* stream) {
  if (w < 2 || w > 14 || tw_count(nw) < 1 || tw_w1(w, nw) > 14) return -1;
  const long long n = nb * tw_count(nw);
  if (n == 0) return 0;
  msm_table_build_kernel<<<(unsigned)((n + 127) / 128), 128, 0, (cudaStream_t)stream>>>((const G1Affine*)bases_affine,
                                                                                       nb, w, nw, (G1Jac*)out_jac);
  return finish_launch("msm_table_build");
}
void bsc_msm_table_build_host(const void* bases_affine, long long nb, int w, int nw, void* out_jac) {
  const G1Affine* bases = (const G1Affine*)bases_affine;
  G1Jac* out = (G1Jac*)out_jac;
  const int cnt = tw_count(nw);
  for (long long t = 0; t < nb * cnt; t++) {
    const int i = (int)(t % cnt);
    const int H = tw_half(w, nw, i);
    G1Jac J = g1j_from_affine(bases[t / cnt]);
    for (int k = 0; k < tw_shift(w, nw, i); k++) J = g1j_dbl(J);
    G1Jac acc = J;
    G1Jac* dst = out + (size_t)(t / cnt) * tw_entries(w, nw) + tw_offset(w, nw, i);
    dst[0] = acc;
    for (int k = 1; k < H; k++) {
      acc = g1j_add(acc, J);
      dst[k] = acc;
    }
  }
}
int bsc_msm_table_dc(const long long* Q, int ldq, const int* sel, int P, const int* pcount, const int* sidx, const int* bidx,
                     const int* og, int G, int T, const void* table_affine, int w, int nw, void* out_jac,
                     long long ostride_p, void* stream);
int bsc_msm_table(const long long* Q, int ldq, const int* sel, int P, const int* sidx, const int* bidx, const int* og,
                  int G, int T, const void* table_affine, int w, int nw, void* out_jac, long long ostride_p,
                  void* stream) {
  return bsc_msm_table_dc(Q, ldq, sel, P, nullptr, sidx, bidx, og, G, T, table_affine, w, nw, out_jac, ostride_p, stream);
}
// pcount (may be null): device-resident number of live rows (<= P); rows >= *pcount are skipped
int bsc_msm_table_dc(const long long* Q, int ldq, const int* sel, int P, const int* pcount, const int* sidx, const int* bidx,
                     const int* og, int G, int T, const void* table_affine, int w, int nw, void* out_jac,
                     long long ostride_p, void* stream) {
  if (T > MSM_TMAX) return -1;
  const long long n = (long long)G * P;
  if (n == 0) return 0;
  static const int minb = getenv("BSC_MSM_MINB") ? atoi(getenv("BSC_MSM_MINB")) : 4;
  const unsigned grid = (unsigned)((n + 127) / 128);
  if (minb >= 8)
    msm_table_kernel<8><<<grid, 128, 0, (cudaStream_t)stream>>>(Q, ldq, sel, P, sidx, bidx, og, G, T, (const G1Affine*)table_affine, w, nw, (G1Jac*)out_jac, ostride_p, pcount);
  else if (minb >= 6)
    msm_table_kernel<6><<<grid, 128, 0, (cudaStream_t)stream>>>(Q, ldq, sel, P, sidx, bidx, og, G, T, (const G1Affine*)table_affine, w, nw, (G1Jac*)out_jac, ostride_p, pcount);
  else
    msm_table_kernel<4><<<grid, 128, 0, (cudaStream_t)stream>>>(Q, ldq, sel, P, sidx, bidx, og, G, T, (const G1Affine*)table_affine, w, nw, (G1Jac*)out_jac, ostride_p, pcount);
  return finish_launch("msm_table");
}
void bsc_msm_table_host(const long long* Q, int ldq, const int* sel, int P, const int* sidx, const int* bidx,
                        const int* og, int G, int T, const void* table_affine, int w, int nw, void* out_jac,
                        long long ostride_p) {
  G1Jac* out = (G1Jac*)out_jac;
  for (int g = 0; g < G; g++)
    for (int p = 0; p < P; p++) {
      const long long* row = Q + (size_t)(sel ? sel[p] : p) * ldq + sidx[g];
      out[(size_t)p * ostride_p + (og ? og[g] : g)] =
          g1_table_msm<MSM_TMAX>((const G1Affine*)table_affine, bidx[g], (const int64_t*)row, T, w, nw);
    }
}

int bsc_g1_comb_mul(const void* scalars_u256, int N, const void* table_affine, int w, int nw, void* out_jac,
                    void* stream) {
  if (N == 0) return 0;
  if (w < 2 || w > 12 || nw * w < 257 || nw > 48) return -1;
  const long long threads = (long long)N * COMB_LANES;
  g1_comb_mul_kernel<<<(unsigned)((threads + 127) / 128), 128, 0, (cudaStream_t)stream>>>((const U256*)scalars_u256, N,
                                                                         (const G1Affine*)table_affine, w, nw,
                                                                         (G1Jac*)out_jac);
  return finish_launch("g1_comb_mul");
}

int bsc_g1_fixed_base_mul(const void* scalars_u256, int N, const void* base_affine, void* out_jac, void* stream) {
  G1Affine b;
  BSC_CHECK(cudaMemcpy(&b, base_affine, sizeof(b), cudaMemcpyDefault));
  g1_fixed_base_mul_kernel<<<(N + 63) / 64, 64, 0, (cudaStream_t)stream>>>((const U256*)scalars_u256, N, b,
                                                                            (G1Jac*)out_jac);
  return finish_launch("g1_fixed_base_mul");
}

int bsc_g1_normalize(const void* in_jac, void* out_affine, long long N, void* stream) {
  if (N == 0) return 0;
  if (N <= 8192) {
    g1_normalize_kernel<1><<<(unsigned)((N + 127) / 128), 128, 0, (cudaStream_t)stream>>>((const G1Jac*)in_jac,
                                                                                          (G1Affine*)out_affine, N);
  } else {
    const long long nt = (N + 7) / 8;
    g1_normalize_kernel<8><<<(unsigned)((nt + 127) / 128), 128, 0, (cudaStream_t)stream>>>((const G1Jac*)in_jac,
                                                                                           (G1Affine*)out_affine, N);
  }
  return finish_launch("g1_normalize");
}

int bsc_g1_affine_to_bytes(const void* in_affine, unsigned char* out, long long N, void* stream) {
  if (N == 0) return 0;
  g1_affine_to_bytes_kernel<<<(unsigned)((N + 127) / 128), 128, 0, (cudaStream_t)stream>>>((const G1Affine*)in_affine,
                                                                                          out, N);
  return finish_launch("g1_affine_to_bytes");
}

int bsc_g1_bytes_to_affine(const unsigned char* in, void* out_affine, long long N, void* stream) {
  if (N == 0) return 0;
  g1_bytes_to_affine_kernel<<<(unsigned)((N + 127) / 128), 128, 0, (cudaStream_t)stream>>>(in, (G1Affine*)out_affine,
                                                                                          N);
  return finish_launch("g1_bytes_to_affine");
}

int bsc_g1_sum_rows_dc(const void* pts_affine, const int* sel, int nsel, const int* nsel_dev, long long G, void* out_jac,
                       void* stream) {
  if (G == 0) return 0;
  g1_sum_rows_kernel<<<(unsigned)((G + 127) / 128), 128, 0, (cudaStream_t)stream>>>((const G1Affine*)pts_affine, sel,
                                                                                   nsel, G, (G1Jac*)out_jac, nsel_dev);
  return finish_launch("g1_sum_rows");
}
int bsc_g1_sum_rows(const void* pts_affine, const int* sel, int nsel, long long G, void* out_jac, void* stream) {
  return bsc_g1_sum_rows_dc(pts_affine, sel, nsel, nullptr, G, out_jac, stream);
}

// nsel_dev (may be null): device-resident row count (<= nsel)
int bsc_g1_sum_rows_jac_dc(const void* pts_jac, const int* sel, int nsel, const int* nsel_dev, long long G, void* out_jac,
                           void* stream) {
  if (G == 0) return 0;
  g1_sum_rows_jac_kernel<<<(unsigned)((G + 127) / 128), 128, 0, (cudaStream_t)stream>>>((const G1Jac*)pts_jac, sel, nsel,
                                                                                       G, (G1Jac*)out_jac, nsel_dev);
  return finish_launch("g1_sum_rows_jac");
}
int bsc_g1_sum_rows_jac(const void* pts_jac, const int* sel, int nsel, long long G, void* out_jac, void* stream) {
  return bsc_g1_sum_rows_jac_dc(pts_jac, sel, nsel, nullptr, G, out_jac, stream);
}
// (workspace for the partial sums: grown on demand, one per device, kept for the life of the process; the first call
// on a device - or one with a larger P * ceil(G/128) than any before - allocates, so make it outside stream capture)
int bsc_g1_sum_cols_jac(const void* pts_jac, int P, int G, void* out_jac, void* stream) {
  if (P == 0) return 0;
  static std::mutex mu;
  static G1Jac* ws[64] = {nullptr};
  static size_t ws_n[64] = {0};
  int dev = 0;
  BSC_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return -2;
  const int S = (G + 127) / 128, per = (G + S - 1) / S;
  const size_t need = (size_t)P * S;
  G1Jac* part;
  {
    std::lock_guard<std::mutex> lk(mu);
    if (ws_n[dev] < need) {
      if (ws[dev]) BSC_CHECK(cudaFree(ws[dev]));
      ws[dev] = nullptr;
      ws_n[dev] = 0;
      BSC_CHECK(cudaMalloc((void**)&ws[dev], need * sizeof(G1Jac)));
      ws_n[dev] = need;
    }
    part = ws[dev];
  }
  g1_sum_cols_jac_part_kernel<<<dim3(P, S), 128, 0, (cudaStream_t)stream>>>((const G1Jac*)pts_jac, G, per, part);
  g1_sum_cols_jac_final_kernel<<<P, 32, 0, (cudaStream_t)stream>>>(part, S, (G1Jac*)out_jac);
  return finish_launch("g1_sum_cols_jac", 2);
}
void bsc_g1_sum_rows_jac_host(const void* pts_jac, const int* sel, int nsel, long long G, void* out_jac) {
  const G1Jac* pts = (const G1Jac*)pts_jac;
  G1Jac* out = (G1Jac*)out_jac;
  for (long long g = 0; g < G; g++) {
    G1Jac acc = g1j_inf();
    for (int q = 0; q < nsel; q++) acc = g1j_add(acc, pts[(size_t)sel[q] * G + g]);
    out[g] = acc;
  }
}
void bsc_g1_sum_cols_jac_host(const void* pts_jac, int P, int G, void* out_jac) {
  const G1Jac* pts = (const G1Jac*)pts_jac;
  G1Jac* out = (G1Jac*)out_jac;
  for (int p = 0; p < P; p++) {
    G1Jac acc = g1j_inf();
    for (int g = 0; g < G; g++) acc = g1j_add(acc, pts[(size_t)p * G + g]);
    out[p] = acc;
  }
}

int bsc_g1_sum_cols(const void* pts_affine, int P, int G, void* out_jac, void* stream) {
  if (P == 0) return 0;
  g1_sum_cols_kernel<<<P, 128, 0, (cudaStream_t)stream>>>((const G1Affine*)pts_affine, G, (G1Jac*)out_jac);
  return finish_launch("g1_sum_cols");
}


// ---- host mirrors (CPU plumbing path + oracles; same templates as the kernels) --------------------
void bsc_msm_groups_host(const long long* Q, int ldq, const int* sel, int P, const int* sidx, const int* bidx,
                         const int* og, int G, int T, const void* bases_affine, void* out_jac, long long ostride_p) {
  const G1Affine* bases = (const G1Affine*)bases_affine;
  G1Jac* out = (G1Jac*)out_jac;
  for (int g = 0; g < G; g++)
    for (int p = 0; p < P; p++) {
      const long long* row = Q + (size_t)(sel ? sel[p] : p) * ldq + sidx[g];
      uint64_t maxmag = 0;
      for (int t = 0; t < T; t++) maxmag |= row[t] < 0 ? (uint64_t)0 - (uint64_t)row[t] : (uint64_t)row[t];
      int nbits = 0;
      while (nbits < 64 && (maxmag >> nbits)) nbits++;
      out[(size_t)p * ostride_p + (og ? og[g] : g)] =
          g1_straus_i64<MSM_TMAX>(bases + bidx[g], (const int64_t*)row, T, nbits);
    }
}
void bsc_g1_normalize_host(const void* in_jac, void* out_affine, long long N) {
  const G1Jac* in = (const G1Jac*)in_jac;
  G1Affine* out = (G1Affine*)out_affine;
  for (long long i = 0; i < N; i++) out[i] = g1j_to_affine(in[i]);
}
void bsc_g1_sum_rows_host(const void* pts_affine, const int* sel, int nsel, long long G, void* out_jac) {
  const G1Affine* pts = (const G1Affine*)pts_affine;
  G1Jac* out = (G1Jac*)out_jac;
  for (long long g = 0; g < G; g++) {
    G1Jac acc = g1j_inf();
    for (int q = 0; q < nsel; q++) acc = g1j_add_affine(acc, pts[(size_t)sel[q] * G + g]);
    out[g] = acc;
  }
}
void bsc_g1_sum_cols_host(const void* pts_affine, int P, int G, void* out_jac) {
  const G1Affine* pts = (const G1Affine*)pts_affine;
  G1Jac* out = (G1Jac*)out_jac;
  for (int p = 0; p < P; p++) {
    G1Jac acc = g1j_inf();
    for (int g = 0; g < G; g++) acc = g1j_add_affine(acc, pts[(size_t)p * G + g]);
    out[p] = acc;
  }
}
void bsc_g1_fixed_base_mul_host(const void* scalars_u256, int N, const void* base_affine, void* out_jac) {
  const U256* k = (const U256*)scalars_u256;
  G1Jac* out = (G1Jac*)out_jac;
  for (int i = 0; i < N; i++) out[i] = g1_mul_u256(*(const G1Affine*)base_affine, k[i]);
}

}  // extern "C"
