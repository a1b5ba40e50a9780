// Counter-based DP noise, fixed-point quantisation and exact-integer Shamir sharing (host+device).
//
// What it reproduces from the reference:
//  * noise: presampled table samples[100, d] ~ sigma*N(0,1), getNoise(it) = (-1/batch)*samples[it%100]
//    (ML/Pytorch/client_obj.py:59-67, 97-98; ML/code/logistic_model.py:79-83, 108-109).  Here any
//    table entry is regenerated on demand from (seed, peer, slot, coord) with Philox4x32-10, so the
//    "RequestNoise" RPC (DistSys/main.go:1592-1660) needs no data movement at all.
//  * quantisation: int64(x * 10^PRECISION), truncation toward zero (DistSys/kyber.go:698-710).
//  * Shamir: x_s = s - 10, y = sum_j a_j x^j over the integers (DistSys/kyber.go:579-609) — evaluated
//    exactly in int64 instead of the reference's float pow()+truncate; recovery (kyber.go:809-857,
//    float QR least squares + round) is replaced by exact Newton divided differences in 128-bit ints.
#pragma once
#include <stdint.h>
#include <math.h>
#include "field.cuh"

namespace bsc {

constexpr int POLY_SIZE = 10;      // DistSys/main.go:46
constexpr int MAX_SHARES = 64;     // TOTAL_SHARES = ceil(2*POLY_SIZE/na)*na <= 64 for na <= 32
constexpr int SHARE_X_OFFSET = 10; // x = s - 10, DistSys/kyber.go:588

struct Philox4 {
  uint32_t x, y, z, w;
};

HD uint32_t mulhi32(uint32_t a, uint32_t b) { return (uint32_t)(((uint64_t)a * b) >> 32); }

HD Philox4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 10; r++) {
    uint32_t h0 = mulhi32(0xD2511F53u, c0), l0 = 0xD2511F53u * c0;
    uint32_t h1 = mulhi32(0xCD9E8D57u, c2), l1 = 0xCD9E8D57u * c2;
    uint32_t n0 = h1 ^ c1 ^ k0, n1 = l1, n2 = h0 ^ c3 ^ k1, n3 = l0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  Philox4 o;
  o.x = c0; o.y = c1; o.z = c2; o.w = c3;
  return o;
}

// Four standard normals for coordinates 4*cb .. 4*cb+3 of table row `slot` of peer `peer`.
HD void gauss4(uint64_t seed, uint32_t peer, uint32_t slot, uint32_t cb, float out[4]) {
  Philox4 r = philox4x32_10(cb, slot, peer, 0x4e6f6973u, (uint32_t)seed, (uint32_t)(seed >> 32));
  const float inv32 = 2.3283064365386963e-10f;  // 2^-32
  float u1 = ((float)r.x + 1.0f) * inv32;
  float u2 = (float)r.y * inv32;
  float u3 = ((float)r.z + 1.0f) * inv32;
  float u4 = (float)r.w * inv32;
  if (u1 > 1.0f) u1 = 1.0f;
  if (u3 > 1.0f) u3 = 1.0f;
  float r1 = sqrtf(-2.0f * logf(u1));
  float r2 = sqrtf(-2.0f * logf(u3));
  const float twopi = 6.283185307179586f;
  out[0] = r1 * cosf(twopi * u2);
  out[1] = r1 * sinf(twopi * u2);
  out[2] = r2 * cosf(twopi * u4);
  out[3] = r2 * sinf(twopi * u4);
}
HD float gauss1(uint64_t seed, uint32_t peer, uint32_t slot, uint32_t coord) {
  float g[4];
  gauss4(seed, peer, slot, coord >> 2, g);
  return g[coord & 3];
}

// kyber.go:698-710 — Go's int64(float64) conversion truncates toward zero.
HD int64_t quantize_fixed(float x, double scale) { return (int64_t)((double)x * scale); }

// Exact Horner evaluation at x = s - 10 for s in [0, nshares).  Wrapping int64 arithmetic:
// additive homomorphism holds mod 2^64, and true values stay < 2^63 for clipped updates.
HD int64_t shamir_eval(const int64_t a[POLY_SIZE], int s) {
  const int64_t x = (int64_t)s - SHARE_X_OFFSET;
  uint64_t acc = 0;
#pragma unroll
  for (int j = POLY_SIZE - 1; j >= 0; j--) acc = acc * (uint64_t)x + (uint64_t)a[j];
  return (int64_t)acc;
}

// num / den for a small non-zero den, assuming the division is exact (the caller verifies q*den == num).
// Device code has no 128-bit divide instruction — a software __int128 division is hundreds of
// instructions, and share recovery does 45 of them per chunk — so the quotient is formed as
// (num >> tz) * inverse(odd part of den) mod 2^128: the 64-bit inverse comes from five Newton steps and one
// more step widens it to 128 bits.
HD __int128 exact_div_small(__int128 num, int64_t den) {
  const bool neg = den < 0;
  uint64_t d = neg ? (uint64_t)0 - (uint64_t)den : (uint64_t)den;
  if (d == 0) return 0;   // duplicate share indices: the caller's q*den == num check reports it
  int tz = 0;
  while ((d & 1) == 0) {
    d >>= 1;
    tz++;
  }
  uint64_t inv = d;                       // d*d = 1 (mod 8): three correct bits
#pragma unroll
  for (int it = 0; it < 5; it++) inv *= 2 - d * inv;   // 6, 12, 24, 48, 96 bits (mod 2^64)
  const unsigned __int128 d128 = d;
  unsigned __int128 inv128 = inv;
  inv128 *= (unsigned __int128)2 - d128 * inv128;       // 128 bits
  const __int128 shifted = num >> tz;                   // arithmetic shift; exact when 2^tz divides num
  const __int128 q = (__int128)((unsigned __int128)shifted * inv128);
  return neg ? -q : q;
}

// Recover the POLY_SIZE integer coefficients from POLY_SIZE shares (xs[i] = share index s_i,
// ys[i] = aggregated share).  Exact: Newton divided differences of an integer polynomial at
// integer nodes are integers, so every division below is exact.  Returns false if a division
// leaves a remainder (corrupted / inconsistent shares).
HDN inline bool shamir_recover(const int* xs, const int64_t* ys, int64_t out[POLY_SIZE]) {
  __int128 dd[POLY_SIZE];
  for (int i = 0; i < POLY_SIZE; i++) dd[i] = (__int128)ys[i];
  bool exact = true;
  for (int k = 1; k < POLY_SIZE; k++) {
    for (int i = POLY_SIZE - 1; i >= k; i--) {
      __int128 num = dd[i] - dd[i - 1];
      int64_t den = (int64_t)(xs[i] - xs[i - k]);
      const __int128 q = exact_div_small(num, den);
      if (q * den != num) exact = false;
      dd[i] = q;
    }
  }
  // Newton -> monomial: p(x) = dd[0] + (x-x0)(dd[1] + (x-x1)(...))
  __int128 c[POLY_SIZE];
  for (int i = 0; i < POLY_SIZE; i++) c[i] = 0;
  c[0] = dd[POLY_SIZE - 1];
  int deg = 0;
  for (int k = POLY_SIZE - 2; k >= 0; k--) {
    const __int128 xk = (__int128)(xs[k] - SHARE_X_OFFSET);
    // c(x) = c(x) * (x - xk) + dd[k]
    deg++;
    for (int i = deg; i >= 1; i--) c[i] = c[i - 1] - xk * c[i];
    c[0] = dd[k] - xk * c[0];
  }
  for (int i = 0; i < POLY_SIZE; i++) out[i] = (int64_t)c[i];
  return exact;
}

// Recovery as one exact matrix product.  The share abscissae are fixed small integers, so the inverse of the
// Vandermonde system is a constant rational matrix: coef_j = (sum_i M[j][i] * y_i) / D with integer M (Lagrange
// basis coefficients over the common denominator D; for ten consecutive nodes D = 9! and |M| < 2^37).  Each
// coefficient is an independent 10-term dot product in 128-bit integers and one exact division by a constant
// (multiplication by the precomputed inverse of D's odd part mod 2^128), so a chunk's recovery spreads over ten
// threads instead of one thread walking the 45 dependent exact divisions of shamir_recover.  The host builds
// the matrix with big integers (ops/secagg.py: recover_matrix) and refuses configurations whose dot products
// could leave 128 bits; those fall back to shamir_recover.  Replaces kyber.go:809-857 (float QR + round).
struct RecoverMat {
  long long M[POLY_SIZE * POLY_SIZE];   // row j = coefficient, column i = share xs[i]
  long long D;                          // common denominator, > 0
  unsigned long long inv_lo, inv_hi;    // (D >> tz)^-1 mod 2^128
  int tz;                               // trailing zero bits of D
  int xs[POLY_SIZE];                    // share indices the matrix was built for
};

HD bool recover_coef_mat(const RecoverMat& R, const int64_t* y, int j, int64_t& coef) {
  __int128 acc = 0;
#pragma unroll
  for (int i = 0; i < POLY_SIZE; i++) acc += (__int128)R.M[j * POLY_SIZE + i] * (__int128)y[i];
  const __int128 sh = acc >> R.tz;
  const unsigned __int128 inv = ((unsigned __int128)R.inv_hi << 64) | (unsigned __int128)R.inv_lo;
  const __int128 q = (__int128)((unsigned __int128)sh * inv);
  coef = (int64_t)q;
  return q * (__int128)R.D == acc;
}

}  // namespace bsc
