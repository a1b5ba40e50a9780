// Latency-oriented Ed25519 scalar multiplication for the per-block VRF evaluation (host+device).
//
// The VRF of every peer is evaluated on the previous block hash (DistSys/vrf.go:54-100), so one
// ~253-bit scalar multiplication per peer sits on the critical path of every block and only ~100
// threads exist to hide latency behind.  The generic 8x32 Montgomery field (field.cuh) is a serial
// carry chain; here the field is 2^255-19 in radix 2^25.5 (ten signed limbs, 64-bit products, lazy
// carries) whose 100 partial products are independent, which is what a lone warp needs to keep the
// integer pipe busy.  Scalar recoding is signed radix-16 (64 digits in [-8, 8]) over a shared table of
// 1H..8H: 252 doublings + <= 64 additions.  Results are canonical compressed points, so they are
// bit-identical to the generic-field implementation in ed25519.cuh (tests compare the two).
#pragma once
#include <stdint.h>
#include "field.cuh"

namespace bsc {

struct F25 {
  int32_t v[10];
};

HD F25 f25_zero() {
  F25 r;
#pragma unroll
  for (int i = 0; i < 10; i++) r.v[i] = 0;
  return r;
}
HD F25 f25_one() {
  F25 r = f25_zero();
  r.v[0] = 1;
  return r;
}
HD F25 f25_add(const F25& a, const F25& b) {
  F25 r;
#pragma unroll
  for (int i = 0; i < 10; i++) r.v[i] = a.v[i] + b.v[i];
  return r;
}
HD F25 f25_sub(const F25& a, const F25& b) {
  F25 r;
#pragma unroll
  for (int i = 0; i < 10; i++) r.v[i] = a.v[i] - b.v[i];
  return r;
}
HD F25 f25_neg(const F25& a) {
  F25 r;
#pragma unroll
  for (int i = 0; i < 10; i++) r.v[i] = -a.v[i];
  return r;
}

// interleaved carry chain; even limbs hold 26 bits, odd limbs 25 bits
HD F25 f25_carry(int64_t h[10]) {
  int64_t c;
  c = (h[0] + ((int64_t)1 << 25)) >> 26; h[1] += c; h[0] -= c << 26;
  c = (h[4] + ((int64_t)1 << 25)) >> 26; h[5] += c; h[4] -= c << 26;
  c = (h[1] + ((int64_t)1 << 24)) >> 25; h[2] += c; h[1] -= c << 25;
  c = (h[5] + ((int64_t)1 << 24)) >> 25; h[6] += c; h[5] -= c << 25;
  c = (h[2] + ((int64_t)1 << 25)) >> 26; h[3] += c; h[2] -= c << 26;
  c = (h[6] + ((int64_t)1 << 25)) >> 26; h[7] += c; h[6] -= c << 26;
  c = (h[3] + ((int64_t)1 << 24)) >> 25; h[4] += c; h[3] -= c << 25;
  c = (h[7] + ((int64_t)1 << 24)) >> 25; h[8] += c; h[7] -= c << 25;
  c = (h[4] + ((int64_t)1 << 25)) >> 26; h[5] += c; h[4] -= c << 26;
  c = (h[8] + ((int64_t)1 << 25)) >> 26; h[9] += c; h[8] -= c << 26;
  c = (h[9] + ((int64_t)1 << 24)) >> 25; h[0] += c * 19; h[9] -= c << 25;
  c = (h[0] + ((int64_t)1 << 25)) >> 26; h[1] += c; h[0] -= c << 26;
  F25 r;
#pragma unroll
  for (int i = 0; i < 10; i++) r.v[i] = (int32_t)h[i];
  return r;
}

// h = f*g (dbl = false) or 2*f*g (dbl = true); inputs bounded by 1.65*2^26, output by 1.01*2^25/2^26
// A real function on the device (see HDNI in field.cuh): one warp per SM runs this kernel, so every
// instruction-cache miss is fully exposed; inlined, one window step is ~7k instructions (118 KB).
template <bool DBL>
HD F25 f25_mul_inl(const F25& f, const F25& g) {
  int32_t g19[10], f2[10];
#pragma unroll
  for (int i = 0; i < 10; i++) {
    g19[i] = 19 * g.v[i];
    f2[i] = 2 * f.v[i];
  }
  int64_t h[10];
#pragma unroll
  for (int k = 0; k < 10; k++) h[k] = 0;
#pragma unroll
  for (int i = 0; i < 10; i++) {
#pragma unroll
    for (int j = 0; j < 10; j++) {
      const int32_t fi = ((i & 1) && (j & 1)) ? f2[i] : f.v[i];
      const bool wrap = (i + j) >= 10;
      const int32_t gj = wrap ? g19[j] : g.v[j];
      h[wrap ? i + j - 10 : i + j] += (int64_t)fi * gj;
    }
  }
  if (DBL) {
#pragma unroll
    for (int k = 0; k < 10; k++) h[k] += h[k];
  }
  return f25_carry(h);
}
// out-of-line instance for code with many call sites (point formulas, inversion chains)
template <bool DBL>
HDNI F25 f25_mul_t(const F25& f, const F25& g) {
  return f25_mul_inl<DBL>(f, g);
}
HD F25 f25_mul(const F25& f, const F25& g) { return f25_mul_t<false>(f, g); }
HD F25 f25_sq(const F25& f) { return f25_mul_t<false>(f, f); }
HD F25 f25_sq2(const F25& f) { return f25_mul_t<true>(f, f); }

HD int64_t ld3(const uint8_t* p) { return (int64_t)p[0] | ((int64_t)p[1] << 8) | ((int64_t)p[2] << 16); }
HD int64_t ld4(const uint8_t* p) { return ld3(p) | ((int64_t)p[3] << 24); }

HD F25 f25_frombytes(const uint8_t s[32]) {
  int64_t h[10];
  h[0] = ld4(s);
  h[1] = ld3(s + 4) << 6;
  h[2] = ld3(s + 7) << 5;
  h[3] = ld3(s + 10) << 3;
  h[4] = ld3(s + 13) << 2;
  h[5] = ld4(s + 16);
  h[6] = ld3(s + 20) << 7;
  h[7] = ld3(s + 23) << 5;
  h[8] = ld3(s + 26) << 4;
  h[9] = (ld3(s + 29) & 8388607) << 2;
  int64_t c;
  c = (h[9] + ((int64_t)1 << 24)) >> 25; h[0] += c * 19; h[9] -= c << 25;
  c = (h[1] + ((int64_t)1 << 24)) >> 25; h[2] += c; h[1] -= c << 25;
  c = (h[3] + ((int64_t)1 << 24)) >> 25; h[4] += c; h[3] -= c << 25;
  c = (h[5] + ((int64_t)1 << 24)) >> 25; h[6] += c; h[5] -= c << 25;
  c = (h[7] + ((int64_t)1 << 24)) >> 25; h[8] += c; h[7] -= c << 25;
  c = (h[0] + ((int64_t)1 << 25)) >> 26; h[1] += c; h[0] -= c << 26;
  c = (h[2] + ((int64_t)1 << 25)) >> 26; h[3] += c; h[2] -= c << 26;
  c = (h[4] + ((int64_t)1 << 25)) >> 26; h[5] += c; h[4] -= c << 26;
  c = (h[6] + ((int64_t)1 << 25)) >> 26; h[7] += c; h[6] -= c << 26;
  c = (h[8] + ((int64_t)1 << 25)) >> 26; h[9] += c; h[8] -= c << 26;
  F25 r;
#pragma unroll
  for (int i = 0; i < 10; i++) r.v[i] = (int32_t)h[i];
  return r;
}

// canonical little-endian encoding (full reduction mod 2^255-19)
HD void f25_tobytes(uint8_t s[32], const F25& f) {
  int32_t h[10];
#pragma unroll
  for (int i = 0; i < 10; i++) h[i] = f.v[i];
  int32_t q = (19 * h[9] + ((int32_t)1 << 24)) >> 25;
  q = (h[0] + q) >> 26; q = (h[1] + q) >> 25; q = (h[2] + q) >> 26; q = (h[3] + q) >> 25; q = (h[4] + q) >> 26;
  q = (h[5] + q) >> 25; q = (h[6] + q) >> 26; q = (h[7] + q) >> 25; q = (h[8] + q) >> 26; q = (h[9] + q) >> 25;
  h[0] += 19 * q;
  int32_t c;
  c = h[0] >> 26; h[1] += c; h[0] -= c << 26;
  c = h[1] >> 25; h[2] += c; h[1] -= c << 25;
  c = h[2] >> 26; h[3] += c; h[2] -= c << 26;
  c = h[3] >> 25; h[4] += c; h[3] -= c << 25;
  c = h[4] >> 26; h[5] += c; h[4] -= c << 26;
  c = h[5] >> 25; h[6] += c; h[5] -= c << 25;
  c = h[6] >> 26; h[7] += c; h[6] -= c << 26;
  c = h[7] >> 25; h[8] += c; h[7] -= c << 25;
  c = h[8] >> 26; h[9] += c; h[8] -= c << 26;
  c = h[9] >> 25; h[9] -= c << 25;
  s[0] = (uint8_t)(h[0] >> 0); s[1] = (uint8_t)(h[0] >> 8); s[2] = (uint8_t)(h[0] >> 16);
  s[3] = (uint8_t)((h[0] >> 24) | (h[1] << 2));
  s[4] = (uint8_t)(h[1] >> 6); s[5] = (uint8_t)(h[1] >> 14);
  s[6] = (uint8_t)((h[1] >> 22) | (h[2] << 3));
  s[7] = (uint8_t)(h[2] >> 5); s[8] = (uint8_t)(h[2] >> 13);
  s[9] = (uint8_t)((h[2] >> 21) | (h[3] << 5));
  s[10] = (uint8_t)(h[3] >> 3); s[11] = (uint8_t)(h[3] >> 11);
  s[12] = (uint8_t)((h[3] >> 19) | (h[4] << 6));
  s[13] = (uint8_t)(h[4] >> 2); s[14] = (uint8_t)(h[4] >> 10); s[15] = (uint8_t)(h[4] >> 18);
  s[16] = (uint8_t)(h[5] >> 0); s[17] = (uint8_t)(h[5] >> 8); s[18] = (uint8_t)(h[5] >> 16);
  s[19] = (uint8_t)((h[5] >> 24) | (h[6] << 1));
  s[20] = (uint8_t)(h[6] >> 7); s[21] = (uint8_t)(h[6] >> 15);
  s[22] = (uint8_t)((h[6] >> 23) | (h[7] << 3));
  s[23] = (uint8_t)(h[7] >> 5); s[24] = (uint8_t)(h[7] >> 13);
  s[25] = (uint8_t)((h[7] >> 21) | (h[8] << 4));
  s[26] = (uint8_t)(h[8] >> 4); s[27] = (uint8_t)(h[8] >> 12);
  s[28] = (uint8_t)((h[8] >> 20) | (h[9] << 6));
  s[29] = (uint8_t)(h[9] >> 2); s[30] = (uint8_t)(h[9] >> 10); s[31] = (uint8_t)(h[9] >> 18);
}

HDN inline F25 f25_sqn(F25 t, int n) {
  for (int i = 0; i < n; i++) t = f25_sq(t);
  return t;
}
// z^(p-2) with the standard 254-squaring / 11-multiplication chain
HDN inline F25 f25_invert(const F25& z) {
  F25 z2 = f25_sq(z);
  F25 t = f25_sqn(z2, 2);
  F25 z9 = f25_mul(t, z);
  F25 z11 = f25_mul(z9, z2);
  t = f25_sq(z11);
  F25 z2_5_0 = f25_mul(t, z9);
  t = f25_sqn(z2_5_0, 5);
  F25 z2_10_0 = f25_mul(t, z2_5_0);
  t = f25_sqn(z2_10_0, 10);
  F25 z2_20_0 = f25_mul(t, z2_10_0);
  t = f25_sqn(z2_20_0, 20);
  t = f25_mul(t, z2_20_0);
  t = f25_sqn(t, 10);
  F25 z2_50_0 = f25_mul(t, z2_10_0);
  t = f25_sqn(z2_50_0, 50);
  F25 z2_100_0 = f25_mul(t, z2_50_0);
  t = f25_sqn(z2_100_0, 100);
  t = f25_mul(t, z2_100_0);
  t = f25_sqn(t, 50);
  t = f25_mul(t, z2_50_0);
  t = f25_sqn(t, 5);
  return f25_mul(t, z11);
}

// Same value as f25_invert through the binary extended Euclid of field.cuh (canonical bytes -> 8x32 limbs ->
// inverse mod 2^255-19 -> limbs): ~15x shorter than the 254-squaring chain when one thread must invert alone, as
// in vrf_noisers_kernel on the block's critical path.  Variable time; only public points' Z are inverted.
HDN inline F25 f25_invert_fast(const F25& z) {
  uint8_t b[32];
  f25_tobytes(b, z);
  const U256 inv = u256_modinv_odd(u256_from_le(b), mod_of<EdP>());
  u256_to_le(b, inv);
  return f25_frombytes(b);
}

// z^((p-5)/8) = z^(2^252 - 3)
HDN inline F25 f25_pow22523_r25(const F25& z) {
  F25 t0 = f25_sq(z);
  F25 t1 = f25_sqn(t0, 2);
  t1 = f25_mul(z, t1);
  t0 = f25_mul(t0, t1);
  t0 = f25_sq(t0);
  t0 = f25_mul(t1, t0);
  t1 = f25_sqn(t0, 5);
  t0 = f25_mul(t1, t0);
  t1 = f25_sqn(t0, 10);
  t1 = f25_mul(t1, t0);
  F25 t2 = f25_sqn(t1, 20);
  t1 = f25_mul(t2, t1);
  t1 = f25_sqn(t1, 10);
  t0 = f25_mul(t1, t0);
  t1 = f25_sqn(t0, 50);
  t1 = f25_mul(t1, t0);
  t2 = f25_sqn(t1, 100);
  t1 = f25_mul(t2, t1);
  t1 = f25_sqn(t1, 50);
  t0 = f25_mul(t1, t0);
  t0 = f25_sqn(t0, 2);
  return f25_mul(t0, z);
}
#if !defined(__CUDA_ARCH__)
// Host only: the same power in radix 2^51 with 128-bit products (25 instead of 100 limb products per
// multiplication).  hashToCurve runs on the host between a block's hash and the launch of the VRF kernel - i.e. on
// the block's critical path - and its two z^(2^252-3) powers (Legendre symbol, square root) were ~100 us of it.
struct F51 {
  uint64_t v[5];
};
inline F51 f51_mul(const F51& a, const F51& b) {
  typedef unsigned __int128 u128;
  const uint64_t M = (1ULL << 51) - 1;
  const uint64_t b1_19 = 19 * b.v[1], b2_19 = 19 * b.v[2], b3_19 = 19 * b.v[3], b4_19 = 19 * b.v[4];
  u128 r0 = (u128)a.v[0] * b.v[0] + (u128)a.v[1] * b4_19 + (u128)a.v[2] * b3_19 + (u128)a.v[3] * b2_19 + (u128)a.v[4] * b1_19;
  u128 r1 = (u128)a.v[0] * b.v[1] + (u128)a.v[1] * b.v[0] + (u128)a.v[2] * b4_19 + (u128)a.v[3] * b3_19 + (u128)a.v[4] * b2_19;
  u128 r2 = (u128)a.v[0] * b.v[2] + (u128)a.v[1] * b.v[1] + (u128)a.v[2] * b.v[0] + (u128)a.v[3] * b4_19 + (u128)a.v[4] * b3_19;
  u128 r3 = (u128)a.v[0] * b.v[3] + (u128)a.v[1] * b.v[2] + (u128)a.v[2] * b.v[1] + (u128)a.v[3] * b.v[0] + (u128)a.v[4] * b4_19;
  u128 r4 = (u128)a.v[0] * b.v[4] + (u128)a.v[1] * b.v[3] + (u128)a.v[2] * b.v[2] + (u128)a.v[3] * b.v[1] + (u128)a.v[4] * b.v[0];
  F51 o;
  uint64_t c;
  c = (uint64_t)(r0 >> 51); o.v[0] = (uint64_t)r0 & M; r1 += c;
  c = (uint64_t)(r1 >> 51); o.v[1] = (uint64_t)r1 & M; r2 += c;
  c = (uint64_t)(r2 >> 51); o.v[2] = (uint64_t)r2 & M; r3 += c;
  c = (uint64_t)(r3 >> 51); o.v[3] = (uint64_t)r3 & M; r4 += c;
  c = (uint64_t)(r4 >> 51); o.v[4] = (uint64_t)r4 & M;
  o.v[0] += 19 * c;
  c = o.v[0] >> 51; o.v[0] &= M; o.v[1] += c;
  return o;
}
inline F51 f51_sqn(F51 t, int n) {
  for (int i = 0; i < n; i++) t = f51_mul(t, t);
  return t;
}
inline F25 f25_pow22523_host51(const F25& zin) {
  const uint64_t M = (1ULL << 51) - 1;
  uint8_t b[32];
  f25_tobytes(b, zin);
  U256 u = u256_from_le(b);
  auto bits = [&u](int lo) -> uint64_t {      // 51 bits of u starting at bit `lo`
    uint64_t r = 0;
    for (int k = 0; k < 51; k++) {
      const int bit = lo + k;
      if (bit < 256) r |= (uint64_t)((u.v[bit >> 5] >> (bit & 31)) & 1u) << k;
    }
    return r;
  };
  F51 z;
  for (int i = 0; i < 5; i++) z.v[i] = bits(51 * i);
  F51 t0 = f51_mul(z, z);
  F51 t1 = f51_sqn(t0, 2);
  t1 = f51_mul(z, t1);
  t0 = f51_mul(t0, t1);
  t0 = f51_mul(t0, t0);
  t0 = f51_mul(t1, t0);
  t1 = f51_sqn(t0, 5);
  t0 = f51_mul(t1, t0);
  t1 = f51_sqn(t0, 10);
  t1 = f51_mul(t1, t0);
  F51 t2 = f51_sqn(t1, 20);
  t1 = f51_mul(t2, t1);
  t1 = f51_sqn(t1, 10);
  t0 = f51_mul(t1, t0);
  t1 = f51_sqn(t0, 50);
  t1 = f51_mul(t1, t0);
  t2 = f51_sqn(t1, 100);
  t1 = f51_mul(t2, t1);
  t1 = f51_sqn(t1, 50);
  t0 = f51_mul(t1, t0);
  t0 = f51_sqn(t0, 2);
  t0 = f51_mul(t0, z);
  // carry once more, pack into 256 bits (value < 2^255 + 2^52), reduce mod p
  uint64_t c;
  for (int pass = 0; pass < 2; pass++) {
    c = t0.v[0] >> 51; t0.v[0] &= M; t0.v[1] += c;
    c = t0.v[1] >> 51; t0.v[1] &= M; t0.v[2] += c;
    c = t0.v[2] >> 51; t0.v[2] &= M; t0.v[3] += c;
    c = t0.v[3] >> 51; t0.v[3] &= M; t0.v[4] += c;
    c = t0.v[4] >> 51; t0.v[4] &= M; t0.v[0] += 19 * c;
  }
  U256 r;
  for (int i = 0; i < 8; i++) r.v[i] = 0;
  for (int i = 0; i < 5; i++)
    for (int k = 0; k < 51; k++) {
      const int bit = 51 * i + k;
      if (bit < 256 && ((t0.v[i] >> k) & 1ULL)) r.v[bit >> 5] |= 1u << (bit & 31);
    }
  const U256 p = mod_of<EdP>();
  while (u256_geq(r, p)) u256_sub(r, r, p);
  u256_to_le(b, r);
  return f25_frombytes(b);
}
#endif
HDN inline F25 f25_pow22523(const F25& z) {
#if !defined(__CUDA_ARCH__)
  return f25_pow22523_host51(z);
#else
  return f25_pow22523_r25(z);
#endif
}
HD bool f25_is_zero(const F25& a) {
  uint8_t s[32];
  f25_tobytes(s, a);
  uint8_t o = 0;
  for (int i = 0; i < 32; i++) o |= s[i];
  return o == 0;
}
HD bool f25_eq(const F25& a, const F25& b) { return f25_is_zero(f25_sub(a, b)); }
HD F25 f25_from_u32(uint32_t x) {
  F25 r = f25_zero();
  r.v[0] = (int32_t)(x & 0x3ffffff);
  r.v[1] = (int32_t)(x >> 26);
  return r;
}

struct EdP3 { F25 X, Y, Z, T; };        // extended
struct EdP1 { F25 X, Y, Z, T; };        // completed ((X:Z), (Y:T))
struct EdCached { F25 YpX, YmX, Z, T2d; };

HD F25 f25_const_2d() {
  // 2d mod p as little-endian bytes of the canonical value
  const uint8_t b[32] = {0x59, 0xf1, 0xb2, 0x26, 0x94, 0x9b, 0xd6, 0xeb, 0x56, 0xb1, 0x83, 0x82, 0x9a, 0x14, 0xe0, 0x00,
                         0x30, 0xd1, 0xf3, 0xee, 0xf2, 0x80, 0x8e, 0x19, 0xe7, 0xfc, 0xdf, 0x56, 0xdc, 0xd9, 0x06, 0x24};
  return f25_frombytes(b);
}

HD EdP3 edr_identity() {
  EdP3 r;
  r.X = f25_zero(); r.Y = f25_one(); r.Z = f25_one(); r.T = f25_zero();
  return r;
}
HD EdP3 edr_p1_to_p3(const EdP1& p) {
  EdP3 r;
  r.X = f25_mul(p.X, p.T); r.Y = f25_mul(p.Y, p.Z); r.Z = f25_mul(p.Z, p.T); r.T = f25_mul(p.X, p.Y);
  return r;
}
// doubling of the projective part (X:Y:Z) of p
HD EdP1 edr_dbl(const F25& X, const F25& Y, const F25& Z) {
  EdP1 r;
  F25 XX = f25_sq(X), YY = f25_sq(Y), B = f25_sq2(Z);
  F25 A = f25_sq(f25_add(X, Y));
  r.Y = f25_add(YY, XX);
  r.Z = f25_sub(YY, XX);
  r.X = f25_sub(A, r.Y);
  r.T = f25_sub(B, r.Z);
  return r;
}
HD EdCached edr_to_cached(const EdP3& p, const F25& d2) {
  EdCached c;
  c.YpX = f25_add(p.Y, p.X); c.YmX = f25_sub(p.Y, p.X); c.Z = p.Z; c.T2d = f25_mul(p.T, d2);
  return c;
}
// p + q (neg = false) or p - q (neg = true)
HD EdP1 edr_add(const EdP3& p, const EdCached& q, bool neg) {
  EdP1 r;
  F25 a = f25_add(p.Y, p.X), b = f25_sub(p.Y, p.X);
  F25 A = f25_mul(a, neg ? q.YmX : q.YpX);
  F25 Bm = f25_mul(b, neg ? q.YpX : q.YmX);
  F25 C = f25_mul(q.T2d, p.T);
  F25 ZZ = f25_mul(p.Z, q.Z);
  F25 D = f25_add(ZZ, ZZ);
  r.X = f25_sub(A, Bm);
  r.Y = f25_add(A, Bm);
  if (neg) {
    r.Z = f25_sub(D, C);
    r.T = f25_add(D, C);
  } else {
    r.Z = f25_add(D, C);
    r.T = f25_sub(D, C);
  }
  return r;
}
HDN inline void edr_compress(uint8_t out[32], const EdP3& p) {
  F25 zi = f25_invert(p.Z);
  F25 x = f25_mul(p.X, zi), y = f25_mul(p.Y, zi);
  uint8_t xb[32];
  f25_tobytes(out, y);
  f25_tobytes(xb, x);
  out[31] ^= (uint8_t)((xb[0] & 1) << 7);
}
// affine (x, y) as 64 canonical little-endian bytes -> extended
HD EdP3 edr_from_affine_bytes(const uint8_t xy[64]) {
  EdP3 r;
  r.X = f25_frombytes(xy);
  r.Y = f25_frombytes(xy + 32);
  r.Z = f25_one();
  r.T = f25_mul(r.X, r.Y);
  return r;
}

HD F25 f25_const_d() {
  const uint8_t b[32] = {0xa3, 0x78, 0x59, 0x13, 0xca, 0x4d, 0xeb, 0x75, 0xab, 0xd8, 0x41, 0x41, 0x4d, 0x0a, 0x70, 0x00,
                         0x98, 0xe8, 0x79, 0x77, 0x79, 0x40, 0xc7, 0x8c, 0x73, 0xfe, 0x6f, 0x2b, 0xee, 0x6c, 0x03, 0x52};
  return f25_frombytes(b);
}
HD F25 f25_const_sqrtm1() {
  const uint8_t b[32] = {0xb0, 0xa0, 0x0e, 0x4a, 0x27, 0x1b, 0xee, 0xc4, 0x78, 0xe4, 0x2f, 0xad, 0x06, 0x18, 0x43, 0x2f,
                         0xa7, 0xd7, 0xfb, 0x3d, 0x99, 0x00, 0x4d, 0x2b, 0x0b, 0xdf, 0xc1, 0x4f, 0x80, 0x24, 0x83, 0x2b};
  return f25_frombytes(b);
}
// x from y and a sign bit (same contract as ed_from_parity_and_y in ed25519.cuh)
HDN inline bool edr_from_parity_and_y(EdP3& r, uint32_t sign, const F25& y) {
  const F25 one = f25_one();
  F25 y2 = f25_sq(y);
  F25 u = f25_sub(y2, one);
  F25 v = f25_add(f25_mul(y2, f25_const_d()), one);
  F25 v3 = f25_mul(f25_sq(v), v);
  F25 v7 = f25_mul(f25_sq(v3), v);
  F25 x = f25_mul(f25_mul(u, v3), f25_pow22523(f25_mul(u, v7)));
  F25 vx2 = f25_mul(v, f25_sq(x));
  if (!f25_eq(vx2, u)) {
    if (!f25_is_zero(f25_add(vx2, u))) return false;
    x = f25_mul(x, f25_const_sqrtm1());
  }
  uint8_t xb[32];
  f25_tobytes(xb, x);
  if ((uint32_t)(xb[0] & 1) != sign) x = f25_mul(f25_neg(x), one);   // negate, then re-carry the limbs
  r.X = x;
  r.Y = y;
  r.Z = one;
  r.T = f25_mul(x, y);
  return true;
}
// Elligator2 + cofactor clearing in the fast field (same map as ed_elligator2_x8 in ed25519.cuh)
HDN inline EdP3 edr_elligator2_x8(const uint8_t h[32]) {
  uint8_t t[32];
  for (int i = 0; i < 32; i++) t[i] = h[i];
  const uint32_t bit = t[31] >> 7;
  t[31] &= 127;
  const F25 one = f25_one();
  const F25 A = f25_from_u32(486662);
  F25 r = f25_frombytes(t);
  F25 den = f25_add(f25_sq2(r), one);
  F25 v = f25_mul(f25_neg(A), f25_invert_fast(den));
  F25 v2 = f25_sq(v);
  F25 e = f25_add(f25_add(f25_mul(v2, v), v), f25_mul(v2, A));
  e = f25_mul(e, one);   // re-carry the three-term sum before it is used as a multiplication input
  // Legendre symbol e^((p-1)/2) = (e^((p-5)/8))^4 * e^2
  F25 chi = f25_mul(f25_sqn(f25_pow22523(e), 2), f25_sq(e));
  const bool minus1 = !f25_eq(chi, one) && !f25_is_zero(chi);
  if (minus1) v = f25_mul(f25_sub(f25_neg(v), A), one);
  F25 y = f25_mul(f25_sub(v, one), f25_invert_fast(f25_mul(f25_add(v, one), one)));
  EdP3 p;
  if (!edr_from_parity_and_y(p, bit, y)) p = edr_identity();
  EdP1 c = edr_dbl(p.X, p.Y, p.Z);
  for (int k = 0; k < 2; k++) {
    F25 X = f25_mul(c.X, c.T), Y = f25_mul(c.Y, c.Z), Z = f25_mul(c.Z, c.T);
    c = edr_dbl(X, Y, Z);
  }
  return edr_p1_to_p3(c);
}
// affine x || y (canonical little-endian) of an extended point
HDN inline void edr_to_affine_bytes(uint8_t out[64], const EdP3& p) {
  F25 zi = f25_invert_fast(p.Z);
  f25_tobytes(out, f25_mul(p.X, zi));
  f25_tobytes(out + 32, f25_mul(p.Y, zi));
}

// table[k] = (k+1) * H in cached form, k = 0..7
HDN inline void edr_build_table(EdCached table[8], const EdP3& H, const F25& d2) {
  table[0] = edr_to_cached(H, d2);
  EdP3 acc = H;
  for (int k = 1; k < 8; k++) {
    acc = edr_p1_to_p3(edr_add(acc, table[0], false));
    table[k] = edr_to_cached(acc, d2);
  }
}

// scalar (32 little-endian bytes, < 2^255) times the table's base point
HDN inline EdP3 edr_scalarmult(const uint8_t a[32], const EdCached* table) {
  signed char e[64];
  for (int i = 0; i < 32; i++) {
    e[2 * i] = (signed char)(a[i] & 15);
    e[2 * i + 1] = (signed char)((a[i] >> 4) & 15);
  }
  int carry = 0;
  for (int i = 0; i < 63; i++) {
    e[i] = (signed char)(e[i] + carry);
    carry = (e[i] + 8) >> 4;
    e[i] = (signed char)(e[i] - (carry << 4));
  }
  e[63] = (signed char)(e[63] + carry);
  EdP3 r = edr_identity();
  for (int i = 63; i >= 0; i--) {
    // r = 16 r : three projective doublings and a final one back to extended coordinates
    EdP1 t = edr_dbl(r.X, r.Y, r.Z);
#pragma unroll 1
    for (int k = 0; k < 3; k++) {
      F25 X = f25_mul(t.X, t.T), Y = f25_mul(t.Y, t.Z), Z = f25_mul(t.Z, t.T);
      t = edr_dbl(X, Y, Z);
    }
    r = edr_p1_to_p3(t);
    const int dgt = e[i];
    if (dgt != 0) {
      const int m = dgt < 0 ? -dgt : dgt;
      r = edr_p1_to_p3(edr_add(r, table[m - 1], dgt < 0));
    }
  }
  return r;
}

}  // namespace bsc
