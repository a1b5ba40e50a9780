// 256-bit Montgomery prime-field arithmetic shared by host (oracle / keygen / ledger) and device
// (MSM, Schnorr, VRF kernels).  Eight 32-bit limbs, little-endian, R = 2^256.
//
// Replaces the reference's amd64 assembly field (lib/dedis/kyber/pairing/bn256/gfp.s:5-95,
// mul.h, mul_bmi2.h: 4x64-bit limbs with MULX) with a limb width that maps to the sm_100a
// integer pipe (IMAD.WIDE / IADD3 with carry).  The same source compiles for the CPU so every
// device result has a bit-exact host oracle.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define HD __host__ __device__ __forceinline__
#define HDN __host__ __device__
// fe_mul is deliberately a real function on the device: fully inlined, a mixed point addition is ~5k
// SASS instructions (80 KB) and the MSM kernels stall on instruction fetch (ncu: stall_no_instruction is
// the top reason); as a call the hot loop fits the instruction cache.
#define HDNI __host__ __device__ __noinline__
#else
#define HD inline
#define HDN
#define HDNI
#endif

namespace bsc {

struct U256 {
  uint32_t v[8];
};

HD bool u256_is_zero(const U256& a) {
  uint32_t o = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) o |= a.v[i];
  return o == 0;
}
HD bool u256_eq(const U256& a, const U256& b) {
  uint32_t o = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) o |= a.v[i] ^ b.v[i];
  return o == 0;
}
// r = a + b, returns carry
HD uint32_t u256_add(U256& r, const U256& a, const U256& b) {
  uint64_t c = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    c += (uint64_t)a.v[i] + b.v[i];
    r.v[i] = (uint32_t)c;
    c >>= 32;
  }
  return (uint32_t)c;
}
// r = a - b, returns borrow (1 if a < b)
HD uint32_t u256_sub(U256& r, const U256& a, const U256& b) {
  int64_t c = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    c += (int64_t)a.v[i] - (int64_t)b.v[i];
    r.v[i] = (uint32_t)c;
    c >>= 32;  // arithmetic shift keeps the borrow as -1
  }
  return (uint32_t)(c & 1);
}

// ---- modulus traits -------------------------------------------------------------------------
// BN256 base field p = 36u^4+36u^3+24u^2+6u+1, u = 1868033^3
// (lib/dedis/kyber/pairing/bn256/constants.go:12-16, 43-56).
struct BnP {
  HD static constexpr uint32_t m(int i) {
    constexpr uint32_t t[8] = {0x5e089667u, 0x185cac6cu, 0x20b5b59eu, 0xee5b88d1u,
                               0x6184dc21u, 0xaa6fecb8u, 0x4aa387f9u, 0x8fb501e3u};
    return t[i];
  }
  HD static constexpr uint32_t r2(int i) {
    constexpr uint32_t t[8] = {0x7e444f56u, 0x9c21c3ffu, 0xb2efb0c2u, 0x409ed151u,
                               0x80fb1651u, 0x0c6dc37bu, 0x2c2380b7u, 0x7c36e0e6u};
    return t[i];
  }
  HD static constexpr uint32_t r1(int i) {
    constexpr uint32_t t[8] = {0xa1f76999u, 0xe7a35393u, 0xdf4a4a61u, 0x11a4772eu,
                               0x9e7b23deu, 0x55901347u, 0xb55c7806u, 0x704afe1cu};
    return t[i];
  }
  static constexpr uint32_t n0 = 0x7f17daa9u;
};
// BN256 group order (constants.go:18-20) — scalar field of G1/G2.
struct BnN {
  HD static constexpr uint32_t m(int i) {
    constexpr uint32_t t[8] = {0x57ac7261u, 0x1a2ef45bu, 0xf82b3924u, 0x2e8d8e12u,
                               0x6184dc21u, 0xaa6fecb8u, 0x4aa387f9u, 0x8fb501e3u};
    return t[i];
  }
  HD static constexpr uint32_t r2(int i) {
    constexpr uint32_t t[8] = {0x2affbc35u, 0xb5f03013u, 0x0792e95du, 0x85a1f7dau,
                               0xa6ee4895u, 0x26841e5fu, 0x765aefd5u, 0x3d8f6c73u};
    return t[i];
  }
  HD static constexpr uint32_t r1(int i) {
    constexpr uint32_t t[8] = {0xa8538d9fu, 0xe5d10ba4u, 0x07d4c6dbu, 0xd17271edu,
                               0x9e7b23deu, 0x55901347u, 0xb55c7806u, 0x704afe1cu};
    return t[i];
  }
  static constexpr uint32_t n0 = 0x2d284e5fu;
};
// Curve25519 base field 2^255-19 (VRF, coniks ECVRF; DistSys/vrf.go:16-32 uses it via coniks-go).
struct EdP {
  HD static constexpr uint32_t m(int i) {
    constexpr uint32_t t[8] = {0xffffffedu, 0xffffffffu, 0xffffffffu, 0xffffffffu,
                               0xffffffffu, 0xffffffffu, 0xffffffffu, 0x7fffffffu};
    return t[i];
  }
  HD static constexpr uint32_t r2(int i) {
    constexpr uint32_t t[8] = {0x000005a4u, 0, 0, 0, 0, 0, 0, 0};
    return t[i];
  }
  HD static constexpr uint32_t r1(int i) {
    constexpr uint32_t t[8] = {0x00000026u, 0, 0, 0, 0, 0, 0, 0};
    return t[i];
  }
  static constexpr uint32_t n0 = 0x286bca1bu;
};
// Ed25519 group order l = 2^252 + 27742317777372353535851937790883648493.
struct EdL {
  HD static constexpr uint32_t m(int i) {
    constexpr uint32_t t[8] = {0x5cf5d3edu, 0x5812631au, 0xa2f79cd6u, 0x14def9deu,
                               0x00000000u, 0x00000000u, 0x00000000u, 0x10000000u};
    return t[i];
  }
  HD static constexpr uint32_t r2(int i) {
    constexpr uint32_t t[8] = {0x449c0f01u, 0xa40611e3u, 0x68859347u, 0xd00e1ba7u,
                               0x17f5be65u, 0xceec73d2u, 0x7c309a3du, 0x0399411bu};
    return t[i];
  }
  HD static constexpr uint32_t r1(int i) {
    constexpr uint32_t t[8] = {0x8d98951du, 0xd6ec3174u, 0x737dcf70u, 0xc6ef5bf4u,
                               0xfffffffeu, 0xffffffffu, 0xffffffffu, 0x0fffffffu};
    return t[i];
  }
  static constexpr uint32_t n0 = 0x12547e1bu;
};

template <class M>
HD U256 mod_of() {
  U256 r;
#pragma unroll
  for (int i = 0; i < 8; i++) r.v[i] = M::m(i);
  return r;
}

// Field element in Montgomery form (value * R mod m), always fully reduced (< m).
template <class M>
struct Fe {
  U256 x;
};

template <class M>
HD Fe<M> fe_zero() {
  Fe<M> r;
#pragma unroll
  for (int i = 0; i < 8; i++) r.x.v[i] = 0;
  return r;
}
template <class M>
HD Fe<M> fe_one() {
  Fe<M> r;
#pragma unroll
  for (int i = 0; i < 8; i++) r.x.v[i] = M::r1(i);
  return r;
}
template <class M>
HD bool fe_is_zero(const Fe<M>& a) { return u256_is_zero(a.x); }
template <class M>
HD bool fe_eq(const Fe<M>& a, const Fe<M>& b) { return u256_eq(a.x, b.x); }

template <class M>
HD Fe<M> fe_add(const Fe<M>& a, const Fe<M>& b) {
  Fe<M> r;
  uint32_t carry = u256_add(r.x, a.x, b.x);
  U256 t;
  const U256 m = mod_of<M>();
  uint32_t borrow = u256_sub(t, r.x, m);
  // take t when the 257-bit sum >= m
  if (carry || !borrow) r.x = t;
  return r;
}
template <class M>
HD Fe<M> fe_sub(const Fe<M>& a, const Fe<M>& b) {
  Fe<M> r;
  uint32_t borrow = u256_sub(r.x, a.x, b.x);
  if (borrow) {
    const U256 m = mod_of<M>();
    u256_add(r.x, r.x, m);
  }
  return r;
}
template <class M>
HD Fe<M> fe_neg(const Fe<M>& a) {
  Fe<M> r;
  if (u256_is_zero(a.x)) return a;
  const U256 m = mod_of<M>();
  u256_sub(r.x, m, a.x);
  return r;
}
template <class M>
HD Fe<M> fe_dbl(const Fe<M>& a) { return fe_add(a, a); }

#if defined(__CUDA_ARCH__)
// Device multiplier.  The running sum is kept as two 8-limb arrays, `even` (bit offset 0) and `odd` (bit
// offset 32): a product a[j]*b_i with j even lands on a 64-bit-aligned limb pair of `even`, with j odd on a
// pair of `odd`, so every 32x32 product is one mad.lo.cc/madc.hi.cc pair on an aligned register pair, which
// ptxas turns into a single IMAD.WIDE.U32(.X) with carry in/out: ~250 SASS instructions per product instead
// of ~510 for the plain CIOS loop below (+21% measured fe_mul throughput; the rest is the IMAD.WIDE pipe).
// After each limb of b the value is divided by 2^32 by swapping the roles of the two arrays.  The moduli here
// use all 256 bits, so the sum can exceed the 288 bits the two arrays hold; `ovf` carries that bit (weight
// 2^288, i.e. 2^256 after the shift) into the next step and into the final conditional subtraction.
__device__ __forceinline__ void eo_mul_n(uint32_t* acc, const uint32_t* a, uint32_t bi) {
#pragma unroll
  for (int j = 0; j < 8; j += 2)
    asm("mul.lo.u32 %0, %2, %3; mul.hi.u32 %1, %2, %3;" : "=r"(acc[j]), "=r"(acc[j + 1]) : "r"(a[j]), "r"(bi));
}
__device__ __forceinline__ void eo_cmad_n(uint32_t* acc, const uint32_t* a, uint32_t bi) {
  asm volatile("mad.lo.cc.u32 %0, %2, %3, %0; madc.hi.cc.u32 %1, %2, %3, %1;"
               : "+r"(acc[0]), "+r"(acc[1]) : "r"(a[0]), "r"(bi));
#pragma unroll
  for (int j = 2; j < 8; j += 2)
    asm volatile("madc.lo.cc.u32 %0, %2, %3, %0; madc.hi.cc.u32 %1, %2, %3, %1;"
                 : "+r"(acc[j]), "+r"(acc[j + 1]) : "r"(a[j]), "r"(bi));
}
// odd = (odd >> 64) + a*bi + carry-in, top limb also takes the overflow bit of the previous step
__device__ __forceinline__ void eo_madc_n_rshift(uint32_t* odd, const uint32_t* a, uint32_t bi, uint32_t ovf) {
#pragma unroll
  for (int j = 0; j < 6; j += 2)
    asm volatile("madc.lo.cc.u32 %0, %2, %3, %4; madc.hi.cc.u32 %1, %2, %3, %5;"
                 : "=r"(odd[j]), "=r"(odd[j + 1]) : "r"(a[j]), "r"(bi), "r"(odd[j + 2]), "r"(odd[j + 3]));
  asm volatile("madc.lo.cc.u32 %0, %2, %3, 0; madc.hi.u32 %1, %2, %3, %4;"
               : "=r"(odd[6]), "=r"(odd[7]) : "r"(a[6]), "r"(bi), "r"(ovf));
}
template <class M>
__device__ __forceinline__ void eo_mad_redc(uint32_t* even, uint32_t* odd, const uint32_t* a, uint32_t bi,
                                            const uint32_t* mod, uint32_t& ovf, bool first) {
  if (first) {
    eo_mul_n(odd, a + 1, bi);
    eo_mul_n(even, a, bi);
  } else {
    asm volatile("add.cc.u32 %0, %0, %1;" : "+r"(even[0]) : "r"(odd[1]));
    eo_madc_n_rshift(odd, a + 1, bi, ovf);
    eo_cmad_n(even, a, bi);
    asm volatile("addc.u32 %0, %0, 0;" : "+r"(odd[7]));
  }
  const uint32_t mi = even[0] * M::n0;
  eo_cmad_n(odd, mod + 1, mi);
  asm volatile("addc.u32 %0, 0, 0;" : "=r"(ovf));
  eo_cmad_n(even, mod, mi);
  asm volatile("addc.cc.u32 %0, %0, 0; addc.u32 %1, %1, 0;" : "+r"(odd[7]), "+r"(ovf));
}
#endif

// Montgomery product a*b*R^-1 mod m (host: CIOS, 32-bit limbs, 64-bit partial products; device: the even/odd
// IMAD.WIDE form above — both return the canonical representative, so host and device stay bit-identical).
template <class M>
HDNI Fe<M> fe_mul(const Fe<M>& a, const Fe<M>& b) {
#if defined(__CUDA_ARCH__)
  uint32_t even[8], odd[8], mod[8];
  uint32_t ovf = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) mod[i] = M::m(i);
#pragma unroll
  for (int i = 0; i < 8; i += 2) {
    eo_mad_redc<M>(even, odd, a.x.v, b.x.v[i], mod, ovf, i == 0);
    eo_mad_redc<M>(odd, even, a.x.v, b.x.v[i + 1], mod, ovf, false);
  }
  uint32_t top;
  asm volatile("add.cc.u32 %0, %0, %1;" : "+r"(even[0]) : "r"(odd[1]));
#pragma unroll
  for (int i = 1; i < 7; i++) asm volatile("addc.cc.u32 %0, %0, %1;" : "+r"(even[i]) : "r"(odd[i + 1]));
  asm volatile("addc.cc.u32 %0, %0, 0; addc.u32 %1, %2, 0;" : "+r"(even[7]), "=r"(top) : "r"(ovf));
  Fe<M> r;
#pragma unroll
  for (int i = 0; i < 8; i++) r.x.v[i] = even[i];
  U256 d;
  const U256 m = mod_of<M>();
  const uint32_t borrow = u256_sub(d, r.x, m);
  if (top || !borrow) r.x = d;
  return r;
#else
  uint32_t t[10];
#pragma unroll
  for (int i = 0; i < 10; i++) t[i] = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    uint64_t c = 0;
    const uint32_t bi = b.x.v[i];
#pragma unroll
    for (int j = 0; j < 8; j++) {
      uint64_t s = (uint64_t)a.x.v[j] * bi + t[j] + c;
      t[j] = (uint32_t)s;
      c = s >> 32;
    }
    uint64_t s = (uint64_t)t[8] + c;
    t[8] = (uint32_t)s;
    t[9] = (uint32_t)(s >> 32);
    const uint32_t q = t[0] * M::n0;
    s = (uint64_t)q * M::m(0) + t[0];
    c = s >> 32;
#pragma unroll
    for (int j = 1; j < 8; j++) {
      s = (uint64_t)q * M::m(j) + t[j] + c;
      t[j - 1] = (uint32_t)s;
      c = s >> 32;
    }
    s = (uint64_t)t[8] + c;
    t[7] = (uint32_t)s;
    t[8] = t[9] + (uint32_t)(s >> 32);
  }
  Fe<M> r;
#pragma unroll
  for (int i = 0; i < 8; i++) r.x.v[i] = t[i];
  U256 d;
  const U256 m = mod_of<M>();
  uint32_t borrow = u256_sub(d, r.x, m);
  if (t[8] || !borrow) r.x = d;
  return r;
#endif
}
template <class M>
HD Fe<M> fe_sqr(const Fe<M>& a) { return fe_mul(a, a); }

// plain integer (< 2^256, any value) -> Montgomery form
template <class M>
HD Fe<M> fe_from_u256(const U256& a) {
  Fe<M> t, r2;
  t.x = a;
#pragma unroll
  for (int i = 0; i < 8; i++) r2.x.v[i] = M::r2(i);
  // a may be >= m; fe_mul only needs a*b < m*R which holds for a < R, b < m
  return fe_mul(t, r2);
}
template <class M>
HD U256 fe_to_u256(const Fe<M>& a) {
  Fe<M> one;
#pragma unroll
  for (int i = 0; i < 8; i++) one.x.v[i] = 0;
  one.x.v[0] = 1;
  return fe_mul(a, one).x;
}
template <class M>
HD Fe<M> fe_from_u64(uint64_t a) {
  U256 t;
#pragma unroll
  for (int i = 0; i < 8; i++) t.v[i] = 0;
  t.v[0] = (uint32_t)a;
  t.v[1] = (uint32_t)(a >> 32);
  return fe_from_u256<M>(t);
}
template <class M>
HD Fe<M> fe_from_i64(int64_t a) {
  if (a >= 0) return fe_from_u64<M>((uint64_t)a);
  return fe_neg(fe_from_u64<M>((uint64_t)0 - (uint64_t)a));
}

// a^e for a 256-bit plain exponent (MSB-first square-and-multiply).
template <class M>
HDN Fe<M> fe_pow(const Fe<M>& a, const U256& e) {
  Fe<M> r = fe_one<M>();
  bool started = false;
  for (int i = 255; i >= 0; i--) {
    if (started) r = fe_sqr(r);
    if ((e.v[i >> 5] >> (i & 31)) & 1) {
      r = started ? fe_mul(r, a) : a;
      started = true;
    }
  }
  return r;
}
// ---- modular inversion -----------------------------------------------------------------------------------
// The reference inverts by Fermat, a^(m-2) (gfp.go:30-48): ~380 dependent 256-bit multiplications, i.e. ~0.3 ms
// for ONE inversion on one GPU thread - and an inversion sits on every Jacobian -> wire-bytes conversion the
// block's tail waits for (commitments, the approvals' T = v*G) and on every VRF output (on the block's chain).
// Here: the Bernstein-Yang "divsteps" gcd in 30-bit batches.  A batch of 30 division steps looks only at the low
// 32 bits of (f, g) and yields a 2x2 transition matrix of 32-bit entries, which is then applied to the full
// 9 x 30-bit signed limbs of (f, g) and, modulo m, to the cofactors (d, e): ~20 batches of cheap 32-bit work
// (IMAD.WIDE + shifts, the GPU's native width) instead of 380 Montgomery products.  Variable time: nothing secret is
// inverted with it (projective Z coordinates of public points).  Same structure as the widely used signed-30
// modular-inverse formulation of that algorithm; written from the paper's recurrences, tested against Python's
// pow(a, -1, m) and against the Fermat form (tests/test_crypto_host.py).
HD bool u256_geq(const U256& a, const U256& b) {
#pragma unroll
  for (int i = 7; i >= 0; i--) {
    if (a.v[i] != b.v[i]) return a.v[i] > b.v[i];
  }
  return true;
}
struct S30 {
  int32_t v[9];   // value = sum v[i] * 2^(30 i), limbs signed, |v[i]| < 2^30 after carrying (top limb free)
};
HD S30 s30_from_u256(const U256& a) {
  S30 r;
  const uint32_t M30 = 0x3FFFFFFFu;
  uint64_t acc = 0;
  int bits = 0, k = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    acc |= (uint64_t)a.v[i] << bits;
    bits += 32;
    while (bits >= 30 && k < 9) {
      r.v[k++] = (int32_t)(acc & M30);
      acc >>= 30;
      bits -= 30;
    }
  }
  r.v[8] = (int32_t)acc;   // 256 - 240 = 16 bits
  return r;
}
HD U256 s30_to_u256(const S30& a) {   // a must be non-negative and < 2^256
  U256 r;
  uint64_t acc = 0;
  int bits = 0, k = 0;
#pragma unroll
  for (int i = 0; i < 9; i++) {
    acc |= (uint64_t)(uint32_t)a.v[i] << bits;
    bits += 30;
    if (bits >= 32 && k < 8) {
      r.v[k++] = (uint32_t)acc;
      acc >>= 32;
      bits -= 32;
    }
  }
  if (k < 8) r.v[k] = (uint32_t)acc;
  return r;
}
struct T2x2 {
  int32_t u, v, q, r;
};
HD int ctz32_nz(uint32_t x) {
#if defined(__CUDA_ARCH__)
  return __ffs((int)x) - 1;
#else
  return __builtin_ctz(x);
#endif
}
// 30 division steps on the low words.  eta = -delta.  On return 2^30 * [f', g'] = t * [f, g].
HD int32_t divsteps_30_var(int32_t eta, uint32_t f0, uint32_t g0, T2x2& t) {
  uint32_t u = 1, v = 0, q = 0, r = 1;
  uint32_t f = f0, g = g0;
  int i = 30;
  for (;;) {
    const int zeros = ctz32_nz(g | (0xFFFFFFFFu << i));
    g >>= zeros;
    u <<= zeros;
    v <<= zeros;
    eta -= zeros;
    i -= zeros;
    if (i == 0) break;
    // f and g are odd here
    if (eta < 0) {
      uint32_t tmp;
      eta = -eta;
      tmp = f; f = g; g = 0u - tmp;
      tmp = u; u = q; q = 0u - tmp;
      tmp = v; v = r; r = 0u - tmp;
    }
    // cancel as many low bits of g as eta and the remaining step count allow (up to 8 at once):
    // w = -g / f mod 2^limit, with 1/f mod 2^8 from three Newton steps (f odd)
    int limit = (eta + 1) > i ? i : (eta + 1);
    if (limit > 8) limit = 8;
    const uint32_t m = (0xFFFFFFFFu >> (32 - limit));
    uint32_t finv = f;                    // f * f = 1 mod 8
    finv *= 2u - f * finv;                // 6 bits
    finv *= 2u - f * finv;                // 12 bits
    const uint32_t w = ((0u - g) * finv) & m;
    g += f * w;
    q += u * w;
    r += v * w;
  }
  t.u = (int32_t)u;
  t.v = (int32_t)v;
  t.q = (int32_t)q;
  t.r = (int32_t)r;
  return eta;
}
// [d, e] <- t * [d, e] / 2^30 mod m, both kept in (-2m, m)
HD void s30_update_de(S30& d, S30& e, const T2x2& t, const S30& mod, uint32_t mod_inv30) {
  const int32_t M30 = 0x3FFFFFFF;
  const int32_t u = t.u, v = t.v, q = t.q, r = t.r;
  const int32_t sd = d.v[8] >> 31, se = e.v[8] >> 31;
  int32_t md = (u & sd) + (v & se);
  int32_t me = (q & sd) + (r & se);
  int32_t di = d.v[0], ei = e.v[0];
  int64_t cd = (int64_t)u * di + (int64_t)v * ei;
  int64_t ce = (int64_t)q * di + (int64_t)r * ei;
  // choose the multiples of m that make the low 30 bits vanish
  md -= (int32_t)((mod_inv30 * (uint32_t)cd + (uint32_t)md) & (uint32_t)M30);
  me -= (int32_t)((mod_inv30 * (uint32_t)ce + (uint32_t)me) & (uint32_t)M30);
  cd += (int64_t)mod.v[0] * md;
  ce += (int64_t)mod.v[0] * me;
  cd >>= 30;
  ce >>= 30;
#pragma unroll
  for (int i = 1; i < 9; i++) {
    di = d.v[i];
    ei = e.v[i];
    cd += (int64_t)u * di + (int64_t)v * ei;
    ce += (int64_t)q * di + (int64_t)r * ei;
    cd += (int64_t)mod.v[i] * md;
    ce += (int64_t)mod.v[i] * me;
    d.v[i - 1] = (int32_t)cd & M30;
    cd >>= 30;
    e.v[i - 1] = (int32_t)ce & M30;
    ce >>= 30;
  }
  d.v[8] = (int32_t)cd;
  e.v[8] = (int32_t)ce;
}
// [f, g] <- t * [f, g] / 2^30 (exact)
HD void s30_update_fg(S30& f, S30& g, const T2x2& t) {
  const int32_t M30 = 0x3FFFFFFF;
  const int32_t u = t.u, v = t.v, q = t.q, r = t.r;
  int32_t fi = f.v[0], gi = g.v[0];
  int64_t cf = (int64_t)u * fi + (int64_t)v * gi;
  int64_t cg = (int64_t)q * fi + (int64_t)r * gi;
  cf >>= 30;
  cg >>= 30;
#pragma unroll
  for (int i = 1; i < 9; i++) {
    fi = f.v[i];
    gi = g.v[i];
    cf += (int64_t)u * fi + (int64_t)v * gi;
    cg += (int64_t)q * fi + (int64_t)r * gi;
    f.v[i - 1] = (int32_t)cf & M30;
    cf >>= 30;
    g.v[i - 1] = (int32_t)cg & M30;
    cg >>= 30;
  }
  f.v[8] = (int32_t)cf;
  g.v[8] = (int32_t)cg;
}
// a^-1 mod m for odd m < 2^256 and 0 <= a < m with gcd(a, m) = 1 (plain integers); returns 0 for a = 0.
HDN inline U256 u256_modinv_odd(const U256& a, const U256& m) {
  U256 zero;
#pragma unroll
  for (int i = 0; i < 8; i++) zero.v[i] = 0;
  if (u256_is_zero(a)) return zero;
  const S30 mod = s30_from_u256(m);
  // m^-1 mod 2^30 by Newton iteration
  uint32_t minv = m.v[0];
  minv *= 2u - m.v[0] * minv;
  minv *= 2u - m.v[0] * minv;
  minv *= 2u - m.v[0] * minv;
  minv *= 2u - m.v[0] * minv;
  minv &= 0x3FFFFFFFu;
  S30 d, e, f = mod, g = s30_from_u256(a);
#pragma unroll
  for (int i = 0; i < 9; i++) d.v[i] = e.v[i] = 0;
  e.v[0] = 1;
  int32_t eta = -1;
  for (int it = 0; it < 40; it++) {     // 25 batches always suffice for 256-bit inputs; 40 is a guard
    T2x2 t;
    eta = divsteps_30_var(eta, (uint32_t)f.v[0], (uint32_t)g.v[0], t);
    s30_update_de(d, e, t, mod, minv);
    s30_update_fg(f, g, t);
    int32_t nz = 0;
#pragma unroll
    for (int i = 0; i < 9; i++) nz |= g.v[i];
    if (nz == 0) break;
  }
  // now f = +-1 and d = +-a^-1 with d in (-2m, m): fold the sign of f in and bring d into [0, m)
  const int32_t fneg = f.v[8] >> 31;
  int64_t c = 0;
  const int32_t M30 = 0x3FFFFFFF;
  // d <- (d < 0 ? d + m : d)
  int32_t add = d.v[8] >> 31;
#pragma unroll
  for (int i = 0; i < 9; i++) {
    c += (int64_t)d.v[i] + (mod.v[i] & add);
    d.v[i] = (i < 8) ? ((int32_t)c & M30) : (int32_t)c;
    if (i < 8) c >>= 30;
  }
  // d <- (f < 0 ? -d : d)
  c = 0;
#pragma unroll
  for (int i = 0; i < 9; i++) {
    c += (int64_t)((d.v[i] ^ fneg) - fneg);
    d.v[i] = (i < 8) ? ((int32_t)c & M30) : (int32_t)c;
    if (i < 8) c >>= 30;
  }
  // twice: d may still be in (-2m, 0)
  for (int rep = 0; rep < 2; rep++) {
    add = d.v[8] >> 31;
    c = 0;
#pragma unroll
    for (int i = 0; i < 9; i++) {
      c += (int64_t)d.v[i] + (mod.v[i] & add);
      d.v[i] = (i < 8) ? ((int32_t)c & M30) : (int32_t)c;
      if (i < 8) c >>= 30;
    }
  }
  // d in [0, 2m) possible after the first fold: subtract m once if needed
  U256 out = s30_to_u256(d);
  if (u256_geq(out, m)) u256_sub(out, out, m);
  return out;
}

// Inverse of a Montgomery-form element.  inv(0) = 0.  (x = aR; modinv gives a^-1 R^-1; two Montgomery products
// with R^2 bring it back to a^-1 R.)
template <class M>
HDN Fe<M> fe_inv(const Fe<M>& a) {
  Fe<M> t, r2;
  t.x = u256_modinv_odd(a.x, mod_of<M>());
#pragma unroll
  for (int i = 0; i < 8; i++) r2.x.v[i] = M::r2(i);
  return fe_mul(fe_mul(t, r2), r2);
}
// The reference's Fermat form a^(m-2) (gfp.go:30-48); kept as the oracle fe_inv is tested against.
template <class M>
HDN Fe<M> fe_inv_fermat(const Fe<M>& a) {
  U256 e = mod_of<M>();
  U256 two;
  for (int i = 0; i < 8; i++) two.v[i] = 0;
  two.v[0] = 2;
  u256_sub(e, e, two);
  return fe_pow(a, e);
}

// big-endian 32-byte <-> U256
HD U256 u256_from_be(const uint8_t* b) {
  U256 r;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    const uint8_t* q = b + 28 - 4 * i;
    r.v[i] = ((uint32_t)q[0] << 24) | ((uint32_t)q[1] << 16) | ((uint32_t)q[2] << 8) | q[3];
  }
  return r;
}
HD void u256_to_be(uint8_t* b, const U256& a) {
#pragma unroll
  for (int i = 0; i < 8; i++) {
    uint8_t* q = b + 28 - 4 * i;
    q[0] = (uint8_t)(a.v[i] >> 24);
    q[1] = (uint8_t)(a.v[i] >> 16);
    q[2] = (uint8_t)(a.v[i] >> 8);
    q[3] = (uint8_t)(a.v[i]);
  }
}
HD U256 u256_from_le(const uint8_t* b) {
  U256 r;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    const uint8_t* q = b + 4 * i;
    r.v[i] = ((uint32_t)q[3] << 24) | ((uint32_t)q[2] << 16) | ((uint32_t)q[1] << 8) | q[0];
  }
  return r;
}
HD void u256_to_le(uint8_t* b, const U256& a) {
#pragma unroll
  for (int i = 0; i < 8; i++) {
    uint8_t* q = b + 4 * i;
    q[0] = (uint8_t)(a.v[i]);
    q[1] = (uint8_t)(a.v[i] >> 8);
    q[2] = (uint8_t)(a.v[i] >> 16);
    q[3] = (uint8_t)(a.v[i] >> 24);
  }
}

}  // namespace bsc
