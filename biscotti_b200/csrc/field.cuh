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
// Fermat inversion a^(m-2) (the reference does the same, gfp.go:30-48).  inv(0) = 0.
template <class M>
HDN Fe<M> fe_inv(const Fe<M>& a) {
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
