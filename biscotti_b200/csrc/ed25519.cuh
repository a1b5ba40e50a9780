// Ed25519 group arithmetic (host+device) for the ECVRF used in committee / noiser selection.
//
// The reference calls github.com/coniks-sys/coniks-go/crypto/vrf (Ed25519 + Elligator2 + SHAKE256;
// not vendored) from DistSys/vrf.go:16-100.  This is a from-scratch implementation of the same
// construction on the shared 8x32-bit Montgomery field (field.cuh): extended twisted-Edwards
// coordinates, unified addition, Elligator2 hash-to-curve, cofactor clearing.
#pragma once
#include "field.cuh"

namespace bsc {

typedef Fe<EdP> Fq;   // base field 2^255-19
typedef Fe<EdL> Fs;   // scalar field (group order l)

struct EdPoint {  // extended coordinates, x = X/Z, y = Y/Z, T = XY/Z
  Fq X, Y, Z, T;
};

HD Fq ed_const_d() {
  U256 t = {{0x135978a3u, 0x75eb4dcau, 0x4141d8abu, 0x00700a4du, 0x7779e898u, 0x8cc74079u, 0x2b6ffe73u, 0x52036ceeu}};
  return fe_from_u256<EdP>(t);
}
HD Fq ed_const_2d() {
  U256 t = {{0x26b2f159u, 0xebd69b94u, 0x8283b156u, 0x00e0149au, 0xeef3d130u, 0x198e80f2u, 0x56dffce7u, 0x2406d9dcu}};
  return fe_from_u256<EdP>(t);
}
HD Fq ed_const_sqrtm1() {
  U256 t = {{0x4a0ea0b0u, 0xc4ee1b27u, 0xad2fe478u, 0x2f431806u, 0x3dfbd7a7u, 0x2b4d0099u, 0x4fc1df0bu, 0x2b832480u}};
  return fe_from_u256<EdP>(t);
}

HD EdPoint ed_identity() {
  EdPoint p;
  p.X = fe_zero<EdP>();
  p.Y = fe_one<EdP>();
  p.Z = fe_one<EdP>();
  p.T = fe_zero<EdP>();
  return p;
}

// add-2008-hwcd-3 (a = -1): complete, also valid for doubling
HDN inline EdPoint ed_add(const EdPoint& p, const EdPoint& q, const Fq& d2) {
  Fq A = fe_mul(fe_sub(p.Y, p.X), fe_sub(q.Y, q.X));
  Fq B = fe_mul(fe_add(p.Y, p.X), fe_add(q.Y, q.X));
  Fq C = fe_mul(fe_mul(p.T, d2), q.T);
  Fq D = fe_dbl(fe_mul(p.Z, q.Z));
  Fq E = fe_sub(B, A), F = fe_sub(D, C), G = fe_add(D, C), H = fe_add(B, A);
  EdPoint r;
  r.X = fe_mul(E, F);
  r.Y = fe_mul(G, H);
  r.T = fe_mul(E, H);
  r.Z = fe_mul(F, G);
  return r;
}
// dbl-2008-hwcd (a = -1)
HDN inline EdPoint ed_dbl(const EdPoint& p) {
  Fq A = fe_sqr(p.X), B = fe_sqr(p.Y);
  Fq C = fe_dbl(fe_sqr(p.Z));
  Fq D = fe_neg(A);
  Fq t = fe_add(p.X, p.Y);
  Fq E = fe_sub(fe_sub(fe_sqr(t), A), B);
  Fq G = fe_add(D, B), F = fe_sub(G, C), H = fe_sub(D, B);
  EdPoint r;
  r.X = fe_mul(E, F);
  r.Y = fe_mul(G, H);
  r.T = fe_mul(E, H);
  r.Z = fe_mul(F, G);
  return r;
}
HDN inline EdPoint ed_mul(const EdPoint& p, const U256& k) {
  const Fq d2 = ed_const_2d();
  EdPoint acc = ed_identity();
  bool started = false;
  for (int i = 255; i >= 0; i--) {
    if (started) acc = ed_dbl(acc);
    if ((k.v[i >> 5] >> (i & 31)) & 1) {
      acc = started ? ed_add(acc, p, d2) : p;
      started = true;
    }
  }
  return acc;
}

// 32-byte little-endian y with the sign of x in bit 255
HDN inline void ed_compress(uint8_t out[32], const EdPoint& p) {
  Fq zi = fe_inv(p.Z);
  U256 x = fe_to_u256(fe_mul(p.X, zi));
  U256 y = fe_to_u256(fe_mul(p.Y, zi));
  u256_to_le(out, y);
  out[31] |= (uint8_t)((x.v[0] & 1) << 7);
}
// x from y and a sign bit; false if y is not on the curve
HDN inline bool ed_from_parity_and_y(EdPoint& r, uint32_t sign, const Fq& y) {
  const Fq one = fe_one<EdP>();
  Fq y2 = fe_sqr(y);
  Fq u = fe_sub(y2, one);
  Fq v = fe_add(fe_mul(y2, ed_const_d()), one);
  // x = sqrt(u/v): candidate = u v^3 (u v^7)^((p-5)/8)
  Fq v3 = fe_mul(fe_sqr(v), v);
  Fq v7 = fe_mul(fe_sqr(v3), v);
  U256 e = {{0xfffffffdu, 0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu, 0x0fffffffu}};
  Fq x = fe_mul(fe_mul(u, v3), fe_pow(fe_mul(u, v7), e));
  Fq vx2 = fe_mul(v, fe_sqr(x));
  if (!fe_eq(vx2, u)) {
    if (!fe_eq(vx2, fe_neg(u))) return false;
    x = fe_mul(x, ed_const_sqrtm1());
  }
  U256 xp = fe_to_u256(x);
  if ((xp.v[0] & 1) != sign) x = fe_neg(x);
  r.X = x;
  r.Y = y;
  r.Z = one;
  r.T = fe_mul(x, y);
  return true;
}
HDN inline bool ed_decompress(EdPoint& r, const uint8_t in[32]) {
  uint8_t t[32];
  for (int i = 0; i < 32; i++) t[i] = in[i];
  const uint32_t sign = t[31] >> 7;
  t[31] &= 127;
  return ed_from_parity_and_y(r, sign, fe_from_u256<EdP>(u256_from_le(t)));
}
HDN inline EdPoint ed_basepoint() {
  // y = 4/5, x positive (even)
  EdPoint b;
  Fq y = fe_mul(fe_from_u64<EdP>(4), fe_inv(fe_from_u64<EdP>(5)));
  ed_from_parity_and_y(b, 0, y);
  return b;
}

// Elligator2: 32-byte representative -> Edwards point (sign from bit 255), then x8 to clear the
// cofactor.  hashToCurve of the coniks VRF = this applied to SHAKE256(m)[:32].
HDN inline EdPoint ed_elligator2_x8(const uint8_t h[32]) {
  uint8_t t[32];
  for (int i = 0; i < 32; i++) t[i] = h[i];
  const uint32_t bit = t[31] >> 7;
  t[31] &= 127;
  const Fq one = fe_one<EdP>();
  const Fq A = fe_from_u64<EdP>(486662);
  Fq r = fe_from_u256<EdP>(u256_from_le(t));
  Fq den = fe_add(fe_dbl(fe_sqr(r)), one);          // 1 + 2 r^2
  Fq v = fe_neg(fe_mul(A, fe_inv(den)));             // -A / (1 + 2 r^2)
  Fq v2 = fe_sqr(v);
  Fq e = fe_add(fe_add(fe_mul(v2, v), v), fe_mul(v2, A));  // v^3 + A v^2 + v
  // Legendre symbol via e^((p-1)/2)
  U256 half = {{0xfffffff6u, 0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu, 0x3fffffffu}};
  Fq chi = fe_pow(e, half);
  const bool minus1 = !fe_eq(chi, one) && !fe_is_zero(chi);
  if (minus1) v = fe_sub(fe_neg(v), A);
  // Montgomery u -> Edwards y = (u - 1)/(u + 1)
  Fq y = fe_mul(fe_sub(v, one), fe_inv(fe_add(v, one)));
  EdPoint p;
  if (!ed_from_parity_and_y(p, bit, y)) p = ed_identity();
  p = ed_dbl(ed_dbl(ed_dbl(p)));
  return p;
}

}  // namespace bsc
