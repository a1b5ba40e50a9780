// BN256 G1 (y^2 = x^3 + 3 over Fp) group law, host+device.
//
// Behavioural parity with lib/dedis/kyber/pairing/bn256/curve.go:9-231 (Jacobian coordinates,
// generator (1,-2), wire format 64-byte big-endian affine x||y with all-zero = infinity,
// point.go:109-167).  The formulas are the textbook a=0 Jacobian ones (dbl-2009-l / madd-2007-bl /
// add-2007-bl); the scalar-multiplication strategy (interleaved Straus over small signed scalars,
// batched affine normalisation) is this framework's own — the reference does one MSB-first
// double-and-add per term (curve.go:186-200).
#pragma once
#include "field.cuh"

namespace bsc {

typedef Fe<BnP> Fp;
typedef Fe<BnN> Fr;

struct G1Affine {  // Montgomery-form coordinates; infinity encoded as (0,0)
  Fp x, y;
};
struct G1Jac {  // infinity <=> z == 0
  Fp x, y, z;
};

HD bool g1a_is_inf(const G1Affine& p) { return fe_is_zero(p.x) && fe_is_zero(p.y); }
HD bool g1j_is_inf(const G1Jac& p) { return fe_is_zero(p.z); }

HD G1Jac g1j_inf() {
  G1Jac r;
  r.x = fe_zero<BnP>();
  r.y = fe_one<BnP>();
  r.z = fe_zero<BnP>();
  return r;
}
HD G1Affine g1a_inf() {
  G1Affine r;
  r.x = fe_zero<BnP>();
  r.y = fe_zero<BnP>();
  return r;
}
HD G1Affine g1a_neg(const G1Affine& p) {
  G1Affine r;
  r.x = p.x;
  r.y = fe_neg(p.y);
  return r;
}
HD G1Jac g1j_from_affine(const G1Affine& p) {
  if (g1a_is_inf(p)) return g1j_inf();
  G1Jac r;
  r.x = p.x;
  r.y = p.y;
  r.z = fe_one<BnP>();
  return r;
}
HD G1Affine g1_generator() {
  G1Affine g;
  g.x = fe_one<BnP>();
  g.y = fe_neg(fe_from_u64<BnP>(2));
  return g;
}

HDN inline G1Jac g1j_dbl(const G1Jac& p) {
  if (g1j_is_inf(p)) return p;
  Fp A = fe_sqr(p.x);
  Fp B = fe_sqr(p.y);
  Fp C = fe_sqr(B);
  Fp t = fe_add(p.x, B);
  Fp D = fe_sub(fe_sub(fe_sqr(t), A), C);
  D = fe_dbl(D);
  Fp E = fe_add(fe_dbl(A), A);
  Fp F = fe_sqr(E);
  G1Jac r;
  r.x = fe_sub(F, fe_dbl(D));
  Fp C8 = fe_dbl(fe_dbl(fe_dbl(C)));
  r.z = fe_dbl(fe_mul(p.y, p.z));
  r.y = fe_sub(fe_mul(E, fe_sub(D, r.x)), C8);
  return r;
}

// p (Jacobian) + q (affine)
HDN inline G1Jac g1j_add_affine(const G1Jac& p, const G1Affine& q) {
  if (g1a_is_inf(q)) return p;
  if (g1j_is_inf(p)) return g1j_from_affine(q);
  Fp Z1Z1 = fe_sqr(p.z);
  Fp U2 = fe_mul(q.x, Z1Z1);
  Fp S2 = fe_mul(fe_mul(q.y, p.z), Z1Z1);
  Fp H = fe_sub(U2, p.x);
  Fp rr = fe_sub(S2, p.y);
  if (fe_is_zero(H)) {
    if (fe_is_zero(rr)) return g1j_dbl(p);
    return g1j_inf();
  }
  rr = fe_dbl(rr);
  Fp HH = fe_sqr(H);
  Fp I = fe_dbl(fe_dbl(HH));
  Fp J = fe_mul(H, I);
  Fp V = fe_mul(p.x, I);
  G1Jac r;
  r.x = fe_sub(fe_sub(fe_sqr(rr), J), fe_dbl(V));
  Fp YJ = fe_mul(p.y, J);
  r.y = fe_sub(fe_mul(rr, fe_sub(V, r.x)), fe_dbl(YJ));
  Fp t = fe_add(p.z, H);
  r.z = fe_sub(fe_sub(fe_sqr(t), Z1Z1), HH);
  return r;
}

// general Jacobian + Jacobian
HDN inline G1Jac g1j_add(const G1Jac& p, const G1Jac& q) {
  if (g1j_is_inf(p)) return q;
  if (g1j_is_inf(q)) return p;
  Fp Z1Z1 = fe_sqr(p.z);
  Fp Z2Z2 = fe_sqr(q.z);
  Fp U1 = fe_mul(p.x, Z2Z2);
  Fp U2 = fe_mul(q.x, Z1Z1);
  Fp S1 = fe_mul(fe_mul(p.y, q.z), Z2Z2);
  Fp S2 = fe_mul(fe_mul(q.y, p.z), Z1Z1);
  Fp H = fe_sub(U2, U1);
  Fp rr = fe_sub(S2, S1);
  if (fe_is_zero(H)) {
    if (fe_is_zero(rr)) return g1j_dbl(p);
    return g1j_inf();
  }
  rr = fe_dbl(rr);
  Fp I = fe_sqr(fe_dbl(H));
  Fp J = fe_mul(H, I);
  Fp V = fe_mul(U1, I);
  G1Jac r;
  r.x = fe_sub(fe_sub(fe_sqr(rr), J), fe_dbl(V));
  Fp SJ = fe_mul(S1, J);
  r.y = fe_sub(fe_mul(rr, fe_sub(V, r.x)), fe_dbl(SJ));
  Fp t = fe_add(p.z, q.z);
  r.z = fe_mul(fe_sub(fe_sub(fe_sqr(t), Z1Z1), Z2Z2), H);
  return r;
}

// Normalise given zinv = 1/z (caller supplies the inverse so it can be batched).
HD G1Affine g1j_to_affine_with_inv(const G1Jac& p, const Fp& zinv) {
  if (g1j_is_inf(p)) return g1a_inf();
  Fp zi2 = fe_sqr(zinv);
  G1Affine r;
  r.x = fe_mul(p.x, zi2);
  r.y = fe_mul(fe_mul(p.y, zi2), zinv);
  return r;
}
HDN inline G1Affine g1j_to_affine(const G1Jac& p) {
  if (g1j_is_inf(p)) return g1a_inf();
  return g1j_to_affine_with_inv(p, fe_inv(p.z));
}

// 64-byte wire format (plain big-endian x||y; zeros = infinity) — point.go:109-167.
HD void g1a_to_bytes(uint8_t* out, const G1Affine& p) {
  if (g1a_is_inf(p)) {
    for (int i = 0; i < 64; i++) out[i] = 0;
    return;
  }
  u256_to_be(out, fe_to_u256(p.x));
  u256_to_be(out + 32, fe_to_u256(p.y));
}
HD G1Affine g1a_from_bytes(const uint8_t* in) {
  G1Affine r;
  r.x = fe_from_u256<BnP>(u256_from_be(in));
  r.y = fe_from_u256<BnP>(u256_from_be(in + 32));
  return r;
}
HDN inline bool g1a_on_curve(const G1Affine& p) {
  if (g1a_is_inf(p)) return true;
  Fp y2 = fe_sqr(p.y);
  Fp x3 = fe_add(fe_mul(fe_sqr(p.x), p.x), fe_from_u64<BnP>(3));
  return fe_eq(y2, x3);
}

// k * P for a plain 256-bit scalar, MSB-first double-and-add (used by host keygen and the oracle;
// the device MSM does not call this).
HDN inline G1Jac g1_mul_u256(const G1Affine& p, const U256& k) {
  G1Jac acc = g1j_inf();
  bool started = false;
  for (int i = 255; i >= 0; i--) {
    if (started) acc = g1j_dbl(acc);
    if ((k.v[i >> 5] >> (i & 31)) & 1) {
      acc = g1j_add_affine(acc, p);
      started = true;
    }
  }
  return acc;
}

// Interleaved (Straus) multi-scalar multiplication with signed 64-bit scalars:
//   sum_t s[t] * base[t],  t < T.  nbits = bit length of max |s[t]| (caller computes).
// One shared doubling chain; a negative scalar adds the negated base, which equals the
// reference's SetInt64 -> reduce-mod-order behaviour (kyber.go:541-552).
template <int TMAX>
HDN inline G1Jac g1_straus_i64(const G1Affine* base, const int64_t* s, int T, int nbits) {
  uint64_t mag[TMAX];
  uint32_t negmask = 0;
  for (int t = 0; t < T; t++) {
    int64_t v = s[t];
    if (v < 0) {
      negmask |= 1u << t;
      mag[t] = (uint64_t)0 - (uint64_t)v;
    } else {
      mag[t] = (uint64_t)v;
    }
  }
  G1Jac acc = g1j_inf();
  for (int b = nbits - 1; b >= 0; b--) {
    acc = g1j_dbl(acc);
    for (int t = 0; t < T; t++) {
      if ((mag[t] >> b) & 1) {
        G1Affine q = base[t];
        if ((negmask >> t) & 1) q.y = fe_neg(q.y);
        acc = g1j_add_affine(acc, q);
      }
    }
  }
  return acc;
}

}  // namespace bsc
