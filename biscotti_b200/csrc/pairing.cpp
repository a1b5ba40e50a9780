// BN256 G2, GT and the optimal-ate pairing (host runtime).
//
// Covers the part of the reference's vendored curve that sits off the per-block hot path
// (lib/dedis/kyber/pairing/bn256/{gfp2,gfp6,gfp12,twist,optate}.go, point.go:190-587): G2 points for the
// second half of the SRS (DistSys/publicKey.go:42-57, commitKey.json "Skey"), and the pairing behind
// verifySecret (DistSys/kyber.go:650-673), which the reference defines but never calls (TODO at kyber.go:476).
//
// Own formulation, not a transcription: the tower is Fp2 = Fp[i]/(i^2+1), Fp6 = Fp2[v]/(v^3-xi),
// Fp12 = Fp6[w]/(w^2-v) with xi = 3+i, all Frobenius constants are derived at start-up from xi (no tables of
// magic limbs), the Miller loop keeps the G2 accumulator affine (one shared Fp inversion per step — ~130
// inversions per pairing is nothing on a host core) and multiplies by the sparse line
// yP - lam*xP*w + (lam*xT - yT)*w^3 directly, and a product of pairings shares one final exponentiation.
// Byte-level oracle: biscotti_b200/oracle/bn256_pairing.py (flat Fp2[w]/(w^6-xi), generic exponent).
#include <string.h>
#include <vector>
#include "g1.cuh"

namespace bsc {

// ---------------------------------------------------------------------------------------------------
// Fp2
struct Fp2 {
  Fp a, b;  // a + b*i
};
static inline Fp2 f2_zero() { return {fe_zero<BnP>(), fe_zero<BnP>()}; }
static inline Fp2 f2_one() { return {fe_one<BnP>(), fe_zero<BnP>()}; }
static inline bool f2_is_zero(const Fp2& x) { return fe_is_zero(x.a) && fe_is_zero(x.b); }
static inline bool f2_eq(const Fp2& x, const Fp2& y) { return fe_eq(x.a, y.a) && fe_eq(x.b, y.b); }
static inline Fp2 f2_add(const Fp2& x, const Fp2& y) { return {fe_add(x.a, y.a), fe_add(x.b, y.b)}; }
static inline Fp2 f2_sub(const Fp2& x, const Fp2& y) { return {fe_sub(x.a, y.a), fe_sub(x.b, y.b)}; }
static inline Fp2 f2_neg(const Fp2& x) { return {fe_neg(x.a), fe_neg(x.b)}; }
static inline Fp2 f2_dbl(const Fp2& x) { return f2_add(x, x); }
static inline Fp2 f2_conj(const Fp2& x) { return {x.a, fe_neg(x.b)}; }
static inline Fp2 f2_scale(const Fp2& x, const Fp& k) { return {fe_mul(x.a, k), fe_mul(x.b, k)}; }
static inline Fp2 f2_mul(const Fp2& x, const Fp2& y) {
  const Fp t0 = fe_mul(x.a, y.a), t1 = fe_mul(x.b, y.b);
  const Fp t2 = fe_mul(fe_add(x.a, x.b), fe_add(y.a, y.b));
  return {fe_sub(t0, t1), fe_sub(fe_sub(t2, t0), t1)};
}
static inline Fp2 f2_sqr(const Fp2& x) {
  const Fp t = fe_mul(x.a, x.b);
  return {fe_mul(fe_add(x.a, x.b), fe_sub(x.a, x.b)), fe_add(t, t)};
}
// (a + b i)(3 + i) = (3a - b) + (a + 3b) i
static inline Fp2 f2_mul_xi(const Fp2& x) {
  const Fp a3 = fe_add(fe_dbl(x.a), x.a), b3 = fe_add(fe_dbl(x.b), x.b);
  return {fe_sub(a3, x.b), fe_add(x.a, b3)};
}
static inline Fp2 f2_inv(const Fp2& x) {
  const Fp d = fe_inv(fe_add(fe_sqr(x.a), fe_sqr(x.b)));
  return {fe_mul(x.a, d), fe_neg(fe_mul(x.b, d))};
}

// ---------------------------------------------------------------------------------------------------
// Fp6 = Fp2[v]/(v^3 - xi)
struct Fp6 {
  Fp2 c0, c1, c2;
};
static inline Fp6 f6_zero() { return {f2_zero(), f2_zero(), f2_zero()}; }
static inline Fp6 f6_one() { return {f2_one(), f2_zero(), f2_zero()}; }
static inline Fp6 f6_add(const Fp6& x, const Fp6& y) { return {f2_add(x.c0, y.c0), f2_add(x.c1, y.c1), f2_add(x.c2, y.c2)}; }
static inline Fp6 f6_sub(const Fp6& x, const Fp6& y) { return {f2_sub(x.c0, y.c0), f2_sub(x.c1, y.c1), f2_sub(x.c2, y.c2)}; }
static inline Fp6 f6_neg(const Fp6& x) { return {f2_neg(x.c0), f2_neg(x.c1), f2_neg(x.c2)}; }
static inline Fp6 f6_mul_v(const Fp6& x) { return {f2_mul_xi(x.c2), x.c0, x.c1}; }
static inline Fp6 f6_scale(const Fp6& x, const Fp& k) { return {f2_scale(x.c0, k), f2_scale(x.c1, k), f2_scale(x.c2, k)}; }
static Fp6 f6_mul(const Fp6& x, const Fp6& y) {
  const Fp2 v0 = f2_mul(x.c0, y.c0), v1 = f2_mul(x.c1, y.c1), v2 = f2_mul(x.c2, y.c2);
  const Fp2 t0 = f2_sub(f2_sub(f2_mul(f2_add(x.c1, x.c2), f2_add(y.c1, y.c2)), v1), v2);
  const Fp2 t1 = f2_sub(f2_sub(f2_mul(f2_add(x.c0, x.c1), f2_add(y.c0, y.c1)), v0), v1);
  const Fp2 t2 = f2_sub(f2_sub(f2_mul(f2_add(x.c0, x.c2), f2_add(y.c0, y.c2)), v0), v2);
  return {f2_add(v0, f2_mul_xi(t0)), f2_add(t1, f2_mul_xi(v2)), f2_add(t2, v1)};
}
// x * (s0 + s1 v): the shape of the w-part of a line function
static Fp6 f6_mul_sparse01(const Fp6& x, const Fp2& s0, const Fp2& s1) {
  return {f2_add(f2_mul(x.c0, s0), f2_mul_xi(f2_mul(x.c2, s1))), f2_add(f2_mul(x.c0, s1), f2_mul(x.c1, s0)),
          f2_add(f2_mul(x.c1, s1), f2_mul(x.c2, s0))};
}
static Fp6 f6_inv(const Fp6& x) {
  const Fp2 A = f2_sub(f2_sqr(x.c0), f2_mul_xi(f2_mul(x.c1, x.c2)));
  const Fp2 B = f2_sub(f2_mul_xi(f2_sqr(x.c2)), f2_mul(x.c0, x.c1));
  const Fp2 C = f2_sub(f2_sqr(x.c1), f2_mul(x.c0, x.c2));
  const Fp2 F = f2_add(f2_mul(x.c0, A), f2_mul_xi(f2_add(f2_mul(x.c2, B), f2_mul(x.c1, C))));
  const Fp2 Fi = f2_inv(F);
  return {f2_mul(A, Fi), f2_mul(B, Fi), f2_mul(C, Fi)};
}

// ---------------------------------------------------------------------------------------------------
// Fp12 = Fp6[w]/(w^2 - v); flat view: coefficient of w^k is (k even ? a : b).c[k/2]
struct Fp12 {
  Fp6 a, b;
};
static inline Fp12 f12_one() { return {f6_one(), f6_zero()}; }
static inline Fp12 f12_conj(const Fp12& x) { return {x.a, f6_neg(x.b)}; }
static Fp12 f12_mul(const Fp12& x, const Fp12& y) {
  const Fp6 aa = f6_mul(x.a, y.a), bb = f6_mul(x.b, y.b);
  const Fp6 cr = f6_sub(f6_sub(f6_mul(f6_add(x.a, x.b), f6_add(y.a, y.b)), aa), bb);
  return {f6_add(aa, f6_mul_v(bb)), cr};
}
static Fp12 f12_sqr(const Fp12& x) {
  // (a + bw)^2 = (a^2 + b^2 v) + 2ab w, with a^2 + b^2 v = (a + b)(a + bv) - ab - ab v
  const Fp6 ab = f6_mul(x.a, x.b);
  const Fp6 t = f6_mul(f6_add(x.a, x.b), f6_add(x.a, f6_mul_v(x.b)));
  return {f6_sub(f6_sub(t, ab), f6_mul_v(ab)), f6_add(ab, ab)};
}
static Fp12 f12_inv(const Fp12& x) {
  const Fp6 d = f6_inv(f6_sub(f6_mul(x.a, x.a), f6_mul_v(f6_mul(x.b, x.b))));
  return {f6_mul(x.a, d), f6_neg(f6_mul(x.b, d))};
}
static bool f12_is_one(const Fp12& x) {
  return f2_eq(x.a.c0, f2_one()) && f2_is_zero(x.a.c1) && f2_is_zero(x.a.c2) && f2_is_zero(x.b.c0) &&
         f2_is_zero(x.b.c1) && f2_is_zero(x.b.c2);
}
static inline Fp2* f12_coeff(Fp12& x, int k) {
  Fp6& h = (k & 1) ? x.b : x.a;
  return (k >> 1) == 0 ? &h.c0 : ((k >> 1) == 1 ? &h.c1 : &h.c2);
}
// f * (l0 + (s0 + s1 v) w) with l0 in Fp
static Fp12 f12_mul_line(const Fp12& f, const Fp& l0, const Fp2& s0, const Fp2& s1) {
  const Fp6 fb_s = f6_mul_sparse01(f.b, s0, s1);
  const Fp6 fa_s = f6_mul_sparse01(f.a, s0, s1);
  return {f6_add(f6_scale(f.a, l0), f6_mul_v(fb_s)), f6_add(fa_s, f6_scale(f.b, l0))};
}

// ---------------------------------------------------------------------------------------------------
// Frobenius: (c_k w^k)^p = conj(c_k) * gamma_k * w^k, gamma_k = xi^(k(p-1)/6)
static Fp2 g_gamma[6];
static bool g_pairing_ready = false;
static const uint64_t BN_U = 6518589491078791937ull;   // constants.go:13

static Fp2 f2_pow_u256(const Fp2& x, const U256& e) {
  Fp2 r = f2_one();
  for (int i = 255; i >= 0; i--) {
    r = f2_sqr(r);
    if ((e.v[i >> 5] >> (i & 31)) & 1) r = f2_mul(r, x);
  }
  return r;
}
static void pairing_init() {
  if (g_pairing_ready) return;
  U256 e = mod_of<BnP>();   // (p - 1) / 6 by long division in base 2^32
  e.v[0] -= 1;              // p is odd
  uint64_t rem = 0;
  for (int i = 7; i >= 0; i--) {
    const uint64_t cur = (rem << 32) | e.v[i];
    e.v[i] = (uint32_t)(cur / 6);
    rem = cur % 6;
  }
  const Fp2 xi = {fe_from_u64<BnP>(3), fe_one<BnP>()};
  g_gamma[0] = f2_one();
  g_gamma[1] = f2_pow_u256(xi, e);
  for (int k = 2; k < 6; k++) g_gamma[k] = f2_mul(g_gamma[k - 1], g_gamma[1]);
  g_pairing_ready = true;
}
static Fp12 f12_frob(const Fp12& x) {
  Fp12 r = x;
  for (int k = 0; k < 6; k++) {
    Fp2* c = f12_coeff(r, k);
    *c = f2_mul(f2_conj(*c), g_gamma[k]);
  }
  return r;
}
static Fp12 f12_pow_u64(const Fp12& x, uint64_t e) {
  Fp12 r = f12_one();
  for (int i = 63; i >= 0; i--) {
    r = f12_sqr(r);
    if ((e >> i) & 1) r = f12_mul(r, x);
  }
  return r;
}
static Fp12 f12_pow_u256(const Fp12& x, const U256& e) {
  Fp12 r = f12_one();
  for (int i = 255; i >= 0; i--) {
    r = f12_sqr(r);
    if ((e.v[i >> 5] >> (i & 31)) & 1) r = f12_mul(r, x);
  }
  return r;
}

// ---------------------------------------------------------------------------------------------------
// G2: y^2 = x^3 + 3/xi over Fp2 (twist.go:7-17)
struct G2Affine {
  Fp2 x, y;
  bool inf;
};
struct G2Jac {
  Fp2 x, y, z;  // infinity <=> z == 0
};
static Fp2 twist_b() {
  const Fp2 xi = {fe_from_u64<BnP>(3), fe_one<BnP>()};
  return f2_scale(f2_inv(xi), fe_from_u64<BnP>(3));
}
static G2Affine g2a_inf() { return {f2_zero(), f2_zero(), true}; }
static G2Jac g2j_inf() { return {f2_zero(), f2_one(), f2_zero()}; }
static G2Jac g2j_from_affine(const G2Affine& p) { return p.inf ? g2j_inf() : G2Jac{p.x, p.y, f2_one()}; }
static G2Affine g2j_to_affine(const G2Jac& p) {
  if (f2_is_zero(p.z)) return g2a_inf();
  const Fp2 zi = f2_inv(p.z), zi2 = f2_sqr(zi);
  return {f2_mul(p.x, zi2), f2_mul(f2_mul(p.y, zi2), zi), false};
}
static bool g2a_on_curve(const G2Affine& p) {
  if (p.inf) return true;
  return f2_eq(f2_sqr(p.y), f2_add(f2_mul(f2_sqr(p.x), p.x), twist_b()));
}
static G2Jac g2j_dbl(const G2Jac& p) {
  if (f2_is_zero(p.z)) return p;
  const Fp2 A = f2_sqr(p.x), B = f2_sqr(p.y), C = f2_sqr(B);
  const Fp2 D = f2_dbl(f2_sub(f2_sub(f2_sqr(f2_add(p.x, B)), A), C));
  const Fp2 E = f2_add(f2_dbl(A), A), F = f2_sqr(E);
  G2Jac r;
  r.x = f2_sub(F, f2_dbl(D));
  r.y = f2_sub(f2_mul(E, f2_sub(D, r.x)), f2_dbl(f2_dbl(f2_dbl(C))));
  r.z = f2_dbl(f2_mul(p.y, p.z));
  return r;
}
static G2Jac g2j_add(const G2Jac& p, const G2Jac& q) {
  if (f2_is_zero(p.z)) return q;
  if (f2_is_zero(q.z)) return p;
  const Fp2 z1z1 = f2_sqr(p.z), z2z2 = f2_sqr(q.z);
  const Fp2 u1 = f2_mul(p.x, z2z2), u2 = f2_mul(q.x, z1z1);
  const Fp2 s1 = f2_mul(f2_mul(p.y, q.z), z2z2), s2 = f2_mul(f2_mul(q.y, p.z), z1z1);
  const Fp2 h = f2_sub(u2, u1), rr = f2_dbl(f2_sub(s2, s1));
  if (f2_is_zero(h)) return f2_is_zero(rr) ? g2j_dbl(p) : g2j_inf();
  const Fp2 i = f2_sqr(f2_dbl(h)), j = f2_mul(h, i), v = f2_mul(u1, i);
  G2Jac r;
  r.x = f2_sub(f2_sub(f2_sqr(rr), j), f2_dbl(v));
  r.y = f2_sub(f2_mul(rr, f2_sub(v, r.x)), f2_dbl(f2_mul(s1, j)));
  r.z = f2_mul(f2_sub(f2_sub(f2_sqr(f2_add(p.z, q.z)), z1z1), z2z2), h);
  return r;
}
static G2Jac g2_mul_u256(const G2Affine& p, const U256& k) {
  const G2Jac base = g2j_from_affine(p);
  G2Jac acc = g2j_inf();
  for (int i = 255; i >= 0; i--) {
    acc = g2j_dbl(acc);
    if ((k.v[i >> 5] >> (i & 31)) & 1) acc = g2j_add(acc, base);
  }
  return acc;
}

// wire form (point.go:280-330): 0x01 | x.i | x.1 | y.i | y.1, 32-byte big-endian each; infinity: first byte 0x00
static void g2a_to_bytes(uint8_t* out, const G2Affine& p) {
  memset(out, 0, 129);
  if (p.inf) return;
  out[0] = 1;
  u256_to_be(out + 1, fe_to_u256(p.x.b));
  u256_to_be(out + 33, fe_to_u256(p.x.a));
  u256_to_be(out + 65, fe_to_u256(p.y.b));
  u256_to_be(out + 97, fe_to_u256(p.y.a));
}
static G2Affine g2a_from_bytes(const uint8_t* in) {
  if (in[0] != 1) return g2a_inf();
  G2Affine r;
  r.x.b = fe_from_u256<BnP>(u256_from_be(in + 1));
  r.x.a = fe_from_u256<BnP>(u256_from_be(in + 33));
  r.y.b = fe_from_u256<BnP>(u256_from_be(in + 65));
  r.y.a = fe_from_u256<BnP>(u256_from_be(in + 97));
  r.inf = false;
  return r;
}
static U256 u256_from_hex(const char* s) {
  uint8_t b[32];
  for (int i = 0; i < 32; i++) {
    auto nib = [](char c) { return (uint8_t)(c <= '9' ? c - '0' : (c | 32) - 'a' + 10); };
    b[i] = (uint8_t)((nib(s[2 * i]) << 4) | nib(s[2 * i + 1]));
  }
  return u256_from_be(b);
}
static G2Affine g2_generator() {   // twist.go:20-31, out of Montgomery form
  G2Affine g;
  g.x.a = fe_from_u256<BnP>(u256_from_hex("8f25386f72c9462b81597d65ae2092c4b97792155dcdaad32b8a6dd41792534c"));
  g.x.b = fe_from_u256<BnP>(u256_from_hex("2ecca446ff6f3d4d03c76e9b5c752f28bc37b364cb05ac4a37eb32e1c3245970"));
  g.y.a = fe_from_u256<BnP>(u256_from_hex("274e5747e8cafacc3716cc8699db79b22f0e4ff3c23e898f694420a3be3087a5"));
  g.y.b = fe_from_u256<BnP>(u256_from_hex("2db10ef5233b0fe3962b9ee6a4bbc2b5bde01a54f3513d42df972e128f31bf12"));
  g.inf = false;
  return g;
}

// ---------------------------------------------------------------------------------------------------
// Optimal ate: f_{6u+2,Q}(P) * l_{[6u+2]Q, pi(Q)}(P) * l_{[6u+2]Q + pi(Q), -pi^2(Q)}(P), then ^((p^12-1)/n)
static G2Affine g2_frob(const G2Affine& q) {   // Frobenius of the untwisted point, pulled back to the twist
  return {f2_mul(f2_conj(q.x), g_gamma[2]), f2_mul(f2_conj(q.y), g_gamma[3]), false};
}
// One Miller step: multiplies f by the line through t and q (tangent when q == nullptr) evaluated at p and
// replaces t by the sum.  Untwisting (x, y) -> (x w^2, y w^3) turns the twist slope lam into lam*w.
static void miller_step(Fp12& f, G2Affine& t, const G2Affine* q, const G1Affine& p) {
  Fp2 lam;
  if (q == nullptr) {
    const Fp2 x2 = f2_sqr(t.x);
    lam = f2_mul(f2_add(f2_dbl(x2), x2), f2_inv(f2_dbl(t.y)));
  } else {
    lam = f2_mul(f2_sub(q->y, t.y), f2_inv(f2_sub(q->x, t.x)));
  }
  const Fp2& ox = q ? q->x : t.x;
  const Fp2 x3 = f2_sub(f2_sub(f2_sqr(lam), t.x), ox);
  const Fp2 y3 = f2_sub(f2_mul(lam, f2_sub(t.x, x3)), t.y);
  f = f12_mul_line(f, p.y, f2_neg(f2_scale(lam, p.x)), f2_sub(f2_mul(lam, t.x), t.y));
  t.x = x3;
  t.y = y3;
}
static Fp12 miller(const G2Affine& q, const G1Affine& p) {
  if (q.inf || g1a_is_inf(p)) return f12_one();
  const unsigned __int128 loop = (unsigned __int128)6 * BN_U + 2;
  int top = 127;
  while (!((loop >> top) & 1)) top--;
  Fp12 f = f12_one();
  G2Affine t = q;
  for (int i = top - 1; i >= 0; i--) {
    f = f12_sqr(f);
    miller_step(f, t, nullptr, p);
    if ((loop >> i) & 1) miller_step(f, t, &q, p);
  }
  const G2Affine q1 = g2_frob(q);
  G2Affine q2 = g2_frob(q1);
  q2.y = f2_neg(q2.y);
  miller_step(f, t, &q1, p);
  miller_step(f, t, &q2, p);
  return f;
}
// Easy part by conjugation/Frobenius, hard part (p^4 - p^2 + 1)/n by the u-power addition chain of
// Scott, Benger, Charlemagne, Dominguez Perez, Kachisa (the decomposition every BN library uses).
static Fp12 final_exp(const Fp12& in) {
  Fp12 f = f12_mul(f12_conj(in), f12_inv(in));      // ^(p^6 - 1)
  f = f12_mul(f12_frob(f12_frob(f)), f);            // ^(p^2 + 1)
  const Fp12 fp1 = f12_frob(f), fp2 = f12_frob(fp1), fp3 = f12_frob(fp2);
  const Fp12 fu = f12_pow_u64(f, BN_U), fu2 = f12_pow_u64(fu, BN_U), fu3 = f12_pow_u64(fu2, BN_U);
  const Fp12 fu2p = f12_frob(fu2);
  const Fp12 y0 = f12_mul(f12_mul(fp1, fp2), fp3);
  const Fp12 y1 = f12_conj(f);
  const Fp12 y2 = f12_frob(fu2p);
  const Fp12 y3 = f12_conj(f12_frob(fu));
  const Fp12 y4 = f12_conj(f12_mul(fu, fu2p));
  const Fp12 y5 = f12_conj(fu2);
  const Fp12 y6 = f12_conj(f12_mul(fu3, f12_frob(fu3)));
  Fp12 t0 = f12_mul(f12_mul(f12_sqr(y6), y4), y5);
  Fp12 t1 = f12_mul(f12_mul(y3, y5), t0);
  t0 = f12_mul(t0, y2);
  t1 = f12_sqr(f12_mul(f12_sqr(t1), t0));
  t0 = f12_mul(t1, y1);
  t1 = f12_mul(t1, y0);
  return f12_mul(f12_sqr(t0), t1);
}

// GT wire form: coefficients of w^0..w^5, each i-part then 1-part, 32-byte big-endian (384 bytes)
static void f12_to_bytes(uint8_t* out, const Fp12& x) {
  Fp12 c = x;
  for (int k = 0; k < 6; k++) {
    const Fp2* e = f12_coeff(c, k);
    u256_to_be(out + 64 * k, fe_to_u256(e->b));
    u256_to_be(out + 64 * k + 32, fe_to_u256(e->a));
  }
}
static Fp12 f12_from_bytes(const uint8_t* in) {
  Fp12 r;
  for (int k = 0; k < 6; k++) {
    Fp2* e = f12_coeff(r, k);
    e->b = fe_from_u256<BnP>(u256_from_be(in + 64 * k));
    e->a = fe_from_u256<BnP>(u256_from_be(in + 64 * k + 32));
  }
  return r;
}

struct PairingOnce {
  PairingOnce() { pairing_init(); }
};
static void ensure_init() { static PairingOnce once; }

}  // namespace bsc

using namespace bsc;

extern "C" {

// ---- G2 (129-byte wire form; a first byte of 0x00 is the point at infinity) ---------------------------
void bsc_g2_generator(uint8_t* out129) { g2a_to_bytes(out129, g2_generator()); }
int bsc_g2_on_curve(const uint8_t* in129) {
  if (in129[0] == 0) return 1;
  if (in129[0] != 1) return 0;
  return g2a_on_curve(g2a_from_bytes(in129)) ? 1 : 0;
}
void bsc_g2_add(uint8_t* out129, const uint8_t* a129, const uint8_t* b129) {
  g2a_to_bytes(out129, g2j_to_affine(g2j_add(g2j_from_affine(g2a_from_bytes(a129)), g2j_from_affine(g2a_from_bytes(b129)))));
}
void bsc_g2_neg(uint8_t* out129, const uint8_t* a129) {
  G2Affine a = g2a_from_bytes(a129);
  if (!a.inf) a.y = f2_neg(a.y);
  g2a_to_bytes(out129, a);
}
// scalar: 32-byte big-endian, taken as is (callers reduce mod the group order)
void bsc_g2_mul(uint8_t* out129, const uint8_t* pt129, const uint8_t* k32) {
  g2a_to_bytes(out129, g2j_to_affine(g2_mul_u256(g2a_from_bytes(pt129), u256_from_be(k32))));
}
// PKG2[i] = tau^i * G2, i < d (DistSys/publicKey.go:42-57) -> d * 129 bytes
void bsc_srs_g2_host(const uint8_t* tau32, int d, uint8_t* out) {
  // chained: PKG2[i] = tau * PKG2[i-1], so a short tau (the reference's tau = 2) costs a few point operations per entry
  const U256 tau = u256_from_be(tau32);
  G2Affine cur = g2_generator();
  for (int i = 0; i < d; i++) {
    g2a_to_bytes(out + 129 * (size_t)i, cur);
    if (i + 1 < d) cur = g2j_to_affine(g2_mul_u256(cur, tau));
  }
}

// ---- GT / pairing --------------------------------------------------------------------------------------
void bsc_pairing(const uint8_t* g1_64, const uint8_t* g2_129, uint8_t* out384) {
  ensure_init();
  f12_to_bytes(out384, final_exp(miller(g2a_from_bytes(g2_129), g1a_from_bytes(g1_64))));
}
// 1 iff prod_i e(P_i, Q_i) == 1; one final exponentiation for the whole product
int bsc_pairing_check(const uint8_t* g1s, const uint8_t* g2s, int n) {
  ensure_init();
  Fp12 f = f12_one();
  for (int i = 0; i < n; i++) f = f12_mul(f, miller(g2a_from_bytes(g2s + 129 * (size_t)i), g1a_from_bytes(g1s + 64 * (size_t)i)));
  return f12_is_one(final_exp(f)) ? 1 : 0;
}
void bsc_gt_mul(uint8_t* out384, const uint8_t* a384, const uint8_t* b384) {
  f12_to_bytes(out384, f12_mul(f12_from_bytes(a384), f12_from_bytes(b384)));
}
void bsc_gt_pow(uint8_t* out384, const uint8_t* a384, const uint8_t* k32) {
  f12_to_bytes(out384, f12_pow_u256(f12_from_bytes(a384), u256_from_be(k32)));
}

// verifySecret (DistSys/kyber.go:650-673): e(C, g2) == e(W, tau*g2 - x*g2) * e(B, g2)^y, checked as
// e(C - y*B, g2) * e(-W, tau*g2 - x*g2) == 1.  B is the first SRS point of the chunk's slice (g1 for chunk 0,
// which is the only case the reference's formula covers; chunk c is committed against PKG1[10c..], kyber.go:172-202).
int bsc_kzg_verify_share(const uint8_t* commit64, const uint8_t* witness64, const uint8_t* base64,
                         const uint8_t* g2_129, const uint8_t* tau_g2_129, int x, long long y) {
  ensure_init();
  const G1Affine C = g1a_from_bytes(commit64), W = g1a_from_bytes(witness64), B = g1a_from_bytes(base64);
  if (!g1a_on_curve(C) || !g1a_on_curve(W)) return 0;
  const U256 ys = fe_to_u256(fe_neg(fe_from_i64<BnN>((int64_t)y)));
  const G1Affine lhs = g1j_to_affine(g1j_add_affine(g1_mul_u256(B, ys), C));           // C - y*B
  const U256 xs = fe_to_u256(fe_neg(fe_from_i64<BnN>((int64_t)x)));
  const G2Affine g2 = g2a_from_bytes(g2_129);
  const G2Affine rhs2 = g2j_to_affine(g2j_add(g2_mul_u256(g2, xs), g2j_from_affine(g2a_from_bytes(tau_g2_129))));
  const Fp12 f = f12_mul(miller(g2, lhs), miller(rhs2, g1a_neg(W)));
  return f12_is_one(final_exp(f)) ? 1 : 0;
}

}  // extern "C"
