// Host-side crypto runtime: BN256 G1 byte API, SRS generation, BLAKE2b / BLAKE2Xb, SHAKE256,
// Schnorr signatures.  Everything here shares field.cuh / g1.cuh with the device kernels, so the
// CPU results are the bit-exact oracle for the CUDA path.
//
// Reference parity: DistSys/publicKey.go:28-106 (GenerateKey / GenerateClientKey),
// DistSys/kyber.go:873-933 (SchnorrSign / SchnorrVerify / hashSchnorr),
// lib/dedis/kyber/xof/blake2xb/blake.go:18-85, lib/dedis/kyber/util/random/rand.go (scalar Pick by
// rejection sampling of 256-bit draws), lib/dedis/kyber/pairing/bn256/point.go:109-167 (wire format).
#include <string.h>
#include <stdlib.h>
#include <vector>
#include "g1.cuh"
#include "sha256.cuh"

using namespace bsc;

// ------------------------------------------------------------------------------------------------
// BLAKE2b
// ------------------------------------------------------------------------------------------------
namespace {

const uint64_t B2_IV[8] = {0x6a09e667f3bcc908ULL, 0xbb67ae8584caa73bULL, 0x3c6ef372fe94f82bULL, 0xa54ff53a5f1d36f1ULL,
                           0x510e527fade682d1ULL, 0x9b05688c2b3e6c1fULL, 0x1f83d9abfb41bd6bULL, 0x5be0cd19137e2179ULL};
const uint8_t B2_SIGMA[12][16] = {
    {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15}, {14, 10, 4, 8, 9, 15, 13, 6, 1, 12, 0, 2, 11, 7, 5, 3},
    {11, 8, 12, 0, 5, 2, 15, 13, 10, 14, 3, 6, 7, 1, 9, 4}, {7, 9, 3, 1, 13, 12, 11, 14, 2, 6, 5, 10, 4, 0, 15, 8},
    {9, 0, 5, 7, 2, 4, 10, 15, 14, 1, 11, 12, 6, 8, 3, 13}, {2, 12, 6, 10, 0, 11, 8, 3, 4, 13, 7, 5, 15, 14, 1, 9},
    {12, 5, 1, 15, 14, 13, 4, 10, 0, 7, 6, 3, 9, 2, 8, 11}, {13, 11, 7, 14, 12, 1, 3, 9, 5, 0, 15, 4, 8, 6, 2, 10},
    {6, 15, 14, 9, 11, 3, 0, 8, 12, 2, 13, 7, 1, 4, 10, 5}, {10, 2, 8, 4, 7, 6, 1, 5, 15, 11, 9, 14, 3, 12, 13, 0},
    {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15}, {14, 10, 4, 8, 9, 15, 13, 6, 1, 12, 0, 2, 11, 7, 5, 3}};

inline uint64_t rotr64(uint64_t x, int n) { return (x >> n) | (x << (64 - n)); }
inline uint64_t load64le(const uint8_t* p) {
  uint64_t v = 0;
  for (int i = 7; i >= 0; i--) v = (v << 8) | p[i];
  return v;
}

struct Blake2b {
  uint64_t h[8];
  uint64_t t0, t1;
  uint8_t buf[128];
  size_t fill;
  size_t outlen;
};

void b2_compress(Blake2b& s, const uint8_t* block, bool last) {
  uint64_t m[16], v[16];
  for (int i = 0; i < 16; i++) m[i] = load64le(block + 8 * i);
  for (int i = 0; i < 8; i++) {
    v[i] = s.h[i];
    v[i + 8] = B2_IV[i];
  }
  v[12] ^= s.t0;
  v[13] ^= s.t1;
  if (last) v[14] = ~v[14];
#define B2G(a, b, c, d, x, y)        \
  v[a] = v[a] + v[b] + (x);          \
  v[d] = rotr64(v[d] ^ v[a], 32);    \
  v[c] = v[c] + v[d];                \
  v[b] = rotr64(v[b] ^ v[c], 24);    \
  v[a] = v[a] + v[b] + (y);          \
  v[d] = rotr64(v[d] ^ v[a], 16);    \
  v[c] = v[c] + v[d];                \
  v[b] = rotr64(v[b] ^ v[c], 63);
#pragma GCC unroll 12
  for (int r = 0; r < 12; r++) {   // unrolled: the message schedule becomes compile-time register picks
    const uint8_t* sg = B2_SIGMA[r];
    B2G(0, 4, 8, 12, m[sg[0]], m[sg[1]]);
    B2G(1, 5, 9, 13, m[sg[2]], m[sg[3]]);
    B2G(2, 6, 10, 14, m[sg[4]], m[sg[5]]);
    B2G(3, 7, 11, 15, m[sg[6]], m[sg[7]]);
    B2G(0, 5, 10, 15, m[sg[8]], m[sg[9]]);
    B2G(1, 6, 11, 12, m[sg[10]], m[sg[11]]);
    B2G(2, 7, 8, 13, m[sg[12]], m[sg[13]]);
    B2G(3, 4, 9, 14, m[sg[14]], m[sg[15]]);
  }
#undef B2G
  for (int i = 0; i < 8; i++) s.h[i] ^= v[i] ^ v[i + 8];
}

// cfg = 64-byte parameter block
void b2_init_cfg(Blake2b& s, const uint8_t cfg[64]) {
  for (int i = 0; i < 8; i++) s.h[i] = B2_IV[i] ^ load64le(cfg + 8 * i);
  s.t0 = s.t1 = 0;
  s.fill = 0;
  s.outlen = cfg[0];
  memset(s.buf, 0, sizeof(s.buf));
}
void b2_update(Blake2b& s, const uint8_t* in, size_t n) {
  while (n) {
    if (s.fill == 128) {  // only compress when more data follows (last block is handled in final)
      s.t0 += 128;
      if (s.t0 < 128) s.t1++;
      b2_compress(s, s.buf, false);
      s.fill = 0;
    }
    size_t take = 128 - s.fill;
    if (take > n) take = n;
    memcpy(s.buf + s.fill, in, take);
    s.fill += take;
    in += take;
    n -= take;
  }
}
void b2_final(Blake2b& s, uint8_t* out) {
  s.t0 += s.fill;
  if (s.t0 < s.fill) s.t1++;
  memset(s.buf + s.fill, 0, 128 - s.fill);
  b2_compress(s, s.buf, true);
  uint8_t full[64];
  for (int i = 0; i < 8; i++)
    for (int j = 0; j < 8; j++) full[8 * i + j] = (uint8_t)(s.h[i] >> (8 * j));
  memcpy(out, full, s.outlen);
}
void b2_init_keyed(Blake2b& s, size_t outlen, const uint8_t* key, size_t keylen, uint32_t xof_len) {
  uint8_t cfg[64];
  memset(cfg, 0, 64);
  cfg[0] = (uint8_t)outlen;
  cfg[1] = (uint8_t)keylen;
  cfg[2] = 1;
  cfg[3] = 1;
  cfg[12] = (uint8_t)xof_len;
  cfg[13] = (uint8_t)(xof_len >> 8);
  cfg[14] = (uint8_t)(xof_len >> 16);
  cfg[15] = (uint8_t)(xof_len >> 24);
  b2_init_cfg(s, cfg);
  if (keylen) {
    uint8_t kb[128];
    memset(kb, 0, 128);
    memcpy(kb, key, keylen);
    b2_update(s, kb, 128);
  }
}

// BLAKE2Xb with unknown output length, keyed (golang.org/x/crypto/blake2b NewXOF semantics).
struct Blake2xb {
  Blake2b root;
  uint8_t h0[64];
  uint8_t block[64];
  uint32_t node;
  size_t used;
  bool reading;
};
void b2x_init(Blake2xb& x, const uint8_t* seed, size_t seedlen) {
  uint8_t k[64];
  if (seedlen > 64) {
    Blake2b t;
    b2_init_keyed(t, 64, nullptr, 0, 0);
    b2_update(t, seed, seedlen);
    b2_final(t, k);
    seed = k;
    seedlen = 64;
  }
  b2_init_keyed(x.root, 64, seed, seedlen, 0xFFFFFFFFu);
  x.node = 0;
  x.used = 64;
  x.reading = false;
}
void b2x_write(Blake2xb& x, const uint8_t* m, size_t n) { b2_update(x.root, m, n); }
void b2x_read(Blake2xb& x, uint8_t* out, size_t n) {
  if (!x.reading) {
    b2_final(x.root, x.h0);
    x.reading = true;
  }
  while (n) {
    if (x.used == 64) {
      uint8_t cfg[64];
      memset(cfg, 0, 64);
      cfg[0] = 64;
      cfg[4] = 64;
      cfg[8] = (uint8_t)x.node;
      cfg[9] = (uint8_t)(x.node >> 8);
      cfg[10] = (uint8_t)(x.node >> 16);
      cfg[11] = (uint8_t)(x.node >> 24);
      cfg[12] = cfg[13] = cfg[14] = cfg[15] = 0xFF;
      cfg[17] = 64;
      Blake2b b;
      b2_init_cfg(b, cfg);
      b2_update(b, x.h0, 64);
      b2_final(b, x.block);
      x.node++;
      x.used = 0;
    }
    size_t take = 64 - x.used;
    if (take > n) take = n;
    memcpy(out, x.block + x.used, take);
    x.used += take;
    out += take;
    n -= take;
  }
}

// ---- Keccak / SHAKE256 ---------------------------------------------------------------------------
const uint64_t KECCAK_RC[24] = {0x0000000000000001ULL, 0x0000000000008082ULL, 0x800000000000808aULL, 0x8000000080008000ULL,
                                0x000000000000808bULL, 0x0000000080000001ULL, 0x8000000080008081ULL, 0x8000000000008009ULL,
                                0x000000000000008aULL, 0x0000000000000088ULL, 0x0000000080008009ULL, 0x000000008000000aULL,
                                0x000000008000808bULL, 0x800000000000008bULL, 0x8000000000008089ULL, 0x8000000000008003ULL,
                                0x8000000000008002ULL, 0x8000000000000080ULL, 0x000000000000800aULL, 0x800000008000000aULL,
                                0x8000000080008081ULL, 0x8000000000008080ULL, 0x0000000080000001ULL, 0x8000000080008008ULL};
const int KECCAK_ROT[24] = {1, 3, 6, 10, 15, 21, 28, 36, 45, 55, 2, 14, 27, 41, 56, 8, 25, 43, 62, 18, 39, 61, 20, 44};
const int KECCAK_PIL[24] = {10, 7, 11, 17, 18, 3, 5, 16, 8, 21, 24, 4, 15, 23, 19, 13, 12, 2, 20, 14, 22, 9, 6, 1};
inline uint64_t rotl64(uint64_t x, int n) { return (x << n) | (x >> (64 - n)); }
void keccak_f(uint64_t st[25]) {
  for (int r = 0; r < 24; r++) {
    uint64_t bc[5];
    for (int i = 0; i < 5; i++) bc[i] = st[i] ^ st[i + 5] ^ st[i + 10] ^ st[i + 15] ^ st[i + 20];
    for (int i = 0; i < 5; i++) {
      uint64_t t = bc[(i + 4) % 5] ^ rotl64(bc[(i + 1) % 5], 1);
      for (int j = 0; j < 25; j += 5) st[j + i] ^= t;
    }
    uint64_t t = st[1];
    for (int i = 0; i < 24; i++) {
      int j = KECCAK_PIL[i];
      uint64_t b = st[j];
      st[j] = rotl64(t, KECCAK_ROT[i]);
      t = b;
    }
    for (int j = 0; j < 25; j += 5) {
      for (int i = 0; i < 5; i++) bc[i] = st[j + i];
      for (int i = 0; i < 5; i++) st[j + i] ^= (~bc[(i + 1) % 5]) & bc[(i + 2) % 5];
    }
    st[0] ^= KECCAK_RC[r];
  }
}

}  // namespace

struct Shake256 {
  uint64_t st[25];
  size_t pos;
  bool squeezing;
};
static const size_t SHAKE256_RATE = 136;
void shake256_init(Shake256& s) {
  memset(&s, 0, sizeof(s));
}
void shake256_absorb(Shake256& s, const uint8_t* in, size_t n) {
  for (size_t i = 0; i < n; i++) {
    s.st[s.pos >> 3] ^= (uint64_t)in[i] << (8 * (s.pos & 7));
    if (++s.pos == SHAKE256_RATE) {
      keccak_f(s.st);
      s.pos = 0;
    }
  }
}
void shake256_squeeze(Shake256& s, uint8_t* out, size_t n) {
  if (!s.squeezing) {
    s.st[s.pos >> 3] ^= (uint64_t)0x1F << (8 * (s.pos & 7));
    s.st[(SHAKE256_RATE - 1) >> 3] ^= (uint64_t)0x80 << (8 * ((SHAKE256_RATE - 1) & 7));
    keccak_f(s.st);
    s.pos = 0;
    s.squeezing = true;
  }
  for (size_t i = 0; i < n; i++) {
    if (s.pos == SHAKE256_RATE) {
      keccak_f(s.st);
      s.pos = 0;
    }
    out[i] = (uint8_t)(s.st[s.pos >> 3] >> (8 * (s.pos & 7)));
    s.pos++;
  }
}

// ------------------------------------------------------------------------------------------------
// helpers
// ------------------------------------------------------------------------------------------------
static bool u256_lt(const U256& a, const U256& b) {
  for (int i = 7; i >= 0; i--) {
    if (a.v[i] != b.v[i]) return a.v[i] < b.v[i];
  }
  return false;
}

namespace bsc {
void sha256_fast(const uint8_t* data, size_t n, uint8_t out[32]);   // sha256_ni.cpp
}

// hashSchnorr (kyber.go:928-933): XOF keyed by the point bytes, absorb the message, Pick a scalar.
static Fr hash_schnorr(const uint8_t* msg, size_t n, const uint8_t T64[64]) {
  Blake2xb x;
  b2x_init(x, T64, 64);
  b2x_write(x, msg, n);
  const U256 order = mod_of<BnN>();
  for (;;) {
    uint8_t b[32];
    b2x_read(x, b, 32);
    U256 c = u256_from_be(b);
    if (!u256_is_zero(c) && u256_lt(c, order)) return fe_from_u256<BnN>(c);
  }
}

extern "C" {

void bsc_sha256(const uint8_t* data, size_t n, uint8_t out[32]) { sha256_oneshot(data, n, out); }

void bsc_blake2b(const uint8_t* data, size_t n, const uint8_t* key, size_t keylen, uint8_t* out, size_t outlen) {
  Blake2b s;
  b2_init_keyed(s, outlen, key, keylen, 0);
  b2_update(s, data, n);
  b2_final(s, out);
}
void bsc_blake2xb(const uint8_t* seed, size_t seedlen, const uint8_t* msg, size_t n, uint8_t* out, size_t outlen) {
  Blake2xb x;
  b2x_init(x, seed, seedlen);
  b2x_write(x, msg, n);
  b2x_read(x, out, outlen);
}
void bsc_shake256(const uint8_t* data, size_t n, uint8_t* out, size_t outlen) {
  Shake256 s;
  shake256_init(s);
  shake256_absorb(s, data, n);
  shake256_squeeze(s, out, outlen);
}

// ---- G1 byte API ---------------------------------------------------------------------------------
void bsc_g1_generator(uint8_t out[64]) { g1a_to_bytes(out, g1_generator()); }
int bsc_g1_on_curve(const uint8_t in[64]) { return g1a_on_curve(g1a_from_bytes(in)) ? 1 : 0; }
void bsc_g1_add(uint8_t out[64], const uint8_t a[64], const uint8_t b[64]) {
  G1Jac r = g1j_add_affine(g1j_from_affine(g1a_from_bytes(a)), g1a_from_bytes(b));
  g1a_to_bytes(out, g1j_to_affine(r));
}
void bsc_g1_neg(uint8_t out[64], const uint8_t a[64]) { g1a_to_bytes(out, g1a_neg(g1a_from_bytes(a))); }
// scalar: 32-byte big-endian, reduced mod the group order first (mod.Int semantics)
void bsc_g1_mul(uint8_t out[64], const uint8_t in[64], const uint8_t k_be[32]) {
  U256 k = fe_to_u256(fe_from_u256<BnN>(u256_from_be(k_be)));
  g1a_to_bytes(out, g1j_to_affine(g1_mul_u256(g1a_from_bytes(in), k)));
}
// naive signed-int64 MSM oracle, one double-and-add per term like createCommitment (kyber.go:533-562)
void bsc_g1_msm_i64_host(uint8_t out[64], const uint8_t* bases64, const long long* scalars, int n) {
  G1Jac acc = g1j_inf();
  for (int i = 0; i < n; i++) {
    G1Affine b = g1a_from_bytes(bases64 + 64 * (size_t)i);
    long long s = scalars[i];
    uint64_t mag = s < 0 ? (uint64_t)0 - (uint64_t)s : (uint64_t)s;
    if (s < 0) b = g1a_neg(b);
    U256 k;
    for (int j = 0; j < 8; j++) k.v[j] = 0;
    k.v[0] = (uint32_t)mag;
    k.v[1] = (uint32_t)(mag >> 32);
    acc = g1j_add(acc, g1_mul_u256(b, k));
  }
  g1a_to_bytes(out, g1j_to_affine(acc));
}
// wire bytes <-> device layout (Montgomery affine, 16 x uint32)
void bsc_g1_bytes_to_affine_host(const uint8_t* in, void* out_affine, long long N) {
  G1Affine* o = (G1Affine*)out_affine;
  for (long long i = 0; i < N; i++) o[i] = g1a_from_bytes(in + 64 * i);
}
void bsc_g1_affine_to_bytes_host(const void* in_affine, uint8_t* out, long long N) {
  const G1Affine* a = (const G1Affine*)in_affine;
  for (long long i = 0; i < N; i++) g1a_to_bytes(out + 64 * i, a[i]);
}
void bsc_g1_jac_to_bytes_host(const void* in_jac, uint8_t* out, long long N) {
  const G1Jac* a = (const G1Jac*)in_jac;
  for (long long i = 0; i < N; i++) g1a_to_bytes(out + 64 * i, g1j_to_affine(a[i]));
}
// Jacobian -> 64-byte wire form for the selected rows, with one shared field inversion (Montgomery's trick).
// Used at block assembly: the device hands over the contributors' full commitments in Jacobian form and
// no inversion kernel sits on the block's critical path.  idx may be null (rows 0..n-1).
void bsc_g1_jac_to_bytes_batch_host(const void* in_jac, const int* idx, int n, uint8_t* out) {
  const G1Jac* a = (const G1Jac*)in_jac;
  std::vector<Fp> pre((size_t)n);
  Fp run = fe_one<BnP>();
  for (int i = 0; i < n; i++) {
    const G1Jac& p = a[idx ? idx[i] : i];
    if (!g1j_is_inf(p)) run = fe_mul(run, p.z);
    pre[(size_t)i] = run;
  }
  Fp inv = fe_inv(run);
  for (int i = n - 1; i >= 0; i--) {
    const G1Jac& p = a[idx ? idx[i] : i];
    if (g1j_is_inf(p)) {
      memset(out + 64 * (size_t)i, 0, 64);
      continue;
    }
    const Fp zi = i > 0 ? fe_mul(inv, pre[(size_t)i - 1]) : inv;
    inv = fe_mul(inv, p.z);
    g1a_to_bytes(out + 64 * (size_t)i, g1j_to_affine_with_inv(p, zi));
  }
}
// host mirror of the device Straus kernel (same template) for CPU-side unit tests
void bsc_g1_straus_host(uint8_t out[64], const void* bases_affine, const long long* s, int T) {
  uint64_t maxmag = 0;
  for (int t = 0; t < T; t++) maxmag |= s[t] < 0 ? (uint64_t)0 - (uint64_t)s[t] : (uint64_t)s[t];
  int nbits = 0;
  while (nbits < 64 && (maxmag >> nbits)) nbits++;
  G1Jac r = g1_straus_i64<16>((const G1Affine*)bases_affine, (const int64_t*)s, T, nbits);
  g1a_to_bytes(out, g1j_to_affine(r));
}

// SRS: PKG1[i] = tau^i * G for i < d (DistSys/publicKey.go:28-66).  tau = 2 reproduces the reference's
// hard-coded trapdoor (:36); any other 32-byte big-endian tau gives a proper setup.
void bsc_srs_g1_host(const uint8_t tau_be[32], int d, uint8_t* out) {
  const Fr tau = fe_from_u256<BnN>(u256_from_be(tau_be));
  Fr pw = fe_one<BnN>();
  const G1Affine g = g1_generator();
  const Fr two = fe_from_u64<BnN>(2);
  const bool is_two = fe_eq(tau, two);
  G1Jac run = g1j_from_affine(g);
  for (int i = 0; i < d; i++) {
    if (is_two) {
      g1a_to_bytes(out + 64 * (size_t)i, g1j_to_affine(run));
      run = g1j_dbl(run);
    } else {
      g1a_to_bytes(out + 64 * (size_t)i, g1j_to_affine(g1_mul_u256(g, fe_to_u256(pw))));
      pw = fe_mul(pw, tau);
    }
  }
}
// tau^i as plain little-endian 8x32 words, for the device fixed-base kernel
void bsc_fr_powers(const uint8_t tau_be[32], int d, uint32_t* out_words) {
  const Fr tau = fe_from_u256<BnN>(u256_from_be(tau_be));
  Fr pw = fe_one<BnN>();
  for (int i = 0; i < d; i++) {
    U256 k = fe_to_u256(pw);
    memcpy(out_words + 8 * (size_t)i, k.v, 32);
    pw = fe_mul(pw, tau);
  }
}

// Test hook for the two inversion routines of field.cuh: a^-1 mod m by binary extended Euclid (fe_inv) and by the
// reference's Fermat exponentiation (fe_inv_fermat).  which: 0 = BN256 base field, 1 = BN256 group order,
// 2 = 2^255-19.  Values are 32-byte big-endian plain integers (reduced mod m on input).
void bsc_fe_inv_both(int which, const uint8_t a_be[32], uint8_t out_fast[32], uint8_t out_fermat[32]) {
  const U256 a = u256_from_be(a_be);
  if (which == 0) {
    const Fp x = fe_from_u256<BnP>(a);
    u256_to_be(out_fast, fe_to_u256(fe_inv(x)));
    u256_to_be(out_fermat, fe_to_u256(fe_inv_fermat(x)));
  } else if (which == 1) {
    const Fr x = fe_from_u256<BnN>(a);
    u256_to_be(out_fast, fe_to_u256(fe_inv(x)));
    u256_to_be(out_fermat, fe_to_u256(fe_inv_fermat(x)));
  } else {
    const Fe<EdP> x = fe_from_u256<EdP>(a);
    u256_to_be(out_fast, fe_to_u256(fe_inv(x)));
    u256_to_be(out_fermat, fe_to_u256(fe_inv_fermat(x)));
  }
}

// ---- Schnorr (kyber.go:873-933) ------------------------------------------------------------------
// pub = x*G
void bsc_schnorr_pubkey(const uint8_t x_be[32], uint8_t pub[64]) {
  U256 x = fe_to_u256(fe_from_u256<BnN>(u256_from_be(x_be)));
  g1a_to_bytes(pub, g1j_to_affine(g1_mul_u256(g1_generator(), x)));
}
// sig = c || r, c = H(T || msg), r = v - x*c ; v supplied by the caller (32 random bytes)
void bsc_schnorr_sign(const uint8_t* msg, size_t n, const uint8_t x_be[32], const uint8_t v_be[32], uint8_t sig[64]) {
  const Fr x = fe_from_u256<BnN>(u256_from_be(x_be));
  const Fr v = fe_from_u256<BnN>(u256_from_be(v_be));
  uint8_t T[64];
  g1a_to_bytes(T, g1j_to_affine(g1_mul_u256(g1_generator(), fe_to_u256(v))));
  const Fr c = hash_schnorr(msg, n, T);
  const Fr r = fe_sub(v, fe_mul(x, c));
  u256_to_be(sig, fe_to_u256(c));
  u256_to_be(sig + 32, fe_to_u256(r));
}
// Batched signing split in two so that the scalar multiplications T_i = v_i*G can run on the GPU
// (g1_comb_mul_kernel): (1) derive nonces, (2) given the T_i, finish c_i = H(T_i || msg_i), r_i = v_i - x_i*c_i.
// Nonce: v = SHA-256("bsc-schnorr-nonce" || x || tag) mod order, tag = caller-chosen unique bytes.
void bsc_schnorr_nonces(int n, const uint8_t* tags, int taglen, const uint8_t* x_be, uint32_t* v_words_le,
                        uint8_t* v_be) {
  std::vector<uint8_t> buf(17 + 32 + (size_t)taglen);
  memcpy(buf.data(), "bsc-schnorr-nonce", 17);
  for (int i = 0; i < n; i++) {
    memcpy(buf.data() + 17, x_be + 32 * (size_t)i, 32);
    memcpy(buf.data() + 49, tags + (size_t)taglen * i, (size_t)taglen);
    uint8_t h[32];
    sha256_fast(buf.data(), buf.size(), h);      // SHA-NI when the CPU has it (sha256_ni.cpp)
    const U256 v = fe_to_u256(fe_from_u256<BnN>(u256_from_be(h)));
    memcpy(v_words_le + 8 * (size_t)i, v.v, 32);
    u256_to_be(v_be + 32 * (size_t)i, v);
  }
}
void bsc_schnorr_finish_batch(int n, const uint8_t* msgs, int msglen, const uint8_t* x_be, const uint8_t* v_be,
                              const uint8_t* T, uint8_t* sigs) {
  for (int i = 0; i < n; i++) {
    const Fr x = fe_from_u256<BnN>(u256_from_be(x_be + 32 * (size_t)i));
    const Fr v = fe_from_u256<BnN>(u256_from_be(v_be + 32 * (size_t)i));
    const Fr c = hash_schnorr(msgs + (size_t)msglen * i, (size_t)msglen, T + 64 * (size_t)i);
    const Fr r = fe_sub(v, fe_mul(x, c));
    u256_to_be(sigs + 64 * (size_t)i, fe_to_u256(c));
    u256_to_be(sigs + 64 * (size_t)i + 32, fe_to_u256(r));
  }
}
int bsc_schnorr_verify(const uint8_t* msg, size_t n, const uint8_t pub[64], const uint8_t sig[64]) {
  const U256 order = mod_of<BnN>();
  U256 cu = u256_from_be(sig), ru = u256_from_be(sig + 32);
  if (!u256_lt(cu, order) || !u256_lt(ru, order)) return 0;
  G1Jac T = g1j_add(g1_mul_u256(g1_generator(), ru), g1_mul_u256(g1a_from_bytes(pub), cu));
  uint8_t Tb[64];
  g1a_to_bytes(Tb, g1j_to_affine(T));
  const Fr c = hash_schnorr(msg, n, Tb);
  return u256_eq(fe_to_u256(c), cu) ? 1 : 0;
}

}  // extern "C"
