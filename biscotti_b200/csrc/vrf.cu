// ECVRF (Ed25519 / Elligator2 / SHAKE256) — host prove/verify plus the batched device kernels that
// evaluate every simulated peer's VRF in one launch, and a batched SHA-256 kernel.
//
// Reference: DistSys/vrf.go:16-52 (two coniks VRF keypairs per peer, Prove on the latest block
// hash, 32-byte output + 96-byte proof).  coniks-go itself is not in the reference tree; the
// construction implemented here is  H = 8*Elligator2(SHAKE256(m)),  Gamma = x*H,
// r = SHAKE256(skhr || pk || m) mod l,  s = SHAKE256(G || H || pk || Gamma || rG || rH || m) mod l,
// t = r - s*x,  proof = s || t || Gamma,  output = SHAKE256(Gamma || m)[:32].
#include "common.cuh"
#include "ed25519.cuh"
#include "ed25519_r25.cuh"
#include "sha256.cuh"
#include <string.h>
#include <vector>

// SHAKE256 lives in host_crypto.cpp
struct Shake256 {
  uint64_t st[25];
  size_t pos;
  bool squeezing;
};
void shake256_init(Shake256& s);
void shake256_absorb(Shake256& s, const uint8_t* in, size_t n);
void shake256_squeeze(Shake256& s, uint8_t* out, size_t n);

namespace bsc {

// ---- SHA-512 (host only; Ed25519 secret expansion) ------------------------------------------------
static const uint64_t K512[80] = {
    0x428a2f98d728ae22ULL, 0x7137449123ef65cdULL, 0xb5c0fbcfec4d3b2fULL, 0xe9b5dba58189dbbcULL, 0x3956c25bf348b538ULL,
    0x59f111f1b605d019ULL, 0x923f82a4af194f9bULL, 0xab1c5ed5da6d8118ULL, 0xd807aa98a3030242ULL, 0x12835b0145706fbeULL,
    0x243185be4ee4b28cULL, 0x550c7dc3d5ffb4e2ULL, 0x72be5d74f27b896fULL, 0x80deb1fe3b1696b1ULL, 0x9bdc06a725c71235ULL,
    0xc19bf174cf692694ULL, 0xe49b69c19ef14ad2ULL, 0xefbe4786384f25e3ULL, 0x0fc19dc68b8cd5b5ULL, 0x240ca1cc77ac9c65ULL,
    0x2de92c6f592b0275ULL, 0x4a7484aa6ea6e483ULL, 0x5cb0a9dcbd41fbd4ULL, 0x76f988da831153b5ULL, 0x983e5152ee66dfabULL,
    0xa831c66d2db43210ULL, 0xb00327c898fb213fULL, 0xbf597fc7beef0ee4ULL, 0xc6e00bf33da88fc2ULL, 0xd5a79147930aa725ULL,
    0x06ca6351e003826fULL, 0x142929670a0e6e70ULL, 0x27b70a8546d22ffcULL, 0x2e1b21385c26c926ULL, 0x4d2c6dfc5ac42aedULL,
    0x53380d139d95b3dfULL, 0x650a73548baf63deULL, 0x766a0abb3c77b2a8ULL, 0x81c2c92e47edaee6ULL, 0x92722c851482353bULL,
    0xa2bfe8a14cf10364ULL, 0xa81a664bbc423001ULL, 0xc24b8b70d0f89791ULL, 0xc76c51a30654be30ULL, 0xd192e819d6ef5218ULL,
    0xd69906245565a910ULL, 0xf40e35855771202aULL, 0x106aa07032bbd1b8ULL, 0x19a4c116b8d2d0c8ULL, 0x1e376c085141ab53ULL,
    0x2748774cdf8eeb99ULL, 0x34b0bcb5e19b48a8ULL, 0x391c0cb3c5c95a63ULL, 0x4ed8aa4ae3418acbULL, 0x5b9cca4f7763e373ULL,
    0x682e6ff3d6b2b8a3ULL, 0x748f82ee5defb2fcULL, 0x78a5636f43172f60ULL, 0x84c87814a1f0ab72ULL, 0x8cc702081a6439ecULL,
    0x90befffa23631e28ULL, 0xa4506cebde82bde9ULL, 0xbef9a3f7b2c67915ULL, 0xc67178f2e372532bULL, 0xca273eceea26619cULL,
    0xd186b8c721c0c207ULL, 0xeada7dd6cde0eb1eULL, 0xf57d4f7fee6ed178ULL, 0x06f067aa72176fbaULL, 0x0a637dc5a2c898a6ULL,
    0x113f9804bef90daeULL, 0x1b710b35131c471bULL, 0x28db77f523047d84ULL, 0x32caab7b40c72493ULL, 0x3c9ebe0a15c9bebcULL,
    0x431d67c49c100d4cULL, 0x4cc5d4becb3e42b6ULL, 0x597f299cfc657e2aULL, 0x5fcb6fab3ad6faecULL, 0x6c44198c4a475817ULL};
static inline uint64_t r64(uint64_t x, int n) { return (x >> n) | (x << (64 - n)); }
static void sha512(const uint8_t* data, size_t n, uint8_t out[64]) {
  uint64_t h[8] = {0x6a09e667f3bcc908ULL, 0xbb67ae8584caa73bULL, 0x3c6ef372fe94f82bULL, 0xa54ff53a5f1d36f1ULL,
                   0x510e527fade682d1ULL, 0x9b05688c2b3e6c1fULL, 0x1f83d9abfb41bd6bULL, 0x5be0cd19137e2179ULL};
  size_t total = n + 1 + 16;
  size_t padded = (total + 127) / 128 * 128;
  uint8_t* buf = new uint8_t[padded];
  memset(buf, 0, padded);
  memcpy(buf, data, n);
  buf[n] = 0x80;
  uint64_t bits = (uint64_t)n * 8;
  for (int i = 0; i < 8; i++) buf[padded - 1 - i] = (uint8_t)(bits >> (8 * i));
  for (size_t off = 0; off < padded; off += 128) {
    uint64_t w[80];
    for (int i = 0; i < 16; i++) {
      uint64_t v = 0;
      for (int j = 0; j < 8; j++) v = (v << 8) | buf[off + 8 * i + j];
      w[i] = v;
    }
    for (int i = 16; i < 80; i++) {
      uint64_t s0 = r64(w[i - 15], 1) ^ r64(w[i - 15], 8) ^ (w[i - 15] >> 7);
      uint64_t s1 = r64(w[i - 2], 19) ^ r64(w[i - 2], 61) ^ (w[i - 2] >> 6);
      w[i] = w[i - 16] + s0 + w[i - 7] + s1;
    }
    uint64_t a = h[0], b = h[1], c = h[2], d = h[3], e = h[4], f = h[5], g = h[6], hh = h[7];
    for (int i = 0; i < 80; i++) {
      uint64_t S1 = r64(e, 14) ^ r64(e, 18) ^ r64(e, 41);
      uint64_t ch = (e & f) ^ (~e & g);
      uint64_t t1 = hh + S1 + ch + K512[i] + w[i];
      uint64_t S0 = r64(a, 28) ^ r64(a, 34) ^ r64(a, 39);
      uint64_t mj = (a & b) ^ (a & c) ^ (b & c);
      uint64_t t2 = S0 + mj;
      hh = g; g = f; f = e; e = d + t1; d = c; c = b; b = a; a = t1 + t2;
    }
    h[0] += a; h[1] += b; h[2] += c; h[3] += d; h[4] += e; h[5] += f; h[6] += g; h[7] += hh;
  }
  delete[] buf;
  for (int i = 0; i < 8; i++)
    for (int j = 0; j < 8; j++) out[8 * i + j] = (uint8_t)(h[i] >> (56 - 8 * j));
}

// 64 little-endian bytes mod l -> Montgomery scalar
static Fs sc_reduce64(const uint8_t b[64]) {
  Fs lo = fe_from_u256<EdL>(u256_from_le(b));
  Fs hi = fe_from_u256<EdL>(u256_from_le(b + 32));
  Fs r2;
  for (int i = 0; i < 8; i++) r2.x.v[i] = EdL::r2(i);  // Montgomery form of 2^256 mod l
  return fe_add(lo, fe_mul(hi, r2));
}

static void expand_secret(const uint8_t seed[32], U256& x, uint8_t skhr[32]) {
  uint8_t h[64];
  sha512(seed, 32, h);
  h[0] &= 248;
  h[31] &= 127;
  h[31] |= 64;
  x = u256_from_le(h);
  memcpy(skhr, h + 32, 32);
}

static EdPoint hash_to_curve(const uint8_t* m, size_t n) {
  uint8_t hm[32];
  Shake256 s;
  shake256_init(s);
  shake256_absorb(s, m, n);
  shake256_squeeze(s, hm, 32);
  return ed_elligator2_x8(hm);
}

// ---- device kernels ------------------------------------------------------------------------------
// out[i] = compress(k[i] * base[b(i)]) for 256-bit scalars; base index = i % nbase (nbase = 1: the
// shared H of a block; nbase = N: per-element bases)
__global__ void ed25519_scalarmult_kernel(const U256* __restrict__ k, const EdPoint* __restrict__ bases, int nbase,
                                          int N, unsigned char* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  EdPoint r = ed_mul(bases[i % nbase], k[i]);
  ed_compress(out + 32 * (size_t)i, r);
}

// Latency-optimised form: radix-2^25.5 field, signed radix-16 windows over a shared table of 1H..8H
// built once per CTA.  scalars: N x 32 little-endian bytes; hxy: affine H as 64 little-endian bytes.
__global__ void __launch_bounds__(32) ed25519_scalarmult_r25_kernel(const unsigned char* __restrict__ scalars,
                                                                    const unsigned char* __restrict__ hxy, int N,
                                                                    unsigned char* __restrict__ out) {
  __shared__ EdCached table[8];
  if (threadIdx.x == 0) {
    uint8_t b[64];
    for (int i = 0; i < 64; i++) b[i] = hxy[i];
    edr_build_table(table, edr_from_affine_bytes(b), f25_const_2d());
  }
  __syncthreads();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  uint8_t a[32];
  for (int k = 0; k < 32; k++) a[k] = scalars[32 * (size_t)i + k];
  edr_compress(out + 32 * (size_t)i, edr_scalarmult(a, table));
}

// Quad-parallel form: four lanes cooperate on one scalar multiplication.  Every point operation of the
// window method is two rounds of four independent field multiplications (doubling: XX, YY, 2ZZ, (X+Y)^2
// then the four completed->extended products; addition: A, B, C, ZZ then the same four products), so a
// quad runs each round in the time of one multiplication and exchanges the results with width-4
// shuffles.  The sequential depth drops from ~2400 multiplications to ~630 rounds — this is what the
// per-block critical path sees, because the VRF of block t+1 cannot start before block t is hashed.
__device__ __forceinline__ F25 quad_pick(const F25& r, int src) {
  F25 o;
#pragma unroll
  for (int i = 0; i < 10; i++) o.v[i] = __shfl_sync(0xffffffffu, r.v[i], src, 4);
  return o;
}
__device__ __forceinline__ F25 f25_sel4(int j, const F25& a, const F25& b, const F25& c, const F25& d) {
  // branch-free: lanes of one quad take different operands, and a branchy select diverges inside every warp
  const int32_t m0 = -(int32_t)(j == 0), m1 = -(int32_t)(j == 1), m2 = -(int32_t)(j == 2), m3 = -(int32_t)(j == 3);
  F25 o;
#pragma unroll
  for (int i = 0; i < 10; i++) o.v[i] = (a.v[i] & m0) | (b.v[i] & m1) | (c.v[i] & m2) | (d.v[i] & m3);
  return o;
}

__global__ void __launch_bounds__(128) ed25519_scalarmult_quad_kernel(const unsigned char* __restrict__ scalars,
                                                                      const unsigned char* __restrict__ hxy, int N,
                                                                      unsigned char* __restrict__ out,
                                                                      int32_t* __restrict__ xyz_out) {
  __shared__ EdCached table[9];   // 1H..8H, then the identity (digit 0) so every quad runs the same steps
  if (threadIdx.x == 0) {
    uint8_t b[64];
    for (int i = 0; i < 64; i++) b[i] = hxy[i];
    edr_build_table(table, edr_from_affine_bytes(b), f25_const_2d());
    table[8].YpX = f25_one();
    table[8].YmX = f25_one();
    table[8].Z = f25_one();
    table[8].T2d = f25_zero();
  }
  __syncthreads();
  const int gid = blockIdx.x * blockDim.x + threadIdx.x;
  const int peer = gid >> 2, j = gid & 3;
  const int pi = peer < N ? peer : N - 1;      // idle quads mirror the last peer so shuffles stay full-warp
  // signed radix-16 recoding (every lane of the quad computes the same digits)
  signed char e[64];
  for (int i = 0; i < 32; i++) {
    const unsigned char a = scalars[32 * (size_t)pi + i];
    e[2 * i] = (signed char)(a & 15);
    e[2 * i + 1] = (signed char)((a >> 4) & 15);
  }
  int carry = 0;
  for (int i = 0; i < 63; i++) {
    e[i] = (signed char)(e[i] + carry);
    carry = (e[i] + 8) >> 4;
    e[i] = (signed char)(e[i] - (carry << 4));
  }
  e[63] = (signed char)(e[63] + carry);

  F25 X = f25_zero(), Y = f25_one(), Z = f25_one(), T = f25_zero();   // identity, replicated in the quad
#pragma unroll 1
  for (int i = 63; i >= 0; i--) {
    const int dgt = e[i];
    // steps 0..3 are doublings, step 4 is the table addition.  A zero digit adds the identity instead of
    // skipping the step: control flow must stay warp-uniform because the shuffles below are full-warp.
    const int nsteps = 5;
#pragma unroll 1
    for (int step = 0; step < nsteps; step++) {
      F25 a, b;
      const bool is_add = (step == 4);
      if (!is_add) {
        // lane 2 forms 2*Z*Z directly (operand Z+Z) so that every input of the second round stays inside
        // the radix-2^25.5 multiplication bounds, exactly like the reference formulas' sq2
        const F25 xy = f25_add(X, Y);
        a = f25_sel4(j, X, Y, Z, xy);
        b = f25_sel4(j, X, Y, f25_add(Z, Z), xy);
      } else {
        // Subtracting a table entry = adding its negation (swap Y+X / Y-X, negate 2dT); the choice is made
        // with masks, not branches: digits differ between the quads of a warp.
        const int32_t nm = -(int32_t)(dgt < 0);
        const int mag = (dgt ^ nm) - nm;
        const EdCached& q = table[mag == 0 ? 8 : mag - 1];
        F25 qp, qm, qt;
#pragma unroll
        for (int i = 0; i < 10; i++) {
          qp.v[i] = (q.YpX.v[i] & ~nm) | (q.YmX.v[i] & nm);
          qm.v[i] = (q.YmX.v[i] & ~nm) | (q.YpX.v[i] & nm);
          qt.v[i] = (q.T2d.v[i] ^ nm) - nm;
        }
        const F25 ypx = f25_add(Y, X), ymx = f25_sub(Y, X);
        a = f25_sel4(j, ypx, ymx, qt, Z);
        b = f25_sel4(j, qp, qm, T, q.Z);
      }
      const F25 r = f25_mul_inl<false>(a, b);   // two call sites only: inlined, no call / stack traffic
      const F25 r0 = quad_pick(r, 0), r1 = quad_pick(r, 1), r2 = quad_pick(r, 2), r3 = quad_pick(r, 3);
      F25 cX, cY, cZ, cT;   // completed point ((cX:cZ), (cY:cT))
      if (!is_add) {
        // r0 = XX, r1 = YY, r2 = 2ZZ, r3 = (X+Y)^2
        cY = f25_add(r1, r0);
        cZ = f25_sub(r1, r0);
        cX = f25_sub(r3, cY);
        cT = f25_sub(r2, cZ);
      } else {
        // r0 = A, r1 = B, r2 = C, r3 = ZZ
        const F25 D = f25_add(r3, r3);
        cX = f25_sub(r0, r1);
        cY = f25_add(r0, r1);
        cZ = f25_add(D, r2);
        cT = f25_sub(D, r2);
      }
      // second round: X3 = cX*cT, Y3 = cY*cZ, Z3 = cZ*cT, T3 = cX*cY
      const F25 a2 = f25_sel4(j, cX, cY, cZ, cX);
      const F25 b2 = f25_sel4(j, cT, cZ, cT, cY);
      const F25 s = f25_mul_inl<false>(a2, b2);
      X = quad_pick(s, 0);
      Y = quad_pick(s, 1);
      Z = quad_pick(s, 2);
      T = quad_pick(s, 3);
    }
  }
  if (peer < N && j == 0) {
    if (xyz_out) {
      // projective result; the host normalises all peers with one shared inversion (Montgomery's trick):
      // the ~265 sequential squarings of an in-kernel inversion would be a fifth of this kernel
      int32_t* o = xyz_out + 30 * (size_t)peer;
      for (int i = 0; i < 10; i++) {
        o[i] = X.v[i];
        o[10 + i] = Y.v[i];
        o[20 + i] = Z.v[i];
      }
    } else {
      EdP3 p;
      p.X = X; p.Y = Y; p.Z = Z; p.T = T;
      edr_compress(out + 32 * (size_t)peer, p);
    }
  }
}

// ------------------------------------------------------------------------------------------------------------
// Warp-wide variant: one peer per warp, ONE LIMB PER LANE.
//
// The VRF of every peer is evaluated on the previous block hash, so this kernel is the head of every block's
// dependency chain and there is nothing to overlap it with: what counts is the latency of one scalar multiplication.
// In the quad kernel a lane computes a whole 10x10-limb product (100 dependent-ish 64-bit multiply-adds, ~0.65 us);
// here lane (g, k) of a warp computes limb k of product g: ten multiply-adds, operands fetched with warp shuffles, and
// the carry chain is three parallel passes (each limb hands its carry to the next lane) instead of a ten-step
// sequential one.  Three products run side by side (lanes 0-9, 10-19, 20-29; lanes 30-31 mirror 20-21), and every lane
// keeps limb k of ALL coordinates of the running point (replicated across the three groups), so the additions and
// subtractions between products are plain per-lane integer adds.  A doubling is three such steps (XX YY 2ZZ | (X+Y)^2
// Y3 Z3 | X3 T3 +1 free slot), an addition of a table entry two (its Z*Zq product rides in the free slot of the
// doubling before it): 14 steps per radix-16 digit, ~900 steps per scalar instead of 640 full products.
// Bounds follow the radix-2^25.5 reference formulas: products take operands up to 1.65 * 2^26 / 2^25, the rounded
// three-pass carry returns limbs within 2^25 + 19 / 2^24 + 1.
// ------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void wide_mul3(int32_t a0, int32_t b0, int32_t a1, int32_t b1, int32_t a2, int32_t b2, int g,
                                          int k, int32_t& r0, int32_t& r1, int32_t& r2) {
  const unsigned FULL = 0xffffffffu;
  const int base = g * 10;
  const int32_t a = g == 0 ? a0 : (g == 1 ? a1 : a2);
  const int32_t b = g == 0 ? b0 : (g == 1 ? b1 : b2);
  int64_t acc = 0;
#pragma unroll
  for (int i = 0; i < 10; i++) {
    int32_t fi = __shfl_sync(FULL, a, base + i);
    int j = k - i;
    const bool wrap = j < 0;
    if (wrap) j += 10;
    int32_t gj = __shfl_sync(FULL, b, base + j);
    if (wrap) gj *= 19;
    if ((i & 1) && (j & 1)) fi *= 2;
    acc += (int64_t)fi * (int64_t)gj;
  }
  const int sft = (k & 1) ? 25 : 26;
  const int src = base + (k == 0 ? 9 : k - 1);
  int64_t c = (acc + ((int64_t)1 << (sft - 1))) >> sft;
  acc -= c << sft;
  int64_t cin = __shfl_sync(FULL, c, src);
  if (k == 0) cin *= 19;
  acc += cin;
  int32_t c2 = (int32_t)((acc + ((int64_t)1 << (sft - 1))) >> sft);
  acc -= (int64_t)c2 << sft;
  int32_t cin2 = __shfl_sync(FULL, c2, src);
  if (k == 0) cin2 *= 19;
  int32_t h = (int32_t)acc + cin2;
  const int32_t c3 = (h + (1 << (sft - 1))) >> sft;
  h -= c3 << sft;
  int32_t cin3 = __shfl_sync(FULL, c3, src);
  if (k == 0) cin3 *= 19;
  h += cin3;
  r0 = __shfl_sync(FULL, h, k);
  r1 = __shfl_sync(FULL, h, 10 + k);
  r2 = __shfl_sync(FULL, h, 20 + k);
}

constexpr int WIDE_WARPS = 4;
__global__ void __launch_bounds__(32 * WIDE_WARPS) ed25519_scalarmult_wide_kernel(const unsigned char* __restrict__ scalars,
                                                                                  const unsigned char* __restrict__ hxy, int N,
                                                                                  int32_t* __restrict__ xyz_out) {
  __shared__ int32_t tab[WIDE_WARPS][8][4][10];   // per warp: (j+1)H as (Y+X, Y-X, Z, 2dT), one limb per lane
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int peer = blockIdx.x * WIDE_WARPS + warp;
  const int pi = peer < N ? peer : N - 1;          // idle warps mirror the last peer (control flow stays uniform per warp)
  const int g = lane < 30 ? lane / 10 : 2;
  const int k = lane < 30 ? lane % 10 : lane - 30;
  // limb k of the constants and of H
  int32_t d2k, hxk, hyk, onek = (k == 0) ? 1 : 0;
  {
    uint8_t b[64];
    for (int i = 0; i < 64; i++) b[i] = hxy[i];
    const F25 hx = f25_frombytes(b), hy = f25_frombytes(b + 32), d2 = f25_const_2d();
    hxk = hx.v[k];
    hyk = hy.v[k];
    d2k = d2.v[k];
  }
  int32_t u0, u1, u2;
  // ---- table: P = H (extended), cached(P) = (Y+X, Y-X, Z, 2d T)
  int32_t X = hxk, Y = hyk, Z = onek, T;
  wide_mul3(X, Y, X, Y, X, Y, g, k, T, u1, u2);                 // T = x*y
  int32_t Td;
  wide_mul3(T, d2k, T, d2k, T, d2k, g, k, Td, u1, u2);          // 2d*T
  const int32_t hp = Y + X, hm = Y - X, hz = Z, ht = Td;        // cached H, in registers
  if (lane < 10) {
    tab[warp][0][0][k] = hp;
    tab[warp][0][1][k] = hm;
    tab[warp][0][2][k] = hz;
    tab[warp][0][3][k] = ht;
  }
#pragma unroll 1
  for (int j = 1; j < 8; j++) {
    // P <- P + H (full outputs), then cache it
    int32_t a, bq, c, dd;
    wide_mul3(Y + X, hp, Y - X, hm, T, ht, g, k, a, bq, c);
    wide_mul3(Z, hz, Z, hz, Z, hz, g, k, dd, u1, u2);
    const int32_t D = dd + dd, cX = a - bq, cY = a + bq, cZ = D + c, cT = D - c;
    wide_mul3(cX, cT, cY, cZ, cZ, cT, g, k, X, Y, Z);
    wide_mul3(cX, cY, cX, cY, cX, cY, g, k, T, u1, u2);
    wide_mul3(T, d2k, T, d2k, T, d2k, g, k, Td, u1, u2);
    if (lane < 10) {
      tab[warp][j][0][k] = Y + X;
      tab[warp][j][1][k] = Y - X;
      tab[warp][j][2][k] = Z;
      tab[warp][j][3][k] = Td;
    }
  }
  __syncwarp();
  // ---- signed radix-16 recoding (every lane computes the same digits)
  signed char e[64];
  for (int i = 0; i < 32; i++) {
    const unsigned char a = scalars[32 * (size_t)pi + i];
    e[2 * i] = (signed char)(a & 15);
    e[2 * i + 1] = (signed char)((a >> 4) & 15);
  }
  int carry = 0;
  for (int i = 0; i < 63; i++) {
    e[i] = (signed char)(e[i] + carry);
    carry = (e[i] + 8) >> 4;
    e[i] = (signed char)(e[i] - (carry << 4));
  }
  e[63] = (signed char)(e[63] + carry);
  // ---- main loop
  X = 0;
  Y = onek;
  Z = onek;
  T = 0;
#pragma unroll 1
  for (int i = 63; i >= 0; i--) {
    const int dgt = e[i];
    const int mag = dgt < 0 ? -dgt : dgt;
    int32_t qp = 0, qm = 0, qz = 0, qt = 0;
    if (mag) {   // uniform per warp: one peer per warp
      const int32_t tp = tab[warp][mag - 1][0][k], tm = tab[warp][mag - 1][1][k];
      qz = tab[warp][mag - 1][2][k];
      qt = tab[warp][mag - 1][3][k];
      if (dgt < 0) {
        qp = tm;
        qm = tp;
        qt = -qt;
      } else {
        qp = tp;
        qm = tm;
      }
    }
    int32_t dz = 0;
#pragma unroll 1
    for (int rep = 0; rep < 4; rep++) {
      int32_t XX, YY, ZZ2, A, Y3, Z3, X3, T3;
      wide_mul3(X, X, Y, Y, Z, Z + Z, g, k, XX, YY, ZZ2);
      const int32_t cY = YY + XX, cZ = YY - XX, cT = ZZ2 - cZ, xy = X + Y;
      wide_mul3(xy, xy, cY, cZ, cZ, cT, g, k, A, Y3, Z3);
      const int32_t cX = A - cY;
      // third slot of the last doubling: Z3 * Zq for the addition that follows
      wide_mul3(cX, cT, cX, cY, Z3, qz, g, k, X3, T3, dz);
      X = X3;
      Y = Y3;
      Z = Z3;
      T = T3;
    }
    if (mag) {
      int32_t a, bq, c;
      wide_mul3(Y + X, qp, Y - X, qm, T, qt, g, k, a, bq, c);
      const int32_t D = dz + dz, cX = a - bq, cY = a + bq, cZ = D + c, cT = D - c;
      wide_mul3(cX, cT, cY, cZ, cZ, cT, g, k, X, Y, Z);
    }
  }
  if (peer < N && lane < 10) {
    int32_t* o = xyz_out + 30 * (size_t)peer;
    o[k] = X;
    o[10 + k] = Y;
    o[20 + k] = Z;
  }
}

// batched SHA-256: digest of msgs[i] (each `len` bytes, `stride` apart) -> out[i][32]
__global__ void sha256_batch_kernel(const unsigned char* __restrict__ msgs, int len, long long stride, int N,
                                    unsigned char* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  sha256_oneshot(msgs + (size_t)i * stride, (size_t)len, out + 32 * (size_t)i);
}

}  // namespace bsc

using namespace bsc;

extern "C" {

void bsc_sha512(const uint8_t* data, size_t n, uint8_t out[64]) { sha512(data, n, out); }

// public key (32 bytes) from a 32-byte seed
void bsc_vrf_pubkey(const uint8_t seed[32], uint8_t pk[32]) {
  U256 x;
  uint8_t skhr[32];
  expand_secret(seed, x, skhr);
  ed_compress(pk, ed_mul(ed_basepoint(), x));
}
// expanded secret scalar (little-endian words) for the device kernel
void bsc_vrf_secret_scalar(const uint8_t seed[32], uint32_t out_words[8]) {
  U256 x;
  uint8_t skhr[32];
  expand_secret(seed, x, skhr);
  memcpy(out_words, x.v, 32);
}
// H = hashToCurve(m) in the device struct layout (4 x 8 words, Montgomery form)
void bsc_vrf_hash_to_curve(const uint8_t* m, size_t n, void* out_point, uint8_t out_bytes[32]) {
  EdPoint h = hash_to_curve(m, n);
  if (out_point) memcpy(out_point, &h, sizeof(h));
  if (out_bytes) ed_compress(out_bytes, h);
}
// affine H = hashToCurve(m) as x || y, 32 canonical little-endian bytes each (input of the r25 kernel)
void bsc_vrf_hash_to_curve_xy(const uint8_t* m, size_t n, uint8_t out[64]) {
  // fast-field Elligator2 (bit-identical to the generic-field hash_to_curve; checked in the tests): this
  // runs on the host at the top of every block, ahead of the VRF kernel
  uint8_t hm[32];
  Shake256 s;
  shake256_init(s);
  shake256_absorb(s, m, n);
  shake256_squeeze(s, hm, 32);
  edr_to_affine_bytes(out, edr_elligator2_x8(hm));
}
// generic-field version of the same (oracle for the fast one)
void bsc_vrf_hash_to_curve_xy_generic(const uint8_t* m, size_t n, uint8_t out[64]) {
  EdPoint h = hash_to_curve(m, n);
  Fq zi = fe_inv(h.Z);
  u256_to_le(out, fe_to_u256(fe_mul(h.X, zi)));
  u256_to_le(out + 32, fe_to_u256(fe_mul(h.Y, zi)));
}
// output = SHAKE256(Gamma || m)[:32]
void bsc_vrf_output_from_gamma(const uint8_t gamma[32], const uint8_t* m, size_t n, uint8_t out[32]) {
  Shake256 s;
  shake256_init(s);
  shake256_absorb(s, gamma, 32);
  shake256_absorb(s, m, n);
  shake256_squeeze(s, out, 32);
}

void bsc_vrf_prove(const uint8_t seed[32], const uint8_t* m, size_t n, uint8_t vrf[32], uint8_t proof[96]) {
  U256 x;
  uint8_t skhr[32], pk[32];
  expand_secret(seed, x, skhr);
  const EdPoint B = ed_basepoint();
  ed_compress(pk, ed_mul(B, x));
  const EdPoint H = hash_to_curve(m, n);
  uint8_t hB[32], gammaB[32], gB[32], grB[32], hrB[32];
  ed_compress(hB, H);
  ed_compress(gammaB, ed_mul(H, x));
  ed_compress(gB, B);
  uint8_t rH[64], sH[64];
  Shake256 s;
  shake256_init(s);
  shake256_absorb(s, skhr, 32);
  shake256_absorb(s, pk, 32);
  shake256_absorb(s, m, n);
  shake256_squeeze(s, rH, 64);
  const Fs r = sc_reduce64(rH);
  const U256 rp = fe_to_u256(r);
  ed_compress(grB, ed_mul(B, rp));
  ed_compress(hrB, ed_mul(H, rp));
  shake256_init(s);
  shake256_absorb(s, gB, 32);
  shake256_absorb(s, hB, 32);
  shake256_absorb(s, pk, 32);
  shake256_absorb(s, gammaB, 32);
  shake256_absorb(s, grB, 32);
  shake256_absorb(s, hrB, 32);
  shake256_absorb(s, m, n);
  shake256_squeeze(s, sH, 64);
  const Fs sc = sc_reduce64(sH);
  const Fs xs = fe_from_u256<EdL>(x);
  const Fs t = fe_sub(r, fe_mul(sc, xs));
  u256_to_le(proof, fe_to_u256(sc));
  u256_to_le(proof + 32, fe_to_u256(t));
  memcpy(proof + 64, gammaB, 32);
  bsc_vrf_output_from_gamma(gammaB, m, n, vrf);
}

int bsc_vrf_verify(const uint8_t pk[32], const uint8_t* m, size_t n, const uint8_t vrf[32], const uint8_t proof[96]) {
  EdPoint P, Gamma;
  if (!ed_decompress(P, pk) || !ed_decompress(Gamma, proof + 64)) return 0;
  const U256 sc = u256_from_le(proof), t = u256_from_le(proof + 32);
  const EdPoint B = ed_basepoint();
  const EdPoint H = hash_to_curve(m, n);
  const Fq d2 = ed_const_2d();
  // rG = tG + sP ; rH = tH + s*Gamma
  const EdPoint rG = ed_add(ed_mul(B, t), ed_mul(P, sc), d2);
  const EdPoint rHp = ed_add(ed_mul(H, t), ed_mul(Gamma, sc), d2);
  uint8_t gB[32], hB[32], grB[32], hrB[32], sH[64];
  ed_compress(gB, B);
  ed_compress(hB, H);
  ed_compress(grB, rG);
  ed_compress(hrB, rHp);
  Shake256 s;
  shake256_init(s);
  shake256_absorb(s, gB, 32);
  shake256_absorb(s, hB, 32);
  shake256_absorb(s, pk, 32);
  shake256_absorb(s, proof + 64, 32);
  shake256_absorb(s, grB, 32);
  shake256_absorb(s, hrB, 32);
  shake256_absorb(s, m, n);
  shake256_squeeze(s, sH, 64);
  const U256 s2 = fe_to_u256(sc_reduce64(sH));
  if (!u256_eq(s2, sc)) return 0;
  uint8_t out[32];
  bsc_vrf_output_from_gamma(proof + 64, m, n, out);
  return memcmp(out, vrf, 32) == 0 ? 1 : 0;
}

int bsc_lottery_noisers(const uint8_t* vrf_out, int len, const int* stake, int n, int source, int want, int* noisers);

// All workers' noiser committees in one call: output_p = SHAKE256(Gamma_p || m)[:32], then the stake
// lottery of vrf.go:54-100 seeded by output_p with p itself excluded.  out[p*want + j] = noiser id or -1.
void bsc_noisers_batch(const uint8_t* gammas, const uint8_t* m, size_t mlen, const int* stake, int n,
                       const uint8_t* is_worker, int want, int* out, uint8_t* vrf_out) {
  // ticket list and eligibility are shared by every peer of the block: build them once
  std::vector<int> tickets;
  int distinct = 0;
  for (int id = 0; id < n; id++) {
    for (int k = 0; k < stake[id]; k++) tickets.push_back(id);
    distinct += stake[id] > 0;
  }
  std::vector<int> stamp((size_t)n, -1);
  for (int p = 0; p < n; p++) {
    for (int j = 0; j < want; j++) out[(size_t)p * want + j] = -1;
    uint8_t o[32];
    bsc_vrf_output_from_gamma(gammas + 32 * (size_t)p, m, mlen, o);
    if (vrf_out) memcpy(vrf_out + 32 * (size_t)p, o, 32);
    if ((is_worker && !is_worker[p]) || tickets.empty()) continue;
    // vrf.go:54-100: 2-byte big-endian draws at a byte cursor, SHA-256 re-hash when the bytes run out,
    // duplicates and the peer itself skipped
    const int eligible = distinct - (stake[p] > 0 ? 1 : 0);
    const int need = want < eligible ? want : eligible;
    uint8_t buf[32];
    memcpy(buf, o, 32);
    size_t i = 0;
    int got = 0;
    while (got < need) {
      if (i + 1 >= 32) {
        uint8_t h[32];
        sha256_oneshot(buf, 32, h);
        memcpy(buf, h, 32);
        i = 0;
      }
      const int winner = tickets[((size_t)buf[i] * 256 + buf[i + 1]) % tickets.size()];
      if (winner != p && stamp[(size_t)winner] != p) {
        stamp[(size_t)winner] = p;
        out[(size_t)p * want + got++] = winner;
      }
      i++;
    }
  }
}

// device: gamma[i] = compress(x[i] * H) for all peers
int bsc_ed25519_scalarmult_batch(const void* scalars_u256, const void* bases, int nbase, int N, unsigned char* out,
                                 void* stream) {
  if (N == 0) return 0;
  ed25519_scalarmult_kernel<<<(N + 63) / 64, 64, 0, (cudaStream_t)stream>>>((const U256*)scalars_u256,
                                                                            (const EdPoint*)bases, nbase, N, out);
  return finish_launch("ed25519_scalarmult_batch");
}
int bsc_ed25519_scalarmult_r25(const unsigned char* scalars, const unsigned char* hxy, int N, unsigned char* out,
                               void* stream) {
  if (N == 0) return 0;
  ed25519_scalarmult_r25_kernel<<<(N + 31) / 32, 32, 0, (cudaStream_t)stream>>>(scalars, hxy, N, out);
  return finish_launch("ed25519_scalarmult_r25");
}
int bsc_ed25519_scalarmult_quad(const unsigned char* scalars, const unsigned char* hxy, int N, unsigned char* out,
                                void* stream) {
  if (N == 0) return 0;
  const int threads = 4 * N;
  ed25519_scalarmult_quad_kernel<<<(threads + 127) / 128, 128, 0, (cudaStream_t)stream>>>(scalars, hxy, N, out, nullptr);
  return finish_launch("ed25519_scalarmult_quad");
}
// same, but leaves the results projective: xyz_out[i] = X, Y, Z as 3 x 10 radix-2^25.5 limbs
int bsc_ed25519_scalarmult_quad_xyz(const unsigned char* scalars, const unsigned char* hxy, int N, int* xyz_out,
                                    void* stream) {
  if (N == 0) return 0;
  const int threads = 4 * N;
  ed25519_scalarmult_quad_kernel<<<(threads + 127) / 128, 128, 0, (cudaStream_t)stream>>>(scalars, hxy, N, nullptr,
                                                                                          xyz_out);
  return finish_launch("ed25519_scalarmult_quad_xyz");
}
// warp-wide latency kernel (one peer per warp, one limb per lane); same contract as the quad _xyz entry point
int bsc_ed25519_scalarmult_wide_xyz(const unsigned char* scalars, const unsigned char* hxy, int N, int* xyz_out,
                                    void* stream) {
  if (N == 0) return 0;
  ed25519_scalarmult_wide_kernel<<<(N + WIDE_WARPS - 1) / WIDE_WARPS, 32 * WIDE_WARPS, 0, (cudaStream_t)stream>>>(
      scalars, hxy, N, xyz_out);
  return finish_launch("ed25519_scalarmult_wide_xyz");
}
// The head of a block's chain in one call: hash-to-curve of the previous block hash on the host (into a pinned 64-byte
// buffer), its upload, the scalar multiplication of every peer (wide != 0: warp-wide kernel, else four lanes per peer)
// and, when `event` is given, a record behind the kernel.  h_pinned must stay untouched until the copy has run.
int bsc_vrf_launch(const uint8_t* msg, int len, uint8_t* h_pinned, void* h_dev, const unsigned char* scalars, int N,
                   int* xyz_out, int wide, void* stream, void* event) {
  bsc_vrf_hash_to_curve_xy(msg, (size_t)len, h_pinned);
  BSC_CHECK(cudaMemcpyAsync(h_dev, h_pinned, 64, cudaMemcpyHostToDevice, (cudaStream_t)stream));
  const int rc = wide ? bsc_ed25519_scalarmult_wide_xyz(scalars, (const unsigned char*)h_dev, N, xyz_out, stream)
                      : bsc_ed25519_scalarmult_quad_xyz(scalars, (const unsigned char*)h_dev, N, xyz_out, stream);
  if (rc) return rc;
  if (event) BSC_CHECK(cudaEventRecord((cudaEvent_t)event, (cudaStream_t)stream));
  return 0;
}
// host: compress N projective points with a single field inversion (prefix products)
void bsc_ed25519_compress_batch(const int* xyz, int N, unsigned char* out) {
  std::vector<F25> pre((size_t)N);
  F25 run = f25_one();
  for (int i = 0; i < N; i++) {
    F25 z;
    for (int k = 0; k < 10; k++) z.v[k] = xyz[30 * (size_t)i + 20 + k];
    run = f25_mul(run, z);
    pre[(size_t)i] = run;
  }
  F25 inv = f25_invert(run);
  for (int i = N - 1; i >= 0; i--) {
    F25 x, y, z;
    for (int k = 0; k < 10; k++) {
      x.v[k] = xyz[30 * (size_t)i + k];
      y.v[k] = xyz[30 * (size_t)i + 10 + k];
      z.v[k] = xyz[30 * (size_t)i + 20 + k];
    }
    const F25 zi = i > 0 ? f25_mul(inv, pre[(size_t)i - 1]) : inv;
    inv = f25_mul(inv, z);
    uint8_t xb[32];
    f25_tobytes(out + 32 * (size_t)i, f25_mul(y, zi));
    f25_tobytes(xb, f25_mul(x, zi));
    out[32 * (size_t)i + 31] ^= (uint8_t)((xb[0] & 1) << 7);
  }
}
void bsc_ed25519_scalarmult_r25_host(const unsigned char* scalars, const unsigned char* hxy, int N,
                                     unsigned char* out) {
  EdCached table[8];
  edr_build_table(table, edr_from_affine_bytes(hxy), f25_const_2d());
  for (int i = 0; i < N; i++) edr_compress(out + 32 * (size_t)i, edr_scalarmult(scalars + 32 * (size_t)i, table));
}
void bsc_ed25519_scalarmult_batch_host(const void* scalars_u256, const void* bases, int nbase, int N,
                                       unsigned char* out) {
  const U256* k = (const U256*)scalars_u256;
  const EdPoint* b = (const EdPoint*)bases;
  for (int i = 0; i < N; i++) ed_compress(out + 32 * (size_t)i, ed_mul(b[i % nbase], k[i]));
}
int bsc_sha256_batch(const unsigned char* msgs, int len, long long stride, int N, unsigned char* out, void* stream) {
  if (N == 0) return 0;
  sha256_batch_kernel<<<(N + 63) / 64, 64, 0, (cudaStream_t)stream>>>(msgs, len, stride, N, out);
  return finish_launch("sha256_batch");
}

}  // extern "C"
