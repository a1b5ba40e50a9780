// BLAKE2b / BLAKE2Xb and SHAKE256 in a host+device form, for the kernels that keep the block's
// dependency chain on the GPU (kernels_chain.cu): the Schnorr challenge c = H(T || m)
// (DistSys/kyber.go:928-933, lib/dedis/kyber/xof/blake2xb/blake.go:18-37) and the VRF output
// SHAKE256(Gamma || m) (coniks ECVRF, DistSys/vrf.go:54-100).  host_crypto.cpp keeps its own streaming
// implementations; tests/test_chain_kernels_gpu.py checks that both agree byte for byte.
//
// These are fixed-shape one-shot forms: messages up to one rate block (SHAKE, 136 B) / two 128-byte
// blocks (BLAKE2b key block + message block), which is all the chain needs.
#pragma once
#include <stdint.h>
#include "field.cuh"

namespace bsc {

HD uint64_t hd_rotr64(uint64_t x, int n) { return (x >> n) | (x << (64 - n)); }
HD uint64_t hd_rotl64(uint64_t x, int n) { return (x << n) | (x >> (64 - n)); }

HD uint64_t b2_iv(int i) {
  constexpr uint64_t IV[8] = {0x6a09e667f3bcc908ULL, 0xbb67ae8584caa73bULL, 0x3c6ef372fe94f82bULL, 0xa54ff53a5f1d36f1ULL,
                              0x510e527fade682d1ULL, 0x9b05688c2b3e6c1fULL, 0x1f83d9abfb41bd6bULL, 0x5be0cd19137e2179ULL};
  return IV[i];
}
HD int b2_sigma(int r, int i) {
  constexpr uint8_t S[10][16] = {
      {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15}, {14, 10, 4, 8, 9, 15, 13, 6, 1, 12, 0, 2, 11, 7, 5, 3},
      {11, 8, 12, 0, 5, 2, 15, 13, 10, 14, 3, 6, 7, 1, 9, 4}, {7, 9, 3, 1, 13, 12, 11, 14, 2, 6, 5, 10, 4, 0, 15, 8},
      {9, 0, 5, 7, 2, 4, 10, 15, 14, 1, 11, 12, 6, 8, 3, 13}, {2, 12, 6, 10, 0, 11, 8, 3, 4, 13, 7, 5, 15, 14, 1, 9},
      {12, 5, 1, 15, 14, 13, 4, 10, 0, 7, 6, 3, 9, 2, 8, 11}, {13, 11, 7, 14, 12, 1, 3, 9, 5, 0, 15, 4, 8, 6, 2, 10},
      {6, 15, 14, 9, 11, 3, 0, 8, 12, 2, 13, 7, 1, 4, 10, 5}, {10, 2, 8, 4, 7, 6, 1, 5, 15, 11, 9, 14, 3, 12, 13, 0}};
  return S[r % 10][i];
}

// one BLAKE2b compression: h is updated in place; m = 16 little-endian words of the 128-byte block,
// t = byte counter after this block, last = final-block flag
HDN inline void b2_compress_hd(uint64_t h[8], const uint64_t m[16], uint64_t t, bool last) {
  uint64_t v[16];
#pragma unroll
  for (int i = 0; i < 8; i++) {
    v[i] = h[i];
    v[i + 8] = b2_iv(i);
  }
  v[12] ^= t;
  if (last) v[14] = ~v[14];
#define BSC_B2G(a, b, c, d, x, y)      \
  v[a] = v[a] + v[b] + (x);            \
  v[d] = hd_rotr64(v[d] ^ v[a], 32);   \
  v[c] = v[c] + v[d];                  \
  v[b] = hd_rotr64(v[b] ^ v[c], 24);   \
  v[a] = v[a] + v[b] + (y);            \
  v[d] = hd_rotr64(v[d] ^ v[a], 16);   \
  v[c] = v[c] + v[d];                  \
  v[b] = hd_rotr64(v[b] ^ v[c], 63);
#pragma unroll
  for (int r = 0; r < 12; r++) {   // unrolled: the message schedule becomes compile-time register picks
    BSC_B2G(0, 4, 8, 12, m[b2_sigma(r, 0)], m[b2_sigma(r, 1)]);
    BSC_B2G(1, 5, 9, 13, m[b2_sigma(r, 2)], m[b2_sigma(r, 3)]);
    BSC_B2G(2, 6, 10, 14, m[b2_sigma(r, 4)], m[b2_sigma(r, 5)]);
    BSC_B2G(3, 7, 11, 15, m[b2_sigma(r, 6)], m[b2_sigma(r, 7)]);
    BSC_B2G(0, 5, 10, 15, m[b2_sigma(r, 8)], m[b2_sigma(r, 9)]);
    BSC_B2G(1, 6, 11, 12, m[b2_sigma(r, 10)], m[b2_sigma(r, 11)]);
    BSC_B2G(2, 7, 8, 13, m[b2_sigma(r, 12)], m[b2_sigma(r, 13)]);
    BSC_B2G(3, 4, 9, 14, m[b2_sigma(r, 14)], m[b2_sigma(r, 15)]);
  }
#undef BSC_B2G
#pragma unroll
  for (int i = 0; i < 8; i++) h[i] ^= v[i] ^ v[i + 8];
}

HD uint64_t hd_load64le(const uint8_t* p) {
  uint64_t v = 0;
  for (int i = 7; i >= 0; i--) v = (v << 8) | p[i];
  return v;
}

// BLAKE2Xb(seed = 64-byte key, message = 64 bytes): the root hash h0 (64 bytes as 8 words) that every
// output block is derived from.  Parameter block: digest 64, key 64, fanout 1, depth 1, xof length
// 0xFFFFFFFF (unknown), as golang.org/x/crypto/blake2b NewXOF sets it.
HDN inline void b2x_root_64_64(const uint8_t key[64], const uint8_t msg[64], uint64_t h0[8]) {
  uint64_t h[8];
  for (int i = 0; i < 8; i++) h[i] = b2_iv(i);
  h[0] ^= 0x01014040ULL;                       // digest_length 64 | key_length 64 << 8 | fanout 1 << 16 | depth 1 << 24
  h[1] ^= 0xFFFFFFFFULL << 32;                 // bytes 12..15: xof length
  uint64_t m[16];
  for (int i = 0; i < 8; i++) m[i] = hd_load64le(key + 8 * i);
  for (int i = 8; i < 16; i++) m[i] = 0;
  b2_compress_hd(h, m, 128, false);            // the key block, padded to 128 bytes
  for (int i = 0; i < 8; i++) m[i] = hd_load64le(msg + 8 * i);
  for (int i = 8; i < 16; i++) m[i] = 0;
  b2_compress_hd(h, m, 192, true);
  for (int i = 0; i < 8; i++) h0[i] = h[i];
}
// output block `node` (64 bytes) of the XOF with root h0
HDN inline void b2x_block(const uint64_t h0[8], uint32_t node, uint8_t out[64]) {
  uint64_t h[8];
  for (int i = 0; i < 8; i++) h[i] = b2_iv(i);
  h[0] ^= 0x40ULL | (64ULL << 32);             // digest 64, fanout 0, depth 0, leaf length 64 (bytes 4..7)
  h[1] ^= (uint64_t)node | (0xFFFFFFFFULL << 32);   // node offset (bytes 8..11), xof length (12..15)
  h[2] ^= 64ULL << 8;                          // byte 17: inner length 64 (byte 16 = node depth 0)
  uint64_t m[16];
  for (int i = 0; i < 8; i++) m[i] = h0[i];
  for (int i = 8; i < 16; i++) m[i] = 0;
  b2_compress_hd(h, m, 64, true);
  for (int i = 0; i < 8; i++)
    for (int j = 0; j < 8; j++) out[8 * i + j] = (uint8_t)(h[i] >> (8 * j));
}

// ---- Keccak-f[1600] / SHAKE256 ---------------------------------------------------------------------
HD uint64_t keccak_rc(int i) {
  constexpr uint64_t RC[24] = {0x0000000000000001ULL, 0x0000000000008082ULL, 0x800000000000808aULL, 0x8000000080008000ULL,
                               0x000000000000808bULL, 0x0000000080000001ULL, 0x8000000080008081ULL, 0x8000000000008009ULL,
                               0x000000000000008aULL, 0x0000000000000088ULL, 0x0000000080008009ULL, 0x000000008000000aULL,
                               0x000000008000808bULL, 0x800000000000008bULL, 0x8000000000008089ULL, 0x8000000000008003ULL,
                               0x8000000000008002ULL, 0x8000000000000080ULL, 0x000000000000800aULL, 0x800000008000000aULL,
                               0x8000000080008081ULL, 0x8000000000008080ULL, 0x0000000080000001ULL, 0x8000000080008008ULL};
  return RC[i];
}
HD int keccak_rot(int i) {
  constexpr int R[24] = {1, 3, 6, 10, 15, 21, 28, 36, 45, 55, 2, 14, 27, 41, 56, 8, 25, 43, 62, 18, 39, 61, 20, 44};
  return R[i];
}
HD int keccak_pil(int i) {
  constexpr int P[24] = {10, 7, 11, 17, 18, 3, 5, 16, 8, 21, 24, 4, 15, 23, 19, 13, 12, 2, 20, 14, 22, 9, 6, 1};
  return P[i];
}
HDN inline void keccak_f_hd(uint64_t st[25]) {
  for (int r = 0; r < 24; r++) {
    uint64_t bc[5];
    for (int i = 0; i < 5; i++) bc[i] = st[i] ^ st[i + 5] ^ st[i + 10] ^ st[i + 15] ^ st[i + 20];
    for (int i = 0; i < 5; i++) {
      const uint64_t t = bc[(i + 4) % 5] ^ hd_rotl64(bc[(i + 1) % 5], 1);
      for (int j = 0; j < 25; j += 5) st[j + i] ^= t;
    }
    uint64_t t = st[1];
    for (int i = 0; i < 24; i++) {
      const int j = keccak_pil(i);
      const uint64_t b = st[j];
      st[j] = hd_rotl64(t, keccak_rot(i));
      t = b;
    }
    for (int j = 0; j < 25; j += 5) {
      for (int i = 0; i < 5; i++) bc[i] = st[j + i];
      for (int i = 0; i < 5; i++) st[j + i] ^= (~bc[(i + 1) % 5]) & bc[(i + 2) % 5];
    }
    st[0] ^= keccak_rc(r);
  }
}
// SHAKE256(a || b)[:32] for len(a) + len(b) < 136 (one rate block)
HDN inline void shake256_2part_32(const uint8_t* a, int na, const uint8_t* b, int nb, uint8_t out[32]) {
  uint64_t st[25];
  for (int i = 0; i < 25; i++) st[i] = 0;
  int pos = 0;
  for (int i = 0; i < na; i++, pos++) st[pos >> 3] ^= (uint64_t)a[i] << (8 * (pos & 7));
  for (int i = 0; i < nb; i++, pos++) st[pos >> 3] ^= (uint64_t)b[i] << (8 * (pos & 7));
  st[pos >> 3] ^= (uint64_t)0x1F << (8 * (pos & 7));
  st[135 >> 3] ^= (uint64_t)0x80 << (8 * (135 & 7));
  keccak_f_hd(st);
  for (int i = 0; i < 32; i++) out[i] = (uint8_t)(st[i >> 3] >> (8 * (i & 7)));
}

}  // namespace bsc
