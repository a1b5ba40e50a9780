// SHA-256, host+device.  Used for block hashes (DistSys/block.go:23-28), the stake lottery's
// re-hash chain (DistSys/vrf.go:79-83, 134-140) and the on-device hash of the weight buffer.
#pragma once
#include <stdint.h>
#include <stddef.h>
#include "field.cuh"

namespace bsc {

struct Sha256 {
  uint32_t h[8];
  uint8_t buf[64];
  uint64_t len;   // total bytes absorbed
  uint32_t fill;  // bytes in buf
};

HD uint32_t sha_rotr(uint32_t x, int n) { return (x >> n) | (x << (32 - n)); }

HD uint32_t sha_k(int i) {
  constexpr uint32_t K[64] = {
      0x428a2f98, 0x71374491, 0xb5c0fbcf, 0xe9b5dba5, 0x3956c25b, 0x59f111f1, 0x923f82a4, 0xab1c5ed5,
      0xd807aa98, 0x12835b01, 0x243185be, 0x550c7dc3, 0x72be5d74, 0x80deb1fe, 0x9bdc06a7, 0xc19bf174,
      0xe49b69c1, 0xefbe4786, 0x0fc19dc6, 0x240ca1cc, 0x2de92c6f, 0x4a7484aa, 0x5cb0a9dc, 0x76f988da,
      0x983e5152, 0xa831c66d, 0xb00327c8, 0xbf597fc7, 0xc6e00bf3, 0xd5a79147, 0x06ca6351, 0x14292967,
      0x27b70a85, 0x2e1b2138, 0x4d2c6dfc, 0x53380d13, 0x650a7354, 0x766a0abb, 0x81c2c92e, 0x92722c85,
      0xa2bfe8a1, 0xa81a664b, 0xc24b8b70, 0xc76c51a3, 0xd192e819, 0xd6990624, 0xf40e3585, 0x106aa070,
      0x19a4c116, 0x1e376c08, 0x2748774c, 0x34b0bcb5, 0x391c0cb3, 0x4ed8aa4a, 0x5b9cca4f, 0x682e6ff3,
      0x748f82ee, 0x78a5636f, 0x84c87814, 0x8cc70208, 0x90befffa, 0xa4506ceb, 0xbef9a3f7, 0xc67178f2};
  return K[i];
}

// one compression of a 64-byte block given as 16 big-endian words
HDN inline void sha256_compress(uint32_t h[8], const uint32_t win[16]) {
  uint32_t w[16];
#pragma unroll
  for (int i = 0; i < 16; i++) w[i] = win[i];
  uint32_t a = h[0], b = h[1], c = h[2], d = h[3], e = h[4], f = h[5], g = h[6], hh = h[7];
#pragma unroll
  for (int i = 0; i < 64; i++) {
    uint32_t wi;
    if (i < 16) {
      wi = w[i];
    } else {
      uint32_t w15 = w[(i + 1) & 15], w2 = w[(i + 14) & 15];
      uint32_t s0 = sha_rotr(w15, 7) ^ sha_rotr(w15, 18) ^ (w15 >> 3);
      uint32_t s1 = sha_rotr(w2, 17) ^ sha_rotr(w2, 19) ^ (w2 >> 10);
      wi = w[i & 15] + s0 + w[(i + 9) & 15] + s1;
      w[i & 15] = wi;
    }
    uint32_t S1 = sha_rotr(e, 6) ^ sha_rotr(e, 11) ^ sha_rotr(e, 25);
    uint32_t ch = (e & f) ^ (~e & g);
    uint32_t t1 = hh + S1 + ch + sha_k(i) + wi;
    uint32_t S0 = sha_rotr(a, 2) ^ sha_rotr(a, 13) ^ sha_rotr(a, 22);
    uint32_t mj = (a & b) ^ (a & c) ^ (b & c);
    uint32_t t2 = S0 + mj;
    hh = g; g = f; f = e; e = d + t1; d = c; c = b; b = a; a = t1 + t2;
  }
  h[0] += a; h[1] += b; h[2] += c; h[3] += d; h[4] += e; h[5] += f; h[6] += g; h[7] += hh;
}

HD void sha256_init(Sha256& s) {
  s.h[0] = 0x6a09e667; s.h[1] = 0xbb67ae85; s.h[2] = 0x3c6ef372; s.h[3] = 0xa54ff53a;
  s.h[4] = 0x510e527f; s.h[5] = 0x9b05688c; s.h[6] = 0x1f83d9ab; s.h[7] = 0x5be0cd19;
  s.len = 0;
  s.fill = 0;
}
HDN inline void sha256_block_bytes(uint32_t h[8], const uint8_t* p) {
  uint32_t w[16];
#pragma unroll
  for (int i = 0; i < 16; i++)
    w[i] = ((uint32_t)p[4 * i] << 24) | ((uint32_t)p[4 * i + 1] << 16) | ((uint32_t)p[4 * i + 2] << 8) | p[4 * i + 3];
  sha256_compress(h, w);
}
HDN inline void sha256_update(Sha256& s, const uint8_t* data, size_t n) {
  s.len += n;
  while (n) {
    if (s.fill == 0 && n >= 64) {
      sha256_block_bytes(s.h, data);
      data += 64;
      n -= 64;
      continue;
    }
    uint32_t take = 64 - s.fill;
    if (take > n) take = (uint32_t)n;
    for (uint32_t i = 0; i < take; i++) s.buf[s.fill + i] = data[i];
    s.fill += take;
    data += take;
    n -= take;
    if (s.fill == 64) {
      sha256_block_bytes(s.h, s.buf);
      s.fill = 0;
    }
  }
}
HDN inline void sha256_final(Sha256& s, uint8_t out[32]) {
  uint64_t bits = s.len * 8;
  s.buf[s.fill++] = 0x80;
  if (s.fill > 56) {
    while (s.fill < 64) s.buf[s.fill++] = 0;
    sha256_block_bytes(s.h, s.buf);
    s.fill = 0;
  }
  while (s.fill < 56) s.buf[s.fill++] = 0;
  for (int i = 0; i < 8; i++) s.buf[56 + i] = (uint8_t)(bits >> (56 - 8 * i));
  sha256_block_bytes(s.h, s.buf);
  for (int i = 0; i < 8; i++) {
    out[4 * i] = (uint8_t)(s.h[i] >> 24);
    out[4 * i + 1] = (uint8_t)(s.h[i] >> 16);
    out[4 * i + 2] = (uint8_t)(s.h[i] >> 8);
    out[4 * i + 3] = (uint8_t)(s.h[i]);
  }
}
HDN inline void sha256_oneshot(const uint8_t* data, size_t n, uint8_t out[32]) {
  Sha256 s;
  sha256_init(s);
  sha256_update(s, data, n);
  sha256_final(s, out);
}

}  // namespace bsc
