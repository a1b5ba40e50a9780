// SHA-256 compression with the x86 SHA extensions, selected at run time.
//
// The block hash covers the full weight vector (DistSys/block.go:23-28: 63 kB at MNIST, 2 MB for the
// CIFAR CNN) and sits on the critical path of every block: the next committee lottery needs it.  The
// portable compression in sha256.cuh runs at ~150 MB/s; SHA-NI reaches >1.5 GB/s.  ledger.cpp calls
// bsc_sha256_fast, which falls back to the portable code when the CPU lacks the extension.
#include <stddef.h>
#include <stdint.h>
#include <string.h>
#include "sha256.cuh"

#if defined(__x86_64__) || defined(__i386__)
#include <immintrin.h>
#define BSC_HAVE_X86 1
#endif

namespace bsc {

#ifdef BSC_HAVE_X86
__attribute__((target("sha,sse4.1,ssse3"))) static void sha256_blocks_ni(uint32_t state[8], const uint8_t* data,
                                                                          size_t nblocks) {
  static const uint32_t K[64] = {
      0x428a2f98, 0x71374491, 0xb5c0fbcf, 0xe9b5dba5, 0x3956c25b, 0x59f111f1, 0x923f82a4, 0xab1c5ed5, 0xd807aa98,
      0x12835b01, 0x243185be, 0x550c7dc3, 0x72be5d74, 0x80deb1fe, 0x9bdc06a7, 0xc19bf174, 0xe49b69c1, 0xefbe4786,
      0x0fc19dc6, 0x240ca1cc, 0x2de92c6f, 0x4a7484aa, 0x5cb0a9dc, 0x76f988da, 0x983e5152, 0xa831c66d, 0xb00327c8,
      0xbf597fc7, 0xc6e00bf3, 0xd5a79147, 0x06ca6351, 0x14292967, 0x27b70a85, 0x2e1b2138, 0x4d2c6dfc, 0x53380d13,
      0x650a7354, 0x766a0abb, 0x81c2c92e, 0x92722c85, 0xa2bfe8a1, 0xa81a664b, 0xc24b8b70, 0xc76c51a3, 0xd192e819,
      0xd6990624, 0xf40e3585, 0x106aa070, 0x19a4c116, 0x1e376c08, 0x2748774c, 0x34b0bcb5, 0x391c0cb3, 0x4ed8aa4a,
      0x5b9cca4f, 0x682e6ff3, 0x748f82ee, 0x78a5636f, 0x84c87814, 0x8cc70208, 0x90befffa, 0xa4506ceb, 0xbef9a3f7,
      0xc67178f2};
  const __m128i shuf = _mm_set_epi64x(0x0c0d0e0f08090a0bULL, 0x0405060700010203ULL);
  // state words (a..h) -> the ABEF / CDGH register layout the sha256rnds2 instruction expects
  __m128i tmp = _mm_loadu_si128((const __m128i*)&state[0]);     // d c b a (high..low = a b c d in memory)
  __m128i st1 = _mm_loadu_si128((const __m128i*)&state[4]);     // e f g h
  tmp = _mm_shuffle_epi32(tmp, 0xB1);                            // c d a b
  st1 = _mm_shuffle_epi32(st1, 0x1B);                            // e f g h reversed
  __m128i st0 = _mm_alignr_epi8(tmp, st1, 8);                    // ABEF
  st1 = _mm_blend_epi16(st1, tmp, 0xF0);                         // CDGH
  while (nblocks--) {
    const __m128i save0 = st0, save1 = st1;
    __m128i m[4];
    for (int i = 0; i < 4; i++) m[i] = _mm_shuffle_epi8(_mm_loadu_si128((const __m128i*)(data + 16 * i)), shuf);
    for (int r = 0; r < 16; r++) {
      // rounds 4r .. 4r+3 use message quad w = m[r & 3]
      __m128i w = m[r & 3];
      __m128i msg = _mm_add_epi32(w, _mm_loadu_si128((const __m128i*)&K[4 * r]));
      st1 = _mm_sha256rnds2_epu32(st1, st0, msg);
      msg = _mm_shuffle_epi32(msg, 0x0E);
      st0 = _mm_sha256rnds2_epu32(st0, st1, msg);
      if (r < 12) {
        // schedule the quad needed four iterations from now: W[t] = s1(W[t-2]) + W[t-7] + s0(W[t-15]) + W[t-16]
        __m128i w0 = m[r & 3], w1 = m[(r + 1) & 3], w2 = m[(r + 2) & 3], w3 = m[(r + 3) & 3];
        __m128i x = _mm_sha256msg1_epu32(w0, w1);
        x = _mm_add_epi32(x, _mm_alignr_epi8(w3, w2, 4));
        m[r & 3] = _mm_sha256msg2_epu32(x, w3);
      }
    }
    st0 = _mm_add_epi32(st0, save0);
    st1 = _mm_add_epi32(st1, save1);
    data += 64;
  }
  tmp = _mm_shuffle_epi32(st0, 0x1B);                            // FEBA
  st1 = _mm_shuffle_epi32(st1, 0xB1);                            // DCHG
  st0 = _mm_blend_epi16(tmp, st1, 0xF0);                         // DCBA
  st1 = _mm_alignr_epi8(st1, tmp, 8);                            // HGFE
  _mm_storeu_si128((__m128i*)&state[0], st0);
  _mm_storeu_si128((__m128i*)&state[4], st1);
}

static bool have_sha_ni() {
  static int cached = -1;
  if (cached < 0) {
    __builtin_cpu_init();
    cached = (__builtin_cpu_supports("sha") && __builtin_cpu_supports("sse4.1") && __builtin_cpu_supports("ssse3")) ? 1 : 0;
  }
  return cached == 1;
}
#endif

void sha256_fast(const uint8_t* data, size_t n, uint8_t out[32]) {
#ifdef BSC_HAVE_X86
  if (have_sha_ni()) {
    uint32_t h[8] = {0x6a09e667, 0xbb67ae85, 0x3c6ef372, 0xa54ff53a, 0x510e527f, 0x9b05688c, 0x1f83d9ab, 0x5be0cd19};
    const size_t full = n / 64;
    sha256_blocks_ni(h, data, full);
    uint8_t tail[128];
    const size_t rem = n - full * 64;
    memset(tail, 0, sizeof(tail));
    memcpy(tail, data + full * 64, rem);
    tail[rem] = 0x80;
    const size_t tb = rem + 9 <= 64 ? 1 : 2;
    const uint64_t bits = (uint64_t)n * 8;
    for (int i = 0; i < 8; i++) tail[tb * 64 - 1 - i] = (uint8_t)(bits >> (8 * i));
    sha256_blocks_ni(h, tail, tb);
    for (int i = 0; i < 8; i++) {
      out[4 * i] = (uint8_t)(h[i] >> 24);
      out[4 * i + 1] = (uint8_t)(h[i] >> 16);
      out[4 * i + 2] = (uint8_t)(h[i] >> 8);
      out[4 * i + 3] = (uint8_t)h[i];
    }
    return;
  }
#endif
  sha256_oneshot(data, n, out);
}

}  // namespace bsc

extern "C" {
void bsc_sha256_fast(const uint8_t* data, size_t n, uint8_t out[32]) { bsc::sha256_fast(data, n, out); }
int bsc_sha256_has_ni() {
#ifdef BSC_HAVE_X86
  return bsc::have_sha_ni() ? 1 : 0;
#else
  return 0;
#endif
}
}
