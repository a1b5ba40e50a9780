// Kernels that keep a block's dependency chain on the GPU.
//
// A Biscotti block is a serial chain: previous block hash -> every peer's VRF -> noiser lottery -> DP noise ->
// Multi-Krum -> approvals -> who contributes -> share aggregation -> reconstruction -> new weights -> block
// hash.  The reference walks that chain with RPC round trips between OS processes (DistSys/main.go:1494-1589,
// krum.go:227-365, main.go:2046-2155).  Here every link whose input is produced on the device is a kernel
// that reads that input on the device, so the host reads results once per block, at the block's tail:
//
//   vrf_noisers_kernel      Gamma_p (projective, from ed25519_scalarmult_quad_kernel) -> canonical bytes ->
//                           VRF output SHAKE256(Gamma || m) -> stake lottery of the noisers (vrf.go:54-100)
//   block_plan_kernel       Krum verdict -> approved set, contributors ("first NUM_SAMPLES/2 shares",
//                           main.go:360), and the compacted local row lists the aggregation kernels consume
//   schnorr_nonce_kernel /  the verifiers' approvals of every Krum candidate (krum.go:338-358,
//   schnorr_finish_kernel   kyber.go:873-896): nonce derivation, and c = H(T || m), r = v - x c once the
//                           comb kernel has produced T = v G — packed next to the commitment bytes so that one
//                           row per peer is what travels between GPUs and to the host
//
// Each kernel has a host restatement (bsc_noisers_batch, engine._plan_host, bsc_schnorr_nonces /
// bsc_schnorr_finish_batch) that the GPU tests compare against.
#include "common.cuh"
#include "g1.cuh"
#include "sha256.cuh"
#include "hashes_hd.cuh"
#include "ed25519_r25.cuh"
#include <string.h>

namespace bsc {

// ---------------------------------------------------------------------------------------------------------
// VRF output + noiser lottery, one thread per local peer.
//   xyz      : [n][30] projective Gamma of every peer (X, Y, Z as 10 radix-2^25.5 limbs each)
//   msg      : the VRF input (previous block hash), mlen <= 64
//   stake    : [n] lottery tickets per peer (0 = absent / dead)
//   is_worker: [n] only workers fetch noise
//   peer_ids : [P] global ids of this rank's peers
//   out      : [P][want] noiser ids, -1 = none
//   vrf_out  : optional [P][32]
// dynamic smem: (n + 1) ints (prefix sums of the stake = the ticket list in closed form)
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) vrf_noisers_kernel(const int32_t* __restrict__ xyz, const uint8_t* __restrict__ msg,
                                                          int mlen, const int* __restrict__ stake, int n,
                                                          const uint8_t* __restrict__ is_worker,
                                                          const int* __restrict__ peer_ids, int P, int want,
                                                          int* __restrict__ out, uint8_t* __restrict__ vrf_out) {
  extern __shared__ int pre[];   // pre[i] = tickets of ids < i ; pre[n] = total
  __shared__ int s_distinct;
  if (threadIdx.x == 0) {
    int run = 0, distinct = 0;
    for (int i = 0; i < n; i++) {
      pre[i] = run;
      run += stake[i];
      distinct += stake[i] > 0;
    }
    pre[n] = run;
    s_distinct = distinct;
  }
  __syncthreads();
  const int lp = blockIdx.x * blockDim.x + threadIdx.x;
  if (lp >= P) return;
  const int p = peer_ids ? peer_ids[lp] : lp;
  for (int j = 0; j < want; j++) out[(size_t)lp * want + j] = -1;
  const bool need_draw = (!is_worker || is_worker[p]) && pre[n] > 0;
  if (!need_draw && !vrf_out) return;
  // canonical compressed Gamma: y with the sign of x in the top bit
  F25 X, Y, Z;
  for (int k = 0; k < 10; k++) {
    X.v[k] = xyz[30 * (size_t)p + k];
    Y.v[k] = xyz[30 * (size_t)p + 10 + k];
    Z.v[k] = xyz[30 * (size_t)p + 20 + k];
  }
  const F25 zi = f25_invert_fast(Z);
  uint8_t gam[32], xb[32];
  f25_tobytes(gam, f25_mul(Y, zi));
  f25_tobytes(xb, f25_mul(X, zi));
  gam[31] ^= (uint8_t)((xb[0] & 1) << 7);
  uint8_t m[64];
  for (int i = 0; i < mlen; i++) m[i] = msg[i];
  uint8_t buf[32];
  shake256_2part_32(gam, 32, m, mlen, buf);
  if (vrf_out)
    for (int i = 0; i < 32; i++) vrf_out[32 * (size_t)lp + i] = buf[i];
  if (!need_draw) return;
  // vrf.go:54-100: 2-byte big-endian draws at a byte cursor that advances by one, SHA-256 re-hash of the
  // buffer when the bytes run out, the peer itself and duplicates skipped
  const int total = pre[n];
  const int eligible = s_distinct - (stake[p] > 0 ? 1 : 0);
  const int need = want < eligible ? want : eligible;
  int got = 0, i = 0;
  int* mine = out + (size_t)lp * want;
  while (got < need) {
    if (i + 1 >= 32) {
      uint8_t h[32];
      sha256_oneshot(buf, 32, h);
      for (int k = 0; k < 32; k++) buf[k] = h[k];
      i = 0;
    }
    const int ticket = (int)(((unsigned)buf[i] * 256u + buf[i + 1]) % (unsigned)total);
    int lo = 0, hi = n - 1;      // largest id with pre[id] <= ticket and stake > 0
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (pre[mid] <= ticket) lo = mid; else hi = mid - 1;
    }
    const int winner = lo;
    bool dup = (winner == p);
    for (int k = 0; k < got; k++) dup = dup || (mine[k] == winner);
    if (!dup) mine[got++] = winner;
    i++;
  }
}

// ---------------------------------------------------------------------------------------------------------
// Krum verdict -> the block's plan.  Single CTA, thread per Krum candidate (nk <= 1024).
//   accept     : [nk] verdict of the defence (Krum accept set / RONI mask / all ones)
//   share_rank : [nk] position of each candidate in this block's share-arrival order (a permutation)
//   local_row  : [nk] row of the candidate in this rank's tensors, -1 if another rank holds it
//   approved_ok: enough verifiers are alive to sign (>= floor(nv/2), main.go:1686)
//   trigger    : miners act on the first `trigger` share sets (NUM_SAMPLES/2, main.go:360)
// outputs (all int32):
//   counts[0..3] = {approved, contributors, local contributors, local approved}
//   contrib_pos[k] = Krum position of the k-th contributor (ascending id)
//   sel[k]         = local row of the k-th local contributor
//   appr_local[k]  = local row of the k-th local approved peer
//   csel[k]        = index of the k-th local contributor inside appr_local
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024) block_plan_kernel(const unsigned char* __restrict__ accept,
                                                          const int* __restrict__ share_rank,
                                                          const int* __restrict__ local_row, int nk, int approved_ok,
                                                          int trigger, int* __restrict__ counts,
                                                          int* __restrict__ contrib_pos, int* __restrict__ sel,
                                                          int* __restrict__ appr_local, int* __restrict__ csel) {
  __shared__ unsigned char s_appr[1024], s_con[1024];
  __shared__ short s_rank[1024];
  __shared__ int s_row[1024];
  const int i = threadIdx.x;
  if (i < nk) {
    s_appr[i] = (accept[i] != 0 && approved_ok) ? 1 : 0;
    s_rank[i] = (short)share_rank[i];
    s_row[i] = local_row[i];
  }
  __syncthreads();
  if (i < nk) {
    int earlier = 0;
    for (int j = 0; j < nk; j++) earlier += (s_appr[j] && s_rank[j] < s_rank[i]);
    s_con[i] = (s_appr[i] && earlier < trigger) ? 1 : 0;
  }
  __syncthreads();
  if (i < nk) {
    int na = 0, nc = 0, nla = 0, nlc = 0;
    for (int j = 0; j < i; j++) {
      na += s_appr[j];
      nc += s_con[j];
      nla += (s_appr[j] && s_row[j] >= 0);
      nlc += (s_con[j] && s_row[j] >= 0);
    }
    if (s_con[i]) contrib_pos[nc] = i;
    if (s_appr[i] && s_row[i] >= 0) appr_local[nla] = s_row[i];
    if (s_con[i] && s_row[i] >= 0) {
      sel[nlc] = s_row[i];
      csel[nlc] = nla;
    }
    if (i == nk - 1) {
      counts[0] = na + s_appr[i];
      counts[1] = nc + s_con[i];
      counts[2] = nlc + (s_con[i] && s_row[i] >= 0);
      counts[3] = nla + (s_appr[i] && s_row[i] >= 0);
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// Approvals.  One thread per (candidate, signer) pair; pair = cand * ns + j.
//   nonce v = SHA-256("bsc-schnorr-nonce" || x_signer || salt[64] || it || peer || signer) mod order
// (salt = per-run salt || previous block hash, so a key never meets the same nonce on two different chains)
// ---------------------------------------------------------------------------------------------------------
struct SchnorrSalt {
  uint8_t b[64];
};
__global__ void __launch_bounds__(128) schnorr_nonce_kernel(const uint8_t* __restrict__ secrets_be, SchnorrSalt salt,
                                                            long long it, const int* __restrict__ cand_ids, int nc,
                                                            const int* __restrict__ signers, int ns,
                                                            U256* __restrict__ k_out) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nc * ns) return;
  const int p = cand_ids[t / ns], v = signers[t % ns];
  uint8_t buf[17 + 32 + 80];
  const char* tag = "bsc-schnorr-nonce";
  for (int i = 0; i < 17; i++) buf[i] = (uint8_t)tag[i];
  for (int i = 0; i < 32; i++) buf[17 + i] = secrets_be[32 * (size_t)v + i];
  for (int i = 0; i < 64; i++) buf[49 + i] = salt.b[i];
  for (int i = 0; i < 8; i++) buf[113 + i] = (uint8_t)((unsigned long long)it >> (8 * i));
  for (int i = 0; i < 4; i++) buf[121 + i] = (uint8_t)((unsigned)p >> (8 * i));
  for (int i = 0; i < 4; i++) buf[125 + i] = (uint8_t)((unsigned)v >> (8 * i));
  uint8_t h[32];
  sha256_oneshot(buf, sizeof(buf), h);
  k_out[t] = fe_to_u256(fe_from_u256<BnN>(u256_from_be(h)));
}

HD bool u256_less(const U256& a, const U256& b) {
  for (int i = 7; i >= 0; i--)
    if (a.v[i] != b.v[i]) return a.v[i] < b.v[i];
  return false;
}

// hashSchnorr (kyber.go:928-933) for a 64-byte message: BLAKE2Xb keyed by the point bytes, scalar picked by
// rejection sampling of successive 32-byte reads (util/random/rand.go)
HDN inline Fr hash_schnorr_64(const uint8_t msg[64], const uint8_t T64[64]) {
  uint64_t h0[8];
  b2x_root_64_64(T64, msg, h0);
  const U256 order = mod_of<BnN>();
  for (uint32_t node = 0;; node++) {
    uint8_t blk[64];
    b2x_block(h0, node, blk);
    for (int half = 0; half < 2; half++) {
      const U256 c = u256_from_be(blk + 32 * half);
      if (!u256_is_zero(c) && u256_less(c, order)) return fe_from_u256<BnN>(c);
    }
  }
}

//   T        : [nc*ns][64] wire bytes of v*G
//   k        : [nc*ns] nonces (plain integers)
//   cbytes   : [*][64] commitment bytes, row = cand_rows[cand]
//   pack     : [*][pack_stride] per-peer output row: 64 B commitment, then ns_max signatures of 64 B
__global__ void __launch_bounds__(128) schnorr_finish_kernel(const uint8_t* __restrict__ T, const U256* __restrict__ k,
                                                             const uint8_t* __restrict__ secrets_be,
                                                             const uint8_t* __restrict__ cbytes,
                                                             const int* __restrict__ cand_rows, int nc,
                                                             const int* __restrict__ signers, int ns,
                                                             uint8_t* __restrict__ pack, int pack_stride) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nc * ns) return;
  const int cand = t / ns, j = t % ns;
  const int row = cand_rows[cand];
  if (row < 0) return;
  uint8_t msg[64], Tb[64];
  for (int i = 0; i < 64; i++) {
    msg[i] = cbytes[64 * (size_t)row + i];
    Tb[i] = T[64 * (size_t)t + i];
  }
  const Fr x = fe_from_u256<BnN>(u256_from_be(secrets_be + 32 * (size_t)signers[j]));
  const Fr v = fe_from_u256<BnN>(k[t]);
  const Fr c = hash_schnorr_64(msg, Tb);
  const Fr r = fe_sub(v, fe_mul(x, c));
  uint8_t* dst = pack + (size_t)row * pack_stride;
  if (j == 0)
    for (int i = 0; i < 64; i++) dst[i] = msg[i];
  u256_to_be(dst + 64 + 64 * j, fe_to_u256(c));
  u256_to_be(dst + 64 + 64 * j + 32, fe_to_u256(r));
}

// rows without approvals (no verifier alive, approvals disabled): the pack row is just the commitment
__global__ void pack_commitments_kernel(const uint8_t* __restrict__ cbytes, int P, uint8_t* __restrict__ pack,
                                        int pack_stride) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= P * 16) return;
  const int row = t >> 4, q = t & 15;
  reinterpret_cast<uint32_t*>(pack + (size_t)row * pack_stride)[q] = reinterpret_cast<const uint32_t*>(cbytes + 64 * (size_t)row)[q];
}

}  // namespace bsc

using namespace bsc;

extern "C" {

int bsc_vrf_noisers(const int* xyz, const unsigned char* msg, int mlen, const int* stake, int n,
                    const unsigned char* is_worker, const int* peer_ids, int P, int want, int* out,
                    unsigned char* vrf_out, void* stream) {
  if (P <= 0) return 0;
  if (mlen < 0 || mlen > 64 || n <= 0 || want < 0) return -1;
  const size_t smem = (size_t)(n + 1) * sizeof(int);
  if (smem > 48 * 1024) return -2;
  vrf_noisers_kernel<<<(P + 127) / 128, 128, smem, (cudaStream_t)stream>>>(xyz, msg, mlen, stake, n, is_worker, peer_ids, P,
                                                                           want, out, vrf_out);
  return finish_launch("vrf_noisers");
}

int bsc_block_plan(const unsigned char* accept, const int* share_rank, const int* local_row, int nk, int approved_ok,
                   int trigger, int* counts, int* contrib_pos, int* sel, int* appr_local, int* csel, void* stream) {
  if (nk < 1 || nk > 1024) return -1;
  block_plan_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(accept, share_rank, local_row, nk, approved_ok, trigger, counts,
                                                          contrib_pos, sel, appr_local, csel);
  return finish_launch("block_plan");
}

// secrets_be: [n_peers][32] big-endian signing keys (device); salt64: 64 host bytes; k_out: [nc*ns] U256 (device)
int bsc_schnorr_nonces_dev(const unsigned char* secrets_be, const unsigned char* salt64, long long it, const int* cand_ids,
                           int nc, const int* signers, int ns, void* k_out, void* stream) {
  if (nc * ns <= 0) return 0;
  SchnorrSalt s;
  memcpy(s.b, salt64, 64);
  schnorr_nonce_kernel<<<(nc * ns + 127) / 128, 128, 0, (cudaStream_t)stream>>>(secrets_be, s, it, cand_ids, nc, signers, ns,
                                                                               (U256*)k_out);
  return finish_launch("schnorr_nonces_dev");
}

int bsc_schnorr_finish_dev(const unsigned char* T, const void* k, const unsigned char* secrets_be, const unsigned char* cbytes,
                           const int* cand_rows, int nc, const int* signers, int ns, unsigned char* pack, int pack_stride,
                           void* stream) {
  if (nc * ns <= 0) return 0;
  if (pack_stride < 64 * (1 + ns)) return -1;
  schnorr_finish_kernel<<<(nc * ns + 127) / 128, 128, 0, (cudaStream_t)stream>>>(T, (const U256*)k, secrets_be, cbytes,
                                                                                 cand_rows, nc, signers, ns, pack, pack_stride);
  return finish_launch("schnorr_finish_dev");
}

int bsc_pack_commitments(const unsigned char* cbytes, int P, unsigned char* pack, int pack_stride, void* stream) {
  if (P <= 0) return 0;
  if (pack_stride < 64 || (pack_stride & 3)) return -1;
  pack_commitments_kernel<<<(P * 16 + 255) / 256, 256, 0, (cudaStream_t)stream>>>(cbytes, P, pack, pack_stride);
  return finish_launch("pack_commitments");
}

// host mirror of hash_schnorr_64 + finish, for tests: sig = c || r
void bsc_schnorr_finish64_host(const unsigned char* T64, const unsigned char* k_words, const unsigned char* x_be,
                               const unsigned char* msg64, unsigned char* sig) {
  U256 kk;
  memcpy(kk.v, k_words, 32);
  const Fr x = fe_from_u256<BnN>(u256_from_be(x_be));
  const Fr v = fe_from_u256<BnN>(kk);
  const Fr c = hash_schnorr_64(msg64, T64);
  const Fr r = fe_sub(v, fe_mul(x, c));
  u256_to_be(sig, fe_to_u256(c));
  u256_to_be(sig + 32, fe_to_u256(r));
}

}  // extern "C"
