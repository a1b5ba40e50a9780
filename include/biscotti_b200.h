/* biscotti_b200 C ABI - every extern "C" entry point of libbiscotti_b200.so.
 *
 * Conventions: device pointers are plain pointers into CUDA global memory; `stream` is a cudaStream_t
 * passed as void*; functions returning int return 0 on success, a cudaError_t on a launch failure, or a
 * negative value for an argument error.  `*_host` functions are CPU mirrors built from the same
 * host/device headers as the kernels (bit-exact oracles and the CPU plumbing path).  Point layouts:
 * affine = 16 x uint32 (x, y in Montgomery form), Jacobian = 24 x uint32, wire = 64 big-endian bytes.
 * Generated from the sources by tools/gen_header.py; comments are the ones beside each definition.
 */
#ifndef BISCOTTI_B200_H
#define BISCOTTI_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* --------------------------------------------------------------------------------------------
 * kernels_ml.cu - Local steps, DP noise, Krum tail, evaluation.  Reference: ML/Pytorch/client.py:38-65, 136-172; ML/code/logistic_model.py:92-140; ML/Pytorch/client_obj.py:97-143; DistSys/main.go:1592-1660; DistSys/krum.go:77-166.
 * -------------------------------------------------------------------------------------------- */
// test hook: 1 = keep K1 on the SIMT kernel (the tensor-core kernel is compared against it and against fp32 PyTorch)
void bsc_softmax_force_simt(int on);

unsigned long long bsc_launch_count();

void bsc_launch_count_reset();

int bsc_local_softmax_step(const float* w, const float* X, const int* y, const int* idx, float* U, float* loss, int P, int Ns, int B, int F, int C, int ldu, float clip, void* stream);

int bsc_local_logistic_step(const float* w, const float* X, const float* y, const int* idx, float* U, int P, int Ns, int B, int F, int ldu, float alpha, float lambda, float own_noise_scale, unsigned long long seed, const int* peer_ids, int slot, void* stream);

int bsc_noise_mix(const float* U, const int* noisers, int nn, unsigned long long seed, int slot, float noise_scale, float* Un, void* hi, void* lo, float* norms, int P, int d, int ldu, int ldh, void* stream);

// Same, with the bf16 planes and norms of peer p written to row dst_row[p] (skipped when negative): the Krum
// candidates come out compacted in Krum order and the Gram kernel reads them without a gather.
int bsc_noise_mix_rows(const float* U, const int* noisers, int nn, unsigned long long seed, int slot, float noise_scale, float* Un, void* hi, void* lo, float* norms, int P, int d, int ldu, int ldh, const int* dst_row, void* stream);

// idx[p][0..B) = B distinct rows in [0, Ns) for peer peer_ids[p] (or p) at iteration `it`; Ns <= 1024, B <= Ns.
int bsc_sample_minibatch(unsigned long long seed, int it, const int* peer_ids, int P, int Ns, int B, int* idx, void* stream);

int bsc_noise_row(float* out, int d, unsigned long long seed, int peer, int slot, float scale, void* stream);

// host restatement of the same generator (CPU path + oracle)
void bsc_noise_row_host(float* out, int d, unsigned long long seed, int peer, int slot, float scale);

int bsc_gram_simt(const float* A, float* Gpart, int n, int d, int lda, int ksplit, void* stream);

int bsc_krum_select(const float* Gpart, int nsplit, int ldg, const float* norms, int n, int clip, float* scores, unsigned char* accept, float* Dout, void* stream);

// split count used by bsc_krum_scores_static for a given dimension (so a caller can reproduce the same
// reduction order with the allocating entry points)
int bsc_krum_static_splits(int d);

// No counterpart in the reference (its Krum runs in NumPy on the host, client_obj.py:114-143); this is
// the allocation-free variant of bsc_gram_simt + bsc_krum_select for callers that cannot call
// cudaMalloc.  Inputs: A [n][lda] fp32 updates (first d columns used), norms [n] squared row norms.
// Outputs: scores [n], accept [n] (1 = in the Multi-Krum accept set).  Limits: 3 <= n <= 128 peers, any
// d >= 1; larger n is rejected with -1 and a message on stderr.  Nothing is allocated or freed, here or
// in any library it calls.  The first call on each device looks up the device address of the scratch
// array (cudaGetSymbolAddress, which may load the CUDA module), so make that first call before starting
// any stream capture.  All calls on a device share the one scratch array: a call rewrites every word it
// later reads (nothing has to start at zero, and contents left by the previous call are never
// consumed), but two calls whose kernels overlap on the GPU — different streams, or different host
// threads — race on it; serialise them on one stream.
int bsc_krum_scores_static(const float* A, const float* norms, int n, int d, int lda, int clip, float* scores, unsigned char* accept, void* stream);

// Multi-CTA Krum tail (scores over a grid of row-groups, then the ranking).  rows / norms may be null
// (identity mapping / norms from the Gram diagonal).  Bit-identical to bsc_krum_select.
int bsc_krum_select_mc(const float* Gpart, int nsplit, int ldg, const float* norms, const int* rows, int n, int clip, float* scores, unsigned char* accept, float* Dout, void* stream);

// Same tail for the fused multi-GPU path: the scores kernel first waits (bounded) until every rank has
// published SLOT_GRAM for `epoch` in my_flags, i.e. all partial Grams have landed in Gpart.
int bsc_krum_select_mc_wait(const float* Gpart, int nsplit, int ldg, const int* rows, int n, int clip, const void* my_flags, int world, unsigned long long epoch, float* scores, unsigned char* accept, void* stream);

int bsc_eval_softmax(const float* Wc, const float* base, const float* X, const int* y, int* wrong, int M, int N, int F, int C, int ldw, void* stream);

int bsc_eval_logistic(const float* Wc, const float* base, const float* X, const float* y, int* wrong, int M, int N, int F, int ldw, void* stream);

/* --------------------------------------------------------------------------------------------
 * gram_tc.cu - tcgen05/TMA Gram kernel for Multi-Krum.  Reference: ML/Pytorch/client_obj.py:132-143.
 * -------------------------------------------------------------------------------------------- */
// G partials of hi/lo planes [rows][ldh] (ldh % 64 == 0, 16-byte aligned base), first n rows valid.
// Gpart: [ksplit][ldg][ldg] with ldg = ceil(n/128)*128; every split must own >= 1 k-block.
int bsc_gram_tc(const void* hi, const void* lo, int n, int rows, int ldh, int ksplit, float* Gpart, int ldg, void* stream);

// fp32 rows -> zero-padded bf16x3 planes [rows_pad][ldk] (+ bias vector); see split_planes_kernel.
int bsc_split_planes(const float* src, const float* base, int outer, int inner, int inner_pad, long long ld_outer, int ld_inner, int cols, void* hi, void* lo, int ldk, int rows_pad, float* bias_out, int bias_off, void* stream);

// wrong[m] += #{n < N : argmax_c (A[n] . B[m*16 + c] + bias[m*16 + c]) != y[n]} for m < M.  A planes: [a_rows_pad][ldk]
// (a_rows_pad % 128 == 0, >= N), B planes: [b_rows_pad][ldk] (b_rows_pad % 128 == 0, >= 16*M).  Caller zero-fills wrong.
int bsc_eval_tc(const void* a_hi, const void* a_lo, int a_rows_pad, const void* b_hi, const void* b_lo, int b_rows_pad, int ldk, const float* bias, const int* y, int* wrong, int N, int M, void* stream);

/* --------------------------------------------------------------------------------------------
 * kernels_secagg.cu - Fixed-point, Shamir shares, aggregation, reconstruction + SGD step.  Reference: DistSys/kyber.go:244-287, 579-609, 698-757, 809-857; DistSys/honest.go:346-502.
 * -------------------------------------------------------------------------------------------- */
int bsc_quantize(const float* U, long long* Q, int P, int d, int ldu, int dpad, double scale, void* stream);

int bsc_shares_eval(const long long* Q, long long* Y, int P, int nchunk, int dpad, int S, void* stream);

// nsel_dev (may be null): device-resident contributor count, see share_sum_kernel
int bsc_share_sum_dc(const float* U, const int* sel, int nsel, const int* nsel_dev, int d, int ldu, int c0, int nc, int S, double scale, long long* Ysum, void* stream);

int bsc_share_sum(const float* U, const int* sel, int nsel, int d, int ldu, int c0, int nc, int S, double scale, long long* Ysum, void* stream);

// gate (may be null): the step is applied only if *gate >= gate_min (device-resident contributor count)
int bsc_reconstruct_step_gated(const long long* Ysum, const int* xs, const unsigned char* avail, int S, int c0, int nc, int d, double scale, double* w64, float* w32, double* dsum, int* bad, const int* gate, int gate_min, void* stream);

int bsc_reconstruct_step(const long long* Ysum, const int* xs, const unsigned char* avail, int S, int c0, int nc, int d, double scale, double* w64, float* w32, double* dsum, int* bad, void* stream);

// Matrix form of bsc_reconstruct_step_gated; `rm` points at a host RecoverMat (see secagg_math.cuh).
int bsc_reconstruct_mat(const long long* Ysum, const void* rm, const unsigned char* avail, int S, int c0, int nc, int d, double scale, double* w64, float* w32, double* dsum, int* bad, const int* gate, int gate_min, void* stream);

// host restatement: coefficients of nc chunks through the matrix; returns the number of inexact coefficients
int bsc_recover_mat_host(const long long* Ysum, const void* rm, int S, int nc, long long* coef_out);

int bsc_recover_mat_sizeof();

int bsc_sum_rows_i64_dc(const long long* Q, int ldq, const int* sel, int nsel, const int* nsel_dev, int ncols, long long* out, void* stream);

int bsc_plain_sum_step_dc(const float* U, const int* sel, int nsel, const int* nsel_dev, int d, int ldu, double* w64, float* w32, double* dsum, const int* gate, int gate_min, void* stream);

int bsc_plain_sum_step(const float* U, const int* sel, int nsel, int d, int ldu, double* w64, float* w32, double* dsum, void* stream);

// ---- host restatements (CPU plumbing path and oracles; same arithmetic, same header) -------------
void bsc_quantize_host(const float* U, long long* Q, int P, int d, int ldu, int dpad, double scale);

void bsc_shares_eval_host(const long long* Q, long long* Y, int P, int nchunk, int dpad, int S);

// returns number of chunks whose recovery was inexact or inconsistent
int bsc_shamir_recover_host(const long long* Ysum, const int* xs, const unsigned char* avail, int S, int nc, long long* coef_out);

/* --------------------------------------------------------------------------------------------
 * kernels_msm.cu - BN256 G1 MSM for commitments and witnesses.  Reference: DistSys/kyber.go:172-202, 456-482, 533-562, 579-646; DistSys/publicKey.go:28-66.
 * -------------------------------------------------------------------------------------------- */
int bsc_msm_groups(const long long* Q, int ldq, const int* sel, int P, const int* sidx, const int* bidx, const int* og, int G, int T, const void* bases_affine, void* out_jac, long long ostride_p, void* stream);

int bsc_msm_table_build(const void* bases_affine, long long nb, int w, int nw, void* out_jac, void* stream);

void bsc_msm_table_build_host(const void* bases_affine, long long nb, int w, int nw, void* out_jac);

int bsc_msm_table(const long long* Q, int ldq, const int* sel, int P, const int* sidx, const int* bidx, const int* og, int G, int T, const void* table_affine, int w, int nw, void* out_jac, long long ostride_p, void* stream);

// pcount (may be null): device-resident number of live rows (<= P); rows >= *pcount are skipped
int bsc_msm_table_dc(const long long* Q, int ldq, const int* sel, int P, const int* pcount, const int* sidx, const int* bidx, const int* og, int G, int T, const void* table_affine, int w, int nw, void* out_jac, long long ostride_p, void* stream);

void bsc_msm_table_host(const long long* Q, int ldq, const int* sel, int P, const int* sidx, const int* bidx, const int* og, int G, int T, const void* table_affine, int w, int nw, void* out_jac, long long ostride_p);

int bsc_g1_comb_mul(const void* scalars_u256, int N, const void* table_affine, int w, int nw, void* out_jac, void* stream);

int bsc_g1_fixed_base_mul(const void* scalars_u256, int N, const void* base_affine, void* out_jac, void* stream);

int bsc_g1_normalize(const void* in_jac, void* out_affine, long long N, void* stream);

int bsc_g1_affine_to_bytes(const void* in_affine, unsigned char* out, long long N, void* stream);

int bsc_g1_bytes_to_affine(const unsigned char* in, void* out_affine, long long N, void* stream);

int bsc_g1_sum_rows_dc(const void* pts_affine, const int* sel, int nsel, const int* nsel_dev, long long G, void* out_jac, void* stream);

int bsc_g1_sum_rows(const void* pts_affine, const int* sel, int nsel, long long G, void* out_jac, void* stream);

// nsel_dev (may be null): device-resident row count (<= nsel)
int bsc_g1_sum_rows_jac_dc(const void* pts_jac, const int* sel, int nsel, const int* nsel_dev, long long G, void* out_jac, void* stream);

int bsc_g1_sum_rows_jac(const void* pts_jac, const int* sel, int nsel, long long G, void* out_jac, void* stream);

// (workspace for the partial sums: grown on demand, one per device, kept for the life of the process; the first call
// on a device - or one with a larger P * ceil(G/128) than any before - allocates, so make it outside stream capture)
int bsc_g1_sum_cols_jac(const void* pts_jac, int P, int G, void* out_jac, void* stream);

void bsc_g1_sum_rows_jac_host(const void* pts_jac, const int* sel, int nsel, long long G, void* out_jac);

void bsc_g1_sum_cols_jac_host(const void* pts_jac, int P, int G, void* out_jac);

int bsc_g1_sum_cols(const void* pts_affine, int P, int G, void* out_jac, void* stream);

// ---- host mirrors (CPU plumbing path + oracles; same templates as the kernels) --------------------
void bsc_msm_groups_host(const long long* Q, int ldq, const int* sel, int P, const int* sidx, const int* bidx, const int* og, int G, int T, const void* bases_affine, void* out_jac, long long ostride_p);

void bsc_g1_normalize_host(const void* in_jac, void* out_affine, long long N);

void bsc_g1_sum_rows_host(const void* pts_affine, const int* sel, int nsel, long long G, void* out_jac);

void bsc_g1_sum_cols_host(const void* pts_affine, int P, int G, void* out_jac);

void bsc_g1_fixed_base_mul_host(const void* scalars_u256, int N, const void* base_affine, void* out_jac);

/* --------------------------------------------------------------------------------------------
 * vrf.cu - ECVRF, batched Ed25519 scalar multiplication, batched SHA-256, noiser lottery.  Reference: DistSys/vrf.go:16-100.
 * -------------------------------------------------------------------------------------------- */
void bsc_sha512(const uint8_t* data, size_t n, uint8_t out[64]);

// public key (32 bytes) from a 32-byte seed
void bsc_vrf_pubkey(const uint8_t seed[32], uint8_t pk[32]);

// expanded secret scalar (little-endian words) for the device kernel
void bsc_vrf_secret_scalar(const uint8_t seed[32], uint32_t out_words[8]);

// H = hashToCurve(m) in the device struct layout (4 x 8 words, Montgomery form)
void bsc_vrf_hash_to_curve(const uint8_t* m, size_t n, void* out_point, uint8_t out_bytes[32]);

// affine H = hashToCurve(m) as x || y, 32 canonical little-endian bytes each (input of the r25 kernel)
void bsc_vrf_hash_to_curve_xy(const uint8_t* m, size_t n, uint8_t out[64]);

// generic-field version of the same (oracle for the fast one)
void bsc_vrf_hash_to_curve_xy_generic(const uint8_t* m, size_t n, uint8_t out[64]);

// output = SHAKE256(Gamma || m)[:32]
void bsc_vrf_output_from_gamma(const uint8_t gamma[32], const uint8_t* m, size_t n, uint8_t out[32]);

void bsc_vrf_prove(const uint8_t seed[32], const uint8_t* m, size_t n, uint8_t vrf[32], uint8_t proof[96]);

int bsc_vrf_verify(const uint8_t pk[32], const uint8_t* m, size_t n, const uint8_t vrf[32], const uint8_t proof[96]);

// All workers' noiser committees in one call: output_p = SHAKE256(Gamma_p || m)[:32], then the stake
// lottery of vrf.go:54-100 seeded by output_p with p itself excluded.  out[p*want + j] = noiser id or -1.
void bsc_noisers_batch(const uint8_t* gammas, const uint8_t* m, size_t mlen, const int* stake, int n, const uint8_t* is_worker, int want, int* out, uint8_t* vrf_out);

// device: gamma[i] = compress(x[i] * H) for all peers
int bsc_ed25519_scalarmult_batch(const void* scalars_u256, const void* bases, int nbase, int N, unsigned char* out, void* stream);

int bsc_ed25519_scalarmult_r25(const unsigned char* scalars, const unsigned char* hxy, int N, unsigned char* out, void* stream);

int bsc_ed25519_scalarmult_quad(const unsigned char* scalars, const unsigned char* hxy, int N, unsigned char* out, void* stream);

// same, but leaves the results projective: xyz_out[i] = X, Y, Z as 3 x 10 radix-2^25.5 limbs
int bsc_ed25519_scalarmult_quad_xyz(const unsigned char* scalars, const unsigned char* hxy, int N, int* xyz_out, void* stream);

// warp-wide latency kernel (one peer per warp, one limb per lane); same contract as the quad _xyz entry point
int bsc_ed25519_scalarmult_wide_xyz(const unsigned char* scalars, const unsigned char* hxy, int N, int* xyz_out, void* stream);

// The head of a block's chain in one call: hash-to-curve of the previous block hash on the host (into a pinned 64-byte
// buffer), its upload, the scalar multiplication of every peer (wide != 0: warp-wide kernel, else four lanes per peer)
// and, when `event` is given, a record behind the kernel.  h_pinned must stay untouched until the copy has run.
int bsc_vrf_launch(const uint8_t* msg, int len, uint8_t* h_pinned, void* h_dev, const unsigned char* scalars, int N, int* xyz_out, int wide, void* stream, void* event);

// host: compress N projective points with a single field inversion (prefix products)
void bsc_ed25519_compress_batch(const int* xyz, int N, unsigned char* out);

void bsc_ed25519_scalarmult_r25_host(const unsigned char* scalars, const unsigned char* hxy, int N, unsigned char* out);

void bsc_ed25519_scalarmult_batch_host(const void* scalars_u256, const void* bases, int nbase, int N, unsigned char* out);

int bsc_sha256_batch(const unsigned char* msgs, int len, long long stride, int N, unsigned char* out, void* stream);

/* --------------------------------------------------------------------------------------------
 * host_crypto.cpp - Host crypto runtime (G1 byte API, SRS, hashes, Schnorr).  Reference: DistSys/kyber.go:873-933; DistSys/publicKey.go; lib/dedis/kyber/pairing/bn256/point.go:109-167; lib/dedis/kyber/xof/blake2xb/blake.go.
 * -------------------------------------------------------------------------------------------- */
void bsc_sha256(const uint8_t* data, size_t n, uint8_t out[32]);

void bsc_blake2b(const uint8_t* data, size_t n, const uint8_t* key, size_t keylen, uint8_t* out, size_t outlen);

void bsc_blake2xb(const uint8_t* seed, size_t seedlen, const uint8_t* msg, size_t n, uint8_t* out, size_t outlen);

void bsc_shake256(const uint8_t* data, size_t n, uint8_t* out, size_t outlen);

// ---- G1 byte API ---------------------------------------------------------------------------------
void bsc_g1_generator(uint8_t out[64]);

int bsc_g1_on_curve(const uint8_t in[64]);

void bsc_g1_add(uint8_t out[64], const uint8_t a[64], const uint8_t b[64]);

void bsc_g1_neg(uint8_t out[64], const uint8_t a[64]);

// scalar: 32-byte big-endian, reduced mod the group order first (mod.Int semantics)
void bsc_g1_mul(uint8_t out[64], const uint8_t in[64], const uint8_t k_be[32]);

// naive signed-int64 MSM oracle, one double-and-add per term like createCommitment (kyber.go:533-562)
void bsc_g1_msm_i64_host(uint8_t out[64], const uint8_t* bases64, const long long* scalars, int n);

// wire bytes <-> device layout (Montgomery affine, 16 x uint32)
void bsc_g1_bytes_to_affine_host(const uint8_t* in, void* out_affine, long long N);

void bsc_g1_affine_to_bytes_host(const void* in_affine, uint8_t* out, long long N);

void bsc_g1_jac_to_bytes_host(const void* in_jac, uint8_t* out, long long N);

// Jacobian -> 64-byte wire form for the selected rows, with one shared field inversion (Montgomery's trick).
// Used at block assembly: the device hands over the contributors' full commitments in Jacobian form and
// no inversion kernel sits on the block's critical path.  idx may be null (rows 0..n-1).
void bsc_g1_jac_to_bytes_batch_host(const void* in_jac, const int* idx, int n, uint8_t* out);

// host mirror of the device Straus kernel (same template) for CPU-side unit tests
void bsc_g1_straus_host(uint8_t out[64], const void* bases_affine, const long long* s, int T);

// SRS: PKG1[i] = tau^i * G for i < d (DistSys/publicKey.go:28-66).  tau = 2 reproduces the reference's
// hard-coded trapdoor (:36); any other 32-byte big-endian tau gives a proper setup.
void bsc_srs_g1_host(const uint8_t tau_be[32], int d, uint8_t* out);

// tau^i as plain little-endian 8x32 words, for the device fixed-base kernel
void bsc_fr_powers(const uint8_t tau_be[32], int d, uint32_t* out_words);

// Test hook for the two inversion routines of field.cuh: a^-1 mod m by binary extended Euclid (fe_inv) and by the
// reference's Fermat exponentiation (fe_inv_fermat).  which: 0 = BN256 base field, 1 = BN256 group order,
// 2 = 2^255-19.  Values are 32-byte big-endian plain integers (reduced mod m on input).
void bsc_fe_inv_both(int which, const uint8_t a_be[32], uint8_t out_fast[32], uint8_t out_fermat[32]);

// ---- Schnorr (kyber.go:873-933) ------------------------------------------------------------------
// pub = x*G
void bsc_schnorr_pubkey(const uint8_t x_be[32], uint8_t pub[64]);

// sig = c || r, c = H(T || msg), r = v - x*c ; v supplied by the caller (32 random bytes)
void bsc_schnorr_sign(const uint8_t* msg, size_t n, const uint8_t x_be[32], const uint8_t v_be[32], uint8_t sig[64]);

// Batched signing split in two so that the scalar multiplications T_i = v_i*G can run on the GPU
// (g1_comb_mul_kernel): (1) derive nonces, (2) given the T_i, finish c_i = H(T_i || msg_i), r_i = v_i - x_i*c_i.
// Nonce: v = SHA-256("bsc-schnorr-nonce" || x || tag) mod order, tag = caller-chosen unique bytes.
void bsc_schnorr_nonces(int n, const uint8_t* tags, int taglen, const uint8_t* x_be, uint32_t* v_words_le, uint8_t* v_be);

void bsc_schnorr_finish_batch(int n, const uint8_t* msgs, int msglen, const uint8_t* x_be, const uint8_t* v_be, const uint8_t* T, uint8_t* sigs);

int bsc_schnorr_verify(const uint8_t* msg, size_t n, const uint8_t pub[64], const uint8_t sig[64]);

/* --------------------------------------------------------------------------------------------
 * ledger.cpp - Ledger and stake lottery.  Reference: DistSys/block.go, blockData.go, update.go, blockchain.go, honest.go:631-653, vrf.go:54-182.
 * -------------------------------------------------------------------------------------------- */
void* bsc_chain_new(int d);

void bsc_chain_free(void* c);

int bsc_chain_len(void* c);

int bsc_chain_dim(void* c);

long long bsc_chain_latest_iteration(void* c);

// ---- block builder -------------------------------------------------------------------------------
void* bsc_block_new(long long iteration, const double* w, int d);

void bsc_block_free(void* b);

void bsc_block_add_update(void* bp, int source, int iteration, int accepted, const uint8_t* commitment, int clen, const double* delta, int dlen, const double* noise, int nlen, const double* noised, int ndlen, const uint8_t* sigs, int nsigs, int siglen);

// Batch form for secure-aggregation blocks: n accepted updates that carry only (source, commitment,
// signatures) — honest.go:391-440 appends exactly such commitment-only updates.  sigs may be null.
void bsc_block_add_commit_updates(void* bp, int n, const int* sources, int iteration, const uint8_t* commitments, int clen, const uint8_t* sigs, int nsigs_each, int siglen);

void bsc_block_set_stake(void* bp, const int* ids, const int* vals, int n);

// Seal against the chain tip and append (NewBlock + AddBlock, block.go:30-44, blockchain.go:12-16).
// timestamp < 0 => reference rule: 0 for an empty block, wall-clock seconds otherwise.
// Consumes the builder.  Returns the new chain length.
int bsc_chain_append(void* cp, void* bp, long long timestamp);

// The whole tail of a secure-aggregation block in one call (createBlockSecAgg + AddBlock, honest.go:391-440,
// block.go:30-44): the contributors' rows are picked out of the per-peer pack (64-byte commitment followed by
// the verifiers' 64-byte approvals, row = peer id), stake[source] += stake_unit, the block is sealed against
// the tip, appended and its hash returned.  `stake` (dense, node id = position) is updated in place and
// recorded in the block.  n == 0 mints the empty block (timestamp rule as in bsc_chain_append).
int bsc_chain_append_packed(void* cp, long long iteration, const double* w, int d, const int* sources, int n, const uint8_t* pack, long long pack_stride, int nsigs, int* stake, int nstake, int stake_unit, long long timestamp, uint8_t out_hash[32]);

// Deterministic stand-in for a network arrival order: a Fisher-Yates shuffle of 0..n-1 driven by splitmix64
// from `seed` (unbiased bounded draws by rejection).  Every rank computes the same permutation.
void bsc_seeded_perm(unsigned long long seed, int n, int* out);

// ---- a block's plan in one call --------------------------------------------------------------------------------
// Everything about block `it` that follows from the previous hash, the stake map and who is alive - what every peer
// of the reference derives for itself in prepareForNextIteration / getRoleNames (main.go:486-565, 1059-1187) and what
// the verifiers' and miners' arrival order decides (krum.go:296-314, 368-388; main.go:360, 1686, 2079):
//   verifiers (sorted), miners; workers = alive non-committee peers;
//   krum_set  = the first krum_thresh workers of the seeded "verify" arrival order, sorted by id; with -rs, a
//               permutation seeded by the iteration keeps num_samples of them;
//   share_rank[i] = position of candidate i in the seeded "shares" arrival order;
//   signers   = live verifiers; approved_ok = #signers >= floor(nv/2) and > 0;
//   shares_avail / xs = which of the S shares live miners hold, and the first ten of them.
// seed_verify / seed_shares / seed_rs are the 64-bit seeds of the three permutations (ops/chain.py: seed64).
// Arrays are caller-allocated with room for n entries (64 for the committees, S for shares_avail, 10 for xs).
// Returns 0; out_counts = {nv, nm, nworkers, nk, ns, approved_ok, enough_shares, clip}.
int bsc_plan_block(const uint8_t* prev_hash, int hash_len, const int* stake, const uint8_t* alive, int n, int want_v, int want_m, int krum_thresh, int num_samples, int random_sampling, unsigned long long seed_verify, unsigned long long seed_shares, unsigned long long seed_rs, double adversary_frac, int shares_per_miner, int S, int* verifiers, int* miners, int* workers, int* krum_set, int* share_rank, int* signers, uint8_t* shares_avail, int* xs, uint8_t* is_worker, int* stake_masked, int* out_counts);

// getBlock(it) = Blocks[it+1] with the consistency check (blockchain.go:77-96): index or -1/-2
int bsc_chain_block_index(void* cp, long long iteration);

void bsc_chain_hash(void* cp, int index, uint8_t out[32]);

void bsc_chain_prev_hash(void* cp, int index, uint8_t out[32], int* len);

void bsc_chain_weights(void* cp, int index, double* out);

long long bsc_chain_block_info(void* cp, int index, long long* iteration, int* nupdates, int* nstake);

void bsc_chain_block_stake(void* cp, int index, int* ids, int* vals);

int bsc_chain_update_info(void* cp, int index, int u, int* source, int* accepted, uint8_t commitment[64], int* nsigs);

// recompute every hash and the prev-hash links; 1 = consistent
int bsc_chain_verify(void* cp);

// serialisation (RegisterPeer chain transfer main.go:420-436; doubles as checkpoint format)
size_t bsc_chain_serialize(void* cp, uint8_t* out, size_t cap);

void* bsc_chain_deserialize(const uint8_t* in, size_t n);

// single-block wire form (RegisterBlock gossip, main.go:398-409)
size_t bsc_chain_block_serialize(void* cp, int index, uint8_t* out, size_t cap);

// Append or replace from a received block.  Return codes:
//   1 appended, 2 replaced an empty block by a non-empty one (honest.go:631-653), 0 duplicate/ignored,
//  -1 malformed, -2 does not link to our chain, -3 hash mismatch
int bsc_chain_accept_block(void* cp, const uint8_t* in, size_t n);

const char* bsc_chain_dump(void* cp, size_t* len);

int bsc_chain_equal(void* a, void* b);

const char* bsc_go_float(double v);

// roles: verifiers first, then miners continue from the same byte cursor (vrf.go:103-182)
int bsc_lottery_roles(const uint8_t* input, int inlen, const int* stake, int n, int nverifiers, int nminers, int* verifiers, int* miners, int* nv_out, int* nm_out);

// noisers from a peer's private VRF output, excluding itself (vrf.go:54-100)
int bsc_lottery_noisers(const uint8_t* vrf_out, int len, const int* stake, int n, int source, int want, int* noisers);

/* --------------------------------------------------------------------------------------------
 * exec.cu - 
 * -------------------------------------------------------------------------------------------- */
// Runs cmds[0..n) in order.  Returns 0, or (index + 1) * 100000 + |rc| of the first failing command.
int bsc_exec(const void* cmds_, int n, void* const* streams, int nstreams);

int bsc_cmd_sizeof();

// Events the command lists record and wait on (timing disabled: cheapest form).
void* bsc_event_create();

void bsc_event_destroy(void* e);

int bsc_event_record(void* e, void* stream);

int bsc_event_synchronize(void* e);

int bsc_event_query(void* e);

int bsc_stream_wait_event(void* stream, void* e);

/* --------------------------------------------------------------------------------------------
 * fused.cu - 
 * -------------------------------------------------------------------------------------------- */
// ---- symmetric memory plumbing (cudaMalloc + CUDA IPC; handles travel over torch.distributed) ----------
int bsc_symm_alloc(size_t bytes, void** out);

int bsc_symm_free(void* p);

int bsc_ipc_get_handle(void* p, unsigned char out[64]);

int bsc_ipc_open_handle(const unsigned char in[64], void** out);

int bsc_ipc_close_handle(void* p);

int bsc_symm_signal(void* const* flag_ptrs, int world, int rank, int slot, unsigned long long epoch, void* stream);

int bsc_symm_wait(const void* my_flags, int world, int slot, unsigned long long epoch, void* stream);

// dst_ptrs[q]: rank q's gather buffer [world][bytes_per_rank] (peer-mapped); src: this rank's block (16-byte aligned,
// bytes_per_rank a multiple of 16).  On return of the kernel every rank's block for `epoch` is in the local buffer.
int bsc_symm_allgather_rows(void* const* dst_ptrs, void* const* flag_ptrs, const void* my_flags, int world, int rank, const void* src, long long bytes_per_rank, int slot, unsigned long long epoch, void* stream);

// Fused all-gather + Gram.  plane_hi/lo[q]: rank q's [Sr][ldh] bf16 planes (peer-mapped), gpart[q]: rank q's
// [world][ldg][ldg] fp32 reduction buffer (slot r receives rank r's reduced partial), flags[q]: rank q's flag
// array, lpart: local scratch [ksplit][npairs][128*128] fp32, done_counter: local uint32[1 + npairs] (zeroed once;
// the counters only grow), seq: 1-based count of launches of this kernel on this rank.
// world*Sr must be a multiple of 128; Sr a multiple of 8 that divides 128, or a multiple of 128 (more than 128 peers
// per rank).
int bsc_krum_gram_fused(void* const* plane_hi, void* const* plane_lo, void* const* gpart, void* const* flags, int world, int rank, int Sr, int ldh, int ksplit, int ldg, float* lpart, unsigned int* done_counter, unsigned int seq, unsigned long long epoch, void* stream);

int bsc_krum_select_fused(const float* gpart_local, int nsplit, int ldg, const int* rows, int n, int clip, const void* my_flags, int world, unsigned long long epoch, float* scores, unsigned char* accept, void* stream);

// Fused reduce-scatter + reconstruct + SGD step + all-gather for this rank's chunk range [c0, c0+nc).
//   rm          : host RecoverMat (secagg_math.cuh) for the abscissae in use; use_mat = 0 keeps only rm->xs and
//                 recovers by divided differences
//   local_ready : local uint64 word (zeroed once), done_counter: local uint32 (zeroed once)
//   gate        : device int, may be null: the step is applied only when *gate >= gate_min
// When the kernel retires, every rank's new weight slice has landed in this rank's w64 / w32.
int bsc_secagg_fused(void* const* ypart, void* const* w64, void* const* w32, void* const* flags, int world, int rank, int S, int d, int c0, int nc, const void* rm, int use_mat, const unsigned char* avail_host, double scale, unsigned int* done_counter, unsigned long long* local_ready, int* bad, const int* gate, int gate_min, unsigned long long epoch, void* stream);

// first cross-GPU wait that timed out since the last call (0 = none); clears the record
int bsc_symm_error();

/* --------------------------------------------------------------------------------------------
 * kernels_chain.cu - 
 * -------------------------------------------------------------------------------------------- */
int bsc_vrf_noisers(const int* xyz, const unsigned char* msg, int mlen, const int* stake, int n, const unsigned char* is_worker, const int* peer_ids, int P, int want, int* out, unsigned char* vrf_out, void* stream);

int bsc_block_plan(const unsigned char* accept, const int* share_rank, const int* local_row, int nk, int approved_ok, int trigger, int* counts, int* contrib_pos, int* sel, int* appr_local, int* csel, void* stream);

// secrets_be: [n_peers][32] big-endian signing keys (device); salt64: 64 host bytes; k_out: [nc*ns] U256 (device)
int bsc_schnorr_nonces_dev(const unsigned char* secrets_be, const unsigned char* salt64, long long it, const int* cand_ids, int nc, const int* signers, int ns, void* k_out, void* stream);

int bsc_schnorr_finish_dev(const unsigned char* T, const void* k, const unsigned char* secrets_be, const unsigned char* cbytes, const int* cand_rows, int nc, const int* signers, int ns, unsigned char* pack, int pack_stride, void* stream);

int bsc_pack_commitments(const unsigned char* cbytes, int P, unsigned char* pack, int pack_stride, void* stream);

// host mirror of hash_schnorr_64 + finish, for tests: sig = c || r
void bsc_schnorr_finish64_host(const unsigned char* T64, const unsigned char* k_words, const unsigned char* x_be, const unsigned char* msg64, unsigned char* sig);

/* --------------------------------------------------------------------------------------------
 * kernels_cnn.cu - 
 * -------------------------------------------------------------------------------------------- */
// test hook: 1 = keep the one-conv models on the generic kernels (the shape-specialised ones are compared against them)
void bsc_cnn_force_generic(int on);

// test hook: 1 = conv weight gradient on the SIMT kernel instead of the tensor-core one
void bsc_cnn_force_wgrad_simt(int on);

// scratch: act [P*B*J] floats, dl [P*B*C], partial [P*nslots] with nslots = ceil(J/256) + OC, coef [P],
// fpart [ceil(J/1024)][P*B][C]
long long bsc_cnn1_scratch_floats(int P, int B, int IC, int OC, int K, int pad, int H, int W, int C);

// One-conv-layer CNN local step for P peers.  w: flat global weights; X: [P][B][IC][H][W]; y: [P][B];
// U: [P][ldu] out (= -clip(grad)); loss: [P] (may be null); scratch: see bsc_cnn1_scratch_floats.
int bsc_cnn1_local_step(const float* w, const float* X, const int* y, float* U, float* loss, float* scratch, int P, int B, int IC, int OC, int K, int pad, int H, int W, int C, int ldu, float clip, void* stream);

// Error count of one weight vector over N images with the one-conv model (K14 for the CNN families; the reference
// runs the torch module, ML/Pytorch/client.py:136-172).  wrong: int32[1] out (zeroed here); scratch:
// bsc_cnn1_eval_scratch_floats floats.  Only the two reference shapes with 10 classes are supported (-1 otherwise: the
// caller falls back to the framework's convolution).
long long bsc_cnn1_eval_scratch_floats(int N, int IC, int OC, int K, int pad, int H, int W, int C);

int bsc_cnn1_eval_errors(const float* w, const float* X, const int* y, int* wrong, float* scratch, int N, int IC, int OC, int K, int pad, int H, int W, int C, void* stream);

// scratch (floats): pool1 [PB*n1], dp1 [PB*n1], act [PB*J], dl [PB*C], partial [P*nslots], coef [P],
// arg1 bytes [PB*n1], arg2 bytes [PB*J] (each rounded up to whole floats)
long long bsc_cnn2_scratch_floats(int P, int B, int IC, int H, int W, int OC1, int K1, int pad1, int OC2, int K2, int pad2, int C);

// Local step of the two-convolution model for P peers.  w: flat global weights in named_parameters order
// (conv0.weight, conv0.bias, conv1.weight, conv1.bias, fc.weight, fc.bias); X: [P][B][IC][H][W]; y: [P][B];
// U: [P][ldu] out (= -clip(grad)); loss: [P] or null; scratch: bsc_cnn2_scratch_floats floats.
int bsc_cnn2_local_step(const float* w, const float* X, const int* y, float* U, float* loss, float* scratch, int P, int B, int IC, int H, int W, int OC1, int K1, int pad1, int OC2, int K2, int pad2, int C, int ldu, float clip, void* stream);

/* --------------------------------------------------------------------------------------------
 * pairing.cpp - 
 * -------------------------------------------------------------------------------------------- */
// ---- G2 (129-byte wire form; a first byte of 0x00 is the point at infinity) ---------------------------
void bsc_g2_generator(uint8_t* out129);

int bsc_g2_on_curve(const uint8_t* in129);

void bsc_g2_add(uint8_t* out129, const uint8_t* a129, const uint8_t* b129);

void bsc_g2_neg(uint8_t* out129, const uint8_t* a129);

// scalar: 32-byte big-endian, taken as is (callers reduce mod the group order)
void bsc_g2_mul(uint8_t* out129, const uint8_t* pt129, const uint8_t* k32);

// PKG2[i] = tau^i * G2, i < d (DistSys/publicKey.go:42-57) -> d * 129 bytes
void bsc_srs_g2_host(const uint8_t* tau32, int d, uint8_t* out);

// ---- GT / pairing --------------------------------------------------------------------------------------
void bsc_pairing(const uint8_t* g1_64, const uint8_t* g2_129, uint8_t* out384);

// 1 iff prod_i e(P_i, Q_i) == 1; one final exponentiation for the whole product
int bsc_pairing_check(const uint8_t* g1s, const uint8_t* g2s, int n);

void bsc_gt_mul(uint8_t* out384, const uint8_t* a384, const uint8_t* b384);

void bsc_gt_pow(uint8_t* out384, const uint8_t* a384, const uint8_t* k32);

// verifySecret (DistSys/kyber.go:650-673): e(C, g2) == e(W, tau*g2 - x*g2) * e(B, g2)^y, checked as
// e(C - y*B, g2) * e(-W, tau*g2 - x*g2) == 1.  B is the first SRS point of the chunk's slice (g1 for chunk 0,
// which is the only case the reference's formula covers; chunk c is committed against PKG1[10c..], kyber.go:172-202).
int bsc_kzg_verify_share(const uint8_t* commit64, const uint8_t* witness64, const uint8_t* base64, const uint8_t* g2_129, const uint8_t* tau_g2_129, int x, long long y);

/* --------------------------------------------------------------------------------------------
 * sha256_ni.cpp - 
 * -------------------------------------------------------------------------------------------- */
void bsc_sha256_fast(const uint8_t* data, size_t n, uint8_t out[32]);

int bsc_sha256_has_ni();

#ifdef __cplusplus
}
#endif
#endif /* BISCOTTI_B200_H */
