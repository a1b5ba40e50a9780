// Host ledger runtime: blocks, chain, canonical serialisation, hashing, PrintChain-style dump,
// longest-chain adoption / fork repair helpers, and the stake lottery.
//
// Behavioural parity with the reference:
//   Block / BlockData / Update fields        DistSys/block.go:14-20, blockData.go:10-14, update.go:13-22
//   Hash = SHA256(prev || decimal(ts) || enc) DistSys/block.go:23-28  (enc: see encode_block_data —
//                                             a documented canonical encoding instead of Go gob)
//   empty blocks carry Timestamp 0            DistSys/block.go:32-38
//   genesis = iteration -1, zero weights      DistSys/block.go:46-53
//   getBlock(it) = Blocks[it+1] + check       DistSys/blockchain.go:77-96
//   PrintChain dump                           DistSys/blockchain.go:43-53, blockData.go:24-27, update.go:24-37
//   better-block replacement rule             DistSys/honest.go:631-653
//   lottery                                   DistSys/vrf.go:54-182 (bit-exact: it is the consensus rule)
#include <stdint.h>
#include <string.h>
#include <stdio.h>
#include <math.h>
#include <time.h>
#include <string>
#include <vector>
#include <algorithm>
#include <map>
#include <charconv>
#include "sha256.cuh"

using namespace bsc;

namespace bsc {
void sha256_fast(const uint8_t* data, size_t n, uint8_t out[32]);   // sha256_ni.cpp (SHA-NI when available)
}

namespace {

struct Update {
  int32_t source = 0;
  int32_t iteration = 0;
  uint8_t accepted = 0;
  std::string commitment;               // 64-byte G1 wire bytes (may be empty)
  std::vector<double> delta;            // only in the plain (-sa=false) path
  std::vector<double> noise;            // optional
  std::vector<double> noised;           // optional
  std::vector<std::string> signatures;  // Schnorr c||r blobs
};

struct Block {
  int64_t timestamp = 0;
  int64_t iteration = -1;
  std::vector<double> w;
  std::vector<Update> updates;
  std::string prev_hash;
  std::string hash;
  std::map<int, int> stake;
};

struct Chain {
  std::vector<Block> blocks;
};

void put_u32(std::string& o, uint32_t v) {
  for (int i = 0; i < 4; i++) o.push_back((char)(v >> (8 * i)));
}
void put_u64(std::string& o, uint64_t v) {
  for (int i = 0; i < 8; i++) o.push_back((char)(v >> (8 * i)));
}
void put_f64s(std::string& o, const std::vector<double>& v) {
  put_u64(o, v.size());
  const size_t off = o.size();
  o.resize(off + 8 * v.size());
  memcpy(&o[off], v.data(), 8 * v.size());  // little-endian hosts only (x86-64 / aarch64)
}
void put_bytes(std::string& o, const std::string& b) {
  put_u32(o, (uint32_t)b.size());
  o += b;
}

// Canonical BlockData encoding ("BSCD1"): the hash preimage component that stands in for
// gob(BlockData) (blockData.go:31-41).  Field order follows the Go struct order.
void encode_block_data_to(std::string& o, const Block& b) {
  o += "BSCD1";
  put_u64(o, (uint64_t)b.iteration);
  put_f64s(o, b.w);
  put_u32(o, (uint32_t)b.updates.size());
  for (const Update& u : b.updates) {
    put_u32(o, (uint32_t)u.source);
    put_u32(o, (uint32_t)u.iteration);
    put_f64s(o, u.delta);
    put_bytes(o, u.commitment);
    put_f64s(o, u.noise);
    put_f64s(o, u.noised);
    o.push_back((char)u.accepted);
    put_u32(o, (uint32_t)u.signatures.size());
    for (const std::string& s : u.signatures) put_bytes(o, s);
  }
}
std::string encode_block_data(const Block& b) {
  std::string o;
  encode_block_data_to(o, b);
  return o;
}

void set_hash(Block& b) {
  std::string pre;
  pre.reserve(128 + 8 * b.w.size() + 512 * b.updates.size());   // one buffer: no second copy of the weights
  pre += b.prev_hash;
  pre += std::to_string((long long)b.timestamp);
  encode_block_data_to(pre, b);
  uint8_t h[32];
  sha256_fast((const uint8_t*)pre.data(), pre.size(), h);
  b.hash.assign((const char*)h, 32);
}

// full block wire encoding (chain transfer / checkpoint)
std::string encode_block(const Block& b) {
  std::string o = "BSCB1";
  put_u64(o, (uint64_t)b.timestamp);
  put_bytes(o, b.prev_hash);
  put_bytes(o, b.hash);
  put_u32(o, (uint32_t)b.stake.size());
  for (auto& kv : b.stake) {
    put_u32(o, (uint32_t)kv.first);
    put_u32(o, (uint32_t)kv.second);
  }
  put_bytes(o, encode_block_data(b));
  return o;
}

struct Reader {
  const uint8_t* p;
  size_t n, off = 0;
  bool ok = true;
  bool need(size_t k) {
    if (off + k > n) ok = false;
    return ok;
  }
  uint32_t u32() {
    if (!need(4)) return 0;
    uint32_t v = 0;
    for (int i = 0; i < 4; i++) v |= (uint32_t)p[off + i] << (8 * i);
    off += 4;
    return v;
  }
  uint64_t u64() {
    if (!need(8)) return 0;
    uint64_t v = 0;
    for (int i = 0; i < 8; i++) v |= (uint64_t)p[off + i] << (8 * i);
    off += 8;
    return v;
  }
  std::string bytes() {
    uint32_t k = u32();
    if (!need(k)) return std::string();
    std::string s((const char*)p + off, k);
    off += k;
    return s;
  }
  std::vector<double> f64s() {
    uint64_t k = u64();
    std::vector<double> v;
    if (k > (n - off) / 8) {
      ok = false;
      return v;
    }
    v.resize(k);
    memcpy(v.data(), p + off, 8 * k);
    off += 8 * k;
    return v;
  }
  bool tag(const char* t) {
    size_t k = strlen(t);
    if (!need(k) || memcmp(p + off, t, k) != 0) {
      ok = false;
      return false;
    }
    off += k;
    return true;
  }
};

bool decode_block_data(Reader& r, Block& b) {
  if (!r.tag("BSCD1")) return false;
  b.iteration = (int64_t)r.u64();
  b.w = r.f64s();
  uint32_t nu = r.u32();
  for (uint32_t i = 0; i < nu && r.ok; i++) {
    Update u;
    u.source = (int32_t)r.u32();
    u.iteration = (int32_t)r.u32();
    u.delta = r.f64s();
    u.commitment = r.bytes();
    u.noise = r.f64s();
    u.noised = r.f64s();
    if (r.need(1)) u.accepted = r.p[r.off++];
    uint32_t ns = r.u32();
    for (uint32_t j = 0; j < ns && r.ok; j++) u.signatures.push_back(r.bytes());
    b.updates.push_back(std::move(u));
  }
  return r.ok;
}

bool decode_block(Reader& r, Block& b) {
  if (!r.tag("BSCB1")) return false;
  b.timestamp = (int64_t)r.u64();
  b.prev_hash = r.bytes();
  b.hash = r.bytes();
  uint32_t ns = r.u32();
  for (uint32_t i = 0; i < ns && r.ok; i++) {
    int k = (int)r.u32();
    int v = (int)r.u32();
    b.stake[k] = v;
  }
  std::string bd = r.bytes();
  Reader r2{(const uint8_t*)bd.data(), bd.size()};
  return r.ok && decode_block_data(r2, b);
}

// Go fmt %v for float64 = strconv 'g' with the shortest round-trip digits; the %e form is used when
// exp < -4 || exp >= 6 (shortest mode fixes the decision precision at 6), e.g. 1e+06, 1.5e-05.
std::string go_float(double v) {
  if (v != v) return "NaN";
  if (isinf(v)) return v > 0 ? "+Inf" : "-Inf";
  if (v == 0) return signbit(v) ? "-0" : "0";
  char buf[64];
  auto res = std::to_chars(buf, buf + sizeof(buf), v, std::chars_format::scientific);
  std::string s(buf, res.ptr);  // d.ddddde[+-]XX shortest
  std::string sign;
  if (s[0] == '-') {
    sign = "-";
    s = s.substr(1);
  }
  size_t epos = s.find('e');
  std::string mant = s.substr(0, epos);
  int exp = atoi(s.c_str() + epos + 1);
  std::string digits;
  for (char c : mant)
    if (c != '.') digits.push_back(c);
  if (exp < -4 || exp >= 6) {
    std::string o = sign + digits.substr(0, 1);
    if (digits.size() > 1) o += "." + digits.substr(1);
    char eb[16];
    snprintf(eb, sizeof(eb), "e%c%02d", exp < 0 ? '-' : '+', abs(exp));
    return o + eb;
  }
  std::string o = sign;
  if (exp < 0) {
    o += "0.";
    o.append((size_t)(-exp - 1), '0');
    o += digits;
  } else if ((int)digits.size() <= exp + 1) {
    o += digits;
    o.append((size_t)(exp + 1 - (int)digits.size()), '0');
  } else {
    o += digits.substr(0, (size_t)exp + 1) + "." + digits.substr((size_t)exp + 1);
  }
  return o;
}

std::string hex(const std::string& b) {
  static const char* d = "0123456789abcdef";
  std::string o;
  for (unsigned char c : b) {
    o.push_back(d[c >> 4]);
    o.push_back(d[c & 15]);
  }
  return o;
}

std::string array_to_string(const std::vector<double>& a) {
  std::string o = "[";
  for (size_t i = 0; i < a.size(); i++) {
    if (i) o += ",";
    o += go_float(a[i]);
  }
  return o + "]";
}

// bn256.G1 point String(): "bn256.G1:(x, y)" with 64-hex-digit coordinates; the point at infinity
// prints as (0, 1) after MakeAffine (curve.go:202-224, gfp.go:19-21).
std::string point_string(const std::string& c) {
  std::string x(32, '\0'), y(32, '\0');
  bool inf = true;
  if (c.size() == 64) {
    for (unsigned char ch : c) inf = inf && ch == 0;
    if (!inf) {
      x = c.substr(0, 32);
      y = c.substr(32, 32);
    }
  }
  if (inf) y[31] = 1;
  return "bn256.G1:(" + hex(x) + ", " + hex(y) + ")";
}

std::string dump_chain(const Chain& c) {
  std::string o;
  for (const Block& b : c.blocks) {
    o += "Prev. hash: " + hex(b.prev_hash) + "\n";
    o += "Data: Iteration: " + std::to_string((long long)b.iteration) + ", GlobalW: " + array_to_string(b.w) +
         ", deltas: [";
    for (size_t i = 0; i < b.updates.size(); i++) {
      const Update& u = b.updates[i];
      o += "{Iteration:" + std::to_string(u.iteration) + ", Commitment:" + point_string(u.commitment) +
           ", Deltas:" + array_to_string(u.delta) + "}";
      if (i + 1 != b.updates.size()) o += " ,";
    }
    o += "]\n";
    o += "Hash: " + hex(b.hash) + "\n\n";
  }
  return o;
}

Block genesis(int d) {
  Block b;
  b.timestamp = 0;
  b.iteration = -1;
  b.w.assign((size_t)d, 0.0);
  set_hash(b);
  return b;
}

thread_local std::string g_scratch;

}  // namespace

extern "C" {

void* bsc_chain_new(int d) {
  Chain* c = new Chain();
  c->blocks.push_back(genesis(d));
  return c;
}
void bsc_chain_free(void* c) { delete (Chain*)c; }
int bsc_chain_len(void* c) { return (int)((Chain*)c)->blocks.size(); }
int bsc_chain_dim(void* c) { return (int)((Chain*)c)->blocks.back().w.size(); }
long long bsc_chain_latest_iteration(void* c) { return ((Chain*)c)->blocks.back().iteration; }

// ---- block builder -------------------------------------------------------------------------------
void* bsc_block_new(long long iteration, const double* w, int d) {
  Block* b = new Block();
  b->iteration = iteration;
  b->w.assign(w, w + d);
  return b;
}
void bsc_block_free(void* b) { delete (Block*)b; }
void bsc_block_add_update(void* bp, int source, int iteration, int accepted, const uint8_t* commitment, int clen,
                          const double* delta, int dlen, const double* noise, int nlen, const double* noised,
                          int ndlen, const uint8_t* sigs, int nsigs, int siglen) {
  Block* b = (Block*)bp;
  Update u;
  u.source = source;
  u.iteration = iteration;
  u.accepted = (uint8_t)accepted;
  if (commitment && clen > 0) u.commitment.assign((const char*)commitment, (size_t)clen);
  if (delta && dlen > 0) u.delta.assign(delta, delta + dlen);
  if (noise && nlen > 0) u.noise.assign(noise, noise + nlen);
  if (noised && ndlen > 0) u.noised.assign(noised, noised + ndlen);
  for (int i = 0; i < nsigs; i++) u.signatures.emplace_back((const char*)sigs + (size_t)i * siglen, (size_t)siglen);
  b->updates.push_back(std::move(u));
}
// Batch form for secure-aggregation blocks: n accepted updates that carry only (source, commitment,
// signatures) — honest.go:391-440 appends exactly such commitment-only updates.  sigs may be null.
void bsc_block_add_commit_updates(void* bp, int n, const int* sources, int iteration, const uint8_t* commitments,
                                  int clen, const uint8_t* sigs, int nsigs_each, int siglen) {
  Block* b = (Block*)bp;
  b->updates.reserve(b->updates.size() + (size_t)n);
  for (int i = 0; i < n; i++) {
    Update u;
    u.source = sources[i];
    u.iteration = iteration;
    u.accepted = 1;
    if (commitments && clen > 0) u.commitment.assign((const char*)commitments + (size_t)i * clen, (size_t)clen);
    for (int k = 0; k < nsigs_each; k++)
      u.signatures.emplace_back((const char*)sigs + ((size_t)i * nsigs_each + k) * siglen, (size_t)siglen);
    b->updates.push_back(std::move(u));
  }
}
void bsc_block_set_stake(void* bp, const int* ids, const int* vals, int n) {
  Block* b = (Block*)bp;
  b->stake.clear();
  for (int i = 0; i < n; i++) b->stake[ids[i]] = vals[i];
}
// Seal against the chain tip and append (NewBlock + AddBlock, block.go:30-44, blockchain.go:12-16).
// timestamp < 0 => reference rule: 0 for an empty block, wall-clock seconds otherwise.
// Consumes the builder.  Returns the new chain length.
int bsc_chain_append(void* cp, void* bp, long long timestamp) {
  Chain* c = (Chain*)cp;
  Block* b = (Block*)bp;
  if (timestamp < 0) timestamp = b->updates.empty() ? 0 : (long long)time(nullptr);
  b->timestamp = timestamp;
  b->prev_hash = c->blocks.back().hash;
  set_hash(*b);
  c->blocks.push_back(std::move(*b));
  delete b;
  return (int)c->blocks.size();
}

// The whole tail of a secure-aggregation block in one call (createBlockSecAgg + AddBlock, honest.go:391-440,
// block.go:30-44): the contributors' rows are picked out of the per-peer pack (64-byte commitment followed by
// the verifiers' 64-byte approvals, row = peer id), stake[source] += stake_unit, the block is sealed against
// the tip, appended and its hash returned.  `stake` (dense, node id = position) is updated in place and
// recorded in the block.  n == 0 mints the empty block (timestamp rule as in bsc_chain_append).
int bsc_chain_append_packed(void* cp, long long iteration, const double* w, int d, const int* sources, int n,
                            const uint8_t* pack, long long pack_stride, int nsigs, int* stake, int nstake, int stake_unit,
                            long long timestamp, uint8_t out_hash[32]) {
  Chain* c = (Chain*)cp;
  Block* b = new Block();
  b->iteration = iteration;
  b->w.assign(w, w + d);
  b->updates.reserve((size_t)n);
  for (int i = 0; i < n; i++) {
    Update u;
    u.source = sources[i];
    u.iteration = (int)iteration;
    u.accepted = 1;
    const uint8_t* row = pack + (size_t)sources[i] * (size_t)pack_stride;
    u.commitment.assign((const char*)row, 64);
    for (int k = 0; k < nsigs; k++) u.signatures.emplace_back((const char*)row + 64 + 64 * (size_t)k, (size_t)64);
    b->updates.push_back(std::move(u));
    if (sources[i] >= 0 && sources[i] < nstake) stake[sources[i]] += stake_unit;
  }
  for (int i = 0; i < nstake; i++) b->stake[i] = stake[i];
  bsc_chain_append(cp, b, timestamp);
  memcpy(out_hash, c->blocks.back().hash.data(), 32);
  return (int)c->blocks.size();
}

// Deterministic stand-in for a network arrival order: a Fisher-Yates shuffle of 0..n-1 driven by splitmix64
// from `seed` (unbiased bounded draws by rejection).  Every rank computes the same permutation.
void bsc_seeded_perm(unsigned long long seed, int n, int* out) {
  auto next = [&seed]() {
    unsigned long long z = (seed += 0x9E3779B97F4A7C15ULL);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
  };
  for (int i = 0; i < n; i++) out[i] = i;
  for (int i = n - 1; i > 0; i--) {
    const unsigned long long bound = (unsigned long long)i + 1;
    const unsigned long long lim = ~0ULL - (~0ULL % bound);
    unsigned long long r;
    do {
      r = next();
    } while (r >= lim);
    const int j = (int)(r % bound);
    const int t = out[i];
    out[i] = out[j];
    out[j] = t;
  }
}

int bsc_lottery_roles(const uint8_t* input, int inlen, const int* stake, int n, int nverifiers, int nminers,
                      int* verifiers, int* miners, int* nv_out, int* nm_out);

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
int bsc_plan_block(const uint8_t* prev_hash, int hash_len, const int* stake, const uint8_t* alive, int n, int want_v,
                   int want_m, int krum_thresh, int num_samples, int random_sampling, unsigned long long seed_verify,
                   unsigned long long seed_shares, unsigned long long seed_rs, double adversary_frac, int shares_per_miner,
                   int S, int* verifiers, int* miners, int* workers, int* krum_set, int* share_rank, int* signers,
                   uint8_t* shares_avail, int* xs, uint8_t* is_worker, int* stake_masked, int* out_counts) {
  std::vector<int> st((size_t)n);
  long long total = 0;
  for (int i = 0; i < n; i++) {
    st[(size_t)i] = alive[i] ? stake[i] : 0;
    stake_masked[i] = st[(size_t)i];
    total += st[(size_t)i];
    is_worker[i] = 0;
  }
  for (int k = 0; k < 8; k++) out_counts[k] = 0;
  for (int s = 0; s < S; s++) shares_avail[s] = 0;
  if (total == 0) return 0;
  int nv = 0, nm = 0;
  bsc_lottery_roles(prev_hash, hash_len, st.data(), n, want_v, want_m, verifiers, miners, &nv, &nm);
  std::sort(verifiers, verifiers + nv);     // getRoleNames sorts verifiers (main.go:529-565)
  std::vector<char> committee((size_t)n, 0);
  for (int i = 0; i < nv; i++) committee[(size_t)verifiers[i]] = 1;
  for (int i = 0; i < nm; i++) committee[(size_t)miners[i]] = 1;
  int nw = 0;
  for (int p = 0; p < n; p++)
    if (!committee[(size_t)p] && alive[p]) {
      workers[nw++] = p;
      is_worker[p] = 1;
    }
  out_counts[0] = nv;
  out_counts[1] = nm;
  out_counts[2] = nw;
  if (nv == 0 || nm == 0 || nw == 0) return 0;
  std::vector<int> perm((size_t)n);
  bsc_seeded_perm(seed_verify, nw, perm.data());
  int nk = nw < krum_thresh ? nw : krum_thresh;
  for (int i = 0; i < nk; i++) krum_set[i] = workers[perm[(size_t)i]];
  std::sort(krum_set, krum_set + nk);
  if (random_sampling && nk > num_samples) {
    std::vector<int> keep((size_t)num_samples);
    bsc_seeded_perm(seed_rs, nk, perm.data());
    for (int i = 0; i < num_samples; i++) keep[(size_t)i] = krum_set[perm[(size_t)i]];
    std::sort(keep.begin(), keep.end());
    nk = num_samples;
    for (int i = 0; i < nk; i++) krum_set[i] = keep[(size_t)i];
  }
  bsc_seeded_perm(seed_shares, nk, perm.data());
  for (int r = 0; r < nk; r++) share_rank[perm[(size_t)r]] = r;
  int ns = 0;
  for (int i = 0; i < nv; i++)
    if (alive[verifiers[i]]) signers[ns++] = verifiers[i];
  int have = 0, live_miners = 0;
  for (int j = 0; j < nm; j++) {
    if (!alive[miners[j]]) continue;
    live_miners++;
    for (int s = j * shares_per_miner; s < (j + 1) * shares_per_miner && s < S; s++) shares_avail[s] = 1;
  }
  int nx = 0;
  for (int s = 0; s < S; s++)
    if (shares_avail[s]) {
      have++;
      if (nx < 10) xs[nx++] = s;
    }
  out_counts[3] = nk;
  out_counts[4] = ns;
  out_counts[5] = (ns > 0 && ns >= nv / 2) ? 1 : 0;
  out_counts[6] = (have >= 10 && live_miners > 0) ? 1 : 0;
  out_counts[7] = (int)(adversary_frac * (double)nk);     // floor
  return 0;
}

// getBlock(it) = Blocks[it+1] with the consistency check (blockchain.go:77-96): index or -1/-2
int bsc_chain_block_index(void* cp, long long iteration) {
  Chain* c = (Chain*)cp;
  if ((long long)c->blocks.size() >= iteration + 2 && iteration + 1 >= 0) {
    if (c->blocks[(size_t)(iteration + 1)].iteration != iteration) return -2;
    return (int)(iteration + 1);
  }
  return -1;
}
void bsc_chain_hash(void* cp, int index, uint8_t out[32]) {
  Chain* c = (Chain*)cp;
  const Block& b = index < 0 ? c->blocks.back() : c->blocks[(size_t)index];
  memcpy(out, b.hash.data(), 32);
}
void bsc_chain_prev_hash(void* cp, int index, uint8_t out[32], int* len) {
  Chain* c = (Chain*)cp;
  const Block& b = index < 0 ? c->blocks.back() : c->blocks[(size_t)index];
  *len = (int)b.prev_hash.size();
  memcpy(out, b.prev_hash.data(), b.prev_hash.size());
}
void bsc_chain_weights(void* cp, int index, double* out) {
  Chain* c = (Chain*)cp;
  const Block& b = index < 0 ? c->blocks.back() : c->blocks[(size_t)index];
  memcpy(out, b.w.data(), 8 * b.w.size());
}
long long bsc_chain_block_info(void* cp, int index, long long* iteration, int* nupdates, int* nstake) {
  Chain* c = (Chain*)cp;
  const Block& b = index < 0 ? c->blocks.back() : c->blocks[(size_t)index];
  *iteration = b.iteration;
  *nupdates = (int)b.updates.size();
  *nstake = (int)b.stake.size();
  return b.timestamp;
}
void bsc_chain_block_stake(void* cp, int index, int* ids, int* vals) {
  Chain* c = (Chain*)cp;
  const Block& b = index < 0 ? c->blocks.back() : c->blocks[(size_t)index];
  int i = 0;
  for (auto& kv : b.stake) {
    ids[i] = kv.first;
    vals[i] = kv.second;
    i++;
  }
}
int bsc_chain_update_info(void* cp, int index, int u, int* source, int* accepted, uint8_t commitment[64], int* nsigs) {
  Chain* c = (Chain*)cp;
  const Block& b = index < 0 ? c->blocks.back() : c->blocks[(size_t)index];
  if (u < 0 || u >= (int)b.updates.size()) return -1;
  const Update& up = b.updates[(size_t)u];
  *source = up.source;
  *accepted = up.accepted;
  *nsigs = (int)up.signatures.size();
  memset(commitment, 0, 64);
  memcpy(commitment, up.commitment.data(), up.commitment.size() < 64 ? up.commitment.size() : 64);
  return (int)up.delta.size();
}

// recompute every hash and the prev-hash links; 1 = consistent
int bsc_chain_verify(void* cp) {
  Chain* c = (Chain*)cp;
  for (size_t i = 0; i < c->blocks.size(); i++) {
    Block t = c->blocks[i];
    set_hash(t);
    if (t.hash != c->blocks[i].hash) return 0;
    if (i > 0 && c->blocks[i].prev_hash != c->blocks[i - 1].hash) return 0;
  }
  return 1;
}

// serialisation (RegisterPeer chain transfer main.go:420-436; doubles as checkpoint format)
size_t bsc_chain_serialize(void* cp, uint8_t* out, size_t cap) {
  Chain* c = (Chain*)cp;
  std::string o = "BSCC1";
  put_u32(o, (uint32_t)c->blocks.size());
  for (const Block& b : c->blocks) put_bytes(o, encode_block(b));
  if (out && cap >= o.size()) memcpy(out, o.data(), o.size());
  return o.size();
}
void* bsc_chain_deserialize(const uint8_t* in, size_t n) {
  Reader r{in, n};
  if (!r.tag("BSCC1")) return nullptr;
  uint32_t nb = r.u32();
  Chain* c = new Chain();
  for (uint32_t i = 0; i < nb && r.ok; i++) {
    std::string bb = r.bytes();
    Reader rb{(const uint8_t*)bb.data(), bb.size()};
    Block b;
    if (!decode_block(rb, b)) r.ok = false;
    c->blocks.push_back(std::move(b));
  }
  if (!r.ok || c->blocks.empty()) {
    delete c;
    return nullptr;
  }
  return c;
}
// single-block wire form (RegisterBlock gossip, main.go:398-409)
size_t bsc_chain_block_serialize(void* cp, int index, uint8_t* out, size_t cap) {
  Chain* c = (Chain*)cp;
  const Block& b = index < 0 ? c->blocks.back() : c->blocks[(size_t)index];
  std::string o = encode_block(b);
  if (out && cap >= o.size()) memcpy(out, o.data(), o.size());
  return o.size();
}
// Append or replace from a received block.  Return codes:
//   1 appended, 2 replaced an empty block by a non-empty one (honest.go:631-653), 0 duplicate/ignored,
//  -1 malformed, -2 does not link to our chain, -3 hash mismatch
int bsc_chain_accept_block(void* cp, const uint8_t* in, size_t n) {
  Chain* c = (Chain*)cp;
  Reader r{in, n};
  Block b;
  if (!decode_block(r, b)) return -1;
  Block t = b;
  set_hash(t);
  if (t.hash != b.hash) return -3;
  const long long idx = b.iteration + 1;
  if (idx == (long long)c->blocks.size()) {
    if (b.prev_hash != c->blocks.back().hash) return -2;
    c->blocks.push_back(std::move(b));
    return 1;
  }
  if (idx > 0 && idx < (long long)c->blocks.size()) {
    Block& mine = c->blocks[(size_t)idx];
    if (mine.hash == b.hash) return 0;
    const bool better = b.prev_hash == mine.prev_hash && mine.updates.empty() && !b.updates.empty();
    if (better && idx + 1 == (long long)c->blocks.size()) {
      mine = std::move(b);
      return 2;
    }
    return 0;
  }
  return -2;
}

const char* bsc_chain_dump(void* cp, size_t* len) {
  g_scratch = dump_chain(*(Chain*)cp);
  *len = g_scratch.size();
  return g_scratch.data();
}
int bsc_chain_equal(void* a, void* b) {
  Chain* x = (Chain*)a;
  Chain* y = (Chain*)b;
  if (x->blocks.size() != y->blocks.size()) return 0;
  for (size_t i = 0; i < x->blocks.size(); i++)
    if (x->blocks[i].hash != y->blocks[i].hash) return 0;
  return 1;
}
const char* bsc_go_float(double v) {
  g_scratch = go_float(v);
  return g_scratch.c_str();
}

// ---- stake lottery (DistSys/vrf.go:54-182) ---------------------------------------------------------
// tickets: node id repeated stake[id] times, ids 0..n-1.  Draw = big-endian 2 bytes at cursor i mod
// #tickets, cursor advances by one byte, SHA-256 re-hash of the current buffer when i >= len-1,
// duplicates (and `exclude`) skipped.  `cursor_io`/`buf` carry the state from the verifier draw into
// the miner draw exactly like getVRFRoles does.
// The ticket list in closed form: pre[id] = tickets of ids < id, pre[n] = total.  ticket t belongs to the largest id
// with pre[id] <= t (ids without stake share their successor's offset and are never the largest).  Same draws as the
// reference's materialised list, without O(total stake) work per draw - stake grows by 5 per contribution per block.
struct Tickets {
  std::vector<long long> pre;
  int n = 0;
  int distinct = 0;
  explicit Tickets(const int* stake, int n_) : pre((size_t)n_ + 1, 0), n(n_) {
    for (int i = 0; i < n_; i++) {
      pre[(size_t)i + 1] = pre[(size_t)i] + (stake[i] > 0 ? stake[i] : 0);
      distinct += stake[i] > 0;
    }
  }
  long long total() const { return pre[(size_t)n]; }
  bool has(int id) const { return id >= 0 && id < n && pre[(size_t)id + 1] > pre[(size_t)id]; }
  int owner(long long t) const {
    int lo = 0, hi = n - 1;
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (pre[(size_t)mid] <= t) lo = mid; else hi = mid - 1;
    }
    return lo;
  }
};

static int lottery_draw(std::vector<uint8_t>& buf, size_t& i, const Tickets& tk, int want, int exclude, int* out) {
  // the reference loops forever if `want` exceeds the distinct eligible ids; bound it instead
  const int distinct = tk.distinct - (tk.has(exclude) ? 1 : 0);
  if (want > distinct) want = distinct;
  int got = 0;
  const unsigned long long total = (unsigned long long)tk.total();
  while (got < want) {
    if (i + 1 >= buf.size()) {
      uint8_t h[32];
      sha256_oneshot(buf.data(), buf.size(), h);
      buf.assign(h, h + 32);
      i = 0;
    }
    const int winner = tk.owner((long long)((((unsigned long long)buf[i] * 256ull + buf[i + 1])) % total));
    bool dup = winner == exclude;
    for (int k = 0; k < got; k++) dup = dup || out[k] == winner;
    if (!dup) out[got++] = winner;
    i++;
  }
  return got;
}

// roles: verifiers first, then miners continue from the same byte cursor (vrf.go:103-182)
int bsc_lottery_roles(const uint8_t* input, int inlen, const int* stake, int n, int nverifiers, int nminers,
                      int* verifiers, int* miners, int* nv_out, int* nm_out) {
  Tickets tk(stake, n);
  if (tk.total() == 0) return -1;
  std::vector<uint8_t> buf(input, input + inlen);
  size_t i = 0;
  *nv_out = lottery_draw(buf, i, tk, nverifiers, -1, verifiers);
  *nm_out = lottery_draw(buf, i, tk, nminers, -1, miners);
  return 0;
}
// noisers from a peer's private VRF output, excluding itself (vrf.go:54-100)
int bsc_lottery_noisers(const uint8_t* vrf_out, int len, const int* stake, int n, int source, int want, int* noisers) {
  Tickets tk(stake, n);
  if (tk.total() == 0) return 0;
  std::vector<uint8_t> buf(vrf_out, vrf_out + len);
  size_t i = 0;
  return lottery_draw(buf, i, tk, want, source, noisers);
}

}  // extern "C"
