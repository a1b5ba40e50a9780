"""Host restatements of the device-resident chain links, the protocol knobs that ride on them (-rs sampling,
DP_IN_MODEL, fork repair) and the constant-matrix share recovery — CPU tier."""
import ctypes
import hashlib

import numpy as np
import torch

from biscotti_b200.config import BiscottiConfig
from biscotti_b200.ops import chain as chain_ops, secagg
from biscotti_b200.ops._call import lib, ptr
from biscotti_b200.protocol.engine import BiscottiEngine


def _cfg(**kw):
    base = dict(num_nodes=10, dataset="creditcard", max_iterations=3, shard_size=100, test_size=300)
    base.update(kw)
    return BiscottiConfig(**base)


def test_seeded_perm_is_a_permutation_and_depends_on_the_seed():
    a = chain_ops.seeded_perm(chain_ops.seed64(1, 5, "verify"), 97)
    b = chain_ops.seeded_perm(chain_ops.seed64(1, 5, "verify"), 97)
    c = chain_ops.seeded_perm(chain_ops.seed64(1, 6, "verify"), 97)
    assert sorted(a.tolist()) == list(range(97)) and a.tolist() == b.tolist() and a.tolist() != c.tolist()


def test_block_plan_host_semantics():
    """approved = accepted (when enough verifiers live); contributors = the approved among the first `trigger` of the
    share-arrival order; lists come out in ascending candidate order; local lists index this rank's rows."""
    accept = [1, 0, 1, 1, 1, 1]
    rank = [5, 0, 1, 4, 2, 3]            # arrival order of the candidates: 1, 2, 4, 5, 3, 0
    rows = [0, -1, 1, -1, 2, 3]          # candidates 1 and 3 live on another rank
    counts, cpos, sel, appr, csel = chain_ops.block_plan_host(accept, rank, rows, True, 3)
    assert counts == [5, 3, 3, 4]
    assert cpos == [2, 4, 5]             # the three earliest approved arrivals, ascending
    assert appr == [0, 1, 2, 3] and sel == [1, 2, 3] and csel == [1, 2, 3]
    assert chain_ops.block_plan_host(accept, rank, rows, False, 3)[0] == [0, 0, 0, 0]   # no quorum of verifiers


def test_random_sampling_keeps_num_samples_candidates():
    """-rs (krum.go:310, 368-388): Krum waits for N-nv-na updates, then keeps NUM_SAMPLES of them by a permutation
    seeded with the iteration; the sample is the same on every rank and changes from block to block."""
    cfg = _cfg(num_nodes=20, perc_samples=40, random_sampling=True, max_iterations=3)
    assert cfg.krum_thresh == 14 and cfg.num_samples == 8
    a, b = BiscottiEngine(cfg), BiscottiEngine(cfg, rank=1, world=2)
    sets = []
    for _ in range(3):
        pa, pb = a._host_plan(), b._host_plan()
        assert pa.krum_set == pb.krum_set and len(pa.krum_set) == 8 and pa.clip == 4
        sets.append(tuple(pa.krum_set))
        r = a.step()
        assert r.n_krum == 8
        b.chain, b.iteration, b.stake = a.chain, a.iteration, list(a.stake)
    assert len(set(sets)) > 1
    plain = BiscottiEngine(_cfg(num_nodes=20, perc_samples=40, max_iterations=3))
    assert len(plain._host_plan().krum_set) == 8          # without -rs: the first NUM_SAMPLES arrivals


def test_dp_in_model_puts_own_noise_into_the_committed_delta():
    """DP_IN_MODEL (honest.go:172-179): on the PyTorch path the peer's own noise becomes part of Delta."""
    base = dict(num_nodes=12, dataset="lfw_softmax", shard_size=40, test_size=60, max_iterations=1, commitments=False)
    a = BiscottiEngine(BiscottiConfig(**base))
    b = BiscottiEngine(BiscottiConfig(dp_in_model=True, **base))
    a._local_and_commit(None, 0, timed=False)
    b._local_and_commit(None, 0, timed=False)
    diff = (b.U - a.U)[:, : a.d]
    assert float(diff.abs().max()) > 0
    from biscotti_b200.ops import ml
    own = ml.noise_row(a.d, a.cfg.seed ^ 0x6F776E6E, a.local_ids[3], 0, a.cfg.noise_scale)
    assert torch.allclose(diff[3], own, atol=1e-6)
    ra, rb = a.step(), b.step()
    assert ra.block_hash != rb.block_hash


def test_recover_matrix_equals_divided_differences():
    rng = np.random.RandomState(0)
    for xs in ([0, 1, 2, 3, 4, 5, 6, 7, 8, 9], [0, 1, 2, 3, 4, 5, 6, 14, 15, 16], [7, 8, 9, 10, 11, 12, 13, 14, 15, 16]):
        rm = secagg.recover_matrix(xs)
        assert rm is not None and rm.D > 0
        nc, S = 40, 21
        Q = torch.from_numpy(rng.randint(-2**24, 2**24, size=(3, nc * 10)).astype(np.int64))
        Y = secagg.shares_eval(Q, S).sum(dim=0).contiguous()
        coef = torch.empty(nc, 10, dtype=torch.int64)
        assert lib().bsc_recover_mat_host(ptr(Y), ctypes.byref(rm), S, nc, ptr(coef)) == 0
        assert torch.equal(coef.reshape(-1), Q.sum(dim=0))
        ref = torch.empty(nc, 10, dtype=torch.int64)
        xs_t = torch.tensor(xs, dtype=torch.int32)
        lib().bsc_shamir_recover_host(ptr(Y), ptr(xs_t), None, S, nc, ptr(ref))
        assert torch.equal(coef, ref)
        Y[3, xs[2]] += 1                                     # a corrupted share is reported as inexact
        assert lib().bsc_recover_mat_host(ptr(Y), ctypes.byref(rm), S, nc, ptr(coef)) > 0


def test_approval_nonces_never_repeat_across_runs_or_forks():
    """The nonce tag binds the run (seed, data, flags) and the previous block hash: two runs that reuse the same keys
    on different data, or two forks of one run, never sign with the same T = v*G."""
    from biscotti_b200.crypto.schnorr_batch import SchnorrBatchSigner

    a = BiscottiEngine(_cfg(max_iterations=1))
    b = BiscottiEngine(_cfg(max_iterations=1, epsilon=1.0))          # same seed => same keys, different run
    assert a.keys[0].schnorr_secret == b.keys[0].schnorr_secret and a.run_salt != b.run_salt
    s = SchnorrBatchSigner("cpu")
    h1, h2 = hashlib.sha256(b"fork-1").digest(), hashlib.sha256(b"fork-2").digest()
    sk = a.keys[0].schnorr_secret
    nonces = set()
    for salt in (a.run_salt + h1, a.run_salt + h2, b.run_salt + h1):
        h = s.launch([sk], [chain_ops.approval_tag(salt, 3, 5, 0)])
        nonces.add(bytes(h["vbe"][:32]))
    assert len(nonces) == 3


def test_fork_repair_replaces_an_empty_block_reached_from_the_engine():
    """A peer cut off at height k mints the deterministic empty block on its block deadline (main.go:2326-2355);
    when the partition heals and it is offered the real block of height k, evaluateBlockQuality
    (honest.go:631-653) replaces the empty one, and the peer continues on the common chain."""
    a, b = BiscottiEngine(_cfg(max_iterations=5)), BiscottiEngine(_cfg(max_iterations=5))
    for _ in range(2):
        a.step()
        b.step()
    ra = a.step()                                   # block 2 as the rest of the network sees it
    assert not ra.empty
    rb = b.mint_empty_block()                       # b saw nothing before its deadline
    assert rb.empty and b.chain != a.chain and len(b.chain) == len(a.chain)
    assert b.offer_block(a.chain.serialize_block(-1)) == 2
    assert b.chain == a.chain and b.iteration == a.iteration and b.stake == a.stake
    assert np.allclose(b.w64.numpy(), a.w64.numpy())
    assert b.offer_block(a.chain.serialize_block(-1)) == 0      # offered again: nothing to repair
    a.step()
    b.step()
    assert a.chain == b.chain and a.chain.verify()


def test_synthetic_data_models_and_poisoned_shards():
    """Both generators give per-peer shards of the reference's shapes; a poisoner's shard holds nothing but source-class
    rows carrying the target label (mnist_bad, parse_mnist.py:295-301), and poisoners are the ids above
    ceil(N * (1 - po)) (main.go:836-844).  On "lowrank" data rows of one class are close to each other and far from
    the other classes, which is what lets Multi-Krum see a poisoner; on "blobs" data they are not."""
    import torch

    from biscotti_b200.config import BiscottiConfig
    from biscotti_b200.data.synthetic import make_data, poisoner_ids
    from biscotti_b200.ops import krum as krum_ops, ml

    for model in ("blobs", "lowrank"):
        cfg = BiscottiConfig(num_nodes=20, dataset="mnist", shard_size=60, test_size=200, poison_fraction=0.3, data_model=model)
        assert poisoner_ids(cfg) == {15, 16, 17, 18, 19}
        d = make_data(cfg, list(range(20)))
        assert d.X.shape == (20, 60, 784) and d.y.shape == (20, 60) and d.X_test.shape == (200, 784)
        assert bool(d.poisoned[15]) and not bool(d.poisoned[14])
        assert (d.y[15:] == 7).all() and (d.y_attack == 1).all()
        assert len(set(d.y[0].tolist())) > 3                       # honest shards mix the classes
    # lowrank: the poisoned rows are class-1 rows (nearest class mean is that of class 1)
    cfg = BiscottiConfig(num_nodes=20, dataset="mnist", shard_size=60, test_size=400, poison_fraction=0.3, data_model="lowrank")
    d = make_data(cfg, list(range(20)))
    means = torch.stack([d.X_test[d.y_test == c].mean(0) for c in range(10)])
    nearest = torch.cdist(d.X[15], means).argmin(dim=1)
    assert (nearest == 1).float().mean() > 0.8
    # ... and with no DP noise Multi-Krum rejects exactly the poisoners' updates at the zero model
    w = torch.zeros(cfg.dim)
    idx = torch.arange(10, dtype=torch.int32).repeat(20, 1)
    U, _ = ml.softmax_step(w, d.X, d.y, idx, 10, 100.0)
    _, acc, _ = krum_ops.krum_reference(U[:, : cfg.dim], 10)
    assert not acc[15:].any() and int(acc.sum()) == 10
