"""Device-resident chain links (csrc/kernels_chain.cu, the *_dc kernel variants, reconstruct_mat_kernel) against
their host restatements, and the pipelined engine schedule against the in-line one."""
import ctypes
import hashlib

import numpy as np
import pytest
import torch

from biscotti_b200.config import BiscottiConfig
from biscotti_b200.crypto import primitives as cp
from biscotti_b200.ops import chain as chain_ops, secagg, ml
from biscotti_b200.ops._call import lib, ptr
from biscotti_b200.protocol.engine import BiscottiEngine

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda:0")


def test_vrf_noisers_kernel_matches_host_lottery():
    """Inversion + compression + SHAKE256 + stake lottery in one kernel vs coniks-style prove() and the host lottery."""
    n = 45
    keys = [cp.VRFKey(hashlib.sha256(b"nz%d" % i).digest()) for i in range(n)]
    msg = hashlib.sha256(b"prev-block-hash").digest()
    stake = [10 + 5 * (i % 5) for i in range(n)]
    stake[11] = 0                                              # a dead peer holds no tickets
    is_worker = np.ones(n, dtype=np.uint8)
    is_worker[[2, 9, 30]] = 0
    bv = cp.BatchedVRF(keys, DEV)
    xyz, ev, _ = bv.launch_dev(msg)
    ev.wait()
    local = list(range(7, 40))                                 # a rank's contiguous slice of the peers
    out = torch.full((len(local), 3), -7, dtype=torch.int32, device=DEV)
    vout = torch.zeros(len(local), 32, dtype=torch.uint8, device=DEV)
    chain_ops.vrf_noisers(xyz, torch.tensor(list(msg), dtype=torch.uint8, device=DEV), 32,
                          torch.tensor(stake, dtype=torch.int32, device=DEV), torch.from_numpy(is_worker).to(DEV),
                          torch.tensor(local, dtype=torch.int32, device=DEV), 3, out, vout)
    out, vout = out.cpu().numpy(), vout.cpu().numpy()
    for r, p in enumerate(local):
        vrf = keys[p].prove(msg)[0]
        assert bytes(vout[r]) == vrf
        want = cp.lottery_noisers(vrf, stake, p, 3) if is_worker[p] else []
        assert [q for q in out[r].tolist() if q >= 0] == want, (p, out[r], want)


def test_wide_vrf_kernel_equals_quad_kernel_and_generic_prove():
    """One-limb-per-lane Ed25519 scalar multiplication (the head of every block's chain) vs the quad kernel and the
    generic-field prove(): identical Gamma for every peer, including tiny and near-order scalars."""
    import ctypes

    n = 67
    keys = [cp.VRFKey(hashlib.sha256(b"w%d" % i).digest()) for i in range(n)]
    L = lib()
    for tag in (b"blk-a", b"blk-b"):
        msg = hashlib.sha256(tag).digest()
        bw, bq = cp.BatchedVRF(keys, DEV), cp.BatchedVRF(keys, DEV, kernel="quad4")
        for edge, val in ((0, 1), (1, 2), (2, 8), (3, 2**252 + 27742317777372353535851937790883648492), (4, 2**255 - 1)):
            w = np.frombuffer(int(val).to_bytes(32, "little"), dtype=np.int32).copy()
            bw.scalars[edge] = torch.from_numpy(w).to(DEV)
            bq.scalars[edge] = torch.from_numpy(w).to(DEV)
        xw, ew, _ = bw.launch_dev(msg)
        xq, eq, _ = bq.launch_dev(msg)
        ew.synchronize()
        eq.synchronize()
        gw, gq = np.empty((n, 32), dtype=np.uint8), np.empty((n, 32), dtype=np.uint8)
        hw, hq = xw.cpu().numpy().copy(), xq.cpu().numpy().copy()
        L.bsc_ed25519_compress_batch(ctypes.c_void_p(hw.ctypes.data), n, ctypes.c_void_p(gw.ctypes.data))
        L.bsc_ed25519_compress_batch(ctypes.c_void_p(hq.ctypes.data), n, ctypes.c_void_p(gq.ctypes.data))
        assert (gw == gq).all()
        assert bytes(gw[9]) == keys[9].prove(msg)[1][64:]          # proof = c || t || Gamma


def test_block_plan_kernel_matches_host():
    rng = np.random.RandomState(3)
    for nk, trigger, ok in ((70, 35, True), (5, 2, True), (140, 70, True), (33, 10, False), (1, 1, True)):
        accept = (rng.rand(nk) < 0.6).astype(np.uint8)
        rank = rng.permutation(nk).astype(np.int32)
        rows = np.where(rng.rand(nk) < 0.5, np.arange(nk), -1).astype(np.int32)
        rows[rows >= 0] = np.arange((rows >= 0).sum())
        out = dict(counts=torch.full((8,), -1, dtype=torch.int32, device=DEV),
                   contrib_pos=torch.zeros(nk, dtype=torch.int32, device=DEV), sel=torch.zeros(nk, dtype=torch.int32, device=DEV),
                   appr_local=torch.zeros(nk, dtype=torch.int32, device=DEV), csel=torch.zeros(nk, dtype=torch.int32, device=DEV))
        chain_ops.block_plan(torch.from_numpy(accept).to(DEV), torch.from_numpy(rank).to(DEV), torch.from_numpy(rows).to(DEV),
                             ok, trigger, out)
        counts, cpos, sel, appr, csel = chain_ops.block_plan_host(accept.tolist(), rank.tolist(), rows.tolist(), ok, trigger)
        got = out["counts"].cpu().tolist()
        assert got[:4] == counts
        assert out["contrib_pos"].cpu().tolist()[: counts[1]] == cpos
        assert out["sel"].cpu().tolist()[: counts[2]] == sel and out["csel"].cpu().tolist()[: counts[2]] == csel
        assert out["appr_local"].cpu().tolist()[: counts[3]] == appr


def test_device_approvals_match_host_and_verify():
    """nonce kernel + comb kernel + finish kernel (BLAKE2Xb challenge, r = v - x c) vs the host signer, byte for byte."""
    from biscotti_b200.crypto.schnorr_batch import ApprovalSigner

    n = 12
    secrets = [hashlib.sha256(b"sk%d" % i).digest() for i in range(n)]
    signer = ApprovalSigner(DEV, secrets)
    salt = hashlib.sha256(b"run").digest() + hashlib.sha256(b"prev").digest()
    cand = [1, 4, 5, 9, 11]
    rows = [0, -1, 2, 3, 5]                     # candidate 4 belongs to another rank
    signers = [2, 7, 8]
    cbytes = torch.randint(0, 256, (6, 64), dtype=torch.uint8, device=DEV)
    pack = torch.zeros(6, 64 * 4, dtype=torch.uint8, device=DEV)
    signer.launch(salt, 17, torch.tensor(cand, dtype=torch.int32, device=DEV), torch.tensor(rows, dtype=torch.int32, device=DEV),
                  len(cand), torch.tensor(signers, dtype=torch.int32, device=DEV), len(signers), cbytes, pack, after=None,
                  serial=True)
    pack_h, cb_h = pack.cpu().numpy(), cbytes.cpu().numpy()
    mine = [(p, r) for p, r in zip(cand, rows) if r >= 0]
    host = signer.sign_host(salt, 17, [p for p, _ in mine], signers, [cb_h[r].tobytes() for _, r in mine])
    for i, (p, r) in enumerate(mine):
        assert pack_h[r, :64].tobytes() == cb_h[r].tobytes()
        for j, v in enumerate(signers):
            sig = pack_h[r, 64 + 64 * j: 128 + 64 * j].tobytes()
            assert sig == host[i * len(signers) + j]
            assert cp.schnorr_verify(cb_h[r].tobytes(), cp.schnorr_pubkey(secrets[v]), sig)
    assert not pack_h[1].any() and not pack_h[4].any()       # rows nobody signed stay untouched


def test_reconstruct_mat_kernel_equals_divided_difference_kernel_and_honours_the_gate():
    d, S, P = 7850, 21, 9
    nc = secagg.nchunks(d)
    torch.manual_seed(5)
    U = torch.zeros(P, ml.pad_to(d, 4), device=DEV)
    U[:, :d] = torch.randn(P, d, device=DEV) * 0.3
    sel = torch.tensor([0, 2, 3, 7, 8, 1, 1, 1, 1], dtype=torch.int32, device=DEV)
    cnt = torch.tensor([5], dtype=torch.int32, device=DEV)
    Y = secagg.share_sum(U, sel, d, S, count=cnt)                       # only the first five entries are live
    assert torch.equal(Y, secagg.share_sum(U, sel[:5].contiguous(), d, S))
    for xs in (list(range(10)), [0, 1, 2, 3, 4, 5, 6, 14, 15, 16]):
        xs_t = torch.tensor(xs, dtype=torch.int32, device=DEV)
        av = torch.ones(S, dtype=torch.uint8, device=DEV)
        w_ref = torch.randn(d, dtype=torch.float64, device=DEV)
        w_mat, w32 = w_ref.clone(), torch.zeros(d, device=DEV)
        _, bad_ref = secagg.reconstruct_step(Y, xs_t, av, d, w_ref, None)
        bad = torch.zeros(4, dtype=torch.int32, device=DEV)
        secagg.reconstruct_step_dev(Y, xs, xs_t, av, d, w_mat, w32, bad=bad, gate=cnt, gate_min=2)
        assert torch.equal(w_mat, w_ref) and torch.equal(w32, w_ref.float())
        assert int(bad_ref.sum()) == 0 and int(bad.sum()) == 0
        closed = w_mat.clone()
        secagg.reconstruct_step_dev(Y, xs, xs_t, av, d, closed, None, bad=bad, gate=cnt, gate_min=6)   # 5 < 6: no step
        assert torch.equal(closed, w_mat)
        Yb = Y.clone()
        Yb[11, 3] += 1
        secagg.reconstruct_step_dev(Yb, xs, xs_t, av, d, closed, None, bad=bad, gate=None)
        assert int(bad[:2].sum()) > 0


def test_witness_and_aggregate_kernels_read_their_row_counts_on_the_device():
    from biscotti_b200.ops.msm import CommitmentEngine

    d, S = 60, 21
    eng = CommitmentEngine(d, S, tau=2, device=DEV, window=5)
    torch.manual_seed(2)
    U = torch.zeros(6, ml.pad_to(d, 4), device=DEV)
    U[:, :d] = torch.randn(6, d, device=DEV)
    Q = secagg.quantize(U, d, 4)
    sel = torch.tensor([4, 0, 3, 5, 5, 5], dtype=torch.int32, device=DEV)
    cnt = torch.tensor([3], dtype=torch.int32, device=DEV)
    W = eng.witnesses(Q, sel, affine=False, count=cnt, max_rows=5)
    Wref = eng.witnesses(Q, sel[:3].contiguous(), affine=False)
    assert W.shape[0] == 5 and torch.equal(W[:3], Wref)
    csel = torch.tensor([0, 2, 1, 1], dtype=torch.int32, device=DEV)
    c2 = torch.tensor([2], dtype=torch.int32, device=DEV)
    agg = eng.aggregate(W, csel, affine=True, count=c2)
    ref = eng.aggregate(Wref, csel[:2].contiguous(), affine=True)
    assert torch.equal(agg, ref)


def test_pipelined_schedule_gives_the_in_line_chain():
    """Chain / off-chain / approval streams change when kernels run, not what they compute."""
    base = dict(num_nodes=30, dataset="mnist", max_iterations=5, shard_size=100, test_size=500, msm_window=7)
    a = BiscottiEngine(BiscottiConfig(pipeline=True, **base), device=DEV)
    b = BiscottiEngine(BiscottiConfig(pipeline=False, prefetch=False, **base), device=DEV)
    c = BiscottiEngine(BiscottiConfig(pipeline=True, lookahead=False, **base), device=DEV)
    e = BiscottiEngine(BiscottiConfig(pipeline=True, native_schedule=False, **base), device=DEV)   # Python-issued launches
    checked = 0
    for _ in range(6):
        ra, rb, rc, re_ = a.step(), b.step(), c.step(), e.step()
        assert ra.block_hash == rb.block_hash == rc.block_hash == re_.block_hash and ra.contributors == rb.contributors
        assert abs(ra.test_error - rb.test_error) < 1e-9 and ra.attack_rate == re_.attack_rate
        if not ra.empty:
            assert a.verify_last_aggregate() is True and b.verify_last_aggregate() is True
            checked += 1
    torch.cuda.synchronize()
    assert checked >= 4 and a.chain == b.chain == c.chain and a._side is not None and b._side is None
    assert a._ctx_next is not None and c._ctx_next is None          # a really ran one block ahead
    assert a._ns is not None and e._ns is None and a.chain == e.chain   # native command lists vs launches from Python
    # every block's approvals verify against the verifiers' public keys
    rep = next(r for r in a.reports if not r.empty)
    info = a.chain.block_info(rep.iteration + 1)
    assert all(u["nsigs"] == len(rep.verifiers) for u in info["updates"])


def test_lookahead_is_discarded_when_the_world_changes_between_steps():
    """The next block's chain front is queued a block ahead; if peers die (or come back) before the next step() the
    stale plan must be thrown away and the block replanned - same ledger as an engine that never looked ahead."""
    base = dict(num_nodes=24, dataset="mnist", max_iterations=8, shard_size=60, test_size=200, msm_window=6)
    a = BiscottiEngine(BiscottiConfig(pipeline=True, **base), device=DEV)
    b = BiscottiEngine(BiscottiConfig(pipeline=False, prefetch=False, **base), device=DEV)
    for it in range(7):
        if it == 3:
            for e in (a, b):
                for p in (2, 5, 6, 11, 17):
                    e.alive[p] = False
        if it == 5:
            for e in (a, b):
                e.alive[5] = e.alive[11] = True
        ra, rb = a.step(), b.step()
        assert ra.block_hash == rb.block_hash, it
        assert torch.equal(a.w64, b.w64)
    assert a.chain == b.chain and a.stake == b.stake
