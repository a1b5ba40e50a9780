"""Fixed-point, Shamir, commitment and witness algebra on the CPU mirrors (same C++ as the kernels).

Oracles: exact integer restatements of DistSys/kyber.go:579-609 (share evaluation), :244-287
(aggregation), :809-857 (recovery), :533-562 / :579-646 (commitment, witness = commit(quotient))."""
import ctypes

import numpy as np
import torch

from biscotti_b200.ops import secagg
from biscotti_b200.ops.msm import CommitmentEngine
from biscotti_b200.oracle import bn256 as O


def test_quantize_truncates_toward_zero():
    U = torch.tensor([[-0.01059, 0.01669, 0.00491, -1.99999, 2.99999, 0.0]], dtype=torch.float32)
    Q = secagg.quantize(U, 6, 4)
    # int64(x*10^4) like Go (kyber.go:698-710); float32 inputs are widened to float64 first
    exp = [int(float(v) * 1e4) for v in U[0].double().tolist()]
    assert Q[0, :6].tolist() == exp
    assert Q.shape[1] == 10 and Q[0, 6:].abs().sum() == 0


def test_shares_are_exact_polynomial_evaluations_and_additive():
    g = torch.Generator().manual_seed(0)
    Q = torch.randint(-10**6, 10**6, (5, 30), generator=g, dtype=torch.int64)
    S = 21
    Y = secagg.shares_eval(Q, S)
    for p in range(5):
        for c in range(3):
            a = [int(v) for v in Q[p, 10 * c:10 * c + 10]]
            for s in (0, 7, 10, 20):
                x = s - 10
                assert int(Y[p, c, s]) == sum(a[j] * x ** j for j in range(10))
    # x = 0 share equals the first coefficient
    assert torch.equal(Y[:, :, 10], Q.view(5, 3, 10)[:, :, 0])
    # additive homomorphism: shares of the sum == sum of the shares
    Ysum = Y.sum(0)
    Yof = secagg.shares_eval(Q.sum(0, keepdim=True), S)[0]
    assert torch.equal(Ysum, Yof)


def test_recover_from_any_ten_shares_and_detect_corruption(lib):
    g = torch.Generator().manual_seed(1)
    Q = torch.randint(-10**7, 10**7, (1, 40), generator=g, dtype=torch.int64)
    S = 21
    Y = secagg.shares_eval(Q, S)[0].contiguous()            # [4, 21]
    for xs in ([0, 1, 2, 3, 4, 5, 6, 7, 8, 9], [0, 1, 2, 3, 4, 5, 6, 14, 15, 16], [20, 18, 16, 14, 12, 10, 8, 6, 4, 2]):
        coef = torch.empty(4, 10, dtype=torch.int64)
        xs_t = torch.tensor(xs, dtype=torch.int32)
        bad = lib.bsc_shamir_recover_host(ctypes.c_void_p(Y.data_ptr()), ctypes.c_void_p(xs_t.data_ptr()), None, S, 4,
                                          ctypes.c_void_p(coef.data_ptr()))
        assert bad == 0
        assert torch.equal(coef.view(-1), Q[0])
    Yc = Y.clone()
    Yc[2, 15] += 1
    coef = torch.empty(4, 10, dtype=torch.int64)
    xs_t = torch.arange(10, dtype=torch.int32)
    bad = lib.bsc_shamir_recover_host(ctypes.c_void_p(Yc.data_ptr()), ctypes.c_void_p(xs_t.data_ptr()), None, S, 4,
                                      ctypes.c_void_p(coef.data_ptr()))
    assert bad == 1


def test_reconstruct_step_applies_sum_of_deltas():
    torch.manual_seed(2)
    d = 25
    U = (torch.randn(6, 28) * 0.05).float()
    sel = torch.tensor([0, 2, 5], dtype=torch.int32)
    Ysum = secagg.share_sum(U, sel, d, 21)
    w64 = torch.zeros(d, dtype=torch.float64)
    w32 = torch.zeros(d)
    xs = torch.arange(10, dtype=torch.int32)
    _, bad = secagg.reconstruct_step(Ysum, xs, None, d, w64, w32)
    assert int(bad.sum()) == 0
    exp = secagg.quantize(U[sel.long()].contiguous(), d)[:, :d].sum(0).double() / 1e4
    assert torch.equal(w64, exp)


def _oracle_points(bytes_):
    return [O.from_bytes(bytes_[64 * i:64 * i + 64]) for i in range(len(bytes_) // 64)]


def test_commitments_and_witnesses_match_kzg_algebra():
    d, S, tau = 25, 21, 2
    eng = CommitmentEngine(d, S, tau=tau, device="cpu")
    srs = _oracle_points(eng.srs_bytes)
    assert srs == O.srs(tau, d)
    g = torch.Generator().manual_seed(3)
    Q = torch.zeros(2, 30, dtype=torch.int64)
    Q[:, :d] = torch.randint(-5 * 10**5, 5 * 10**5, (2, d), generator=g)
    C = eng.chunk_commitments(Q)
    Cb = eng.to_bytes(C)
    full = eng.to_bytes(eng.full_commitments(C))
    W = eng.witnesses(Q)
    Wb = eng.to_bytes(W)
    for p in range(2):
        a = [int(v) for v in Q[p]]
        # chunk commitment = MSM over the SRS slice [10c, 10c+10) (kyber.go:172-202)
        acc = None
        for c in range(3):
            hi = min(d, 10 * c + 10)
            cc = O.msm(srs[10 * c:hi], a[10 * c:hi])
            assert Cb[64 * (p * 3 + c):64 * (p * 3 + c) + 64] == O.to_bytes(cc)
            acc = O.add(acc, cc)
        # chunk commitments sum to the full-vector commitment
        assert full[64 * p:64 * p + 64] == O.to_bytes(acc) == O.to_bytes(O.msm(srs, a[:d]))
        # witness = commitment to the quotient (p(X) - p(x)) / (X - x) with the same SRS slice
        for c in (0, 2):
            coeffs = a[10 * c:10 * c + 10]
            for s in (0, 10, 20):
                x = s - 10
                q = [sum(coeffs[j] * x ** (j - k - 1) for j in range(k + 1, 10)) for k in range(9)]
                base = srs[10 * c:10 * c + 9] + [O.GEN] * max(0, 10 * c + 9 - d)
                wq = O.msm(base[:9], q)
                off = 64 * ((p * 3 + c) * S + s)
                assert Wb[off:off + 64] == O.to_bytes(wq)
                # KZG opening with known tau: C - y*tau^(10c)*G == (tau - x) * W
                if 10 * c + 10 <= d:
                    y = sum(coeffs[j] * x ** j for j in range(10))
                    lhs = O.add(O.from_bytes(Cb[64 * (p * 3 + c):64 * (p * 3 + c) + 64]),
                                O.neg(O.mul(O.GEN, y * pow(tau, 10 * c, O.ORDER))))
                    assert lhs == O.mul(wq, (tau - x) % O.ORDER)


def test_miner_aggregation_is_homomorphic():
    d, S = 20, 21
    eng = CommitmentEngine(d, S, tau=2, device="cpu")
    g = torch.Generator().manual_seed(4)
    Q = torch.randint(-10**5, 10**5, (3, 20), generator=g, dtype=torch.int64)
    C = eng.chunk_commitments(Q)
    sel = torch.tensor([0, 2], dtype=torch.int32)
    agg = eng.to_bytes(eng.aggregate(C, sel))
    Csum = eng.to_bytes(eng.chunk_commitments((Q[0] + Q[2]).unsqueeze(0)))
    assert agg == Csum      # aggregateSecret (kyber.go:244-287): sum of commitments commits to the sum


def test_verify_commitment_and_share_opening_helpers():
    d, S = 20, 21
    eng = CommitmentEngine(d, S, tau=2, device="cpu")
    g = torch.Generator().manual_seed(9)
    Q = torch.randint(-10**5, 10**5, (1, 20), generator=g, dtype=torch.int64)
    C = eng.chunk_commitments(Q)
    full = eng.to_bytes(eng.full_commitments(C))
    assert eng.verify_commitment(Q[0], full)
    assert not eng.verify_commitment(Q[0] + 1, full)
    W = eng.witnesses(Q)
    Y = secagg.shares_eval(Q, S)
    Cb, Wb = eng.to_bytes(C), eng.to_bytes(W)
    for c, s in ((0, 3), (1, 10), (1, 20)):
        off = 64 * (c * S + s)
        assert eng.verify_share_trapdoor(Cb[64 * c:64 * c + 64], Wb[off:off + 64], c, s, int(Y[0, c, s]))
        assert not eng.verify_share_trapdoor(Cb[64 * c:64 * c + 64], Wb[off:off + 64], c, s, int(Y[0, c, s]) + 1)


def test_two_width_window_tables_match_the_plain_msm():
    """Fixed-base tables with a wide first window and narrow upper windows (kernels_msm.cu: tw_*) against the table-free
    MSM, on scalars that exercise one window, both windows, the signed-digit carry and the out-of-range fallback."""
    import torch

    from biscotti_b200.ops import secagg
    from biscotti_b200.ops.msm import CommitmentEngine

    d, S = 30, 21
    plain = CommitmentEngine(d, S, tau=2, device="cpu", window=0)
    two = CommitmentEngine(d, S, tau=2, device="cpu", window=6, window2=4, max_scalar_bits=13)
    assert two.nwin == 3 and two.w1 == 4 and two.per_base == 32 + 2 * 8
    Q = torch.zeros(4, plain.dpad, dtype=torch.int64)
    g = torch.Generator().manual_seed(3)
    Q[0] = torch.randint(-31, 32, (plain.dpad,), generator=g)                    # first window only
    Q[1] = torch.randint(-8000, 8000, (plain.dpad,), generator=g)                # all windows
    Q[2, :12] = torch.tensor([32, -32, 33, -33, 31, 511, 512, -512, 8191, -8191, 8192, 1 << 20])   # carries, range edge, fallback
    Q[3] = torch.randint(-(1 << 22), 1 << 22, (plain.dpad,), generator=g)        # mostly out of range: Straus fallback
    Ca, Cb = plain.chunk_commitments(Q), two.chunk_commitments(Q)
    assert plain.to_bytes(Ca) == two.to_bytes(Cb)
    Wa, Wb = plain.witnesses(Q), two.witnesses(Q)
    assert plain.to_bytes(Wa) == two.to_bytes(Wb)
