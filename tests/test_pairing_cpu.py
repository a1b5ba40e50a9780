"""BN256 G2 / GT / optimal-ate pairing (csrc/pairing.cpp) against the big-int oracle, plus verifySecret.

Oracle: biscotti_b200/oracle/bn256_pairing.py (flat Fp12, affine Miller loop, generic final exponent).
Reference behaviour: lib/dedis/kyber/pairing/bn256/{twist,optate,point}.go, DistSys/kyber.go:650-673.
"""
import base64
import json

import torch

from biscotti_b200.crypto import pairing as P
from biscotti_b200.crypto import primitives as cp
from biscotti_b200.ops import secagg
from biscotti_b200.ops.msm import CommitmentEngine
from biscotti_b200.oracle import bn256 as O
from biscotti_b200.oracle import bn256_pairing as Q


def test_g2_group_law_and_wire_format():
    g2 = P.g2_generator()
    assert len(g2) == 129 and g2[0] == 1
    assert g2 == Q.g2_to_bytes(Q.G2_GEN)
    assert P.g2_on_curve(g2)
    bad = bytearray(g2)
    bad[40] ^= 1
    assert not P.g2_on_curve(bytes(bad))
    for k in (1, 2, 3, 0xdeadbeefcafe, O.ORDER - 1):
        assert P.g2_mul(g2, k) == Q.g2_to_bytes(Q.g2_mul(Q.G2_GEN, k))
    assert P.g2_mul(g2, O.ORDER) == b"\x00"                      # infinity marshals to a single zero byte
    assert P.g2_add(g2, P.g2_neg(g2)) == b"\x00"
    assert P.g2_add(g2, b"\x00") == g2
    a, b = P.g2_mul(g2, 77), P.g2_mul(g2, 1000)
    assert P.g2_add(a, b) == P.g2_mul(g2, 1077)
    assert P.g2_add(a, a) == P.g2_mul(g2, 154)


def test_pairing_matches_oracle_bytes():
    g1, g2 = O.to_bytes(O.GEN), P.g2_generator()
    e = P.pair(g1, g2)
    assert e == Q.f12_to_bytes(Q.pairing(O.GEN, Q.G2_GEN))
    p5, q9 = O.mul(O.GEN, 5), Q.g2_mul(Q.G2_GEN, 9)
    assert P.pair(O.to_bytes(p5), Q.g2_to_bytes(q9)) == Q.f12_to_bytes(Q.pairing(p5, q9))


def test_pairing_bilinear_nondegenerate():
    g1, g2 = O.to_bytes(O.GEN), P.g2_generator()
    e = P.pair(g1, g2)
    assert e != P.gt_one()
    assert P.gt_pow(e, O.ORDER) == P.gt_one() or P.gt_mul(P.gt_pow(e, O.ORDER - 1), e) == P.gt_one()
    a, b = 31337, 271828182845
    assert P.pair(O.to_bytes(O.mul(O.GEN, a)), P.g2_mul(g2, b)) == P.gt_pow(e, a * b)
    assert P.gt_mul(P.pair(O.to_bytes(O.mul(O.GEN, a)), g2), P.pair(O.to_bytes(O.mul(O.GEN, b)), g2)) == \
        P.pair(O.to_bytes(O.mul(O.GEN, a + b)), g2)
    # infinity on either side pairs to one
    assert P.pair(bytes(64), g2) == P.gt_one()
    assert P.pair(g1, b"\x00") == P.gt_one()
    # product check: e(aG, Q) * e(-G, aQ) == 1
    assert P.pairing_check([O.to_bytes(O.mul(O.GEN, a)), O.to_bytes(O.neg(O.GEN))], [g2, P.g2_mul(g2, a)])
    assert not P.pairing_check([O.to_bytes(O.mul(O.GEN, a)), O.to_bytes(O.neg(O.GEN))], [g2, P.g2_mul(g2, a + 1)])


def test_verify_secret_pairing_agrees_with_trapdoor_check():
    d, S = 20, 21
    for tau in (2, 0x1234567890abcdef1234567890abcdef):
        eng = CommitmentEngine(d, S, tau=tau, device="cpu")
        g = torch.Generator().manual_seed(11)
        Qs = torch.randint(-10**5, 10**5, (1, d), generator=g, dtype=torch.int64)
        C, W = eng.chunk_commitments(Qs), eng.witnesses(Qs)
        Y = secagg.shares_eval(Qs, S)
        Cb, Wb = eng.to_bytes(C), eng.to_bytes(W)
        for c, s in ((0, 0), (0, 13), (1, 20)):
            off = 64 * (c * S + s)
            cm, wt, y = Cb[64 * c:64 * c + 64], Wb[off:off + 64], int(Y[0, c, s])
            assert eng.verify_share_pairing(cm, wt, c, s, y)
            assert eng.verify_share_trapdoor(cm, wt, c, s, y)
            assert not eng.verify_share_pairing(cm, wt, c, s, y + 1)
            assert not eng.verify_share_pairing(cm, wt, c, (s + 1) % S, y)


def test_verify_secret_reference_form_chunk0():
    """The literal identity of kyber.go:664-671 for chunk 0: e(C, g2) == e(W, tau*g2 - x*g2) * e(g1, g2)^y."""
    d, S, tau = 10, 21, 2
    eng = CommitmentEngine(d, S, tau=tau, device="cpu")
    Qs = torch.tensor([[3, -1, 4, 1, -5, 9, 2, -6, 5, 3]], dtype=torch.int64)
    cm = eng.to_bytes(eng.chunk_commitments(Qs))[:64]
    Wb = eng.to_bytes(eng.witnesses(Qs))
    Y = secagg.shares_eval(Qs, S)
    g1, g2 = O.to_bytes(O.GEN), P.g2_generator()
    s = 17
    x, y = s - 10, int(Y[0, 0, s])
    rhs_g2 = P.g2_add(P.g2_mul(g2, tau), P.g2_neg(P.g2_mul(g2, x)))
    lhs = P.pair(cm, g2)
    rhs = P.gt_mul(P.pair(Wb[64 * s:64 * s + 64], rhs_g2), P.gt_pow(P.pair(g1, g2), y))
    assert lhs == rhs


def test_bootstrap_file_carries_g2_srs(tmp_path):
    cp.write_bootstrap_files(str(tmp_path), n_nodes=3, dims=6, tau=2)
    lines = [json.loads(l) for l in open(tmp_path / "commitKey.json")]
    assert len(lines) == 6
    g2 = P.g2_generator()
    for i, rec in enumerate(lines):
        sk = base64.b64decode(rec["Skey"])
        assert len(sk) == 129 and sk == P.g2_mul(g2, 2 ** i)
        assert base64.b64decode(rec["Pkey"]) == O.to_bytes(O.mul(O.GEN, 2 ** i))
    assert cp.load_commit_key_g2(str(tmp_path / "commitKey.json"))[1] == P.g2_mul(g2, 2)
    # e(PKG1[i], g2) == e(g1, PKG2[i]): the two halves of the SRS are consistent
    assert P.pair(base64.b64decode(lines[3]["Pkey"]), g2) == P.pair(O.to_bytes(O.GEN), base64.b64decode(lines[3]["Skey"]))
