"""Host crypto runtime vs independent oracles (Python big ints, hashlib).

Mirrors the reference's manual crypto sandboxes (kyber-demo/test.go:98-124 Schnorr round trip,
DistSys/test/test.go point unmarshal, vrf_main.go prove/verify + lottery) as real unit tests."""
import ctypes
import hashlib
import random

from biscotti_b200.oracle import bn256 as O
from biscotti_b200.crypto import primitives as cp


def buf(n):
    return ctypes.create_string_buffer(n)


def test_generator_and_scalar_mul(lib):
    g = buf(64)
    lib.bsc_g1_generator(g)
    assert g.raw == O.to_bytes(O.GEN)
    assert lib.bsc_g1_on_curve(g.raw) == 1
    rnd = random.Random(7)
    for _ in range(4):
        k = rnd.getrandbits(256)
        out = buf(64)
        lib.bsc_g1_mul(out, g.raw, k.to_bytes(32, "big"))
        assert out.raw == O.to_bytes(O.mul(O.GEN, k))
        assert lib.bsc_g1_on_curve(out.raw) == 1


def test_add_neg_infinity(lib):
    a = O.mul(O.GEN, 12345)
    out = buf(64)
    neg = buf(64)
    lib.bsc_g1_neg(neg, O.to_bytes(a))
    assert neg.raw == O.to_bytes(O.neg(a))
    lib.bsc_g1_add(out, O.to_bytes(a), neg.raw)
    assert out.raw == b"\x00" * 64          # infinity = 64 zero bytes (point.go:109-167)
    lib.bsc_g1_add(out, O.to_bytes(a), O.to_bytes(a))
    assert out.raw == O.to_bytes(O.add(a, a))


def test_msm_signed_scalars_match_oracle(lib):
    rnd = random.Random(3)
    pts = [O.mul(O.GEN, rnd.getrandbits(200)) for _ in range(10)]
    sc = [rnd.randint(-2**40, 2**40) for _ in range(10)]
    bases = b"".join(O.to_bytes(p) for p in pts)
    arr = (ctypes.c_longlong * 10)(*sc)
    out = buf(64)
    lib.bsc_g1_msm_i64_host(out, bases, arr, 10)
    assert out.raw == O.to_bytes(O.msm(pts, sc))      # negative scalar == reduce mod order (kyber.go:541-552)
    aff = buf(640)
    lib.bsc_g1_bytes_to_affine_host(bases, aff, 10)
    out2 = buf(64)
    lib.bsc_g1_straus_host(out2, aff, arr, 10)
    assert out2.raw == out.raw


def test_srs_powers_of_tau(lib):
    for tau in (2, 987654321):
        assert cp.generate_srs_bytes(6, tau) == b"".join(O.to_bytes(p) for p in O.srs(tau, 6))


def test_hashes(lib):
    m = b"biscotti" * 41
    o = buf(32)
    lib.bsc_sha256(m, len(m), o)
    assert o.raw == hashlib.sha256(m).digest()
    o = buf(64)
    lib.bsc_sha512(m, len(m), o)
    assert o.raw == hashlib.sha512(m).digest()
    lib.bsc_blake2b(m, len(m), None, 0, o, 64)
    assert o.raw == hashlib.blake2b(m).digest()
    key = bytes(range(64))
    lib.bsc_blake2b(m, len(m), key, 64, o, 64)
    assert o.raw == hashlib.blake2b(m, key=key).digest()
    o = buf(200)
    lib.bsc_shake256(m, len(m), o, 200)
    assert o.raw == hashlib.shake_256(m).digest(200)


def test_schnorr_round_trip():
    sk = hashlib.sha256(b"k").digest()
    pk = cp.schnorr_pubkey(sk)
    sig = cp.schnorr_sign(b"commitment-bytes", sk)
    assert len(sig) == 64
    assert cp.schnorr_verify(b"commitment-bytes", pk, sig)
    assert not cp.schnorr_verify(b"commitment-bytez", pk, sig)
    bad = bytearray(sig)
    bad[40] ^= 1
    assert not cp.schnorr_verify(b"commitment-bytes", pk, bytes(bad))


def test_ed25519_rfc8032_public_key():
    seed = bytes.fromhex("9d61b19deffd5a60ba844af492ec2cc44449c5697b326919703bac031cae7f60")
    assert cp.VRFKey(seed).public.hex() == "d75a980182b10ab7d54bfed3c964073a0ee172f3daa62325af021a68f707511a"


def test_vrf_prove_verify_and_batch():
    keys = [cp.VRFKey(hashlib.sha256(bytes([i])).digest()) for i in range(4)]
    msg = hashlib.sha256(b"block").digest()
    outs = []
    for k in keys:
        vrf, proof = k.prove(msg)
        assert len(vrf) == 32 and len(proof) == 96
        assert cp.vrf_verify(k.public, msg, vrf, proof)
        assert not cp.vrf_verify(k.public, msg + b"x", vrf, proof)
        outs.append(vrf)
    assert len(set(outs)) == 4
    assert cp.BatchedVRF(keys).outputs(msg) == outs      # batched Gamma = x*H path gives the same outputs


def _py_lottery(inp, stake, wants, exclude=-1):
    tickets = [i for i, s in enumerate(stake) for _ in range(s)]
    i, res = 0, []
    for want in wants:
        got = []
        while len(got) < want:
            if i >= len(inp) - 1:
                inp = hashlib.sha256(inp).digest()
                i = 0
            w = tickets[(inp[i] * 256 + inp[i + 1]) % len(tickets)]
            if w not in got and w != exclude:
                got.append(w)
            i += 1
        res.append(got)
    return res


def test_lottery_matches_reference_rule():
    # restatement of vrf.go:103-182 (overlapping 2-byte windows, re-hash when the bytes run out)
    for n, nv, nm in ((10, 3, 3), (100, 26, 26), (40, 10, 5)):
        stake = [10 + 5 * (i % 3) for i in range(n)]
        h = hashlib.sha256(f"h{n}".encode()).digest()
        v, m = cp.lottery_roles(h, stake, nv, nm)
        pv, pm = _py_lottery(h, stake, (nv, nm))
        assert (v, m) == (pv, pm)
    out = cp.lottery_noisers(h, stake, 3, 5)
    assert out == _py_lottery(h, stake, (5,), exclude=3)[0]
    assert 3 not in out


def test_role_ids_are_prime_products():
    r = cp.role_ids(6, [1, 2], [2, 3], [5])
    assert r == [1, 2, 6, 3, 1, 5]


def test_p256_vrf_reading_copy_and_hmac_commitments():
    """vrf-reference/: P-256 VRF evaluate / proof_to_hash round trip, rejection of tampering, HMAC commitments."""
    from biscotti_b200.crypto import p256_vrf as V

    assert V._mul(V.G, V.N) is None and (V.G[1] ** 2 - (V.G[0] ** 3 + V.A * V.G[0] + V.B)) % V.P == 0
    key = V.P256VRFKey(b"\x07" * 32)
    index, proof = key.evaluate(b"block-hash")
    assert len(index) == 32 and len(proof) == 129
    assert V.proof_to_hash(key.public, b"block-hash", proof) == index
    assert key.evaluate(b"block-hash")[0] == index                       # a VRF is a function
    assert V.proof_to_hash(key.public, b"other", proof) is None
    other = V.P256VRFKey(b"\x08" * 32)
    assert V.proof_to_hash(other.public, b"block-hash", proof) is None
    bad = bytearray(proof)
    bad[40] ^= 1
    assert V.proof_to_hash(key.public, b"block-hash", bytes(bad)) is None
    assert other.evaluate(b"block-hash")[0] != index
    nonce = bytes(range(16))
    c = V.commit("alice", "app", b"profile", nonce)
    assert len(c) == 32 and V.verify_commitment("alice", "app", c, b"profile", nonce)
    assert not V.verify_commitment("alice", "app", c, b"profile2", nonce)
    assert not V.verify_commitment("alic", "eapp", c, b"profile", nonce)  # length-prefixed fields do not collide


def test_binary_euclid_inverse_matches_fermat_and_python():
    """field.cuh: fe_inv (binary extended Euclid, what the kernels run) == fe_inv_fermat (the reference's a^(m-2),
    gfp.go:30-48) == Python's pow(a, -1, m), on the three moduli in use, including edge values."""
    import ctypes
    import random

    from biscotti_b200 import _native
    from biscotti_b200.oracle import bn256 as O

    L = _native.load()
    mods = {0: O.P, 1: O.ORDER, 2: 2**255 - 19}
    rnd = random.Random(7)
    for which, m in mods.items():
        vals = [1, 2, m - 1, m - 2, 3, 2**128, 2**255 % m, (m + 1) // 2] + [rnd.randrange(1, m) for _ in range(40)]
        for a in vals:
            f, g = ctypes.create_string_buffer(32), ctypes.create_string_buffer(32)
            L.bsc_fe_inv_both(which, int(a).to_bytes(32, "big"), f, g)
            want = pow(a, -1, m)
            assert int.from_bytes(f.raw, "big") == want == int.from_bytes(g.raw, "big"), (which, a)
        f, g = ctypes.create_string_buffer(32), ctypes.create_string_buffer(32)
        L.bsc_fe_inv_both(which, (0).to_bytes(32, "big"), f, g)
        assert int.from_bytes(f.raw, "big") == 0 == int.from_bytes(g.raw, "big")
