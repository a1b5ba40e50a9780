"""P-256 VRF and HMAC commitments — the reference's "reading copy" (vrf-reference/).

The live protocol uses the Ed25519 ECVRF (crypto/primitives.py, csrc/vrf.cu); the reference tree also carries Google
KeyTransparency's P-256 VRF and HMAC commitment scheme "for reading only" (vrf-reference/README.md).  They are kept
here as small host-side Python so the capability exists with the same construction:

  VRF (vrf-reference/crypto/vrf/p256/p256.go:81-222)
      H1: try-and-increment hash to the curve (SHA-512 of counter||m as a compressed x, even y)
      H2: hash to a scalar in [1, N-1]
      Evaluate(m)  -> index = SHA-256(vrf), proof = s(32) || t(32) || vrf(65), vrf = [k]H1(m),
                      s = H2(G, H, [k]G, vrf, [r]G, [r]H), t = r - s*k mod N
      ProofToHash  -> recomputes [t]G + [s]pk and [t]H + [s]vrf and checks H2 gives s back
  Commitments (vrf-reference/crypto/commitments/commitments.go:34-72)
      HMAC-SHA512/256 under a fixed key of prefix || nonce || len(userID) || userID || len(appID) || appID || data

The nonce r of Evaluate is derived from the key and the message (deterministic, RFC 6979 style) instead of
crypto/rand, so results are reproducible; verification is unaffected.
"""
from __future__ import annotations

import hashlib
import hmac
import os

# NIST P-256
P = 0xffffffff00000001000000000000000000000000ffffffffffffffffffffffff
N = 0xffffffff00000000ffffffffffffffffbce6faada7179e84f3b9cac2fc632551
A = P - 3
B = 0x5ac635d8aa3a93e7b3ebbd55769886bc651d06b0cc53b0f63bce3c3e27d2604b
G = (0x6b17d1f2e12c4247f8bce6e563a440f277037d812deb33a0f4a13945d898c296,
     0x4fe342e2fe1a7f9b8ee7eb4a7c0f9e162bce33576b315ececbb6406837bf51f5)


def _add(p, q):
    if p is None:
        return q
    if q is None:
        return p
    if p[0] == q[0]:
        if (p[1] + q[1]) % P == 0:
            return None
        lam = (3 * p[0] * p[0] + A) * pow(2 * p[1], -1, P) % P
    else:
        lam = (q[1] - p[1]) * pow(q[0] - p[0], -1, P) % P
    x = (lam * lam - p[0] - q[0]) % P
    return (x, (lam * (p[0] - x) - p[1]) % P)


def _mul(p, k: int):
    k %= N
    r = None
    while k:
        if k & 1:
            r = _add(r, p)
        p = _add(p, p)
        k >>= 1
    return r


def _marshal(p) -> bytes:          # elliptic.Marshal: uncompressed, 65 bytes
    return b"\x04" + p[0].to_bytes(32, "big") + p[1].to_bytes(32, "big")


def _unmarshal(b: bytes):
    if len(b) != 65 or b[0] != 4:
        return None
    x, y = int.from_bytes(b[1:33], "big"), int.from_bytes(b[33:], "big")
    if (y * y - (x * x * x + A * x + B)) % P:
        return None
    return (x, y)


def _decompress_even(x: int):
    if x >= P:
        return None
    rhs = (x * x * x + A * x + B) % P
    y = pow(rhs, (P + 1) // 4, P)          # P = 3 mod 4
    if y * y % P != rhs:
        return None
    return (x, y if y % 2 == 0 else P - y)


def h1(m: bytes):
    """Hash to a curve point: the first counter whose SHA-512 prefix is the x of a point (p256.go:81-96)."""
    for i in range(100):
        d = hashlib.sha512(i.to_bytes(4, "big") + m).digest()
        pt = _decompress_even(int.from_bytes(d[:32], "big"))
        if pt is not None:
            return pt
    raise ValueError("h1: no point found")


def h2(m: bytes) -> int:
    """Hash to a scalar in [1, N-1] (p256.go:101-115)."""
    i = 0
    while True:
        k = int.from_bytes(hashlib.sha512(i.to_bytes(4, "big") + m).digest()[:32], "big")
        if k < N - 1:
            return k + 1
        i += 1


class P256VRFKey:
    def __init__(self, secret: bytes | None = None):
        self.d = (int.from_bytes(secret or os.urandom(32), "big") % (N - 1)) + 1
        self.public = _mul(G, self.d)

    def evaluate(self, m: bytes):
        """-> (index 32 bytes, proof 129 bytes)."""
        r = h2(b"nonce" + self.d.to_bytes(32, "big") + m)
        H = h1(m)
        vrf = _marshal(_mul(H, self.d))
        s = h2(_marshal(G) + _marshal(H) + _marshal(self.public) + vrf + _marshal(_mul(G, r)) + _marshal(_mul(H, r)))
        t = (r - s * self.d) % N
        return hashlib.sha256(vrf).digest(), s.to_bytes(32, "big") + t.to_bytes(32, "big") + vrf


def proof_to_hash(public, m: bytes, proof: bytes) -> bytes | None:
    """Verify a proof against `public`; returns the 32-byte index or None (p256.go:168-222)."""
    if len(proof) != 129:
        return None
    s, t = int.from_bytes(proof[:32], "big"), int.from_bytes(proof[32:64], "big")
    vrf = proof[64:]
    u = _unmarshal(vrf)
    if u is None:
        return None
    H = h1(m)
    tks_g = _add(_mul(G, t), _mul(public, s))
    tks_h = _add(_mul(H, t), _mul(u, s))
    if tks_g is None or tks_h is None:
        return None
    want = h2(_marshal(G) + _marshal(H) + _marshal(public) + vrf + _marshal(tks_g) + _marshal(tks_h))
    if not hmac.compare_digest(want.to_bytes(32, "big"), proof[:32]):
        return None
    return hashlib.sha256(vrf).digest()


# ---- HMAC commitments -------------------------------------------------------------------------------------------
_FIXED_KEY = bytes([0x19, 0x6e, 0x7e, 0x52, 0x84, 0xa7, 0xef, 0x93, 0x0e, 0xcb, 0x9a, 0x19, 0x78, 0x74, 0x97, 0x55])
_PREFIX = b"Key Transparency Commitment"
COMMITMENT_KEY_LEN = 16


def gen_commitment_key() -> bytes:
    return os.urandom(COMMITMENT_KEY_LEN)


def commit(user_id: str, app_id: str, data: bytes, nonce: bytes) -> bytes:
    mac = hmac.new(_FIXED_KEY, digestmod="sha512_256")
    mac.update(_PREFIX + nonce)
    mac.update(len(user_id.encode()).to_bytes(4, "big") + user_id.encode())
    mac.update(len(app_id.encode()).to_bytes(4, "big") + app_id.encode())
    mac.update(data)
    return mac.digest()


def verify_commitment(user_id: str, app_id: str, commitment: bytes, data: bytes, nonce: bytes) -> bool:
    return hmac.compare_digest(commit(user_id, app_id, data, nonce), commitment)
