"""BN256 G2 / GT / pairing front end (native host code in csrc/pairing.cpp).

Reference surface covered: `pointG2` marshal/add/mul (lib/dedis/kyber/pairing/bn256/point.go:190-400), the G2
half of the SRS (`PKG2`, DistSys/publicKey.go:42-57), `suite.Pair` and `verifySecret`
(DistSys/kyber.go:650-673).  G2 points travel as the reference's wire form: 129 bytes
``0x01|x.i|x.1|y.i|y.1`` or the single byte ``0x00`` for infinity; GT elements as 384 bytes.
"""
from __future__ import annotations

import ctypes

from .._native import load as _load

G2_SIZE = 129
GT_SIZE = 384
ORDER = 65000549695646603732796438742359905742570406053903786389881062969044166799969


def _pad(pt: bytes) -> bytes:
    if pt[:1] == b"\x00":
        return bytes(G2_SIZE)
    if len(pt) != G2_SIZE or pt[0] != 1:
        raise ValueError("malformed G2 point")
    return pt


def _trim(buf) -> bytes:
    raw = bytes(buf.raw)
    return b"\x00" if raw[0] == 0 else raw


def _k(k: int) -> bytes:
    return (k % ORDER).to_bytes(32, "big")


def g2_generator() -> bytes:
    out = ctypes.create_string_buffer(G2_SIZE)
    _load().bsc_g2_generator(out)
    return _trim(out)


def g2_on_curve(pt: bytes) -> bool:
    try:
        return bool(_load().bsc_g2_on_curve(_pad(pt)))
    except ValueError:
        return False


def g2_add(a: bytes, b: bytes) -> bytes:
    out = ctypes.create_string_buffer(G2_SIZE)
    _load().bsc_g2_add(out, _pad(a), _pad(b))
    return _trim(out)


def g2_neg(a: bytes) -> bytes:
    out = ctypes.create_string_buffer(G2_SIZE)
    _load().bsc_g2_neg(out, _pad(a))
    return _trim(out)


def g2_mul(pt: bytes, k: int) -> bytes:
    out = ctypes.create_string_buffer(G2_SIZE)
    _load().bsc_g2_mul(out, _pad(pt), _k(k))
    return _trim(out)


def srs_g2(tau: int, d: int) -> list[bytes]:
    """PKG2[i] = tau^i * G2 for i < d."""
    out = ctypes.create_string_buffer(G2_SIZE * d)
    _load().bsc_srs_g2_host(_k(tau), d, out)
    raw = out.raw
    return [raw[G2_SIZE * i: G2_SIZE * (i + 1)] for i in range(d)]


def pair(g1: bytes, g2: bytes) -> bytes:
    """e(g1, g2) as a 384-byte GT element (suite.Pair)."""
    out = ctypes.create_string_buffer(GT_SIZE)
    _load().bsc_pairing(g1, _pad(g2), out)
    return out.raw


def pairing_check(g1s: list[bytes], g2s: list[bytes]) -> bool:
    """prod_i e(g1s[i], g2s[i]) == 1, with a single final exponentiation."""
    assert len(g1s) == len(g2s)
    return bool(_load().bsc_pairing_check(b"".join(g1s), b"".join(_pad(q) for q in g2s), len(g1s)))


def gt_mul(a: bytes, b: bytes) -> bytes:
    out = ctypes.create_string_buffer(GT_SIZE)
    _load().bsc_gt_mul(out, a, b)
    return out.raw


def gt_pow(a: bytes, k: int) -> bytes:
    out = ctypes.create_string_buffer(GT_SIZE)
    _load().bsc_gt_pow(out, a, _k(k))
    return out.raw


def gt_one() -> bytes:
    return bytes(63) + b"\x01" + bytes(GT_SIZE - 64)


def verify_secret(commitment: bytes, witness: bytes, x: int, y: int, g2: bytes, tau_g2: bytes,
                  base: bytes | None = None) -> bool:
    """verifySecret (kyber.go:650-673): does `witness` open the chunk commitment to value y at point x?

    `base` is the first G1 SRS point of the chunk's slice (PKG1[10c]); None means the generator, i.e. chunk 0,
    which is the case the reference's formula is written for."""
    if base is None:
        from .primitives import g1_generator

        base = g1_generator()
    return bool(_load().bsc_kzg_verify_share(commitment, witness, base, _pad(g2), _pad(tau_g2), int(x), int(y)))
