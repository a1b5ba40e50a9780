"""Pure-Python big-integer BN256 G1 oracle.

Independent restatement (affine formulas, Python ints) of the group the reference uses through
dedis/kyber (lib/dedis/kyber/pairing/bn256/constants.go:12-20, curve.go:9-231, point.go:109-167).
It exists only to check the C++/CUDA field and curve code bit-for-bit; nothing on a hot path calls it.
"""
from __future__ import annotations

P = 65000549695646603732796438742359905742825358107623003571877145026864184071783
ORDER = 65000549695646603732796438742359905742570406053903786389881062969044166799969
B = 3
GEN = (1, P - 2)
INF = None


def is_on_curve(pt) -> bool:
    if pt is None:
        return True
    x, y = pt
    return (y * y - x * x * x - B) % P == 0


def neg(pt):
    if pt is None:
        return None
    return (pt[0], (-pt[1]) % P)


def add(a, b):
    if a is None:
        return b
    if b is None:
        return a
    x1, y1 = a
    x2, y2 = b
    if x1 == x2:
        if (y1 + y2) % P == 0:
            return None
        lam = 3 * x1 * x1 * pow(2 * y1, -1, P) % P
    else:
        lam = (y2 - y1) * pow(x2 - x1, -1, P) % P
    x3 = (lam * lam - x1 - x2) % P
    y3 = (lam * (x1 - x3) - y1) % P
    return (x3, y3)


def mul(pt, k: int):
    k %= ORDER
    acc = None
    addend = pt
    while k:
        if k & 1:
            acc = add(acc, addend)
        addend = add(addend, addend)
        k >>= 1
    return acc


def msm(points, scalars):
    acc = None
    for pt, s in zip(points, scalars):
        acc = add(acc, mul(pt, int(s)))
    return acc


def to_bytes(pt) -> bytes:
    if pt is None:
        return b"\x00" * 64
    return pt[0].to_bytes(32, "big") + pt[1].to_bytes(32, "big")


def from_bytes(b: bytes):
    if b == b"\x00" * 64:
        return None
    return (int.from_bytes(b[:32], "big"), int.from_bytes(b[32:], "big"))


def srs(tau: int, d: int):
    out = []
    pw = 1
    for _ in range(d):
        out.append(mul(GEN, pw))
        pw = pw * tau % ORDER
    return out
