"""Big-int oracle for BN256 G2, GT and the optimal-ate pairing (test oracle for csrc/pairing.cpp).

What it restates: the reference's vendored curve (lib/dedis/kyber/pairing/bn256): the tower
Fp2 = Fp[i]/(i^2+1), xi = i+3 (constants.go:22-41), the sextic twist y^2 = x^3 + 3/xi and its generator
(twist.go:7-31), the 129-byte G2 wire form 0x01|x.i|x.1|y.i|y.1 (point.go:280-330), and the KZG opening
check of DistSys/kyber.go:650-673 (verifySecret).

It is deliberately the *slow, obviously-correct* formulation: Fp12 is the flat ring Fp2[w]/(w^6 - xi), the
Miller loop runs in affine coordinates, and the final exponentiation is a single generic power by
(p^12-1)/n.  The C++ implementation uses the tower, projective-free sparse lines and the u-addition chain;
the tests require both to produce byte-identical GT elements.
"""
from __future__ import annotations

from .bn256 import ORDER, P

U = 6518589491078791937                       # BN parameter (constants.go:13)
XI = (3, 1)                                   # xi = 3 + i
ATE_LOOP = 6 * U + 2

# G2 generator (twist.go:20-31, decoded from Montgomery form); element = (a, b) meaning a + b*i
G2_GEN = (
    (0x8f25386f72c9462b81597d65ae2092c4b97792155dcdaad32b8a6dd41792534c,
     0x2ecca446ff6f3d4d03c76e9b5c752f28bc37b364cb05ac4a37eb32e1c3245970),
    (0x274e5747e8cafacc3716cc8699db79b22f0e4ff3c23e898f694420a3be3087a5,
     0x2db10ef5233b0fe3962b9ee6a4bbc2b5bde01a54f3513d42df972e128f31bf12),
)


# ---- Fp2 -----------------------------------------------------------------------------------------
def f2_add(a, b): return ((a[0] + b[0]) % P, (a[1] + b[1]) % P)
def f2_sub(a, b): return ((a[0] - b[0]) % P, (a[1] - b[1]) % P)
def f2_neg(a): return (-a[0] % P, -a[1] % P)
def f2_mul(a, b): return ((a[0] * b[0] - a[1] * b[1]) % P, (a[0] * b[1] + a[1] * b[0]) % P)
def f2_conj(a): return (a[0], -a[1] % P)
def f2_scale(a, k): return (a[0] * k % P, a[1] * k % P)


def f2_inv(a):
    d = pow(a[0] * a[0] + a[1] * a[1], -1, P)
    return (a[0] * d % P, -a[1] * d % P)


def f2_pow(a, e):
    r = (1, 0)
    while e:
        if e & 1:
            r = f2_mul(r, a)
        a = f2_mul(a, a)
        e >>= 1
    return r


F2_ZERO, F2_ONE = (0, 0), (1, 0)
TWIST_B = f2_scale(f2_inv(XI), 3)             # 3/xi (twist.go:14-17)


# ---- G2 (affine over Fp2; None = infinity) ---------------------------------------------------------
def g2_on_curve(q):
    if q is None:
        return True
    x, y = q
    return f2_mul(y, y) == f2_add(f2_mul(f2_mul(x, x), x), TWIST_B)


def g2_neg(q):
    return None if q is None else (q[0], f2_neg(q[1]))


def g2_add(a, b):
    if a is None:
        return b
    if b is None:
        return a
    if a[0] == b[0]:
        if f2_add(a[1], b[1]) == F2_ZERO:
            return None
        lam = f2_mul(f2_scale(f2_mul(a[0], a[0]), 3), f2_inv(f2_scale(a[1], 2)))
    else:
        lam = f2_mul(f2_sub(b[1], a[1]), f2_inv(f2_sub(b[0], a[0])))
    x = f2_sub(f2_sub(f2_mul(lam, lam), a[0]), b[0])
    return (x, f2_sub(f2_mul(lam, f2_sub(a[0], x)), a[1]))


def g2_mul(q, k: int):
    k %= ORDER
    r = None
    while k:
        if k & 1:
            r = g2_add(r, q)
        q = g2_add(q, q)
        k >>= 1
    return r


def g2_to_bytes(q) -> bytes:
    if q is None:
        return b"\x00"
    (x0, x1), (y0, y1) = q
    return b"\x01" + b"".join(v.to_bytes(32, "big") for v in (x1, x0, y1, y0))


def g2_from_bytes(b: bytes):
    if b[:1] == b"\x00":
        return None
    if b[:1] != b"\x01" or len(b) < 129:
        raise ValueError("malformed G2 point")
    v = [int.from_bytes(b[1 + 32 * i: 33 + 32 * i], "big") for i in range(4)]
    return ((v[1], v[0]), (v[3], v[2]))


# ---- Fp12 = Fp2[w]/(w^6 - xi), element = list of 6 Fp2 coefficients --------------------------------
F12_ONE = [F2_ONE] + [F2_ZERO] * 5


def f12_mul(a, b):
    t = [F2_ZERO] * 11
    for i in range(6):
        if a[i] == F2_ZERO:
            continue
        for j in range(6):
            t[i + j] = f2_add(t[i + j], f2_mul(a[i], b[j]))
    return [f2_add(t[k], f2_mul(t[k + 6], XI)) if k < 5 else t[k] for k in range(6)]


def f12_pow(a, e):
    r = list(F12_ONE)
    while e:
        if e & 1:
            r = f12_mul(r, a)
        a = f12_mul(a, a)
        e >>= 1
    return r


def f12_to_bytes(a) -> bytes:
    """384-byte GT encoding used by the C++ side: coefficients of w^0..w^5, each as (i-part, 1-part) big-endian."""
    return b"".join(c[1].to_bytes(32, "big") + c[0].to_bytes(32, "big") for c in a)


# ---- optimal ate ------------------------------------------------------------------------------------
_G = [f2_pow(XI, k * (P - 1) // 6) for k in range(6)]     # gamma_k = xi^(k(p-1)/6)


def _frob_twist(q):
    """Frobenius of the untwisted point pulled back to the twist."""
    return (f2_mul(f2_conj(q[0]), _G[2]), f2_mul(f2_conj(q[1]), _G[3]))


def _line(t, q, p):
    """Line through twist points t, q (tangent when equal) evaluated at p in E(Fp); returns (line, t+q)."""
    if t[0] == q[0] and t[1] == q[1]:
        lam = f2_mul(f2_scale(f2_mul(t[0], t[0]), 3), f2_inv(f2_scale(t[1], 2)))
    else:
        lam = f2_mul(f2_sub(q[1], t[1]), f2_inv(f2_sub(q[0], t[0])))
    x3 = f2_sub(f2_sub(f2_mul(lam, lam), t[0]), q[0])
    y3 = f2_sub(f2_mul(lam, f2_sub(t[0], x3)), t[1])
    # untwist (x, y) -> (x w^2, y w^3):  l(P) = yP - lam*xP*w + (lam*xT - yT)*w^3
    l = [(p[1] % P, 0), f2_neg(f2_scale(lam, p[0])), F2_ZERO, f2_sub(f2_mul(lam, t[0]), t[1]), F2_ZERO, F2_ZERO]
    return l, (x3, y3)


def miller(q, p):
    if q is None or p is None:
        return list(F12_ONE)
    f = list(F12_ONE)
    t = q
    for bit in bin(ATE_LOOP)[3:]:
        l, t = _line(t, t, p)
        f = f12_mul(f12_mul(f, f), l)
        if bit == "1":
            l, t = _line(t, q, p)
            f = f12_mul(f, l)
    q1 = _frob_twist(q)
    q2 = g2_neg(_frob_twist(q1))
    l, t = _line(t, q1, p)
    f = f12_mul(f, l)
    l, t = _line(t, q2, p)
    return f12_mul(f, l)


def final_exp(f):
    return f12_pow(f, (P ** 12 - 1) // ORDER)


def pairing(p, q):
    """e(p, q) for p in G1 (affine tuple of ints or None), q in G2."""
    return final_exp(miller(q, p))
