"""Secure aggregation ops: quantise, Shamir share evaluation / aggregation, reconstruction + SGD step.

Reference: DistSys/kyber.go:698-743, 579-609, 244-287, 809-857; DistSys/honest.go:391-502.
CUDA path: kernels_secagg.cu.  CPU path: the same C++ arithmetic through the host mirrors.
"""
from __future__ import annotations

import ctypes

import torch

from ._call import lib, ptr, stream, check

POLY_SIZE = 10


def nchunks(d: int) -> int:
    return (d + POLY_SIZE - 1) // POLY_SIZE


def quantize(U: torch.Tensor, d: int, precision: int = 4):
    """Q[p, k] = int64(U[p,k] * 10^precision) (truncation), padded to nchunk*10 columns."""
    P = U.shape[0]
    dpad = nchunks(d) * POLY_SIZE
    Q = torch.empty(P, dpad, dtype=torch.int64, device=U.device)
    scale = float(10 ** precision)
    if U.is_cuda:
        check(lib().bsc_quantize(ptr(U), ptr(Q), P, d, U.stride(0), dpad, ctypes.c_double(scale), stream(U.device)),
              "quantize")
    else:
        lib().bsc_quantize_host(ptr(U), ptr(Q), P, d, U.stride(0), dpad, ctypes.c_double(scale))
    return Q


def shares_eval(Q: torch.Tensor, S: int):
    """Per-peer shares Y[p, c, s] = poly_{p,c}(s - 10)."""
    P, dpad = Q.shape
    nc = dpad // POLY_SIZE
    Y = torch.empty(P, nc, S, dtype=torch.int64, device=Q.device)
    if Q.is_cuda:
        check(lib().bsc_shares_eval(ptr(Q), ptr(Y), P, nc, dpad, S, stream(Q.device)), "shares_eval")
    else:
        lib().bsc_shares_eval_host(ptr(Q), ptr(Y), P, nc, dpad, S)
    return Y


def share_sum(U: torch.Tensor, sel: torch.Tensor, d: int, S: int, precision: int = 4, c0: int = 0, nc: int | None = None,
              out: torch.Tensor | None = None, count: torch.Tensor | None = None):
    """Aggregated shares of the selected rows for chunk range [c0, c0+nc): Ysum [nc, S] int64.
    count (int32 [1], same device): number of live entries of `sel`; on a GPU it is read by the kernel."""
    nc = nchunks(d) - c0 if nc is None else nc
    scale = float(10 ** precision)
    if U.is_cuda:
        Ysum = out if out is not None else torch.empty(nc, S, dtype=torch.int64, device=U.device)
        check(lib().bsc_share_sum_dc(ptr(U), ptr(sel), sel.numel(), ptr(count), d, U.stride(0), c0, nc, S,
                                     ctypes.c_double(scale), ptr(Ysum), stream(U.device)), "share_sum")
        return Ysum
    if count is not None:
        sel = sel[: int(count[0])]
    if sel.numel() == 0:
        Y = torch.zeros(nc, S, dtype=torch.int64)
    else:
        Q = quantize(U[sel.long()].contiguous(), d, precision)
        Y = shares_eval(Q, S)[:, c0:c0 + nc].sum(dim=0)
    if out is not None:
        out.copy_(Y)
        return out
    return Y


def reconstruct_step(Ysum: torch.Tensor, xs: torch.Tensor, avail, d: int, w64: torch.Tensor, w32=None,
                     precision: int = 4, c0: int = 0, want_dsum: bool = False):
    """Recover sum-of-deltas from aggregated shares and apply w += sum (in place).  Returns (dsum, bad)."""
    nc, S = Ysum.shape
    inv = float(10 ** precision)   # de-quantisation divides by 10^p (updateIntToFloat, kyber.go:745-757)
    if Ysum.is_cuda:
        dsum = torch.zeros(d, dtype=torch.float64, device=Ysum.device) if want_dsum else None
        bad = torch.zeros(2, dtype=torch.int32, device=Ysum.device)
        check(lib().bsc_reconstruct_step(ptr(Ysum), ptr(xs), ptr(avail), S, c0, nc, d, ctypes.c_double(inv), ptr(w64),
                                         ptr(w32), ptr(dsum), ptr(bad), stream(Ysum.device)), "reconstruct_step")
        return dsum, bad
    coef = torch.empty(nc, POLY_SIZE, dtype=torch.int64)
    nbad = lib().bsc_shamir_recover_host(ptr(Ysum), ptr(xs), ptr(avail), S, nc, ptr(coef))
    flat = coef.reshape(-1)
    lo, hi = c0 * POLY_SIZE, min(d, (c0 + nc) * POLY_SIZE)
    delta = flat[: hi - lo].double() / inv
    w64[lo:hi] += delta
    if w32 is not None:
        w32[lo:hi] = w64[lo:hi].float()
    dsum = None
    if want_dsum:
        dsum = torch.zeros(d, dtype=torch.float64)
        dsum[lo:hi] = delta
    return dsum, torch.tensor([nbad, nbad], dtype=torch.int32)


# ---------------------------------------------------------------------------------------------
# recovery as a constant matrix (csrc/secagg_math.cuh: RecoverMat)
# ---------------------------------------------------------------------------------------------
class _RecoverMat(ctypes.Structure):
    _fields_ = [("M", ctypes.c_longlong * (POLY_SIZE * POLY_SIZE)), ("D", ctypes.c_longlong),
                ("inv_lo", ctypes.c_ulonglong), ("inv_hi", ctypes.c_ulonglong), ("tz", ctypes.c_int),
                ("xs", ctypes.c_int * POLY_SIZE)]


_recover_cache: dict = {}


def recover_matrix(xs):
    """Exact inverse of the Vandermonde system at the share indices `xs` (x = s - 10) as an integer matrix over a
    common denominator: coef_j = (sum_i M[j][i] y_i) / D.  Built with Python big integers from the Lagrange basis
    (kyber.go:809-857 solves the same system with float QR and rounds).  Returns a ctypes RecoverMat, or None when
    a dot product of int64 shares with row j could exceed 127 bits (the callers then use the divided-difference
    kernel)."""
    key = tuple(int(x) for x in xs)
    if key in _recover_cache:
        return _recover_cache[key]
    from fractions import Fraction
    from math import lcm

    assert len(key) == POLY_SIZE and len(set(key)) == POLY_SIZE
    nodes = [x - 10 for x in key]
    cols = []                                   # cols[i][j] = coefficient j of the Lagrange basis polynomial L_i
    for i, xi in enumerate(nodes):
        poly = [Fraction(1)]
        den = 1
        for k, xk in enumerate(nodes):
            if k == i:
                continue
            poly = [Fraction(0)] + poly         # times x
            for t in range(len(poly) - 1):
                poly[t] -= xk * poly[t + 1]
            den *= xi - xk
        cols.append([c / den for c in poly])
    D = 1
    for col in cols:
        for c in col:
            D = lcm(D, c.denominator)
    M = [[int(cols[i][j] * D) for i in range(POLY_SIZE)] for j in range(POLY_SIZE)]
    ok = D < 2**62 and all(sum(abs(v) for v in row) < 2**63 for row in M)
    rm = None
    if ok:
        rm = _RecoverMat()
        for j in range(POLY_SIZE):
            for i in range(POLY_SIZE):
                rm.M[j * POLY_SIZE + i] = M[j][i]
        tz = (D & -D).bit_length() - 1
        inv = pow(D >> tz, -1, 1 << 128)
        rm.D, rm.tz, rm.inv_lo, rm.inv_hi = D, tz, inv & (2**64 - 1), inv >> 64
        for i in range(POLY_SIZE):
            rm.xs[i] = key[i]
        assert ctypes.sizeof(rm) == lib().bsc_recover_mat_sizeof()
    _recover_cache[key] = rm
    return rm


def reconstruct_step_dev(Ysum, xs_list, xs, avail, d, w64, w32, precision=4, c0=0, bad=None, gate=None, gate_min=0):
    """Device-resident form used on the block's chain: matrix recovery when the abscissae allow it, contributor-count
    gate read on the device (no step when *gate < gate_min), `bad` accumulated into a caller-owned int32[2]."""
    nc, S = Ysum.shape
    inv = float(10 ** precision)
    if not Ysum.is_cuda:
        if gate is not None and int(gate[0]) < gate_min:
            return
        _, b = reconstruct_step(Ysum, xs, avail, d, w64, w32, precision, c0)
        if bad is not None:
            bad[:2] += b.to(bad.dtype)
        return
    rm = recover_matrix(xs_list)
    L = lib()
    if rm is not None:
        check(L.bsc_reconstruct_mat(ptr(Ysum), ctypes.byref(rm), ptr(avail), S, c0, nc, d, ctypes.c_double(inv), ptr(w64),
                                    ptr(w32), None, ptr(bad), ptr(gate), gate_min, stream(Ysum.device)), "reconstruct_mat")
    else:
        check(L.bsc_reconstruct_step_gated(ptr(Ysum), ptr(xs), ptr(avail), S, c0, nc, d, ctypes.c_double(inv), ptr(w64),
                                           ptr(w32), None, ptr(bad), ptr(gate), gate_min, stream(Ysum.device)),
              "reconstruct_step")


def plain_sum_step(U: torch.Tensor, sel: torch.Tensor, d: int, w64: torch.Tensor, w32=None, want_dsum=False,
                   count: torch.Tensor | None = None, gate: torch.Tensor | None = None, gate_min: int = 0):
    """-sa=false / FedSys: w += sum of selected deltas (fp64, fixed order).  count / gate: device-resident row count
    and the contributor count the step is conditional on (no step when gate < gate_min)."""
    if U.is_cuda:
        dsum = torch.zeros(d, dtype=torch.float64, device=U.device) if want_dsum else None
        check(lib().bsc_plain_sum_step_dc(ptr(U), ptr(sel), sel.numel(), ptr(count), d, U.stride(0), ptr(w64), ptr(w32),
                                          ptr(dsum), ptr(gate), gate_min, stream(U.device)), "plain_sum_step")
        return dsum
    if gate is not None and int(gate[0]) < gate_min:
        return None
    if count is not None:
        sel = sel[: int(count[0])]
    acc = torch.zeros(d, dtype=torch.float64)
    for q in sel.tolist():
        acc += U[q, :d].double()
    w64[:d] += acc
    if w32 is not None:
        w32[:d] = w64[:d].float()
    return acc if want_dsum else None
