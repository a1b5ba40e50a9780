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
              out: torch.Tensor | None = None):
    """Aggregated shares of the selected rows for chunk range [c0, c0+nc): Ysum [nc, S] int64."""
    nc = nchunks(d) - c0 if nc is None else nc
    scale = float(10 ** precision)
    if U.is_cuda:
        Ysum = out if out is not None else torch.empty(nc, S, dtype=torch.int64, device=U.device)
        check(lib().bsc_share_sum(ptr(U), ptr(sel), sel.numel(), d, U.stride(0), c0, nc, S, ctypes.c_double(scale),
                                  ptr(Ysum), stream(U.device)), "share_sum")
        return Ysum
    Q = quantize(U[sel.long()].contiguous(), d, precision)
    Y = shares_eval(Q, S)
    return Y[:, c0:c0 + nc].sum(dim=0)


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


def plain_sum_step(U: torch.Tensor, sel: torch.Tensor, d: int, w64: torch.Tensor, w32=None, want_dsum=False):
    """-sa=false / FedSys: w += sum of selected deltas (fp64, fixed order)."""
    if U.is_cuda:
        dsum = torch.zeros(d, dtype=torch.float64, device=U.device) if want_dsum else None
        check(lib().bsc_plain_sum_step(ptr(U), ptr(sel), sel.numel(), d, U.stride(0), ptr(w64), ptr(w32), ptr(dsum),
                                       stream(U.device)), "plain_sum_step")
        return dsum
    acc = torch.zeros(d, dtype=torch.float64)
    for q in sel.tolist():
        acc += U[q, :d].double()
    w64[:d] += acc
    if w32 is not None:
        w32[:d] = w64[:d].float()
    return acc if want_dsum else None
