"""Multi-Krum: Gram matrix, pairwise distances, scores and the accept set.

Reference: ML/Pytorch/client_obj.py:114-143 / ML/code/logistic_validator.py:36-65 (maths) and
DistSys/krum.go:77-166 (clip = floor(0.5 n), sorted-by-id input).  Ties in the final selection are
broken by peer index (numpy's argpartition leaves them unspecified).
"""
from __future__ import annotations

import torch

from ._call import lib, ptr, stream, check


def krum_reference(X: torch.Tensor, clip: int):
    """fp32/fp64 restatement of get_krum_scores + krum.  Returns (scores, accept_mask, D)."""
    n = X.shape[0]
    X = X.double()
    sq = (X * X).sum(dim=1)
    D = sq[:, None] + sq[None, :] - 2.0 * (X @ X.t())
    group = n - clip
    srt, _ = torch.sort(D, dim=1)
    scores = srt[:, 1: max(1, group - 1)].sum(dim=1)
    order = sorted(range(n), key=lambda i: (float(scores[i]), i))
    accept = torch.zeros(n, dtype=torch.bool)
    accept[order[: n - clip]] = True
    return scores, accept, D


def gram_simt(A: torch.Tensor, d: int, ksplit: int = 1):
    """Partial Gram matrices [ksplit, n, n] of A[:, :d] with the SIMT fp32 kernel (baseline path)."""
    n = A.shape[0]
    G = torch.empty(ksplit, n, n, dtype=torch.float32, device=A.device)
    check(lib().bsc_gram_simt(ptr(A), ptr(G), n, d, A.stride(0), ksplit, stream(A.device)), "gram_simt")
    return G


def krum_select(Gpart: torch.Tensor, norms: torch.Tensor, n: int, clip: int, want_D: bool = False):
    """Gpart: [nsplit, ldg, ldg] fp32 partial Grams.  Returns (scores [n], accept uint8 [n], D or None)."""
    nsplit, ldg = Gpart.shape[0], Gpart.shape[1]
    if not Gpart.is_cuda:
        G = Gpart[:, :n, :n].sum(dim=0)
        D = norms[:n, None] + norms[None, :n] - 2.0 * G
        group = n - clip
        srt, _ = torch.sort(D, dim=1)
        scores = srt[:, 1: max(1, group - 1)].sum(dim=1)
        order = sorted(range(n), key=lambda i: (float(scores[i]), i))
        accept = torch.zeros(n, dtype=torch.uint8)
        accept[order[: n - clip]] = 1
        return scores, accept, (D if want_D else None)
    scores = torch.empty(n, dtype=torch.float32, device=Gpart.device)
    accept = torch.empty(n, dtype=torch.uint8, device=Gpart.device)
    D = torch.empty(n, n, dtype=torch.float32, device=Gpart.device) if want_D else None
    L = lib()
    if hasattr(L, "bsc_krum_select_mc"):
        check(L.bsc_krum_select_mc(ptr(Gpart), nsplit, ldg, ptr(norms), None, n, clip, ptr(scores), ptr(accept), ptr(D),
                                   stream(Gpart.device)), "krum_select_mc")
    else:
        check(L.bsc_krum_select(ptr(Gpart), nsplit, ldg, ptr(norms), n, clip, ptr(scores), ptr(accept), ptr(D),
                                stream(Gpart.device)), "krum_select")
    return scores, accept, D


def krum(Un: torch.Tensor, d: int, clip: int, hi=None, lo=None, norms=None, impl: str = "auto"):
    """Full Multi-Krum over the rows of Un[:, :d].  On CUDA uses the tcgen05 Gram kernel when the bf16
    planes are supplied (impl auto/tcgen05), else the SIMT fp32 kernel."""
    n = Un.shape[0]
    if not Un.is_cuda:
        X = Un[:, :d].float()
        G = (X @ X.t())[None]
        nr = (X * X).sum(dim=1) if norms is None else norms
        return krum_select(G, nr, n, clip)
    if norms is None:
        norms = (Un[:, :d] ** 2).sum(dim=1)
    if impl in ("auto", "tcgen05") and hi is not None:
        from . import gram_tc

        if gram_tc.available():
            Gp = gram_tc.gram_tc(hi, lo, n)
            return krum_select(Gp, norms, n, clip)
        if impl == "tcgen05":
            raise RuntimeError("tcgen05 Gram kernel requested but not available")
    ksplit = max(1, min(32, d // 512))
    Gp = gram_simt(Un, d, ksplit)
    return krum_select(Gp, norms, n, clip)
