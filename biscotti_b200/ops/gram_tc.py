"""tcgen05 / TMA Gram kernel wrapper (gram_tc_kernel in csrc/gram_tc.cu)."""
from __future__ import annotations

import torch

from .. import _native
from ._call import ptr, stream, check


def available() -> bool:
    try:
        L = _native.load()
    except Exception:
        return False
    return hasattr(L, "bsc_gram_tc") and torch.cuda.is_available()


def effective_ksplit(kblocks: int, ksplit: int) -> int:
    """Largest split count <= ksplit such that every split owns at least one k-block."""
    ksplit = max(1, min(ksplit, kblocks))
    per = (kblocks + ksplit - 1) // ksplit
    return (kblocks + per - 1) // per


def ksplit_for(n: int, ldh: int) -> int:
    """Split-K factor: spread the k-blocks over the SMs, but keep >= 4 k-blocks per CTA so the
    fixed-order reduction in krum_select stays small."""
    kblocks = (ldh + 63) // 64
    ntile = (n + 127) // 128
    pairs = ntile * (ntile + 1) // 2
    want = max(1, min(kblocks // 4 if kblocks >= 8 else 1, 148 // pairs if pairs <= 148 else 1))
    return effective_ksplit(kblocks, want)


def gram_tc(hi: torch.Tensor, lo: torch.Tensor, n: int, ksplit: int | None = None):
    """Partial Gram matrices [ksplit, ldg, ldg] fp32 of the bf16x3-split rows (hi + lo)."""
    L = _native.load()
    ldh = hi.shape[1]
    kblocks = ldh // 64
    ksplit = effective_ksplit(kblocks, ksplit) if ksplit else ksplit_for(n, ldh)
    ldg = (n + 127) // 128 * 128
    G = torch.empty(ksplit, ldg, ldg, dtype=torch.float32, device=hi.device)
    check(L.bsc_gram_tc(ptr(hi), ptr(lo), n, hi.shape[0], ldh, ksplit, ptr(G), ldg, stream(hi.device)), "gram_tc")
    return G
