"""tcgen05 / TMA Gram kernel wrapper (krum_gram_tc in csrc/gram_tc.cu)."""
from __future__ import annotations

import ctypes

import torch

from .. import _native
from ._call import ptr, stream, check

_state = {}


def available() -> bool:
    try:
        L = _native.load()
    except Exception:
        return False
    return hasattr(L, "bsc_gram_tc") and torch.cuda.is_available()


def ksplit_for(n: int, ldh: int) -> int:
    """Split-K factor: enough CTAs to cover the GPU, at least one 64-wide k-block each."""
    kblocks = (ldh + 63) // 64
    tiles = ((n + 127) // 128) ** 2
    return max(1, min(kblocks, 148 // max(1, tiles)))


def gram_tc(hi: torch.Tensor, lo: torch.Tensor, n: int, ksplit: int | None = None):
    """Partial Gram matrices [ksplit, ldg, ldg] fp32 of the bf16x3-split rows (hi + lo)."""
    L = _native.load()
    ldh = hi.shape[1]
    ksplit = ksplit or ksplit_for(n, ldh)
    ldg = (n + 127) // 128 * 128
    G = torch.empty(ksplit, ldg, ldg, dtype=torch.float32, device=hi.device)
    L.bsc_gram_tc.restype = ctypes.c_int
    check(L.bsc_gram_tc(ptr(hi), ptr(lo), n, hi.shape[0], ldh, ksplit, ptr(G), ldg, stream(hi.device)), "gram_tc")
    return G
