"""Device-resident links of a block's dependency chain (csrc/kernels_chain.cu) and their host restatements.

Reference steps replaced: noiser lottery from the private VRF output (DistSys/vrf.go:54-100), "who got approved
and whose shares arrived first" (DistSys/krum.go:338-358, main.go:360, 1686, 2079), the verifiers' Schnorr
approvals (kyber.go:873-896).  On a GPU each of them reads its input where the previous kernel left it; the
host restatements below are what the CPU plumbing path runs and what the GPU tests compare the kernels with.
"""
from __future__ import annotations

import ctypes
import hashlib

import numpy as np
import torch

from ._call import lib, ptr, stream, check


def seed64(*parts) -> int:
    """64-bit seed from a tuple of run parameters (identical on every rank)."""
    h = hashlib.sha256("-".join(str(p) for p in parts).encode()).digest()
    return int.from_bytes(h[:8], "little")


def seeded_perm(seed: int, n: int) -> np.ndarray:
    """Deterministic permutation of 0..n-1 (ledger.cpp: bsc_seeded_perm) - the stand-in for a network arrival order."""
    out = np.empty(max(1, n), dtype=np.int32)
    lib().bsc_seeded_perm(ctypes.c_ulonglong(seed & (2**64 - 1)), n, out.ctypes.data_as(ctypes.c_void_p))
    return out[:n]


# ---------------------------------------------------------------------------------------------------------
# block plan
# ---------------------------------------------------------------------------------------------------------
COUNT_APPROVED, COUNT_CONTRIB, COUNT_LOCAL_CONTRIB, COUNT_LOCAL_APPROVED = 0, 1, 2, 3


def block_plan_host(accept, share_rank, local_row, approved_ok: bool, trigger: int):
    """Restatement of block_plan_kernel on plain Python lists.  Returns (counts[4], contrib_pos, sel, appr_local, csel)."""
    nk = len(accept)
    appr = [bool(accept[i]) and approved_ok for i in range(nk)]
    order = sorted((share_rank[i], i) for i in range(nk) if appr[i])
    con = [False] * nk
    for _, i in order[:max(0, trigger)]:
        con[i] = True
    contrib_pos = [i for i in range(nk) if con[i]]
    appr_local = [local_row[i] for i in range(nk) if appr[i] and local_row[i] >= 0]
    sel, csel = [], []
    nla = 0
    for i in range(nk):
        if con[i] and local_row[i] >= 0:
            sel.append(local_row[i])
            csel.append(nla)
        if appr[i] and local_row[i] >= 0:
            nla += 1
    counts = [sum(appr), len(contrib_pos), len(sel), len(appr_local)]
    return counts, contrib_pos, sel, appr_local, csel


def block_plan(accept: torch.Tensor, share_rank: torch.Tensor, local_row: torch.Tensor, approved_ok: bool, trigger: int,
               out: dict):
    """accept: uint8 [nk]; share_rank / local_row: int32 [nk].  `out` holds persistent int32 tensors counts[8],
    contrib_pos[>=nk], sel, appr_local, csel [>= local candidates]; they are filled in place."""
    nk = accept.numel()
    if accept.is_cuda:
        check(lib().bsc_block_plan(ptr(accept), ptr(share_rank), ptr(local_row), nk, int(bool(approved_ok)), int(trigger),
                                   ptr(out["counts"]), ptr(out["contrib_pos"]), ptr(out["sel"]), ptr(out["appr_local"]),
                                   ptr(out["csel"]), stream(accept.device)), "block_plan")
        return
    counts, cpos, sel, al, cs = block_plan_host(accept.tolist(), share_rank.tolist(), local_row.tolist(), approved_ok, trigger)
    out["counts"][:4] = torch.tensor(counts, dtype=torch.int32)
    for name, vals in (("contrib_pos", cpos), ("sel", sel), ("appr_local", al), ("csel", cs)):
        if vals:
            out[name][: len(vals)] = torch.tensor(vals, dtype=torch.int32)


# ---------------------------------------------------------------------------------------------------------
# VRF output + noiser lottery on the device
# ---------------------------------------------------------------------------------------------------------
def vrf_noisers(xyz: torch.Tensor, msg_dev: torch.Tensor, mlen: int, stake: torch.Tensor, is_worker: torch.Tensor,
                peer_ids: torch.Tensor, want: int, out: torch.Tensor, vrf_out: torch.Tensor | None = None):
    """xyz int32 [n, 30] projective Gamma of every peer; writes out int32 [P, want] (-1 = none)."""
    n = xyz.shape[0]
    P = peer_ids.numel()
    check(lib().bsc_vrf_noisers(ptr(xyz), ptr(msg_dev), mlen, ptr(stake), n, ptr(is_worker), ptr(peer_ids), P, want, ptr(out),
                                ptr(vrf_out), stream(xyz.device)), "vrf_noisers")
    return out


# ---------------------------------------------------------------------------------------------------------
# approvals
# ---------------------------------------------------------------------------------------------------------
def approval_tag(salt64: bytes, it: int, peer: int, verifier: int) -> bytes:
    """Nonce tag of one approval: per-run salt || previous block hash || (iteration, update, verifier).  Binding the
    chain position means a signing key never meets the same nonce on two different chains (two runs, a fork, a
    resumed run with other flags) - equal nonces over different commitments would leak the key."""
    import struct

    assert len(salt64) == 64
    return salt64 + struct.pack("<qii", it, peer, verifier)
