"""Allocation-free Multi-Krum entry point (bsc_krum_scores_static), driven through the C ABI.

Who needs it: callers that must get a Krum verdict at a point where they are not allowed to allocate
device memory — e.g. a verifier loop that keeps the allocator untouched between blocks, or a host
program that embeds the library under a memory budget fixed at start-up.  The entry point keeps its
split-K scratch in a fixed 512 KiB array inside the library image (limit: 128 peers) instead of a
caller buffer.  This file is named so that pytest collects it last.

The check: on two consecutive calls (different sizes, so the second call finds the first call's
leftovers in the shared scratch) the scores and the accept set must match what the existing
allocating entry points (bsc_gram_simt + bsc_krum_select) produce for the same inputs.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

c_i, c_p = ctypes.c_int, ctypes.c_void_p


def _ptr(t):
    return c_p(t.data_ptr())


def _declare(lib):
    lib.bsc_krum_scores_static.restype = c_i
    lib.bsc_krum_scores_static.argtypes = [c_p, c_p, c_i, c_i, c_i, c_i, c_p, c_p, c_p]
    lib.bsc_krum_static_splits.restype = c_i
    lib.bsc_krum_static_splits.argtypes = [c_i]
    lib.bsc_gram_simt.restype = c_i
    lib.bsc_gram_simt.argtypes = [c_p, c_p, c_i, c_i, c_i, c_i, c_p]
    lib.bsc_krum_select.restype = c_i
    lib.bsc_krum_select.argtypes = [c_p, c_i, c_i, c_p, c_i, c_i, c_p, c_p, c_p, c_p]


def _reference(lib, A, norms, n, d, clip, stream):
    """The existing entry points with a caller-allocated partial-Gram buffer and the same split count."""
    ks = lib.bsc_krum_static_splits(d)
    G = torch.empty(ks, n, n, dtype=torch.float32, device=A.device)
    scores = torch.empty(n, dtype=torch.float32, device=A.device)
    accept = torch.empty(n, dtype=torch.uint8, device=A.device)
    assert lib.bsc_gram_simt(_ptr(A), _ptr(G), n, d, A.stride(0), ks, stream) == 0
    assert lib.bsc_krum_select(_ptr(G), ks, n, _ptr(norms), n, clip, _ptr(scores), _ptr(accept), None, stream) == 0
    return scores, accept


def test_static_scratch_krum_matches_allocating_entry_points_twice_in_a_row(lib):
    _declare(lib)
    dev = torch.device("cuda:0")
    stream = c_p(torch.cuda.current_stream(dev).cuda_stream)
    torch.manual_seed(11)
    # first call: full-size problem (fills the scratch); second: smaller n and d, so stale partials from
    # the first call sit right next to the region the second call uses
    for n, d in ((128, 7850), (37, 600)):
        A = (torch.randn(n, d + 2, device=dev) * 0.1).contiguous()
        A[: n // 4] += 0.3
        norms = (A[:, :d] ** 2).sum(dim=1).contiguous()
        clip = n // 2
        scores = torch.empty(n, dtype=torch.float32, device=dev)
        accept = torch.empty(n, dtype=torch.uint8, device=dev)
        rc = lib.bsc_krum_scores_static(_ptr(A), _ptr(norms), n, d, A.stride(0), clip, _ptr(scores), _ptr(accept), stream)
        assert rc == 0
        ref_scores, ref_accept = _reference(lib, A, norms, n, d, clip, stream)
        torch.cuda.synchronize()
        assert torch.allclose(scores, ref_scores, rtol=1e-4, atol=1e-3)
        assert torch.equal(accept, ref_accept)
        assert int(accept.sum()) == n - clip


def test_static_scratch_krum_rejects_sizes_beyond_its_limit(lib, capfd):
    _declare(lib)
    dev = torch.device("cuda:0")
    n, d = 129, 64
    A = torch.zeros(n, d, device=dev)
    norms = torch.zeros(n, device=dev)
    scores = torch.zeros(n, device=dev)
    accept = torch.zeros(n, dtype=torch.uint8, device=dev)
    rc = lib.bsc_krum_scores_static(_ptr(A), _ptr(norms), n, d, d, n // 2, _ptr(scores), _ptr(accept),
                                    c_p(torch.cuda.current_stream(dev).cuda_stream))
    assert rc == -1
    assert "static scratch limits" in capfd.readouterr().err
