"""Allocation-free Multi-Krum entry point (bsc_krum_scores_static), driven through the C ABI.

Who needs it: callers that must get a Krum verdict at a point where they are not allowed to allocate
device memory — e.g. a verifier loop that keeps the allocator untouched between blocks, or a host
program that embeds the library under a memory budget fixed at start-up.  The entry point keeps its
split-K scratch in a fixed 512 KiB array inside the library image (limit: 128 peers) instead of a
caller buffer.  This file is named so that pytest collects it last.

The checks: on two consecutive calls (different sizes, so the second call finds the first call's
leftovers in the shared scratch) the scores must be bit-identical to what the existing allocating entry
points (bsc_gram_simt + bsc_krum_select) produce for the same inputs, scores and accept set must agree
with a float64 NumPy restatement of the reference's Krum (client_obj.py:114-143), the amount of free
device memory must not move across a call, and a call must be capturable into a CUDA graph (any
allocation or synchronisation inside it would abort the capture).
"""
import ctypes

import numpy as np
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


def _numpy_krum(X, clip):
    """client_obj.py:114-143 in float64: squared distances, per row the sum of the sorted distances [1 : n-clip-1),
    accept set = the n-clip smallest scores."""
    n = len(X)
    sq = np.sum(X ** 2, axis=1)
    dist = sq[:, None] + sq[None] - 2 * X @ X.T
    group = n - clip
    scores = np.array([np.sum(np.sort(dist[i])[1:group - 1]) for i in range(n)])
    return scores, np.sort(np.argpartition(scores, group)[:group])


def _problem(n, d, dev):
    """n - n//2 rows around the origin and n//2 outliers, each pushed away by its own offset, so that the accept set
    (clip = n//2) is exactly the first group whatever the rounding of the scores."""
    A = (torch.randn(n, d + 2, device=dev) * 0.1).contiguous()
    clip = n // 2
    A[n - clip:] += 0.3 * torch.arange(1, clip + 1, device=dev, dtype=torch.float32).unsqueeze(1)
    norms = (A[:, :d] ** 2).sum(dim=1).contiguous()
    return A, norms, clip


def test_static_scratch_krum_matches_allocating_entry_points_twice_in_a_row(lib):
    _declare(lib)
    dev = torch.device("cuda:0")
    stream = c_p(torch.cuda.current_stream(dev).cuda_stream)
    torch.manual_seed(11)
    # first call: full-size problem (fills the scratch); second: smaller n and d, so stale partials from
    # the first call sit right next to the region the second call uses
    for n, d in ((128, 7850), (37, 600)):
        A, norms, clip = _problem(n, d, dev)
        scores = torch.empty(n, dtype=torch.float32, device=dev)
        accept = torch.empty(n, dtype=torch.uint8, device=dev)
        rc = lib.bsc_krum_scores_static(_ptr(A), _ptr(norms), n, d, A.stride(0), clip, _ptr(scores), _ptr(accept), stream)
        assert rc == 0
        ref_scores, ref_accept = _reference(lib, A, norms, n, d, clip, stream)
        torch.cuda.synchronize()
        assert torch.equal(scores, ref_scores)          # same kernels, same split count: same bits
        assert torch.equal(accept, ref_accept)
        assert int(accept.sum()) == n - clip
        np_scores, np_accept = _numpy_krum(A[:, :d].double().cpu().numpy(), clip)
        assert np.allclose(scores.cpu().numpy(), np_scores, rtol=5e-3, atol=5e-2)
        assert np.array_equal(np.flatnonzero(accept.cpu().numpy()), np_accept)
        assert np.array_equal(np_accept, np.arange(n - clip))


def test_static_scratch_krum_allocates_nothing_and_can_be_captured(lib):
    _declare(lib)
    dev = torch.device("cuda:0")
    torch.manual_seed(12)
    n, d = 64, 2048
    A, norms, clip = _problem(n, d, dev)
    scores = torch.empty(n, dtype=torch.float32, device=dev)
    accept = torch.empty(n, dtype=torch.uint8, device=dev)
    stream = c_p(torch.cuda.current_stream(dev).cuda_stream)
    args = (_ptr(A), _ptr(norms), n, d, A.stride(0), clip, _ptr(scores), _ptr(accept))
    assert lib.bsc_krum_scores_static(*args, stream) == 0      # warm-up: symbol lookup, module load
    torch.cuda.synchronize()
    want_scores, want_accept = scores.clone(), accept.clone()
    free0, _ = torch.cuda.mem_get_info(dev)
    assert lib.bsc_krum_scores_static(*args, stream) == 0
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info(dev)
    assert free1 == free0, (free0, free1)
    # a stream capture aborts on cudaMalloc / cudaFree / any synchronisation inside the captured region
    scores.zero_()
    accept.zero_()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        cap = c_p(torch.cuda.current_stream(dev).cuda_stream)
        assert lib.bsc_krum_scores_static(*args, cap) == 0
    graph.replay()
    torch.cuda.synchronize()
    assert torch.equal(scores, want_scores) and torch.equal(accept, want_accept)


def test_static_scratch_krum_rejects_a_row_stride_shorter_than_the_row(lib, capfd):
    _declare(lib)
    dev = torch.device("cuda:0")
    n, d = 16, 64
    A = torch.zeros(n, d, device=dev)
    norms = torch.zeros(n, device=dev)
    scores = torch.zeros(n, device=dev)
    accept = torch.zeros(n, dtype=torch.uint8, device=dev)
    rc = lib.bsc_krum_scores_static(_ptr(A), _ptr(norms), n, d, d - 1, n // 2, _ptr(scores), _ptr(accept),
                                    c_p(torch.cuda.current_stream(dev).cuda_stream))
    assert rc == -1
    assert "smaller than d" in capfd.readouterr().err


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
