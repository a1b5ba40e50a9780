#!/usr/bin/env python
"""Device time of K1 (local softmax step): tensor-core kernel against the SIMT kernel, 100 peers x 10 rows x 784 features.
CUDA events, 20 warm-up launches, shard matrix (188 MB) larger than L2."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from biscotti_b200 import _native  # noqa: E402
from biscotti_b200.ops import ml  # noqa: E402

dev = "cuda:0"
L = _native.load()
P, Ns, B, F, C = 100, 600, 10, 784, 10
torch.manual_seed(0)
w = (torch.randn(C * F + C) * 0.05).to(dev)
X = torch.randn(P, Ns, F, device=dev)
y = torch.randint(0, C, (P, Ns), dtype=torch.int32, device=dev)
idxs = [torch.stack([torch.randperm(Ns)[:B] for _ in range(P)]).int().to(dev) for _ in range(16)]
out = torch.empty(P, C * F + C + 2, device=dev)


def timed(n=64, reps=10):
    """n launches captured in one CUDA graph (so the host's per-call cost does not bound the rate), replayed reps times."""
    import ctypes

    loss = torch.empty(P, device=dev)
    for i in range(20):
        ml.softmax_step(w, X, y, idxs[i % 16], C, 100.0, out=out, ldu=out.stride(0))
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        for i in range(n):
            rc = L.bsc_local_softmax_step(ml.ptr(w), ml.ptr(X), ml.ptr(y), ml.ptr(idxs[i % 16]), ml.ptr(out), ml.ptr(loss), P, Ns, B, F,
                                          C, out.stride(0), ctypes.c_float(100.0), st)
            assert rc == 0
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (n * reps) * 1e3


res = {}
L.bsc_softmax_force_simt(1)
res["simt_us"] = round(timed(), 2)
L.bsc_softmax_force_simt(0)
res["tcgen05_us"] = round(timed(), 2)
print("BENCH_K1 " + json.dumps(res))
