#!/usr/bin/env python
"""Device time of batched model evaluation: tcgen05 (eval_tc_kernel) vs SIMT (eval_softmax_kernel)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from biscotti_b200.ops import ml  # noqa: E402

dev = "cuda:0"
torch.manual_seed(0)
N, F, C = 2000, 784, 10
X = torch.randn(N, F, device=dev)
y = torch.randint(0, C, (N,), dtype=torch.int32, device=dev)
base = torch.randn(C * F + C, device=dev) * 0.1
for M in (1, 70, 280):
    Wc = torch.randn(M, C * F + C + 2, device=dev) * 0.05
    for impl in ("simt", "tc"):
        os.environ["BISCOTTI_EVAL_IMPL"] = impl
        for _ in range(3):
            ml.eval_softmax_errors(Wc, base, X, y, C)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            ml.eval_softmax_errors(Wc, base, X, y, C)
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / 20 * 1e3
        flops = 2.0 * N * F * C * M
        print(f"BENCH_EVAL M={M} N={N} impl={impl}: {us:.1f} us per call ({flops / us / 1e6:.2f} TFLOP/s useful)", flush=True)
