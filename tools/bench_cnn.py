#!/usr/bin/env python
"""Device time of the CNN local-step kernels (mnist_cnn, cifar_cnn, lfw_cnn) for P peers x B images."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from biscotti_b200.models.registry import CNN_ARCH, cnn_param_slices  # noqa: E402
from biscotti_b200.ops import cnn_native  # noqa: E402

P = int(os.environ.get("P", "100"))
B = int(os.environ.get("B", "10"))
kinds = sys.argv[1:] or ["mnist_cnn", "cifar_cnn", "lfw_cnn"]
dev = torch.device("cuda:0")
for kind in kinds:
    arch = CNN_ARCH[kind]
    _, d = cnn_param_slices(arch)
    F = arch["image"][0] * arch["image"][1] * arch["image"][2]
    torch.manual_seed(1)
    w = (torch.randn(d) * 0.02).to(dev)
    X = torch.randn(P, B, F, device=dev)
    y = torch.randint(0, arch["n_classes"], (P, B), dtype=torch.int32, device=dev)
    out = torch.zeros(P, d + 2, device=dev)
    for _ in range(3):
        cnn_native.local_step(kind, w, X, y, 100.0, out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        cnn_native.local_step(kind, w, X, y, 100.0, out)
    e1.record()
    torch.cuda.synchronize()
    print(f"BENCH_CNN {kind} P={P} B={B} d={d}: {e0.elapsed_time(e1) / 10:.3f} ms per local step", flush=True)
    if os.environ.get("PROFILE"):
        from torch.profiler import ProfilerActivity, profile

        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            for _ in range(5):
                cnn_native.local_step(kind, w, X, y, 100.0, out)
            torch.cuda.synchronize()
        rows = sorted(((e.key, e.device_time_total / 5) for e in prof.key_averages() if e.device_time_total > 0), key=lambda kv: -kv[1])
        for k, us in rows[:10]:
            print(f"    {us:9.1f} us  {k[:100]}", flush=True)
