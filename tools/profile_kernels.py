#!/usr/bin/env python
"""Small single-GPU program that launches every hot kernel a few times (for `ncu -k regex:...`)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from biscotti_b200.config import BiscottiConfig  # noqa: E402
from biscotti_b200.ops import cnn_native, ml, secagg  # noqa: E402
from biscotti_b200.parallel.fused import FusedComm  # noqa: E402
from biscotti_b200.protocol.engine import BiscottiEngine  # noqa: E402

dev = torch.device("cuda:0")
# MNIST-100 engine on the fused path (world = 1): local softmax, noise, fused Gram + tail, MSM, secagg, VRF ...
cfg = BiscottiConfig(num_nodes=100, dataset="mnist", max_iterations=20)
eng = BiscottiEngine(cfg, device=dev, comm=FusedComm(0, 1, dev, cfg))
for _ in range(4):
    eng.step()
# CIFAR-200-shaped fused Gram (the bandwidth-relevant case) and CNN local step
cfg2 = BiscottiConfig(num_nodes=128, dataset="cifar")
comm = FusedComm(0, 1, dev, cfg2)
d = cfg2.dim
U = torch.zeros(128, ml.pad_to(d, 4), device=dev)
U[:, :d] = torch.randn(128, d, device=dev) * 0.05
nz = torch.full((128, 1), -1, dtype=torch.int32, device=dev)
for ep in range(1, 4):
    ml.noise_mix(U, nz, 1, 0, 0.0, d, out_hi=comm.hi, out_lo=comm.lo)
    comm.krum_fused(list(range(90)), 45, ep)
    secagg.share_sum(U, torch.arange(35, dtype=torch.int32, device=dev), d, cfg2.total_shares, out=comm.ypart)
    comm.secagg_fused(list(range(10)), [1] * cfg2.total_shares, 4, ep)
w = torch.randn(d, device=dev) * 0.02
Xb = torch.randn(50, 10, 3072, device=dev)
yb = torch.randint(0, 10, (50, 10), dtype=torch.int32, device=dev)
out = torch.zeros(50, ml.pad_to(d, 4), device=dev)
for _ in range(3):
    cnn_native.local_step("cifar_cnn", w, Xb, yb, 100.0, out)
torch.cuda.synchronize()
print("profile_kernels done")
