#!/usr/bin/env python
"""Device time of the per-block VRF evaluation for the scalar-multiplication kernels (diagnostic)."""
import hashlib
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from biscotti_b200.crypto import primitives as cp  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 100
keys = [cp.VRFKey(hashlib.sha256(b"k%d" % i).digest()) for i in range(n)]
msg = hashlib.sha256(b"blk").digest()
for kern in ("wide", "quad4"):
    bv = cp.BatchedVRF(keys, "cuda:0", kernel=kern)
    bv.launch_dev(msg)
    torch.cuda.synchronize()
    ts = []
    for it in range(8):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(bv._dside)
        bv.launch_dev(msg)
        e1.record(bv._dside)
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    print(f"VRF_BENCH kernel={kern} peers={n} device_ms={min(ts[2:]):.3f} all={[round(t, 3) for t in ts]}")
