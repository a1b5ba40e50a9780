#!/usr/bin/env python
"""A/B: wall time per block with and without the next-iteration prefetch (resident shards, 1 GPU)."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from biscotti_b200.config import BiscottiConfig  # noqa: E402
from biscotti_b200.protocol.engine import BiscottiEngine  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 60
for rep in range(2):
    for pf in (False, True):
        cfg = BiscottiConfig(num_nodes=100, dataset="mnist", max_iterations=steps + 20, prefetch=pf)
        eng = BiscottiEngine(cfg, device="cuda:0")
        for _ in range(5):
            eng.step()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(steps):
            eng.step()
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / steps * 1e3
        print(f"prefetch={pf}: {dt:.3f} ms/block  ({1e3 / dt:.1f} blocks/s)  hash={eng.chain.latest_hash().hex()[:12]}", flush=True)
        del eng
        torch.cuda.empty_cache()
