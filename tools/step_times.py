#!/usr/bin/env python
"""Per-block wall time of the first blocks of a run (diagnosis: how long until the steady state of bench.py is reached)."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from biscotti_b200.config import BiscottiConfig  # noqa: E402
from biscotti_b200.protocol.engine import BiscottiEngine  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 60
ring = int(sys.argv[2]) if len(sys.argv) > 2 else 3
cfg = BiscottiConfig(num_nodes=100, dataset="mnist", max_iterations=n + 4, ring_depth=ring)
eng = BiscottiEngine(cfg, device="cuda:0")
torch.cuda.synchronize()
ts = []
sync_at = int(os.environ.get("SYNC_AT", "-1"))
for i in range(n):
    if i == sync_at:
        torch.cuda.synchronize()
    t0 = time.perf_counter()
    eng.step()
    ts.append((time.perf_counter() - t0) * 1e3)
    if i in (sync_at, sync_at + 1):
        print("AFTER_SYNC step", i, {k: round(v, 3) for k, v in eng.reports[-1].phase_ms.items()})
eng.join_streams()
print("ring", ring, "chain", None)
print("STEP_MS " + " ".join(f"{t:.2f}" for t in ts))
import numpy as np  # noqa: E402

U = eng.U[:, : eng.d]
print("mean of blocks 5-24: %.3f ms, 25-59: %.3f ms" % (np.mean(ts[5:25]), np.mean(ts[25:60])))
keys = ["host_queue", "host_wait", "host_tail_py", "block", "h_append", "host_after", "h_plan", "h_front", "h_witness", "total_host"]
for par in (0, 1):
    reps = [r for r in eng.reports[20:] if r.iteration % 2 == par]
    print("PARITY", par, {k: round(float(np.mean([r.phase_ms.get(k, 0.0) for r in reps])), 3) for k in keys},
          "step", round(float(np.mean([ts[r.iteration] for r in reps if r.iteration < len(ts)])), 3))
print("HASH", eng.chain.latest_hash().hex()[:16])
eng.close()
