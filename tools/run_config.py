#!/usr/bin/env python
"""Run one BASELINE.json configuration for a few blocks on cuda:0 and print a JSON summary
(blocks/s, per-phase device ms, Krum / secure-aggregation microseconds, final errors)."""
import argparse
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from biscotti_b200.config import BiscottiConfig  # noqa: E402
from biscotti_b200.protocol.engine import BiscottiEngine  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--dataset", default="mnist")
ap.add_argument("--peers", type=int, default=100)
ap.add_argument("--steps", type=int, default=10)
ap.add_argument("--warmup", type=int, default=2)
ap.add_argument("--poison", type=float, default=0.0)
ap.add_argument("--epsilon", type=float, default=2.0)
ap.add_argument("--shard-size", type=int, default=600)
ap.add_argument("--affine-wire", action="store_true")
ap.add_argument("--msm-table-gb", type=float, default=None)
a = ap.parse_args()

cfg = BiscottiConfig(num_nodes=a.peers, dataset=a.dataset, max_iterations=a.steps + a.warmup + 4, poison_fraction=a.poison,
                     epsilon=a.epsilon, shard_size=a.shard_size, affine_wire=a.affine_wire, msm_table_gb=a.msm_table_gb)
t0 = time.perf_counter()
eng = BiscottiEngine(cfg, device="cuda:0")
torch.cuda.synchronize()
init_s = time.perf_counter() - t0
for _ in range(a.warmup):
    eng.step()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(a.steps):
    eng.step()
torch.cuda.synchronize()
dt = time.perf_counter() - t0
eng._timing = True
eng.step()
r = eng.reports[-1]
out = {"dataset": a.dataset, "peers": a.peers, "d": cfg.dim, "init_s": round(init_s, 2), "ms_per_block": 1e3 * dt / a.steps,
       "blocks_per_s": a.steps / dt, "phase_ms": {k: round(v, 3) for k, v in r.phase_ms.items()},
       "krum_us": round(1e3 * r.phase_ms.get("verify", 0.0), 1),
       "secagg_us": round(1e3 * (r.phase_ms.get("witness", 0) + r.phase_ms.get("miner_aggregate", 0) + r.phase_ms.get("secagg", 0)), 1),
       "contributors": r.n_contributors, "krum_set": r.n_krum, "test_error": r.test_error, "attack_rate": r.attack_rate,
       "bad_shares": sum(x.bad_shares for x in eng.reports), "mem_gb": round(torch.cuda.max_memory_allocated() / 2**30, 2),
       "msm_window": [eng.commit.w, eng.commit.nwin, eng.commit.w1] if eng.commit else None,
       "msm_table_gb": round(eng.commit.table_bytes / 2**30, 2) if eng.commit else None, "chain_ok": eng.chain.verify()}
print("RUN_CONFIG " + json.dumps(out))
