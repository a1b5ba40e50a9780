"""Kernel timeline of steady-state blocks (diagnosis only - numbers under a profiler are never bench values).

    python tools/timeline.py [--peers 100] [--dataset mnist] [--steps 6] [--show 1]

Runs the engine under torch.profiler, then prints, for the last `--show` blocks, every kernel / memcpy with its
stream, start offset and duration (us), so the critical path and the overlap between the chain stream and the
off-chain streams can be read directly.
"""
import argparse
import json
import os
import sys
import tempfile

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from biscotti_b200.config import BiscottiConfig  # noqa: E402
from biscotti_b200.protocol.engine import BiscottiEngine  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--peers", type=int, default=100)
    ap.add_argument("--dataset", default="mnist")
    ap.add_argument("--steps", type=int, default=6)
    ap.add_argument("--show", type=int, default=1)
    ap.add_argument("--sep", type=float, default=None)
    a = ap.parse_args()
    kw = {} if a.sep is None else {"data_sep": a.sep}
    cfg = BiscottiConfig(num_nodes=a.peers, dataset=a.dataset, max_iterations=a.steps + 40, **kw)
    from biscotti_b200.parallel.comm import init_distributed

    rank, local, world = init_distributed("cuda")
    dev = torch.device("cuda", local)
    comm = None
    if world > 1:
        from biscotti_b200.parallel.fused import FusedComm

        comm = FusedComm(rank, world, dev, cfg)
    eng = BiscottiEngine(cfg, device=dev, rank=rank, world=world, comm=comm)
    for _ in range(25):
        eng.step()
    torch.cuda.synchronize()
    from torch.profiler import profile, ProfilerActivity, record_function

    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
        for i in range(a.steps):
            with record_function(f"BLOCK{i}"):
                eng.step()
        torch.cuda.synchronize()
    if rank != 0:
        torch.distributed.barrier()
        return
    path = os.path.join(tempfile.gettempdir(), "bsc_trace.json")
    prof.export_chrome_trace(path)
    ev = json.load(open(path))["traceEvents"]
    blocks = sorted([e for e in ev if e.get("name", "").startswith("BLOCK") and e.get("ph") == "X" and e.get("cat") == "user_annotation"],
                    key=lambda e: e["ts"])
    gpu = [e for e in ev if e.get("ph") == "X" and e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset")]
    gpu.sort(key=lambda e: e["ts"])
    for b in blocks[-a.show:]:
        t0, t1 = b["ts"], b["ts"] + b["dur"]
        print(f"==== {b['name']}: host span {b['dur']:.0f} us")
        for e in gpu:
            if t0 - 50 <= e["ts"] <= t1 + 400:
                nm = e["name"].replace("bsc::", "")[:60]
                print(f"  {e['ts'] - t0:9.1f} +{e['dur']:8.1f}  s{e['args'].get('stream', '?'):<3} {nm}")
    spans = [blocks[i + 1]["ts"] - blocks[i]["ts"] for i in range(len(blocks) - 1)]
    print("block-to-block host period (us):", [round(s) for s in spans])
    if world > 1:
        torch.distributed.barrier()


if __name__ == "__main__":
    main()
