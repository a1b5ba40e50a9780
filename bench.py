#!/usr/bin/env python
"""Headline benchmark: block iterations/sec at 100 peers, MNIST softmax (BASELINE.json).

    python bench.py --gpus N --steps K --warmup W            # this framework
    python bench.py --impl reference ...                     # the unmodified reference, if installable

One "step" is one complete Biscotti block iteration with everything on (local SGD step of every
worker, KZG commitments, VRF noiser lottery + DP noise, Multi-Krum, Shamir shares + witnesses for
every approved worker, miner aggregation, reconstruction, SGD step, block hash + ledger append,
test-error / attack-rate evaluation) — the same work the reference's 39.5 s/block covers.
Total work is fixed at 100 peers, so scaling over N GPUs is strong scaling.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
import threading
import time

BASELINE_BLOCKS_PER_S = 1.0 / 39.5          # BASELINE.md headline: 39.5 s/block, 100 peers, MNIST
REF_UNAVAILABLE = ("reference is a Go 1.10 + CPython 2.7 program (go-python, net/rpc); no go/python2 toolchain in "
                   "this image and it has no setup.py/pyproject, so pip cannot build or install it")


def reference_arm(args):
    print(json.dumps({"impl": "reference", "unavailable": REF_UNAVAILABLE}))
    return 0


def cpu_oracle_arm(args):
    """Same-box anchor: the identical block (same peers, committees, flags, step count) on device="cpu" - PyTorch-CPU local
    steps and Krum, the host restatements of the Shamir / MSM / VRF / Schnorr arithmetic (the CPU mirrors every kernel is
    tested against).  It restates ML/Pytorch/client.py:38-65, client_obj.py:114-143, DistSys/kyber.go:533-646, 809-857 with
    this framework's algorithms (window-free Straus MSM instead of bn256's double-and-add), single process, no sockets - so
    it is a floor for what the reference's own stack would need on this CPU, not the reference itself."""
    import torch

    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from biscotti_b200.config import BiscottiConfig
    from biscotti_b200.protocol.engine import BiscottiEngine

    torch.set_num_threads(max(1, os.cpu_count() or 1))
    cfg = BiscottiConfig(num_nodes=args.peers, dataset=args.dataset, max_iterations=args.steps + args.warmup + 2,
                         prefetch=False)      # every phase is booked under its own block
    eng = BiscottiEngine(cfg, device="cpu")
    for _ in range(args.warmup):
        eng.step()
    t0 = time.perf_counter()
    phases = {}
    for _ in range(args.steps):
        rep = eng.step()
        for k, v in rep.phase_ms.items():
            phases[k] = phases.get(k, 0.0) + v
    dt = time.perf_counter() - t0
    ms = 1e3 * dt / args.steps
    print(json.dumps({
        "metric": "block iterations/sec at 100 peers MNIST", "value": 1000.0 / ms, "unit": "blocks/s", "n_gpus": 0,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": (1000.0 / ms) / BASELINE_BLOCKS_PER_S, "dtype": "fp32 / int64 / BN256 256-bit (CPU)",
        "data": "synthetic (Gaussian-blob MNIST-shaped shards, zero genesis weights)", "impl": "cpu-oracle",
        "config": {"model": f"{args.dataset} d={cfg.dim}", "peers": cfg.num_nodes, "global_batch": cfg.num_nodes * cfg.batch_size,
                   "parallelism": f"single process, {torch.get_num_threads()} CPU threads", "timing": "wall clock"},
        "phase_ms": {k: v / args.steps for k, v in phases.items()},
        "chain_hash": eng.chain.latest_hash().hex(),
        "last_block": {"contributors": rep.n_contributors, "accepted": rep.n_accepted, "krum_set": rep.n_krum,
                       "test_error": rep.test_error, "empty": rep.empty}}))
    return 0


class ClockSampler(threading.Thread):
    """Samples SM clock + throttle reasons through NVML while the timed region runs."""

    def __init__(self, index: int, period: float = 0.025):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop_evt = threading.Event()
        self.ok = False
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.ok = False

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop_evt.wait(self.period)

    def arm(self):
        """Start of the timed region: forget what was sampled during the warm-up."""
        self.samples, self.reasons = [], set()

    def stop(self):
        self._stop_evt.set()
        if self.ok:
            self.join(timeout=2)
        return {"sm_mhz": (statistics.median(self.samples) if self.samples else None), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "cpu-oracle"])
    ap.add_argument("--peers", type=int, default=100)
    ap.add_argument("--dataset", default="mnist")
    ap.add_argument("--comm", default=None, help="fused | nccl (default: fused when available)")
    ap.add_argument("--gram", default="auto")
    ap.add_argument("--data-sep", type=float, default=None,
                    help="class separation of the synthetic data (default: the package default, overlapping classes; 6.0 = "
                         "the well separated blobs round 1 was measured on)")
    ap.add_argument("--profile-phases", action="store_true")
    ap.add_argument("--flush-l2", action="store_true",
                    help="write a 256 MiB buffer between timed iterations (default: off - a block's inputs are far larger "
                         "than L2, see `l2` in the output line)")
    ap.add_argument("--e2e-pipeline", type=int, default=1,
                    help="1: double-buffered H2D + next_inputs (the engine starts the next local step while this block is "
                         "assembled); 0: copy then step")
    args = ap.parse_args()
    if args.impl == "reference":
        return reference_arm(args)
    if args.impl == "cpu-oracle":
        return cpu_oracle_arm(args)

    import torch

    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from biscotti_b200 import _native
    from biscotti_b200.config import BiscottiConfig
    from biscotti_b200.parallel.comm import init_distributed, NcclComm
    from biscotti_b200.protocol.engine import BiscottiEngine

    if not torch.cuda.is_available():
        print(json.dumps({"error": "no CUDA device; bench.py measures the B200 path"}))
        return 1
    rank, local, world = init_distributed("cuda")
    if world == 1:
        torch.cuda.set_device(0)
        local = 0
    dev = torch.device("cuda", local)
    L = _native.load()
    kw = {} if args.data_sep is None else {"data_sep": args.data_sep}
    cfg = BiscottiConfig(num_nodes=args.peers, dataset=args.dataset, max_iterations=args.steps + args.warmup + 8,
                         gram_impl=args.gram, **kw)
    comm = None
    comm_name = "single"
    if world > 1:
        comm_name = args.comm or "fused"
        comm = None
        if comm_name == "fused":
            try:
                from biscotti_b200.parallel.fused import FusedComm

                comm = FusedComm(rank, world, dev, cfg)
            except Exception as e:  # fused path unavailable: say so and use the library baseline
                if rank == 0:
                    print(f"[bench] fused comm unavailable ({e}); using nccl baseline", file=sys.stderr)
                comm_name = "nccl"
        if comm is None:
            comm = NcclComm(rank, world, dev)
    eng = BiscottiEngine(cfg, device=dev, rank=rank, world=world, comm=comm)

    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev) if args.flush_l2 else None   # > 126 MB L2

    def barrier():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize(dev)

    # ---------------- device-timed arm: resident shards, CUDA events per step --------------------
    # The NVML thread starts before the warm-up (library start-up and the first calls' symbol look-ups cost milliseconds of
    # interpreter time); what it sampled before the timed region is dropped at arm().
    sampler = ClockSampler(local)
    sampler.start()
    for _ in range(args.warmup):
        eng.step()
    barrier()
    sampler.arm()
    L.bsc_launch_count_reset()
    evs, wall_seq = [], []
    wall0 = time.perf_counter()
    e_start, e_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e_start.record()
    for _ in range(args.steps):
        if flush is not None:
            flush.zero_()                               # L2 flush between timed iterations (not timed)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        tw = time.perf_counter()
        eng.step()
        wall_seq.append(round((time.perf_counter() - tw) * 1e3, 3))
        e1.record()
        evs.append((e0, e1))
    eng.join_streams()        # the closing event sits behind everything the K steps queued, off-chain work included
    e_end.record()
    barrier()
    wall = time.perf_counter() - wall0
    launches = int(L.bsc_launch_count())
    step_seq = [round(a.elapsed_time(b), 3) for a, b in evs]     # in order: shows where a slow step sits
    step_ms = sorted(step_seq)
    dev_ms = e_start.elapsed_time(e_end)     # ONE interval around the K steps: no untimed gaps between steps
    t = torch.tensor([dev_ms], dtype=torch.float64, device=dev)
    if world > 1:
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
    dev_ms = float(t.item())
    ms_per_step = dev_ms / args.steps

    # ---------------- end-to-end arm: public API, minibatches from pinned host memory ------------
    P, B, F = eng.P, cfg.batch_size, cfg.spec.n_features
    # A ring of distinct host minibatches cut from the peers' shards (one per step, cycling): feeding the SAME minibatch
    # every step would let the model fit it, the updates would quantise to zero and the MSMs would have nothing to do.
    RING = max(1, min(32, cfg.shard_size // B))
    hostXs, hostys = [], []
    Xc, yc = eng.data.X.cpu(), eng.data.y.cpu()
    for r in range(RING):
        hx = torch.empty(max(1, P), B, F).pin_memory()
        hy = torch.empty(max(1, P), B, dtype=torch.int32).pin_memory()
        if P:
            hx.copy_(Xc[:, r * B: (r + 1) * B])
            hy.copy_(yc[:, r * B: (r + 1) * B])
        hostXs.append(hx)
        hostys.append(hy)
    hostX, hosty = hostXs[0], hostys[0]
    feed_i = [0]
    # Optional double-buffered input pipeline (--e2e-pipeline 1): while step i runs, step i+1's minibatch is copied from pinned
    # host memory on a copy stream and handed to the engine as `next_inputs`, which lets it queue that step's local
    # pass behind this step's aggregation.  Every step's H2D copy is inside the timed region either way.
    bufs = [(torch.empty_like(hostX, device=dev), torch.empty_like(hosty, device=dev)) for _ in range(2)]
    copy_stream = torch.cuda.Stream(device=dev)
    res_host = torch.empty(2, dtype=torch.float32).pin_memory()

    def h2d(slot):
        ev = torch.cuda.Event()
        k = feed_i[0] % RING
        feed_i[0] += 1
        with torch.cuda.stream(copy_stream):             # slot's previous reader finished a full step ago
            bufs[slot][0].copy_(hostXs[k], non_blocking=True)
            bufs[slot][1].copy_(hostys[k], non_blocking=True)
            ev.record()
        return (bufs[slot][0], bufs[slot][1], ev)

    prev_rep = [None]

    def note(rep):
        # The step's result - new weights, contributor list, share-check counters, commitment||approval rows - is read back
        # from the device inside step() (pinned D2H copies + one event wait, before the block is hashed and appended).  The
        # error rates of the new weights are computed off the block's chain and read back asynchronously: the loop picks up
        # the previous block's (already landed) so that no extra host wait is added per step.
        res_host[0] = float(rep.iteration)
        p = prev_rep[0]
        if p is not None:
            res_host[1] = p.test_error or 0.0
        prev_rep[0] = rep

    def run_e2e(nsteps):
        if not args.e2e_pipeline:
            dX, dy = bufs[0]
            for _ in range(nsteps):
                k = feed_i[0] % RING
                feed_i[0] += 1
                dX.copy_(hostXs[k], non_blocking=True)           # H2D of this step's inputs
                dy.copy_(hostys[k], non_blocking=True)
                rep = eng.step((dX, dy))
                note(rep)
            return
        torch.cuda.current_stream(dev).synchronize()
        cur = h2d(0)
        torch.cuda.current_stream(dev).wait_event(cur[2])
        for i in range(nsteps):
            nxt = h2d((i + 1) % 2)                               # H2D of the next step's inputs (one copy per step)
            rep = eng.step((cur[0], cur[1]), next_inputs=nxt)
            note(rep)
            cur = nxt

    run_e2e(2)
    barrier()
    e2e_t0 = time.perf_counter()
    run_e2e(args.steps)
    barrier()
    e2e_s = time.perf_counter() - e2e_t0
    te = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        torch.distributed.all_reduce(te, op=torch.distributed.ReduceOp.MAX)
    e2e_s = float(te.item())
    clocks = sampler.stop()

    # chain equality across ranks (the reference's correctness oracle)
    chain_ok = True
    if world > 1:
        hs = comm.all_gather_bytes(eng.chain.latest_hash()) if hasattr(comm, "all_gather_bytes") else [eng.chain.latest_hash()]
        chain_ok = all(h == hs[0] for h in hs)

    phase_ms = None
    if args.profile_phases:          # every rank takes the extra instrumented step (collective kernels inside)
        torch.cuda.synchronize(dev)                  # drain the pipelined schedule before switching to the in-line one
        eng._timing = True
        eng._pre, eng.cfg.prefetch = None, False     # book every phase under this one iteration
        eng.step()
        phase_ms = eng.reports[-1].phase_ms
        eng._timing = False

    table_gb = 0.0
    if eng.commit is not None and eng.commit.c_table is not None:
        table_gb = (eng.commit.c_table.numel() + eng.commit.w_table.numel()) * 4 / 2**30
    shard_mb = eng.data.X.numel() * 4 / 2**20
    if rank == 0:
        value = 1000.0 / ms_per_step
        last = eng.reports[-1]
        out = {
            "metric": "block iterations/sec at 100 peers MNIST",
            "value": value,
            "unit": "blocks/s",
            "n_gpus": world,
            "steps": args.steps,
            "warmup": args.warmup,
            "ms_per_step": ms_per_step,
            "higher_is_better": True,
            "scaling": "strong",
            "vs_baseline": value / BASELINE_BLOCKS_PER_S,
            "dtype": "fp32 (Krum Gram on tcgen05 as bf16x3 split, fp32 accumulate); int64 shares; BN256 256-bit",
            "data": "synthetic (MNIST-shaped shards: ten overlapping Gaussian classes, data_sep=%g; zero genesis weights)" % cfg.data_sep,
            "impl": "ours",
            "config": {"model": "mnist softmax d=7850", "peers": cfg.num_nodes, "global_batch": cfg.num_nodes * B,
                       "seq_len": None, "parallelism": f"peers sharded over {world} GPU(s), comm={comm_name}",
                       "committees": {"noisers": cfg.num_noisers, "verifiers": cfg.num_verifiers, "miners": cfg.num_miners},
                       "perc_samples": cfg.perc_samples, "epsilon": cfg.epsilon, "secure_agg": cfg.secure_agg,
                       "commitments": cfg.commitments,
                       "l2": ("256 MiB flush between timed iterations" if args.flush_l2 else
                              "inputs larger than L2: every block gathers ~0.4 GB of fixed-base window-table entries at "
                              "scalar-dependent (fresh every block) addresses out of %.0f GB of tables, and draws new minibatch "
                              "rows out of %.0f MB of resident shards; L2 is 126 MB"
                              % (table_gb, shard_mb)),
                       "timing": "one CUDA-event interval around the K steps (closing event behind all of the engine's streams), "
                                 "max over ranks; step_ms = per-step event pairs on the caller's stream"},
            "clocks": clocks,
            "e2e": {"value": args.steps / e2e_s, "unit": "blocks/s",
                    "h2d_bytes_per_step": int(hostX.numel() * 4 + hosty.numel() * 4) * world,
                    "d2h_bytes_per_step": int(getattr(eng, "last_d2h_bytes", 0)),
                    "timing": "wall clock between barrier+synchronize, max over ranks"},
            "gpu_launches": launches,
            "wall_ms_per_step": wall * 1e3 / args.steps,
            "chain_equal_across_ranks": chain_ok,
            "chain_hash": eng.chain.latest_hash().hex(),          # same on any N: the ledger does not depend on the sharding
            "host_ms": {k: round(sum(r.phase_ms.get(k, 0.0) for r in eng.reports[-args.steps:]) / args.steps, 4)
                        for k in ("host_queue", "host_wait", "host_tail_py", "block", "h_append", "host_after", "h_local_eval",
                                  "h_plan", "h_plan_host", "h_plan_upload", "h_front", "h_commit_sign", "h_witness", "total_host")},
            "step_ms": {"min": step_ms[0], "median": step_ms[len(step_ms) // 2],
                        "p95": step_ms[min(len(step_ms) - 1, int(0.95 * len(step_ms)))], "max": step_ms[-1],
                        "in_order": step_seq, "host_wall_in_order": wall_seq},
            "last_block": {"contributors": last.n_contributors, "accepted": last.n_accepted, "krum_set": last.n_krum,
                           "test_error": last.test_error, "empty": last.empty},
        }
        if phase_ms is not None:
            out["phase_ms"] = phase_ms
        print(json.dumps(out))
    eng.close()
    if world > 1:
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
