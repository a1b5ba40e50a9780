#!/usr/bin/env python
"""Device-timed microbenchmark of the two fused compute+collective paths against their NCCL baselines
and their rooflines.  Run under torchrun on N GPUs of one box:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/bench_fused.py

Timing: CUDA events on the launching stream, >= 3 warm-up iterations, barrier + synchronize on both sides,
max over ranks.  Roofline of a fused path = max(FLOPs / measured bf16 GEMM peak, bytes that must cross
NVLink per GPU / measured 770 GB/s per direction)."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from biscotti_b200.config import BiscottiConfig  # noqa: E402
from biscotti_b200.ops import gram_tc, krum as krum_ops, ml, secagg  # noqa: E402
from biscotti_b200.parallel.comm import NcclComm, init_distributed  # noqa: E402
from biscotti_b200.parallel.fused import FusedComm  # noqa: E402

NVLINK_GBS = 770.0
try:
    PEAK_TFLOPS = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["bf16_tflops"]
except Exception:
    PEAK_TFLOPS = 1590.0


def timed(fn, iters, world):
    for _ in range(3):
        fn()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    evs = []
    for _ in range(iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        evs.append((e0, e1))
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    ms = sum(a.elapsed_time(b) for a, b in evs) / iters
    t = torch.tensor([ms], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item()) * 1e3   # us


def main():
    rank, local, world = init_distributed("cuda")
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    results = []
    for dataset, peers in (("mnist", 100), ("cifar", 200)):
        cfg = BiscottiConfig(num_nodes=peers, dataset=dataset)
        comm = FusedComm(rank, world, dev, cfg)
        nccl = NcclComm(rank, world, dev)
        d, per = cfg.dim, comm.per
        P = min(per, max(0, peers - rank * per))
        torch.manual_seed(1 + rank)
        U = torch.zeros(max(1, P), ml.pad_to(d, 4), device=dev)
        U[:, :d] = torch.randn(max(1, P), d, device=dev) * 0.05
        noisers = torch.full((max(1, P), 1), -1, dtype=torch.int32, device=dev)
        ml.noise_mix(U, noisers, 1, 0, 0.0, d, out_hi=comm.hi, out_lo=comm.lo)
        Un, hi_l, lo_l, norms_l = ml.noise_mix(U, noisers, 1, 0, 0.0, d)
        krum_set = list(range(int(peers * 0.7)))
        clip = len(krum_set) // 2
        epoch = [0]

        def fused_krum():
            epoch[0] += 1
            comm.krum_fused(krum_set, clip, epoch[0])

        def nccl_krum():
            # baseline: library all-gather of both planes + un-fused tcgen05 Gram + Krum tail
            li = {p: p for p in range(peers)}
            hik = nccl.gather_rows(krum_set, hi_l, li, [0] * peers)
            lok = nccl.gather_rows(krum_set, lo_l, li, [0] * peers)
            nrm = nccl.gather_rows(krum_set, norms_l.unsqueeze(1), li, [0] * peers).squeeze(1)
            Gp = gram_tc.gram_tc(hik, lok, len(krum_set))
            krum_ops.krum_select(Gp, nrm, len(krum_set), clip)

        t_fk = timed(fused_krum, 20, world)
        t_nk = timed(nccl_krum, 20, world)
        # roofline of the fused Krum path
        n_slots = world * comm.Sr
        kcols = comm.ldh / world
        flops = 2.0 * n_slots * n_slots * kcols * 3 / (2.0 if comm.ntile == 1 else 1.0)   # upper-triangular tiles
        pull_bytes = (world - 1) / world * n_slots * kcols * 4.0                             # hi + lo bf16
        push_bytes = (world - 1) * (comm.ntile * comm.ntile) * 128 * 128 * 4.0   # one reduced Gram per source rank
        roof_krum = max(flops / (PEAK_TFLOPS * 1e12), (pull_bytes + push_bytes) / (NVLINK_GBS * 1e9)) * 1e6

        # secure aggregation
        sel = torch.arange(min(P, 35), dtype=torch.int32, device=dev)
        secagg.share_sum(U, sel, d, cfg.total_shares, out=comm.ypart)
        ep2 = [0]

        def fused_sec():
            ep2[0] += 1
            comm.secagg_fused(list(range(10)), [1] * cfg.total_shares, 4, ep2[0])

        w64 = torch.zeros(d, dtype=torch.float64, device=dev)
        w32 = torch.zeros(d, device=dev)
        xs_t = torch.arange(10, dtype=torch.int32, device=dev)

        bad_n = torch.zeros(4, dtype=torch.int32, device=dev)
        av_t = torch.ones(cfg.total_shares, dtype=torch.uint8, device=dev)

        def nccl_sec():
            # baseline: library all-reduce of the share sums + the same matrix-recovery kernel, un-fused
            Y = comm.ypart.clone()
            nccl.all_reduce_i64(Y)
            secagg.reconstruct_step_dev(Y, list(range(10)), xs_t, av_t, d, w64, w32, bad=bad_n)

        t_fs = timed(fused_sec, 20, world)
        t_ns = timed(nccl_sec, 20, world)
        sec_bytes = (world - 1) / world * cfg.nchunk * cfg.total_shares * 8.0 + (world - 1) / world * d * 12.0
        roof_sec = sec_bytes / (NVLINK_GBS * 1e9) * 1e6
        results.append({"dataset": dataset, "peers": peers, "d": d, "world": world, "slots": n_slots, "ksplit": comm.ksplit,
                        "krum_fused_us": round(t_fk, 1), "krum_nccl_us": round(t_nk, 1), "krum_roofline_us": round(roof_krum, 2),
                        "krum_frac_of_roofline": round(roof_krum / t_fk, 4),
                        "krum_nvlink_bytes_per_gpu": int(pull_bytes + push_bytes), "krum_flops_per_gpu": int(flops),
                        "secagg_fused_us": round(t_fs, 1), "secagg_nccl_us": round(t_ns, 1),
                        "secagg_roofline_us": round(roof_sec, 2), "secagg_frac_of_roofline": round(roof_sec / t_fs, 4),
                        "secagg_nvlink_bytes_per_gpu": int(sec_bytes)})
        comm.close()
        torch.cuda.empty_cache()
    if rank == 0:
        print("BENCH_FUSED " + json.dumps(results))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
