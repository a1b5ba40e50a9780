"""Multi-GPU check, run under torchrun on N GPUs of one box (not collected by pytest):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tests/multigpu_check.py

Runs the MNIST engine with the fused peer-memory path and with the NCCL baseline path and verifies the
reference's correctness oracle: every rank ends with a byte-identical chain (localTest.sh:70-85)."""
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from biscotti_b200.config import BiscottiConfig  # noqa: E402
from biscotti_b200.parallel.comm import NcclComm, init_distributed  # noqa: E402
from biscotti_b200.parallel.fused import FusedComm  # noqa: E402
from biscotti_b200.protocol.engine import BiscottiEngine  # noqa: E402


def main():
    rank, local, world = init_distributed("cuda")
    dev = torch.device("cuda", local)
    peers = int(os.environ.get("BSC_PEERS", "100"))
    iters = int(os.environ.get("BSC_ITERS", "6"))
    out = {"world": world}
    if rank == 0:
        # the ledger must not depend on how the peers are sharded: the same run on one GPU, no communicator at all
        cfg = BiscottiConfig(num_nodes=peers, dataset="mnist", max_iterations=iters, shard_size=200, test_size=500)
        solo = BiscottiEngine(cfg, device=dev)
        solo.run(iters)
        torch.cuda.synchronize()
        out["single_gpu_hash"] = solo.chain.latest_hash().hex()[:16]
        del solo
        torch.cuda.empty_cache()
    torch.distributed.barrier()
    for name in ("fused", "nccl"):
        cfg = BiscottiConfig(num_nodes=peers, dataset="mnist", max_iterations=iters, shard_size=200, test_size=500)
        comm = FusedComm(rank, world, dev, cfg) if name == "fused" else NcclComm(rank, world, dev)
        eng = BiscottiEngine(cfg, device=dev, rank=rank, world=world, comm=comm)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        reps = eng.run(iters)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        hashes = comm.all_gather_bytes(eng.chain.latest_hash())
        out[name] = {"chain_equal": all(h == hashes[0] for h in hashes), "ms_per_block": 1e3 * dt / iters,
                     "hash": hashes[0].hex()[:16], "bad": sum(r.bad_shares for r in reps),
                     "empty": sum(r.empty for r in reps), "test_error": reps[-1].test_error,
                     "verify_ok": eng.chain.verify()}
        eng.close()
    if rank == 0:
        print("MULTIGPU_CHECK " + json.dumps(out))
        ok = all(out[k]["chain_equal"] and out[k]["verify_ok"] and out[k]["bad"] == 0 for k in ("fused", "nccl"))
        ok = ok and out["fused"]["hash"] == out["nccl"]["hash"] == out["single_gpu_hash"]
        print("MULTIGPU_CHECK_OK" if ok else "MULTIGPU_CHECK_FAILED")
    torch.distributed.barrier()
    torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
