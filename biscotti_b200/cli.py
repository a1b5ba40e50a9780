"""Command line: the reference's entrypoints re-expressed for one-process-per-GPU execution.

    python -m biscotti_b200 localtest <numNodes> <dataset> [flags]     # DistSys/localTest.sh parity
    python -m biscotti_b200 fedsys    <numNodes> <dataset> [flags]     # FedSys/localTest.sh parity
    python -m biscotti_b200 keygen -n <nodesPerHost> -d <dims> [-h hosts] # keyGeneration parity
    python -m biscotti_b200 vrf-demo                                    # vrf_main.go parity

Flags keep the reference's names (DistSys/main.go:613-647): -na -nv -nn -sa -np -vp -ep -po -ns -rs -c.
`localtest` ends like localTest.sh:61-89: each rank dumps its chain (PrintChain) and the dumps are
compared pairwise; "SUCCESS! Nodes have same blockchain" is printed when they are identical.
Under torchrun the N simulated peers are sharded over the ranks.
"""
from __future__ import annotations

import argparse
import hashlib
import os
import sys


def _bool(s: str) -> bool:
    return str(s).lower() in ("1", "true", "t", "yes")


def add_protocol_flags(ap: argparse.ArgumentParser):
    ap.add_argument("-na", type=int, default=3, help="number of aggregators/miners")
    ap.add_argument("-nv", type=int, default=3, help="number of verifiers")
    ap.add_argument("-nn", type=int, default=2, help="number of noisers")
    ap.add_argument("-sa", type=_bool, default=True, help="secure aggregation")
    ap.add_argument("-np", type=_bool, default=True, help="noising")
    ap.add_argument("-vp", type=_bool, default=True, help="verification")
    ap.add_argument("-ep", type=float, default=2.0, help="epsilon")
    ap.add_argument("-po", type=float, default=0.0, help="fraction of poisoners")
    ap.add_argument("-ns", type=int, default=None, help="percent of updates to collect (default 70; fedsys 35)")
    ap.add_argument("-rs", type=_bool, default=False, help="random sampling of updates before Krum")
    ap.add_argument("-c", type=int, default=0, help="number of colluders")
    ap.add_argument("--defense", default="KRUM", choices=["KRUM", "RONI"])
    ap.add_argument("--iterations", type=int, default=100, help="MAX_ITERATIONS")
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--device", default="auto")
    ap.add_argument("--shard-size", type=int, default=None)
    ap.add_argument("--fail-prob", type=float, default=-0.005)
    ap.add_argument("--tau", type=int, default=2, help="SRS trapdoor (2 = reference compat, 0 = random)")
    ap.add_argument("--comm", default="fused", choices=["nccl", "fused"], help="multi-GPU path: fused peer-memory kernels (default) or the NCCL baseline")
    ap.add_argument("--gram", default="auto")
    ap.add_argument("--msm-table-gb", type=float, default=None, help="HBM budget of the MSM window tables in GiB (default: half the device)")
    ap.add_argument("--print-hash", action="store_true")
    ap.add_argument("--no-dump", action="store_true")
    ap.add_argument("--log", action="store_true", help="emit [peer] log lines on stderr")
    ap.add_argument("--trace", default=None, metavar="FILE",
                    help="instrument every phase (CUDA events) and write a Chrome/Perfetto trace; one file per rank")
    ap.add_argument("--save-chain", default=None)
    ap.add_argument("--checkpoint-every", type=int, default=0, metavar="K",
                    help="with --save-chain: also write the chain every K blocks (atomically), so that a killed run can be resumed")
    ap.add_argument("--partition", action="append", default=[], metavar="PEER:FIRST:LAST",
                    help="cut PEER off for blocks FIRST..LAST (blockNode.sh); may be repeated")
    ap.add_argument("--data-dir", default=None, help="directory of <dataset><i>.npy shard files (default: synthetic data)")
    ap.add_argument("--resume", default=None, help="chain file to adopt before running (restart / rejoin, main.go:1000-1013)")


def config_from_args(a, mode: str):
    from .config import BiscottiConfig

    ns = a.ns if a.ns is not None else (35 if mode == "fedsys" else 70)
    kw = dict(num_nodes=a.num_nodes, dataset=a.dataset, num_miners=a.na, num_verifiers=a.nv, num_noisers=a.nn,
              secure_agg=a.sa, noising=a.np, verification=a.vp, epsilon=a.ep, poison_fraction=a.po, perc_samples=ns,
              random_sampling=a.rs, colluders=a.c, defense=a.defense, max_iterations=a.iterations, seed=a.seed,
              fail_prob=a.fail_prob, srs_tau=a.tau, mode=mode, gram_impl=a.gram, comm_impl=a.comm, data_dir=a.data_dir, msm_table_gb=a.msm_table_gb,
              partitions=tuple(tuple(int(v) for v in x.split(":")) for x in a.partition))
    if a.shard_size:
        kw["shard_size"] = a.shard_size
    return BiscottiConfig(**kw)


def run_localtest(a, mode: str) -> int:
    import torch

    from .parallel.comm import init_distributed, NcclComm
    from .protocol.engine import BiscottiEngine

    dev_type = a.device
    if dev_type == "auto":
        dev_type = "cuda" if torch.cuda.is_available() else "cpu"
    rank, local, world = init_distributed("cuda" if dev_type.startswith("cuda") else "cpu")
    device = torch.device("cuda", local) if dev_type.startswith("cuda") else torch.device("cpu")
    cfg = config_from_args(a, mode)
    comm = None
    if world > 1:
        if cfg.comm_impl == "fused" and device.type == "cuda":
            from .parallel.fused import FusedComm

            comm = FusedComm(rank, world, device, cfg)
        else:
            comm = NcclComm(rank, world, device)
    eng = BiscottiEngine(cfg, device=device, rank=rank, world=world, comm=comm, log_stream=sys.stderr if a.log else None)
    if a.resume:
        # a restarted deployment adopts the saved chain (the ledger is the checkpoint) and continues from its height
        with open(a.resume, "rb") as f:
            adopted = eng.adopt_chain(f.read())
        if rank == 0:
            print(f"resume: adopted={adopted} height={len(eng.chain)} next_iteration={eng.iteration}", file=sys.stderr)
    if a.trace:
        eng._timing = True                     # every phase bracketed by CUDA events (in-line schedule)
        cfg.prefetch = False
    todo = max(0, cfg.max_iterations + 1 - eng.iteration)
    if a.save_chain and a.checkpoint_every > 0:
        for i in range(todo):
            eng.step()
            if rank == 0 and (i + 1) % a.checkpoint_every == 0:
                eng.chain.save(a.save_chain)
        eng.flush_reports()
    else:
        eng.run(todo)
    if a.trace:
        from .utils.trace import write_chrome_trace

        write_chrome_trace(eng.reports, a.trace if world == 1 else f"{a.trace}.rank{rank}", rank)
    dump = eng.print_chain()
    dump_dir = os.environ.get("BISCOTTI_DUMP_DIR")
    if dump_dir:   # one PrintChain file per rank, for the reference-style pairwise `cmp` (deploy/local_test.sh)
        with open(os.path.join(dump_dir, f"chain_rank{rank}.txt"), "w") as f:
            f.write(dump)
    ok = True
    if world > 1:
        digests = comm.all_gather_bytes(hashlib.sha256(dump.encode()).digest())
        ok = all(d == digests[0] for d in digests)
    if rank == 0:
        if not a.no_dump:
            sys.stdout.write(dump)
        if a.print_hash:
            print("chain-hash", hashlib.sha256(dump.encode()).hexdigest())
        if a.save_chain:
            eng.chain.save(a.save_chain)
        last = eng.reports[-1]
        print(f"blocks={len(eng.chain)} final_test_error={last.test_error} attack_rate={last.attack_rate}")
        # localTest.sh:70-89
        print("SUCCESS! Nodes have same blockchain" if ok else "FAILURE! Nodes have different blockchains")
    eng.close()              # unmaps the peers' memory of the fused path (after a barrier)
    if world > 1:
        import torch.distributed as dist

        dist.barrier()
        dist.destroy_process_group()
    return 0 if ok else 1


def run_keygen(a) -> int:
    from .crypto.primitives import write_bootstrap_files

    hosts = None
    if a.hosts:
        with open(a.hosts) as f:
            hosts = [l.strip() for l in f if l.strip()]
    total = a.n * (len(hosts) if hosts else 1)
    write_bootstrap_files(a.out, total, a.d, hosts, seed=a.seed, tau=a.tau, nodes_per_host=a.n)
    print(f"wrote peersfile.txt commitKey.json pKeyG1.json for {total} peers, d={a.d} into {a.out}")
    return 0


def run_vrf_demo(a) -> int:
    """vrf_main.go:75-153: prove / verify and a stake lottery over 100 nodes."""
    from .crypto import primitives as cp

    key = cp.VRFKey(hashlib.sha256(b"vrf-demo").digest())
    msg = hashlib.sha256(b"block-hash").digest()
    vrf, proof = key.prove(msg)
    print("VRF output:", vrf.hex())
    print("Proof verifies:", cp.vrf_verify(key.public, msg, vrf, proof))
    v, m = cp.lottery_roles(msg, [10] * 100, 3, 3)
    nz = cp.lottery_noisers(vrf, [10] * 100, 0, 2)
    print("verifiers", v, "miners", m, "noisers", nz)
    return 0


def run_kzg_demo(a) -> int:
    """kyber-demo/kyber.go + DistSys/kyber.go:289-454 (commented main): share two copies of one vector to three
    miners, aggregate homomorphically, recover the sum, and check a KZG opening."""
    import torch

    from .ops import secagg
    from .ops.msm import CommitmentEngine

    d, S = 25, 21
    eng = CommitmentEngine(d, S, tau=2, device="cpu")
    delta = torch.tensor([[-0.0105, -0.0166, 0.0049] + [0.001 * i for i in range(22)] + [0.0] * 3])
    Q = secagg.quantize(torch.cat([delta, delta]), d, 4)
    print("quantised head:", Q[0, :6].tolist())
    C = eng.chunk_commitments(Q)
    W = eng.witnesses(Q)
    Y = secagg.shares_eval(Q, S)
    per = S // 3
    for m in range(3):
        print(f"miner {m} holds shares {m * per}..{(m + 1) * per - 1} of {Q.shape[1] // 10} chunks")
    Ysum = Y.sum(0).contiguous()
    w = torch.zeros(d, dtype=torch.float64)
    _, bad = secagg.reconstruct_step(Ysum, torch.arange(10, dtype=torch.int32), None, d, w, None)
    print("recovered sum head:", [round(float(v), 4) for v in w[:4]], "inconsistent shares:", int(bad.sum()))
    ok = eng.verify_share_trapdoor(eng.to_bytes(C)[:64], eng.to_bytes(W)[:64], 0, 0, int(Y[0, 0, 0]))
    print("KZG opening of chunk 0 at x=-10 verifies:", ok)
    okp = eng.verify_share_pairing(eng.to_bytes(C)[:64], eng.to_bytes(W)[:64], 0, 0, int(Y[0, 0, 0]))
    print("same opening through the pairing (verifySecret, kyber.go:650-673):", okp)
    return 0 if ok and okp and int(bad.sum()) == 0 else 1


def run_experiment(a) -> int:
    import json

    import torch

    from .eval import experiments as ex

    dev = a.device
    if dev == "auto":
        dev = "cuda:0" if torch.cuda.is_available() else "cpu"
    fn = getattr(ex, a.name)
    kw = {}
    if a.name not in ("committee_capture", "unmasked_probability"):
        kw["device"] = dev
        if a.blocks:
            kw["blocks"] = a.blocks
    rows = fn(**kw)
    if a.csv:
        ex.write_csv(rows if isinstance(rows, list) else [rows], a.csv)
    print(json.dumps(rows, indent=1))
    return 0


def run_make_shards(a) -> int:
    """parse_mnist.py / parse_cifar.py / lfw parsers: [n, F+1] row files -> per-peer shards, test, attack and poisoned
    sets in the layout data/files.py reads.  Without --train the rows are drawn from the synthetic generator, which
    gives a self-contained way to exercise the file-backed path."""
    import numpy as np

    from .config import BiscottiConfig, DATASETS
    from .data.files import write_shards

    spec = DATASETS[a.dataset]
    if a.train:
        tr, te = np.load(a.train), np.load(a.test or a.train)
        X, y, Xt, yt = tr[:, :-1], tr[:, -1], te[:, :-1], te[:, -1]
    else:
        from .data.synthetic import make_data

        cfg = BiscottiConfig(num_nodes=a.n, dataset=a.dataset, seed=a.seed)
        d = make_data(cfg, list(range(a.n)), "cpu")
        X = d.X.reshape(-1, spec.n_features).numpy()
        y = d.y.reshape(-1).numpy().astype(np.float64)
        Xt, yt = d.X_test.numpy(), d.y_test.numpy().astype(np.float64)
    files = write_shards(a.out, a.dataset, X, y, Xt, yt, a.n, source=a.source, target=a.target,
                         standardize=not a.no_standardize, seed=a.seed)
    print(f"wrote {len(files)} files to {a.out} ({a.n} shards of {a.dataset})")
    return 0


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="biscotti_b200", add_help=True)
    sub = ap.add_subparsers(dest="cmd", required=True)
    for name in ("localtest", "fedsys"):
        sp = sub.add_parser(name)
        sp.add_argument("num_nodes", type=int)
        sp.add_argument("dataset")
        add_protocol_flags(sp)
    kg = sub.add_parser("keygen", add_help=False)
    kg.add_argument("-n", type=int, default=10)
    kg.add_argument("-d", type=int, default=25)
    kg.add_argument("-h", dest="hosts", default=None)
    kg.add_argument("--out", default="keys")
    kg.add_argument("--seed", type=int, default=1)
    kg.add_argument("--tau", type=int, default=2)
    ms = sub.add_parser("make-shards")
    ms.add_argument("dataset")
    ms.add_argument("-n", type=int, default=100, help="number of peer shards")
    ms.add_argument("--train", default=None, help=".npy with rows features||label (default: synthetic rows)")
    ms.add_argument("--test", default=None)
    ms.add_argument("--out", default="data")
    ms.add_argument("--source", type=int, default=1)
    ms.add_argument("--target", type=int, default=7)
    ms.add_argument("--no-standardize", action="store_true")
    ms.add_argument("--seed", type=int, default=1)
    sub.add_parser("vrf-demo")
    sub.add_parser("kzg-demo")
    exp = sub.add_parser("experiment")
    exp.add_argument("name", choices=["time_to_accuracy", "cost_breakdown", "fedsys_scale", "committee_scale", "noise_krum", "poison_heatmap",
                                      "privacy_utility", "churn", "committee_capture", "unmasked_probability"])
    exp.add_argument("--device", default="auto")
    exp.add_argument("--blocks", type=int, default=None)
    exp.add_argument("--csv", default=None)
    a = ap.parse_args(argv)
    if a.cmd == "kzg-demo":
        return run_kzg_demo(a)
    if a.cmd == "experiment":
        return run_experiment(a)
    if a.cmd == "localtest":
        return run_localtest(a, "biscotti")
    if a.cmd == "fedsys":
        return run_localtest(a, "fedsys")
    if a.cmd == "keygen":
        return run_keygen(a)
    if a.cmd == "make-shards":
        return run_make_shards(a)
    return run_vrf_demo(a)


if __name__ == "__main__":
    sys.exit(main())
