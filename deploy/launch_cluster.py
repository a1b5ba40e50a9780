#!/usr/bin/env python3
"""Multi-host launcher: one torchrun agent per host, one rank per GPU, peers sharded over all ranks.

Counterpart of the reference's deployment scripts — azure/azure-run/runBiscotti.sh (keygen, build, write the
peers file, copy keys + binary, ssh `deployAzureNodes.sh` per VM, wait) and deployAzureNodes.sh:13-61 (per-VM
launch loop, logs copied back), plus azure-util/get-all-LogFiles.sh.  VM *provisioning* (azure-setup/*.sh: `az`
resource groups, vnets, images) has no counterpart: the hosts file names machines that already exist, exactly
like azure/azure-conf/hosts_sameDC.

    deploy/launch_cluster.py --hosts deploy/hosts.example --gpus-per-host 8 -- 100 mnist -na 3 -nv 3 --comm nccl
    deploy/launch_cluster.py --hosts h.txt --dry-run -- 200 cifar          # print the commands only

Across hosts the collectives go through NCCL (`--comm nccl`); the fused peer-memory kernels (`--comm fused`) map
their symmetric region with CUDA IPC and are therefore for the GPUs of one NVSwitch box.

Every host must see the repository at the same path (shared filesystem or a prior `--sync`).  The first host is
the rendezvous master (the reference's "node 0 must be up first", main.go:962-974).  Logs come back to
--logdir/<host>.log; rank 0's stdout (the chain dump + SUCCESS line) is on the first host's log.
"""
from __future__ import annotations

import argparse
import os
import shlex
import subprocess
import sys
from pathlib import Path

REPO = Path(__file__).resolve().parent.parent


def read_hosts(path: str) -> list[str]:
    with open(path) as f:
        return [l.split("#")[0].strip() for l in f if l.split("#")[0].strip()]


def agent_command(a, hosts: list[str], node_rank: int, app_args: list[str]) -> str:
    run = [a.python, "-m", "torch.distributed.run", f"--nnodes={len(hosts)}", f"--nproc-per-node={a.gpus_per_host}",
           f"--node-rank={node_rank}", f"--master-addr={hosts[0]}", f"--master-port={a.port}",
           "-m", "biscotti_b200", a.mode, *app_args]
    env = f"cd {shlex.quote(a.remote_repo)} && BISCOTTI_DUMP_DIR={shlex.quote(a.remote_logdir)} "
    return f"mkdir -p {shlex.quote(a.remote_logdir)} && " + env + " ".join(shlex.quote(x) for x in run)


def main() -> int:
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--hosts", required=True, help="one host (or user@host) per line; first = master")
    ap.add_argument("--gpus-per-host", type=int, default=8)
    ap.add_argument("--mode", default="localtest", choices=["localtest", "fedsys"])
    ap.add_argument("--port", type=int, default=29500)
    ap.add_argument("--python", default="python")
    ap.add_argument("--remote-repo", default=str(REPO))
    ap.add_argument("--remote-logdir", default="/tmp/biscotti_logs")
    ap.add_argument("--logdir", default="cluster_logs")
    ap.add_argument("--sync", action="store_true", help="rsync the repository (with the built .so) to every host first")
    ap.add_argument("--ssh", default="ssh -o BatchMode=yes -o StrictHostKeyChecking=no")
    ap.add_argument("--timeout", type=int, default=30000, help="seconds, like `timeout 30000 ./DistSys` in deployAzureNodes.sh")
    ap.add_argument("--dry-run", action="store_true")
    ap.add_argument("app", nargs=argparse.REMAINDER, help="-- <numNodes> <dataset> [protocol flags]")
    a = ap.parse_args()
    app_args = [x for x in a.app if x != "--"]
    if len(app_args) < 2:
        ap.error("expected: -- <numNodes> <dataset> [flags]")
    hosts = read_hosts(a.hosts)
    if not hosts:
        ap.error("empty hosts file")
    if not a.dry_run:
        os.makedirs(a.logdir, exist_ok=True)
    ssh = shlex.split(a.ssh)

    if a.sync:
        for h in hosts:
            cmd = ["rsync", "-a", "--delete", "--exclude", ".git", "--exclude", "gpurun_out", f"{REPO}/", f"{h}:{a.remote_repo}/"]
            print("+", " ".join(cmd))
            if not a.dry_run:
                subprocess.run(cmd, check=True)

    procs = []
    for i, h in enumerate(hosts):
        remote = f"timeout {a.timeout} bash -lc {shlex.quote(agent_command(a, hosts, i, app_args))}"
        cmd = [*ssh, h, remote]
        print("+", " ".join(shlex.quote(c) for c in cmd))
        if a.dry_run:
            continue
        log = open(os.path.join(a.logdir, f"{h.replace('@', '_')}.log"), "w")
        procs.append((h, subprocess.Popen(cmd, stdout=log, stderr=subprocess.STDOUT), log))
    if a.dry_run:
        return 0
    rc = 0
    for h, p, log in procs:
        code = p.wait()
        log.close()
        print(f"{h}: exit {code}")
        rc = rc or code
    # per-rank chain dumps (the reference copies every node's log back with scp, deployAzureNodes.sh:60-61)
    for h in hosts:
        subprocess.run(["scp", "-q", f"{h}:{a.remote_logdir}/chain_rank*.txt", a.logdir], check=False)
    dumps = sorted(Path(a.logdir).glob("chain_rank*.txt"))
    if dumps:
        first = dumps[0].read_bytes()
        same = all(d.read_bytes() == first for d in dumps)
        print("SUCCESS! Nodes have same blockchain" if same else "FAILURE! Nodes have different blockchains")
        rc = rc or (0 if same else 1)
    return rc


if __name__ == "__main__":
    sys.exit(main())
