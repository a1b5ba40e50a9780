"""Cross-rank plumbing over torch.distributed (NCCL on GPUs, gloo on CPU).

This is the *baseline* communication path: library collectives around un-fused
kernels — all-gather of the noised updates for Krum (the reference's workers -> verifiers fan-out,
DistSys/main.go:1666-1775), all-reduce of the aggregated integer shares (miner fold + leader gather,
main.go:2157-2235) and a tiny object exchange for commitments (RegisterBlock payload).  The product
path that issues the same transfers from inside the kernels over NVLink is `parallel.fused`.
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist


def init_distributed(device_type: str = "cuda"):
    """One process per GPU; reads RANK / LOCAL_RANK / WORLD_SIZE / MASTER_* from the environment."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        backend = "nccl" if device_type == "cuda" else "gloo"
        if device_type == "cuda":
            torch.cuda.set_device(local)
            dist.init_process_group(backend, rank=rank, world_size=world, device_id=torch.device("cuda", local))
        else:
            dist.init_process_group(backend, rank=rank, world_size=world)
    return rank, local, world


class NcclComm:
    def __init__(self, rank: int, world: int, device):
        self.rank, self.world, self.device = rank, world, torch.device(device)

    def barrier(self):
        if self.world > 1:
            dist.barrier()

    def all_reduce_i64(self, t: torch.Tensor) -> torch.Tensor:
        """Sum of int64 share matrices.  Integer addition is associative, so the result is bit-identical on
        every rank regardless of NCCL's reduction order — this is what keeps the ledgers equal."""
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return t

    def all_reduce_max(self, t: torch.Tensor) -> torch.Tensor:
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t

    def gather_rows(self, ids, local: torch.Tensor, local_index: dict, owner: list) -> torch.Tensor:
        """Rows for global peer ids from wherever they live: all-gather of the (padded) local blocks."""
        if self.world == 1:
            ix = torch.tensor([local_index[p] for p in ids], dtype=torch.long, device=local.device)
            return local.index_select(0, ix).contiguous()
        per = (len(owner) + self.world - 1) // self.world
        pad = torch.zeros(per, *local.shape[1:], dtype=local.dtype, device=local.device)
        pad[: local.shape[0]] = local[:per]
        out = torch.empty(self.world * per, *local.shape[1:], dtype=local.dtype, device=local.device)
        dist.all_gather_into_tensor(out, pad)
        ix = torch.tensor(ids, dtype=torch.long, device=local.device)   # contiguous ownership: global id == row
        return out.index_select(0, ix).contiguous()

    def all_gather_rows_u8(self, local: torch.Tensor) -> torch.Tensor:
        """[per, k] blocks from every rank -> [world*per, k] (the block's commitment || approvals rows, row = peer id)."""
        if self.world == 1:
            return local
        out = torch.empty(self.world * local.shape[0], *local.shape[1:], dtype=local.dtype, device=local.device)
        dist.all_gather_into_tensor(out, local.contiguous())
        return out

    def all_gather_np(self, a):
        """numpy block per rank -> list of every rank's block (host-side plumbing of the CPU path)."""
        import numpy as np

        if self.world == 1:
            return [a]
        t = torch.from_numpy(np.ascontiguousarray(a)).to(self.device)
        out = self.all_gather_rows_u8(t.view(torch.uint8).reshape(1, -1)).cpu().numpy()
        return [out[q].view(a.dtype).reshape(a.shape) for q in range(self.world)]

    def merge_dicts(self, d: dict) -> dict:
        if self.world == 1:
            return d
        objs = [None] * self.world
        dist.all_gather_object(objs, d)
        out = {}
        for o in objs:
            out.update(o)
        return out

    def all_gather_bytes(self, b: bytes) -> list:
        if self.world == 1:
            return [b]
        objs = [None] * self.world
        dist.all_gather_object(objs, b)
        return objs
