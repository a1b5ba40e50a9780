"""Fused compute+collective kernels exercised on one GPU (world = 1: the peer tables all point at the
local buffer, flags are self-signalled).  The multi-rank behaviour is covered by tests/multigpu_check.py
under torchrun on 2+ GPUs."""
import pytest
import torch

from biscotti_b200.config import BiscottiConfig
from biscotti_b200.ops import ml, krum as krum_ops, secagg
from biscotti_b200.parallel.fused import FusedComm
from biscotti_b200.protocol.engine import BiscottiEngine

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda:0")


def test_fused_krum_and_secagg_world1_match_unfused_kernels():
    cfg = BiscottiConfig(num_nodes=100, dataset="mnist")
    comm = FusedComm(0, 1, DEV, cfg)
    d, n = cfg.dim, 100
    torch.manual_seed(0)
    U = torch.zeros(n, ml.pad_to(d, 4), device=DEV)
    U[:, :d] = torch.randn(n, d, device=DEV) * 0.05
    U[:20, :d] += 0.2
    noisers = torch.full((n, 1), -1, dtype=torch.int32, device=DEV)
    Un, hi, lo, norms = ml.noise_mix(U, noisers, 1, 0, 0.0, d, out_hi=comm.hi, out_lo=comm.lo)
    krum_set = sorted(torch.randperm(n)[:70].tolist())
    clip = 35
    scores, acc = comm.krum_fused(krum_set, clip, epoch=1)
    _, aref, _ = krum_ops.krum_reference(U[krum_set, :d].cpu(), clip)
    assert acc.cpu().bool().tolist() == aref.tolist()
    # a second block reuses every buffer and flag (epoch 2)
    ml.noise_mix(U * 0.5, noisers, 1, 0, 0.0, d, out_hi=comm.hi, out_lo=comm.lo)
    _, acc2 = comm.krum_fused(krum_set, clip, epoch=2)
    assert acc2.cpu().bool().tolist() == aref.tolist()
    # secure aggregation
    sel = torch.tensor([1, 5, 9, 33, 70], dtype=torch.int32, device=DEV)
    secagg.share_sum(U, sel, d, cfg.total_shares, out=comm.ypart)
    comm.w64.zero_()
    comm.w32.zero_()
    bad = comm.secagg_fused(list(range(10)), [1] * 21, 4, epoch=1)
    torch.cuda.synchronize()
    assert int(bad.sum()) == 0
    w_before = comm.w64.clone()
    gate = torch.tensor([1], dtype=torch.int32, device=DEV)
    comm.secagg_fused(list(range(10)), [1] * 21, 4, epoch=2, gate=gate, gate_min=2)      # too few contributors: no step
    assert torch.equal(comm.w64, w_before)
    comm.w64.zero_()
    comm.w32.zero_()
    comm.secagg_fused([0, 1, 2, 3, 4, 5, 6, 14, 15, 16], [1] * 7 + [0] * 7 + [1] * 7, 4, epoch=3)   # a miner is down
    w_gap = comm.w64.clone()
    comm.w64.zero_()
    comm.w32.zero_()
    comm.secagg_fused(list(range(10)), [1] * 21, 4, epoch=4)
    w_ref = torch.zeros(d, dtype=torch.float64, device=DEV)
    Ysum = secagg.share_sum(U, sel, d, cfg.total_shares)
    secagg.reconstruct_step(Ysum, torch.arange(10, dtype=torch.int32, device=DEV), None, d, w_ref, None)
    assert torch.equal(comm.w64, w_ref) and torch.equal(w_gap, w_ref)
    assert torch.equal(comm.w32, w_ref.float())
    comm.close()


def test_fused_krum_with_more_than_128_peers_per_rank():
    """200 peers on one rank: 256 slots, two 128-row tiles out of one owner's plane (row-offset TMA boxes)."""
    cfg = BiscottiConfig(num_nodes=200, dataset="mnist")
    comm = FusedComm(0, 1, DEV, cfg)
    assert comm.Sr == 256 and comm.ntile == 2
    d, n = cfg.dim, 200
    torch.manual_seed(1)
    U = torch.zeros(n, ml.pad_to(d, 4), device=DEV)
    U[:, :d] = torch.randn(n, d, device=DEV) * 0.05
    U[150:, :d] += 0.2 * torch.arange(1, 51, device=DEV, dtype=torch.float32).unsqueeze(1)      # fifty outliers
    noisers = torch.full((n, 1), -1, dtype=torch.int32, device=DEV)
    ml.noise_mix(U, noisers, 1, 0, 0.0, d, out_hi=comm.hi, out_lo=comm.lo)
    krum_set = sorted(torch.randperm(n)[:140].tolist())
    clip = 70
    _, acc = comm.krum_fused(krum_set, clip)
    _, aref, _ = krum_ops.krum_reference(U[krum_set, :d].cpu(), clip)
    assert acc.cpu().bool().tolist() == aref.tolist()
    comm.close()


def test_engine_with_fused_comm_world1_is_deterministic():
    cfg = BiscottiConfig(num_nodes=30, dataset="mnist", max_iterations=3, shard_size=100, test_size=300, msm_window=7,
                         data_sep=6.0)
    a = BiscottiEngine(cfg, device=DEV, comm=FusedComm(0, 1, DEV, cfg))
    b = BiscottiEngine(cfg, device=DEV, comm=FusedComm(0, 1, DEV, cfg))
    ra, rb = a.run(), b.run()
    assert a.chain == b.chain and a.chain.verify()
    assert all(r.bad_shares == 0 for r in ra) and any(not r.empty for r in ra)
    assert ra[-1].test_error <= ra[0].test_error
