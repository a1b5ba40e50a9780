"""The Biscotti protocol engine: one object per rank simulating its share of the peers.

What one `step()` does is the reference's block iteration (call stacks §3.3-3.6):
  1. role lottery on the latest block hash            DistSys/main.go:486-565, vrf.go:103-182
  2. local SGD step of every worker                   honest.go:165-200, client_obj.py:73-77
  3. commitment to the quantised update               kyber.go:533-562 (one MSM per chunk, summed)
  4. noiser lottery + DP noise                        vrf.go:54-100, main.go:1592-1660, 1530-1537
  5. Multi-Krum over the first KRUM_UPDATETHRESH      krum.go:227-365, client_obj.py:114-143
  6. Schnorr approvals, >= nv/2 needed                kyber.go:873-896, main.go:1686
  7. Shamir shares + KZG witnesses to the miners      kyber.go:456-482, main.go:1847-1952
  8. miners aggregate, leader reconstructs the sum    kyber.go:244-287, 809-857; honest.go:391-502
  9. w_{t+1} = w_t + sum, stake update, block append  honest.go:391-440, block.go:23-44
 10. flood + test error / attack rate                 main.go:1403-1444, honest.go:141-162

There are no per-peer processes, sockets or wall-clock timeouts: peers are rows of device tensors, message
arrival order is an explicit seeded permutation (so "first NUM_SAMPLES/2 shares win" and Krum's "first
KRUM_UPDATETHRESH updates" are reproducible), and every rank replays the same control flow from the same
block hash, so the ledger is replicated without gossip.

How a block is scheduled on a GPU.  A block is a serial chain (hash -> VRF -> noisers -> noise -> Krum ->
approved / contributors -> share sums -> reconstruction -> new weights -> hash).  Every link whose input is
produced on the device is a kernel that reads it there (ops/chain.py, the *_dc kernel variants): the Krum verdict,
the approved set, the contributor set and the noiser table never visit the host in front of the block's tail.
The host plans what depends only on the previous hash and the stake (roles, arrival orders, candidate set),
queues the chain, queues everything that is off the chain behind it on a second stream (witness MSM, miners'
aggregation, error rates, the next block's local step + commitments) and on a third (the verifiers' approvals of
every candidate), and then waits ONCE, for the tail's device->host copy (weights, counters, contributor list, the
gathered commitment||signature rows).  Block assembly, stake update, hashing and append are one native call.
"""
from __future__ import annotations

import contextlib
import ctypes
import hashlib
import math
import time
from dataclasses import dataclass, field

import numpy as np
import torch

from ..config import BiscottiConfig, DEFAULT_STAKE, STAKE_UNIT, POLY_SIZE, NOISE_TABLE_ITERS
from ..crypto import primitives as cp
from ..crypto.schnorr_batch import ApprovalSigner
from ..data.synthetic import make_data, poisoner_ids
from ..ledger.chain import Blockchain, UpdateRecord
from ..models.registry import Model
from ..ops import ml, krum as krum_ops, secagg, chain as chain_ops
from ..ops._call import lib, ptr, stream as raw_stream, check
from ..ops.msm import CommitmentEngine
from ..utils.log import PeerLogger


@dataclass
class BlockReport:
    iteration: int
    block_hash: bytes
    empty: bool
    verifiers: list
    miners: list
    n_workers: int
    n_krum: int
    n_accepted: int
    n_contributors: int
    loss: float | None = None
    bad_shares: int = 0
    phase_ms: dict = field(default_factory=dict)
    unmasked_updates: int = 0
    contributors: list = field(default_factory=list)
    _test_error: float | None = None
    _attack_rate: float | None = None
    _resolve: object = None        # callable that waits for this block's error-rate kernels and fills the two fields

    # The error rates of the new weights are computed off the block's chain and read back asynchronously; touching
    # them waits for exactly that read-back (the engine itself resolves a report two blocks later at the latest).
    @property
    def test_error(self):
        if self._resolve is not None:
            self._resolve()
        return self._test_error

    @property
    def attack_rate(self):
        if self._resolve is not None:
            self._resolve()
        return self._attack_rate


@dataclass
class _Plan:
    """Everything about a block that follows from the previous hash, the stake and who is alive (host side)."""
    verifiers: list
    miners: list
    workers: list
    prev_hash: bytes
    krum_set: list = field(default_factory=list)
    share_rank: list = field(default_factory=list)
    signers: list = field(default_factory=list)
    approved_ok: bool = False
    trigger: int = 0
    shares_avail: list = field(default_factory=list)
    xs: list = field(default_factory=list)
    enough_shares: bool = False
    clip: int = 0
    n_workers: int = 0
    stake_m: object = None       # int32 [n]: stake with dead peers zeroed (the lottery's ticket counts)
    is_worker: object = None     # uint8 [n]


@dataclass
class _Ctx:
    """One block in flight: its plan, its device-side handles and how far it has been queued."""
    it: int
    par: int
    pl: _Plan
    nk: int
    phases: dict
    alive: list
    stake: list
    timing: bool
    pipelined: bool
    rep: BlockReport | None = None
    empty: bool = False
    native: bool = False          # queued through protocol/schedule.py's command lists
    pre: dict | None = None
    front_queued: bool = False
    pout: dict | None = None
    use_noise: bool = False
    vrf_handle: object = None
    compact: bool = False
    plan_ready: object = None
    sig_event: object = None
    salt64: bytes = b""
    ns: int = 0
    t0: float = 0.0
    stake_m: object = None
    is_worker: object = None
    stake_d: object = None
    isw_d: object = None
    dst_row: object = None
    local_row_d: object = None
    share_rank_d: object = None
    krum_ids_d: object = None
    signers_d: object = None
    xs_t: object = None
    av_t: object = None
    slots_d: object = None
    hash_d: object = None
    noiser_tab: object = None
    noisers_host: object = None
    accept: object = None
    n_lc: object = None
    n_la: object = None
    n_con: object = None


_NP = {torch.int32: np.int32, torch.uint8: np.uint8, torch.int64: np.int64, torch.float32: np.float32}
_SZ = {torch.int32: 4, torch.uint8: 1, torch.int64: 8, torch.float32: 4}


class BiscottiEngine:
    def __init__(self, cfg: BiscottiConfig, device="cpu", rank: int = 0, world: int = 1, comm=None, log_stream=None,
                 external_inputs: bool = False):
        self.cfg = cfg
        self.device = torch.device(device)
        self.rank, self.world, self.comm = rank, world, comm
        n = cfg.num_nodes
        per = (n + world - 1) // world
        self.per = per
        self.local_ids = list(range(rank * per, min(n, (rank + 1) * per)))
        self.owner = [p // per for p in range(n)]
        self.P = len(self.local_ids)
        self.local_index = {p: i for i, p in enumerate(self.local_ids)}
        self.model = Model(cfg)
        self.d = cfg.dim
        self.ldu = ml.pad_to(self.d, 4)
        self.ldh = ml.pad_to(self.d, 64)
        self.S = cfg.total_shares
        self.log = PeerLogger(rank, stream=log_stream)
        self.external_inputs = external_inputs
        # ---- state
        self.chain = Blockchain(self.d)
        self.stake = [DEFAULT_STAKE] * n
        self.alive = [True] * n
        self.iteration = 0
        self._pre = None          # prefetched local step + commitments of the next iteration, see _prefetch_next
        self._ctx_next = None     # the next block, planned and partly queued by the previous step (see _after_block)
        self._ns, self._ns_ok = None, None      # native command-list schedule (protocol/schedule.py), created on first use
        self._faults_applied_for = -1
        self._next_inputs, self._fed_inputs = None, False
        self._eval_waiting: list = []           # reports whose error rates are still being read back
        self._vrf_early = None
        self._main = self._side = self._wit = self._agg = None   # chain / off-chain / witness / aggregation streams (pipelined GPU)
        self._wstream = None                    # alias of _side, kept for callers that look for the witness stream
        self._wevent = None
        # Ring depth of everything a block owns on the device (plan outputs, commitments, share sums, packs, events, command
        # lists): block t reuses the slot of block t - NB, and may only do so once that block's off-chain work (witnesses,
        # miners' aggregates) has read it.  With two slots that wait sat on the chain of every other block.
        self.NB = max(2, int(getattr(cfg, "ring_depth", 3)))
        self._side_done = [None] * self.NB      # per ring slot: off-chain work of that block has finished
        self._eval_done = None
        self.last_aggregate = None
        self.keys = cp.derive_peer_keys(n, cfg.seed)
        self.poisoners = poisoner_ids(cfg)
        self.colluder_ids = set(range(n - cfg.colluders, n)) if cfg.colluders > 0 else set()
        self.data = None if external_inputs else make_data(cfg, self.local_ids, self.device)
        if external_inputs:
            self.data = make_data(cfg, [], self.device)   # test / attack sets only
        w0 = self.model.init_weights()
        if w0.abs().sum() != 0:
            # non-zero genesis (CNNs): record it as the genesis block's GlobalW
            self.chain = Blockchain(self.d)
            self.chain.add_block(-1, w0.numpy(), [], None, timestamp=0)
        self.fused = comm is not None and hasattr(comm, "krum_fused")
        if self.fused:
            # weights live in symmetric peer memory: the owner of a chunk range stores its new slice
            # straight into every rank's copy (secagg_fused_kernel)
            comm.w64.copy_(w0.to(self.device))
            comm.w32.copy_(w0.float().to(self.device))
            self.w64, self.w32 = comm.w64, comm.w32
        else:
            self.w64 = w0.to(self.device)
            self.w32 = self.w64.float()
        self.U = torch.zeros(max(1, self.P), self.ldu, dtype=torch.float32, device=self.device)
        self.commit = None
        if cfg.commitments and cfg.secure_agg and cfg.mode == "biscotti":
            self.commit = CommitmentEngine(self.d, self.S, tau=cfg.srs_tau or self._random_tau(), device=self.device,
                                           window=cfg.msm_window,
                                           table_budget_bytes=None if cfg.msm_table_gb is None else int(cfg.msm_table_gb * (1 << 30)))
        self.vrf = cp.BatchedVRF([k.vrf_noise for k in self.keys], self.device)
        self.signer = None
        if cfg.verification and cfg.sign_approvals and self.commit is not None:
            self.signer = ApprovalSigner(self.device, [k.schnorr_secret for k in self.keys])
        # Nonce salt of the approvals: binds the run (seed, data, every flag) - the previous block hash binds the fork.
        # (Scheduling knobs - prefetch, pipeline, kernel choices - do not change what is signed and stay out of it.)
        sched = {"prefetch", "pipeline", "gram_impl", "comm_impl", "msm_window", "msm_table_gb", "affine_wire", "eval_every", "extra",
                 "max_iterations", "test_size", "lookahead", "native_schedule", "ring_depth"}
        self.run_salt = hashlib.sha256(("bsc-run-salt|" + repr(sorted((k, v) for k, v in cfg.to_dict().items()
                                                                        if k not in sched))).encode()).digest()
        self.nsig_max = cfg.num_verifiers if self.signer is not None else 0
        self.pack_stride = 64 * (1 + self.nsig_max)
        self.reports: list[BlockReport] = []
        self._rng = np.random.RandomState(cfg.seed * 65537 % (2**31))
        self._timing = False
        self._bad_prev = np.zeros(4, dtype=np.int64)
        self._bufs_ready = False

    # ------------------------------------------------------------------------------------------
    def _random_tau(self) -> int:
        h = hashlib.sha256(f"bsc-tau-{self.cfg.seed}".encode()).digest()
        return int.from_bytes(h, "big") % cp_order()

    def _arrival_order(self, ids: list[int], tag: str) -> list[int]:
        """Deterministic stand-in for network arrival order, identical on every rank."""
        perm = chain_ops.seeded_perm(chain_ops.seed64(self.cfg.seed, self.iteration, tag), len(ids))
        return [ids[i] for i in perm]

    def _phase(self, name: str):
        return _Phase(self, name)

    @property
    def pipelined(self) -> bool:
        """Chain / off-chain / approval streams (GPU engines that are not being instrumented)."""
        return (self.device.type == "cuda" and self.cfg.mode != "fedsys" and self.cfg.pipeline and not self._timing)

    def _idx(self, values, dtype=torch.int32) -> torch.Tensor:
        """Small host list -> device tensor through a pinned staging ring (asynchronous H2D, no host sync)."""
        n = len(values)
        if self.device.type != "cuda":
            return torch.tensor(values, dtype=dtype)
        if not hasattr(self, "_stage"):
            self._stage = [torch.empty(4096, dtype=torch.int64).pin_memory() for _ in range(32)]
            self._stage_i = 0
        buf = self._stage[self._stage_i % len(self._stage)]
        self._stage_i += 1
        view = buf.view(dtype)[:n] if n <= buf.view(dtype).numel() else torch.empty(n, dtype=dtype).pin_memory()
        if n:
            view.copy_(torch.as_tensor(values, dtype=dtype))
        return view.to(self.device, non_blocking=True)

    def _idx_pack(self, *specs):
        """Several small host arrays -> device tensors with ONE staging copy.  specs: (values, dtype) pairs; each
        tensor starts on a 16-byte boundary of the shared buffer.  Packing is plain numpy into the pinned buffer."""
        if self.device.type != "cuda":
            return [torch.as_tensor(np.array(v, dtype=_NP[dt])) for v, dt in specs]
        if not hasattr(self, "_pack_stage"):
            self._pack_stage = [torch.empty(1 << 16, dtype=torch.uint8).pin_memory() for _ in range(8)]
            self._pack_np = [b.numpy() for b in self._pack_stage]
            self._pack_i = 0
        k = self._pack_i % len(self._pack_stage)
        self._pack_i += 1
        buf, bnp = self._pack_stage[k], self._pack_np[k]
        offs, off = [], 0
        for v, dt in specs:
            nb = len(v) * _SZ[dt]
            offs.append((off, nb))
            off = (off + nb + 15) & ~15
        if off > buf.numel():
            return [self._idx(list(v), dt) for v, dt in specs]
        for (o, nb), (v, dt) in zip(offs, specs):
            if nb:
                bnp[o: o + nb].view(_NP[dt])[:] = v
        dev_buf = buf[: max(off, 16)].to(self.device, non_blocking=True)
        return [dev_buf[o: o + nb].view(dt) for (o, nb), (v, dt) in zip(offs, specs)]

    def _ensure_bufs(self):
        """Persistent device buffers of everything that crosses a stream boundary (double-buffered by block parity
        where the off-chain stream may still be reading block t while the chain of block t+1 writes)."""
        if self._bufs_ready:
            return
        dev, P = self.device, max(1, self.P)
        nk = max(1, self.cfg.num_nodes)
        i32 = dict(dtype=torch.int32, device=dev)
        self._plan_out = [dict(counts=torch.zeros(8, **i32), contrib_pos=torch.zeros(nk, **i32), sel=torch.zeros(P, **i32),
                               appr_local=torch.zeros(P, **i32), csel=torch.zeros(P, **i32)) for _ in range(self.NB)]
        self._cbytes = [torch.zeros(P, 64, dtype=torch.uint8, device=dev) for _ in range(self.NB)]
        self._pack = [torch.zeros(self.per, self.pack_stride, dtype=torch.uint8, device=dev) for _ in range(self.NB)]
        self._ysum = [torch.zeros(self.cfg.nchunk, self.S, dtype=torch.int64, device=dev) for _ in range(self.NB)]
        self._bad = self.comm.bad if self.fused else torch.zeros(4, dtype=torch.int32, device=dev)
        self._noiser_tab = torch.full((P, max(1, self.cfg.num_noisers)), -1, **i32)
        self._plan_keep = [None] * self.NB
        self._bufs_ready = True

    def _native(self):
        """The native command-list schedule, when this engine's configuration is the standard one (else None: every
        launch is then issued from Python, see protocol/schedule.py for what 'standard' means)."""
        if not self.pipelined or self._main is None:
            return None
        if self._ns_ok is None:
            from .schedule import NativeSchedule

            self._ns_ok = NativeSchedule.supported(self)
            if self._ns_ok:
                torch.cuda.synchronize(self.device)
                self._ns = NativeSchedule(self)
                # Everything built so far (torch, the data, the tables, the command lists) lives as long as the engine: move it
                # out of the garbage collector's young generations, so that a full collection in the middle of a run walks
                # the few objects a block creates instead of the whole heap (a ~1 ms stall of the block loop otherwise).
                import gc

                gc.collect()
                gc.freeze()
        return self._ns

    def _ensure_eval_slots(self):
        if not hasattr(self, "_ev_slots"):
            self._ev_slots = [(torch.zeros(2, dtype=torch.int32, device=self.device),
                               torch.zeros(2, dtype=torch.int32).pin_memory(), torch.cuda.Event()) for _ in range(4)]

    def _attach_eval_resolver(self, rep: BlockReport, cnt_h: torch.Tensor, event):
        """rep's error rates = the two counts that land in `cnt_h` once `event` has passed."""
        it, log = rep.iteration, self.log
        nt, na = self.data.X_test.shape[0], self.data.X_attack.shape[0]

        def resolve():
            event.synchronize()
            c = cnt_h.numpy()
            rep._test_error, rep._attack_rate, rep._resolve = float(c[0] / max(1, nt)), float(c[1] / max(1, na)), None
            log.train_error(it, rep._test_error)
            log.attack_rate(it, rep._attack_rate)

        rep._resolve = resolve
        self._eval_waiting.append(rep)

    def _order_side_after_main(self):
        ev = torch.cuda.Event()
        ev.record()
        self._side.wait_event(ev)
        if self.signer is not None:
            self.signer.stream.wait_event(ev)

    # ------------------------------------------------------------------------------------------
    def select_roles(self):
        cfg = self.cfg
        h = self.chain.latest_hash()
        stake = [s if a else 0 for s, a in zip(self.stake, self.alive)]
        if sum(stake) == 0:
            return [], [], h
        verifiers, miners = cp.lottery_roles(h, stake, cfg.num_verifiers, cfg.num_miners)
        verifiers = sorted(verifiers)  # getRoleNames sorts verifiers (main.go:529-565)
        return verifiers, miners, h

    def sample_minibatch(self, n_local: int, it: int | None = None):
        """Random minibatch without replacement per peer (logistic_model.py:121-125, DataLoader shuffle)."""
        Ns, B = self.cfg.shard_size, self.cfg.batch_size
        n = self.cfg.num_nodes
        it = self.iteration if it is None else it
        if self.device.type == "cuda" and Ns <= 1024 and self.P and not self.external_inputs:
            # one kernel, keyed by (seed, iteration, peer, row): no generator state, no top-k, same rows on any rank
            if not hasattr(self, "_mb_idx"):
                self._mb_idx = torch.empty(self.P, B, dtype=torch.int32, device=self.device)
            check(lib().bsc_sample_minibatch(self.cfg.seed & (2**64 - 1), it, ptr(self.data.peer_ids), self.P, Ns, B,
                                             ptr(self._mb_idx), raw_stream(self.device)), "sample_minibatch")
            return self._mb_idx
        # One seeded draw for ALL n peers (so the result does not depend on how peers are sharded over
        # ranks), generated on the engine's device; each rank keeps the rows of its own peers.
        if not hasattr(self, "_mb_gen"):
            self._mb_gen = torch.Generator(device=self.device)
        self._mb_gen.manual_seed((self.cfg.seed * 1000003 + it * 7919) % (2**62))
        keys = torch.rand(n, Ns, generator=self._mb_gen, device=self.device)
        idx = keys.topk(B, dim=1, largest=False).indices.int()
        lo = self.local_ids[0] if self.local_ids else 0
        return idx[lo: lo + max(1, self.P)].contiguous()

    # ------------------------------------------------------------------------------------------
    def _host_plan(self) -> _Plan:
        """Roles, candidate set, arrival orders and availability of a block - functions of the previous hash, the
        stake map and who is alive, identical on every rank (vrf.go:103-182, krum.go:296-314, 368-388, main.go:360).
        One native call (ledger.cpp: bsc_plan_block)."""
        cfg, it, n = self.cfg, self.iteration, self.cfg.num_nodes
        if not hasattr(self, "_pb"):
            i32 = lambda k: np.zeros(k, dtype=np.int32)   # noqa: E731
            self._pb = dict(verifiers=i32(64), miners=i32(64), workers=i32(n), krum=i32(n), rank=i32(n), signers=i32(64),
                            avail=np.zeros(max(64, self.S), dtype=np.uint8), xs=i32(16), isw=np.zeros(n, dtype=np.uint8),
                            stake_m=i32(n), counts=i32(8), stake=i32(n), alive=np.zeros(n, dtype=np.uint8))
            self._pb_ptr = {k: v.ctypes.data_as(ctypes.c_void_p) for k, v in self._pb.items()}
        pb, pp = self._pb, self._pb_ptr
        pb["stake"][:] = self.stake
        pb["alive"][:] = self.alive
        prev_hash = self.chain.latest_hash()
        lib().bsc_plan_block(prev_hash, len(prev_hash), pp["stake"], pp["alive"], n, cfg.num_verifiers, cfg.num_miners,
                             cfg.krum_thresh, cfg.num_samples, int(cfg.random_sampling),
                             chain_ops.seed64(cfg.seed, it, "verify"), chain_ops.seed64(cfg.seed, it, "shares"),
                             chain_ops.seed64("rs", it), ctypes.c_double(cfg.krum_adversary_frac), cfg.shares_per_miner,
                             self.S, pp["verifiers"], pp["miners"], pp["workers"], pp["krum"], pp["rank"], pp["signers"],
                             pp["avail"], pp["xs"], pp["isw"], pp["stake_m"], pp["counts"])
        nv, nm, nw, nk, ns, ok, enough, clip = (int(x) for x in pb["counts"])
        pl = _Plan(pb["verifiers"][:nv].tolist(), pb["miners"][:nm].tolist(), pb["workers"][:nw].tolist(), prev_hash)
        pl.stake_m, pl.is_worker = pb["stake_m"].copy(), pb["isw"].copy()
        if nv == 0 or nm == 0 or nw == 0:
            return pl
        pl.krum_set = pb["krum"][:nk].tolist()
        pl.clip = clip
        pl.share_rank = pb["rank"][:nk].copy()
        pl.signers = pb["signers"][:ns].tolist()
        pl.approved_ok = bool(ok)
        pl.trigger = max(1, cfg.num_samples // 2) if cfg.share_trigger == "half" else nk
        pl.shares_avail = pb["avail"][: self.S].tolist()
        pl.xs = [s for s, a in enumerate(pl.shares_avail) if a][:POLY_SIZE]
        pl.enough_shares = bool(enough)
        return pl

    # ------------------------------------------------------------------------------------------
    def _local_step(self, inputs, it: int, timed: bool) -> dict:
        """Local SGD step of every local peer from the current global weights (phase 2 of an iteration) into self.U.
        Returns the block's `pre` record; on a pipelined engine `u_ready` marks U complete."""
        cfg, dev = self.cfg, self.device
        slot = it % NOISE_TABLE_ITERS
        phase = self._phase if timed else (lambda name: contextlib.nullcontext())
        out = {"it": it, "Q": None, "chunkC": None, "cbytes": None, "u_ready": None, "c_ready": None,
               "src": None if inputs is None else inputs[0].data_ptr()}
        with phase("local_step"):
            if inputs is not None:
                X, y = inputs
                idx = None
            else:
                X, y = self.data.X, self.data.y
                idx = self.sample_minibatch(self.P, it).to(dev)
            if self.P:
                pid = self.data.peer_ids if not self.external_inputs else torch.tensor(self.local_ids, dtype=torch.int32, device=dev)
                self.model.local_step(self.w32, X, y, idx, self.U, peer_ids=pid, slot=slot)
                if cfg.dp_in_model and cfg.noising and cfg.epsilon > 0 and self.model.kind != "logistic":
                    # DP_IN_MODEL (honest.go:172-179, main.go:155): the peer's own presampled noise goes INTO the delta
                    # that is committed, shared and aggregated (the NumPy logistic path always does this, inside its step)
                    own = pid.view(-1, 1).contiguous()
                    ml.noise_mix(self.U, own, cfg.seed ^ 0x6F776E6E, slot, cfg.noise_scale, self.d, want_planes=False,
                                 in_place=True)
            if dev.type == "cuda" and self.pipelined:
                out["u_ready"] = torch.cuda.Event()
                out["u_ready"].record()
        return out

    def _commit_step(self, out: dict, timed: bool) -> dict:
        """Quantisation, the polynomial commitments of the update and their 64-byte wire form (phase 3): fills Q, chunkC,
        cbytes (device [P, 64]) of the block's `pre` record; `c_ready` marks cbytes complete on a pipelined engine."""
        cfg, dev = self.cfg, self.device
        phase = self._phase if timed else (lambda name: contextlib.nullcontext())
        if self.commit is not None and self.P:
            with phase("commit"):
                Q = secagg.quantize(self.U, self.d, cfg.precision)
                chunkC = self.commit.chunk_commitments(Q, affine=cfg.affine_wire)    # [P, nchunk, 16|24]
                full = self.commit.full_commitments(chunkC)                          # affine [P, 16]
                self._ensure_bufs()
                cb = self._cbytes[out["it"] % self.NB]
                self.commit.to_bytes_into(full, cb)                                  # [P, 64] wire bytes, on the device
                out.update(Q=Q, chunkC=chunkC, cbytes=cb)
                if dev.type == "cuda" and self.pipelined:
                    out["c_ready"] = torch.cuda.Event()
                    out["c_ready"].record()
        return out

    def _local_and_commit(self, inputs, it: int, timed: bool):
        return self._commit_step(self._local_step(inputs, it, timed), timed)

    def _next_inputs_ready(self, stream=None):
        """Inputs of the next block if its local step may be queued now: None = resident shards, False = not available
        (minibatches come from the caller and the next one has not been handed over, or prefetching is off).  `stream`:
        the stream that will read them (default: the current one)."""
        cfg = self.cfg
        if not cfg.prefetch or not cfg.secure_agg or cfg.mode == "fedsys":
            return False   # the plain path still reads U when the block is assembled
        nxt = self._next_inputs
        if nxt is None and (self._fed_inputs or self.external_inputs):
            return False
        if nxt is not None and len(nxt) == 3:      # (X, y, event): the H2D copy was issued on another stream
            (stream or torch.cuda.current_stream(self.device)).wait_event(nxt[2])
            nxt = (nxt[0], nxt[1])
        return nxt

    def _prefetch_next(self, it_next: int):
        """Queue the next iteration's local step and commitment MSM behind this iteration's aggregation kernel.
        They read only the new global weights (already updated in stream order) and the resident shards, so the
        device works on them while the host assembles, signs and hashes the current block.  The reference has no
        such overlap (phases are strictly sequential); the values are identical either way."""
        nxt = self._next_inputs_ready()
        if nxt is False:
            return
        self._pre = self._local_and_commit(nxt, it_next, timed=False)

    def step(self, inputs=None, next_inputs=None) -> BlockReport:
        """Run one block iteration.  `inputs` = (X [P,B,F], y [P,B]) device tensors when the caller feeds
        minibatches (bench end-to-end path); otherwise minibatches are drawn from the resident shards.
        `next_inputs` = the following call's `inputs` (optionally with a CUDA event that marks their H2D copy
        complete), if the caller already has them: a double-buffered input pipeline lets the engine start the
        next local step while this block is being assembled.

        On a GPU the block's chain runs on the engine's own high-priority stream (ordered after the caller's stream on
        entry, and the caller's stream after it on exit), so that the off-chain work queued on the default-priority
        stream - witness MSM, aggregation, the next local step and its commitments - yields SM slots to the kernels the
        block is waiting for."""
        if not self.pipelined:
            return self._step(inputs, next_inputs)
        if self._main is None:
            # chain > approvals > short off-chain kernels > witness MSM (CUDA: a lower number is a higher priority)
            self._main = torch.cuda.Stream(self.device, priority=-3)
            self._side = torch.cuda.Stream(self.device, priority=-1)
            self._wit = torch.cuda.Stream(self.device, priority=0)
            self._agg = torch.cuda.Stream(self.device, priority=-1)
            self._wstream = self._side
        cur = torch.cuda.current_stream(self.device)
        self._main.wait_stream(cur)
        with torch.cuda.stream(self._main):
            rep = self._step(inputs, next_inputs)
        # The caller's stream is ordered after the block's chain (its results: weights, ledger) - not after the off-chain
        # work, which may trail by up to two blocks (the chain itself waits for it, see _queue_chain_front); synchronize()
        # or close() drains everything.
        cur.wait_stream(self._main)
        return rep

    @contextlib.contextmanager
    def _off_chain(self, after: "torch.cuda.Event | None" = None):
        """Queue work off the block's chain: on the second stream of a pipelined engine, in line otherwise."""
        if not self.pipelined:
            yield None
            return
        if after is None:          # default: after everything queued on the chain so far
            after = torch.cuda.Event()
            after.record()
        self._side.wait_event(after)
        with torch.cuda.stream(self._side):
            yield self._side

    def _step(self, inputs=None, next_inputs=None) -> BlockReport:
        cfg = self.cfg
        self._next_inputs, self._fed_inputs = next_inputs, inputs is not None
        if cfg.mode == "fedsys":
            from .fedsys import fedsys_step

            return fedsys_step(self, inputs)
        t0 = time.perf_counter()
        src = None if inputs is None else inputs[0].data_ptr()
        # The previous step may already have planned this block and queued the front of its chain (see _after_block).
        ctx, self._ctx_next = self._ctx_next, None
        if ctx is not None and not self._ctx_valid(ctx, src):
            ctx = None          # somebody changed the world in between (alive set, adopted chain, other inputs): replan
        if ctx is None:
            ctx = self._begin_block()
        ctx.t0 = t0
        self._phases = ctx.phases
        if ctx.empty:
            return self._finish_block(ctx.rep, ctx.phases, t0)        # main.go:1696-1707, 1934-1944, 2099-2117
        if ctx.native:
            ns = self._ns
            if not ctx.front_queued:
                # first block, a replanned block, or no lookahead: queue the whole block now
                self._order_side_after_main()
                ns.queue_sign_T(ctx)
                ns.queue_local_and_eval(ctx.it, None, False, inputs=inputs)
                ns.queue_front(ctx)
                ns.queue_commit_and_sign(ctx)
            ns.queue_back(ctx)
            if not self.fused:
                self._last_ysum = self._ysum[ctx.par]
            # The next block's local step and commitments need only the weights this block's back produces: queue them now,
            # before the host sleeps on this block's read-back, so that they run while the host assembles the block.
            ctx.npre = None
            if cfg.lookahead:
                nxt = self._next_inputs_ready(self._side)
                if nxt is not False:
                    ns.queue_local_and_eval(ctx.it + 1, None, False, inputs=nxt)
                    ctx.npre = ns.queue_commit(ctx.it + 1)
            ctx.phases["host_queue"] = (time.perf_counter() - t0) * 1e3
            host = ns.collect()
            ctx.phases["host_wait"] = (time.perf_counter() - t0) * 1e3 - ctx.phases["host_queue"]
            return self._tail(ctx, host)
        if ctx.pre is None:
            pre, self._pre = self._pre, None
            if not (pre is not None and pre["it"] == ctx.it and pre.get("src") == src):
                with self._off_chain():
                    pre = self._local_and_commit(inputs, ctx.it, timed=True)
            ctx.pre = pre
        if not ctx.front_queued:
            self._queue_approvals(ctx)
            self._queue_chain_front(ctx)
        self._queue_chain_back(ctx)
        return self._tail(ctx)

    # ---- stage 0: everything about the block that follows from the previous hash ----------------------------------
    def _begin_block(self) -> "_Ctx":
        """Fault injection / partitions for this height (prepareForNextIteration, main.go:1059-1187), roles, candidate
        set, arrival orders (host), the VRF launch on the previous hash, and ONE staging copy of the plan to the device."""
        cfg, it, n, dev = self.cfg, self.iteration, self.cfg.num_nodes, self.device
        cuda = dev.type == "cuda"
        if self._faults_applied_for != it:
            self._faults_applied_for = it
            # a failing peer sits this block out and rejoins (main.go:1117-1120)
            if cfg.fail_prob > 0:
                for p in range(1, n):
                    self.alive[p] = self._rng.rand() >= cfg.fail_prob
            # scheduled network partitions (DistSys/blockNode.sh): the peer is cut off for a range of blocks; it holds no
            # role and contributes nothing meanwhile, and is back (with the chain everybody shares) afterwards
            if cfg.partitions:
                cut = {int(p) for p, a, b in cfg.partitions if int(a) <= it <= int(b)}
                was = getattr(self, "_cut", set())
                for p in was - cut:
                    self.alive[p] = True
                for p in cut:
                    self.alive[p] = False
                self._cut = cut
        t_p0 = time.perf_counter()
        ns = self._native()
        pl = ns.plan_block(_Plan) if ns is not None else self._host_plan()
        self._t_plan_host = (time.perf_counter() - t_p0) * 1e3
        krum_set = pl.krum_set
        nk = len(krum_set)
        ctx = _Ctx(it=it, par=it % self.NB, pl=pl, nk=nk, phases={}, alive=list(self.alive), stake=list(self.stake),
                   timing=self._timing, pipelined=self.pipelined)
        n_workers = pl.n_workers if ns is not None else len(pl.workers)
        ctx.rep = BlockReport(it, b"", True, pl.verifiers, pl.miners, n_workers, nk, 0, 0)
        if not pl.verifiers or not pl.miners or not pl.workers or not pl.enough_shares or nk == 0:
            ctx.empty = True
            return ctx
        self._ensure_bufs()
        ctx.pout = self._plan_out[ctx.par]
        prev_hash = pl.prev_hash
        # VRF evaluation of every peer on the previous block hash: launched the moment that block was hashed
        # (_finish_block) on its own high-priority stream; the noise phase consumes it on the device.
        ctx.use_noise = cfg.noising and cfg.num_noisers > 0 and cfg.epsilon > 0
        early, self._vrf_early = self._vrf_early, None
        if ctx.use_noise and early is not None and early[0] == prev_hash:
            ctx.vrf_handle = early[1]
        elif ctx.use_noise:
            ctx.vrf_handle = self.vrf.launch_dev(prev_hash) if cuda else self.vrf.launch(prev_hash)
        # ---- host -> device: the block's plan in one staging copy
        ctx.stake_m, ctx.is_worker = pl.stake_m, pl.is_worker
        ctx.salt64 = self.run_salt + (prev_hash + b"\0" * 32)[:32]
        ctx.ns = len(pl.signers) if (self.signer is not None and pl.approved_ok) else 0
        if ns is not None and nk >= 3:
            ctx.native = True
            t_u0 = time.perf_counter()
            ns.upload_plan(ctx)
            ctx.phases["h_plan_host"] = self._t_plan_host
            ctx.phases["h_plan_upload"] = (time.perf_counter() - t_u0) * 1e3
            return ctx
        kpos = {p: i for i, p in enumerate(krum_set)}
        # Single-rank tcgen05 Krum: noise_mix writes the candidates' bf16 planes and norms compacted and in Krum
        # order (row = position in krum_set), so the Gram kernel needs no gather; the fp32 noised rows are only
        # materialised for the paths that read them (RONI, SIMT Gram, NCCL gathers).
        ctx.compact = (cuda and self.comm is None and cfg.verification and cfg.defense != "RONI"
                       and cfg.gram_impl in ("auto", "tcgen05") and nk >= 3)
        dst_rows = [kpos.get(p, -1) for p in self.local_ids] if ctx.compact else []
        local_row = [self.local_index.get(p, -1) for p in krum_set]
        slots = [self.comm.slot_of(p) for p in krum_set] if self.fused else []
        (ctx.stake_d, ctx.isw_d, ctx.dst_row, ctx.local_row_d, ctx.share_rank_d, ctx.krum_ids_d, ctx.signers_d, ctx.xs_t,
         ctx.av_t, ctx.slots_d, ctx.hash_d) = \
            self._idx_pack((ctx.stake_m, torch.int32), (ctx.is_worker, torch.uint8), (dst_rows, torch.int32),
                           (local_row, torch.int32), (pl.share_rank, torch.int32), (krum_set, torch.int32),
                           (pl.signers, torch.int32), (pl.xs, torch.int32), (pl.shares_avail, torch.uint8),
                           (slots, torch.int32), (np.frombuffer(bytearray(prev_hash), dtype=np.uint8), torch.uint8))
        self._plan_keep[ctx.par] = (ctx.stake_d, ctx.local_row_d, ctx.krum_ids_d, ctx.signers_d)   # read on other streams
        if not ctx.compact:
            ctx.dst_row = None
        if cuda and self.pipelined:
            ctx.plan_ready = torch.cuda.Event()
            ctx.plan_ready.record()
        ctx.salt64 = self.run_salt + (prev_hash + b"\0" * 32)[:32]
        ctx.ns = len(pl.signers) if (self.signer is not None and pl.approved_ok) else 0
        return ctx

    def _ctx_valid(self, ctx: "_Ctx", src) -> bool:
        return (ctx.it == self.iteration and ctx.pl.prev_hash == self.chain.latest_hash() and ctx.alive == self.alive
                and ctx.stake == self.stake and ctx.timing == self._timing and ctx.pipelined == self.pipelined
                and (ctx.pre is None or ctx.pre.get("src") == src))

    # ---- stage 1: approvals (their own stream) -----------------------------------------------------------------------
    def _queue_approvals(self, ctx: "_Ctx"):
        """6 (early). T = v*G, c = H(T || commitment), r = v - x*c depend only on (chain position, update, verifier) and
        on the commitment, not on the Krum verdict: every candidate of this rank is signed now, on the approval stream,
        while the VRF and Krum run; only the rows of the block's updates are ever read."""
        dev, pre, pl = self.device, ctx.pre, ctx.pl
        if dev.type != "cuda" or self.commit is None or not self.P:
            return
        pack = self._pack[ctx.par]
        if ctx.ns:
            ctx.sig_event = self.signer.launch(ctx.salt64, ctx.it, ctx.krum_ids_d, ctx.local_row_d, ctx.nk, ctx.signers_d,
                                               ctx.ns, pre["cbytes"], pack, after=ctx.plan_ready, serial=not self.pipelined,
                                               commit_ready=pre["c_ready"])
        else:
            with self._off_chain(pre["c_ready"]):
                check(lib().bsc_pack_commitments(ptr(pre["cbytes"]), self.P, ptr(pack), pack.stride(0), raw_stream(dev)),
                      "pack_commitments")
                if self.pipelined:
                    ctx.sig_event = torch.cuda.Event()
                    ctx.sig_event.record()

    # ---- stage 2: the chain up to the share sums (writes scratch only, so it may be queued a block ahead) ---------------
    def _queue_chain_front(self, ctx: "_Ctx"):
        cfg, dev, pl, pre, pout = self.cfg, self.device, ctx.pl, ctx.pre, ctx.pout
        cuda = dev.type == "cuda"
        it, nk, krum_set = ctx.it, ctx.nk, pl.krum_set
        slot = it % NOISE_TABLE_ITERS
        prev = self._phases
        self._phases = ctx.phases
        # ---- 4. noisers + noise
        if self.pipelined:
            main = torch.cuda.current_stream(dev)
            if pre["u_ready"] is not None:
                main.wait_event(pre["u_ready"])
            if self._side_done[ctx.par] is not None:
                main.wait_event(self._side_done[ctx.par])   # block it-2's witnesses / aggregation still read this parity's
                                                            # plan lists: stay at most two blocks ahead of the off-chain stream
        with self._phase("noise"):
            nn_cols = max(1, cfg.num_noisers)
            if ctx.use_noise and cuda:
                xyz, vev, _ = ctx.vrf_handle
                vev.wait()
                noiser_tab = self._noiser_tab
                if self.P:
                    chain_ops.vrf_noisers(xyz, ctx.hash_d, len(pl.prev_hash), ctx.stake_d, ctx.isw_d,
                                          self.data.peer_ids if not self.external_inputs else self._idx(self.local_ids),
                                          cfg.num_noisers, noiser_tab)
            elif ctx.use_noise:
                full_tab = self.vrf.noiser_table(ctx.vrf_handle, ctx.stake_m.tolist(), ctx.is_worker, cfg.num_noisers)
                ctx.noisers_host = full_tab
                lo_id = self.local_ids[0] if self.P else 0
                noiser_tab = torch.from_numpy(np.ascontiguousarray(full_tab[lo_id: lo_id + max(1, self.P)]))
            else:
                noiser_tab = torch.full((max(1, self.P), nn_cols), -1, dtype=torch.int32, device=dev)
            ctx.noiser_tab = noiser_tab
            scale = cfg.noise_scale if ctx.use_noise else 0.0
            Un, hi, lo, norms = ml.noise_mix(self.U, noiser_tab, cfg.seed, slot, scale, self.d,
                                             ldh=self.ldh, out_hi=self.comm.hi if self.fused else None,
                                             out_lo=self.comm.lo if self.fused else None, dst_row=ctx.dst_row,
                                             want_un=not ctx.compact)
        # ---- 5. verification -> accept mask on the device
        with self._phase("verify"):
            self._compact = ctx.compact
            if not cfg.verification or nk < 3:
                accept = torch.ones(nk, dtype=torch.uint8, device=dev)
            elif cfg.defense == "RONI":
                accept = self._roni(krum_set, Un)
            else:
                accept = self._krum(krum_set, Un, hi, lo, norms, pl.clip, ctx.slots_d, epoch=it + 1)
            ctx.accept = accept
            # approved set, contributors and the compacted local row lists, still on the device
            chain_ops.block_plan(accept, ctx.share_rank_d, ctx.local_row_d, pl.approved_ok, pl.trigger, pout)
        counts = pout["counts"]
        ctx.n_lc, ctx.n_la, ctx.n_con = counts[2:3], counts[3:4], counts[1:2]
        # ---- 7. share sums of my contributors
        with self._phase("secagg"):
            if cfg.secure_agg:
                out = self.comm.ypart if self.fused else self._ysum[ctx.par]
                secagg.share_sum(self.U, pout["sel"], self.d, self.S, cfg.precision, out=out, count=ctx.n_lc)
        ctx.front_queued = True
        self._phases = prev

    # ---- stage 3: reconstruction + SGD step (the only chain kernels that change the weights) ---------------------------
    def _queue_chain_back(self, ctx: "_Ctx"):
        """8. Gated on the device: fewer than two contributors (main.go:2079) leaves the weights untouched and the tail
        mints the empty block."""
        cfg, dev, pl, pout = self.cfg, self.device, ctx.pl, ctx.pout
        it = ctx.it
        if self.pipelined and self._eval_done is not None:
            torch.cuda.current_stream(dev).wait_event(self._eval_done)     # the previous block's error rates read w
        with self._phase("secagg"):
            if cfg.secure_agg:
                if self.fused:
                    self.comm.secagg_fused(pl.xs, pl.shares_avail, cfg.precision, None, gate=ctx.n_con, gate_min=2)
                else:
                    Ysum = self._ysum[ctx.par]
                    if self.comm is not None:
                        Ysum = self.comm.all_reduce_i64(Ysum)
                    self._last_ysum = Ysum
                    secagg.reconstruct_step_dev(Ysum, pl.xs, ctx.xs_t, ctx.av_t, self.d, self.w64, self.w32, cfg.precision,
                                                bad=self._bad, gate=ctx.n_con, gate_min=2)
            else:
                if self.comm is not None and self.world > 1:
                    rows = self.comm.gather_rows(pl.krum_set, self.U, self.local_index, self.owner)
                    secagg.plain_sum_step(rows, pout["contrib_pos"], self.d, self.w64, self.w32, count=ctx.n_con,
                                          gate=ctx.n_con, gate_min=2)
                else:
                    secagg.plain_sum_step(self.U, pout["sel"], self.d, self.w64, self.w32, count=ctx.n_lc, gate=ctx.n_con,
                                          gate_min=2)

    # ---- stage 4: the block's tail --------------------------------------------------------------------------------------
    def _tail(self, ctx: "_Ctx", host: dict | None = None) -> BlockReport:
        """9. ONE device->host collection (weights, counters, contributor list, gathered commitment||signature rows), then
        block assembly, stake update, hashing and append (one native call), then whatever can be queued for later."""
        cfg, dev, pl, pout, pre, rep = self.cfg, self.device, ctx.pl, ctx.pout, ctx.pre, ctx.rep
        cuda = dev.type == "cuda"
        it, par, krum_set, n = ctx.it, ctx.par, pl.krum_set, cfg.num_nodes
        counts = pout["counts"]
        with self._phase("block"):
            if host is None:
                want = {"w": self.w64, "bad": self._bad, "counts": counts, "cpos": pout["contrib_pos"]}
                if not cfg.secure_agg:
                    want["accept"] = ctx.accept
                if self.colluder_ids and ctx.use_noise and cuda:
                    want["noisers"] = ctx.noiser_tab
                if cuda and self.commit is not None:
                    if ctx.sig_event is not None:
                        torch.cuda.current_stream(dev).wait_event(ctx.sig_event)
                    want["pack"] = self._gather_pack(self._pack[par], par)
                host = self._collect(want)
            cnt = host["counts"]
            n_contrib = int(cnt[1])
            rep.n_accepted = int(host["accept"].sum()) if "accept" in host else int(cnt[0])
            contributors = [int(krum_set[int(i)]) for i in host["cpos"][:n_contrib]] if n_contrib >= 2 else []
            badv = host["bad"].astype(np.int64)
            if badv.shape[0] >= 3 and badv[2]:
                raise RuntimeError(f"rank {self.rank}: a cross-GPU wait timed out (code {int(badv[2])}) in block {it}")
            rep.bad_shares = int((badv[:2] - self._bad_prev[:2]).sum())
            self._bad_prev[: badv.shape[0]] = badv
            # collusion accounting (main.go:1026-1057, 1138-1144)
            if self.colluder_ids and ctx.use_noise:
                noisers_host = ctx.noisers_host
                if noisers_host is None:
                    tab = host["noisers"]
                    if self.comm is not None and self.world > 1:
                        tab = np.concatenate(self.comm.all_gather_np(tab[: self.per] if tab.shape[0] >= self.per
                                                                     else np.pad(tab, ((0, self.per - tab.shape[0]), (0, 0)),
                                                                                 constant_values=-1)))
                    noisers_host = tab if tab.shape[0] >= n else np.pad(tab, ((0, n - tab.shape[0]), (0, 0)), constant_values=-1)
                vcol = any(v in self.colluder_ids for v in pl.verifiers)
                rep.unmasked_updates = sum(
                    1 for p in krum_set if vcol and any(q >= 0 for q in noisers_host[p]) and
                    all(int(q) in self.colluder_ids for q in noisers_host[p] if q >= 0))
            rep.n_contributors = len(contributors)
            rep.contributors = contributors
            updates, packed = [], None
            if contributors:
                if self.commit is not None:
                    pack_host = host.get("pack")
                    nsig = len(pl.signers) if (self.signer is not None and pl.approved_ok) else 0
                    if pack_host is None:          # CPU engine: commitments and approvals of the contributors, now
                        pack_host = self._pack_host(contributors, pl.signers if nsig else [], pre["cbytes"], ctx.salt64, it)
                    packed = (contributors, pack_host, nsig)
                else:
                    for p in contributors:
                        rec = UpdateRecord(source=p, iteration=it, commitment=b"", accepted=True, signatures=[])
                        if not cfg.secure_agg and p in self.local_index and self.world == 1:
                            rec.delta = self.U[self.local_index[p], : self.d].double().cpu().numpy()
                        updates.append(rec)
                        self.stake[p] += STAKE_UNIT
                if not cfg.secure_agg:
                    acc = host["accept"]
                    for i, p in enumerate(krum_set):
                        if not acc[i]:
                            self.stake[p] -= STAKE_UNIT                  # honest.go:362-370 (plain path only)
        agg_box = [None]
        ctx.phases["host_tail_py"] = ctx.phases.pop("block")
        rep = self._finish_block(rep, ctx.phases, ctx.t0, host=host, updates=updates, packed=packed,
                                 then=lambda chain_done, r: self._after_block(ctx, chain_done, r, agg_box))
        self.last_aggregate = agg_box[0] if contributors else None
        return rep

    # ---- stage 5: behind the block ---------------------------------------------------------------------------------------
    def _after_block(self, ctx: "_Ctx", chain_done, rep: BlockReport, agg_box: list):
        """Called once the block is appended and the next block's VRF is under way.  Queues, in the order the next block
        needs them: the next local step (head of the next chain), this block's error rates (they read the new weights and
        must precede the next SGD step), the front of the NEXT block's chain, the next commitments, the next approvals -
        and last this block's witness MSM and miners' aggregation, which nothing waits for."""
        cfg, pout, pre = self.cfg, ctx.pout, ctx.pre
        it = ctx.it
        do_eval = bool(cfg.eval_every) and (it % cfg.eval_every == 0)
        if ctx.native:
            ns = self._ns
            ph = ctx.phases
            tm = time.perf_counter
            t0 = tm()
            npre = getattr(ctx, "npre", None)
            if npre is not None:
                # The next block's local step and commitments were queued before the tail (see _step).  Order of the rest =
                # what the NEXT tail waits for longest first: the next plan and the approvals' T = v*G, this block's error
                # rates (they must read the weights before the next SGD step), the next chain front, the approvals' last kernel.
                t1 = tm()
                nctx = self._begin_block_ahead()
                t2 = tm()
                ahead = nctx is not None and not nctx.empty and nctx.native
                if ahead:
                    ns.queue_sign_T(nctx)
                if do_eval:
                    ns.queue_eval(rep, ctx.par)
                if ahead:
                    nctx.pre = npre
                    t3 = tm()
                    ns.queue_front(nctx)
                    ns.queue_sign_finish(nctx)
                    self._ctx_next = nctx
                    ph["h_front"] = (tm() - t3) * 1e3
                ph["h_local_eval"] = (t1 - t0) * 1e3
                ph["h_plan"] = (t2 - t1) * 1e3
            elif do_eval:
                ns.queue_eval(rep, ctx.par)
            t4 = tm()
            agg_box[0] = ns.queue_witness(ctx)
            ph["h_witness"] = (tm() - t4) * 1e3
            self._agg_event = ns.agg_done[ctx.par]
            self._side_done[ctx.par] = ns.side_done[ctx.par]
            self._wevent = ns.side_done[ctx.par]
            return
        nxt = self._next_inputs_ready() if not self._timing else False
        npre, nctx = None, None
        with self._off_chain(chain_done):
            if nxt is not False:
                npre = self._local_step(nxt, it + 1, timed=False)
            if do_eval:
                self._launch_eval(rep)
                if self.pipelined:
                    self._eval_done = torch.cuda.Event()
                    self._eval_done.record()
        if npre is not None and self.pipelined and cfg.lookahead:
            # plan block it+1 now and queue the front of its chain: VRF -> noisers -> noise -> Krum -> plan -> share sums
            # write scratch only, so nothing is lost if the world changes before the next step() (the context is then
            # discarded); the weights are touched only by the kernels queued inside that step
            nctx = self._begin_block_ahead()
            if nctx is not None and not nctx.empty:
                nctx.pre = npre
                self._queue_chain_front(nctx)
        with self._off_chain(None if npre is None else npre["u_ready"]):
            if npre is not None:
                self._commit_step(npre, timed=False)
        if nctx is not None and not nctx.empty:
            self._queue_approvals(nctx)
            self._ctx_next = nctx
        elif npre is not None:
            self._pre = npre
        with self._off_chain(chain_done):
            if cfg.secure_agg and self.commit is not None and self.P:
                # Every approved worker generates shares + KZG witnesses for the miners (kyber.go:456-482) and the
                # miners add the contributors' witnesses and chunk commitments homomorphically (kyber.go:244-287).
                # Nothing that enters the block depends on them.
                with self._phase("witness"):
                    W = self.commit.witnesses(pre["Q"], pout["appr_local"], affine=cfg.affine_wire, count=ctx.n_la,
                                              max_rows=min(self.P, ctx.nk))
                with self._phase("miner_aggregate"):
                    agg_box[0] = self.commit.aggregate_by_scalars(pre["Q"], pout["sel"], count=ctx.n_lc,
                                                                  affine=cfg.affine_wire)
            if self.pipelined:
                ev = torch.cuda.Event()
                ev.record()
                self._side_done[ctx.par] = ev
                self._wevent = ev

    def _begin_block_ahead(self):
        """_begin_block for the next height, called from inside the current step (self.iteration not advanced yet)."""
        self.iteration += 1
        try:
            return self._begin_block()
        finally:
            self.iteration -= 1

    # ------------------------------------------------------------------------------------------
    def _gather_pack(self, pack: torch.Tensor, par: int) -> torch.Tensor:
        """[per, stride] commitment||signature rows of my peers -> [world*per, stride] rows of every peer (row = id)."""
        if self.comm is None or self.world == 1:
            return pack
        return self.comm.all_gather_rows_u8(pack)

    def _pack_host(self, contributors, signers, cbytes, salt64, it) -> np.ndarray:
        """CPU engine: the contributors' pack rows, built at the tail (host approvals, gloo gather across ranks)."""
        n, st = self.world * self.per, self.pack_stride
        loc = np.zeros((self.per, st), dtype=np.uint8)
        mine = [p for p in contributors if p in self.local_index]
        if mine:
            cb = cbytes.cpu().numpy() if isinstance(cbytes, torch.Tensor) else cbytes
            rows = [self.local_index[p] for p in mine]
            commits = [cb[r].tobytes() for r in rows]
            for r, c in zip(rows, commits):
                loc[r, :64] = np.frombuffer(c, dtype=np.uint8)
            if signers and self.signer is not None:
                sigs = self.signer.sign_host(salt64, it, mine, signers, commits)
                k = len(signers)
                for j, r in enumerate(rows):
                    loc[r, 64: 64 + 64 * k] = np.frombuffer(b"".join(sigs[j * k: (j + 1) * k]), dtype=np.uint8)
        if self.comm is None or self.world == 1:
            out = np.zeros((n, st), dtype=np.uint8)
            out[: self.per] = loc
            return out
        return np.concatenate(self.comm.all_gather_np(loc))

    def _available_shares(self, miners, alive_miners):
        spm = self.cfg.shares_per_miner
        avail = [0] * self.S
        for j, m in enumerate(miners):
            if m in alive_miners:
                for s in range(j * spm, min(self.S, (j + 1) * spm)):
                    avail[s] = 1
        return avail

    def _gather_rows(self, ids: list[int], local: torch.Tensor, ix=None) -> torch.Tensor:
        """Rows of `local` (indexed by local peer position) for global peer ids, in order."""
        if self.comm is None:
            if ix is None:
                ix = self._idx([self.local_index[p] for p in ids], torch.int64)
            return local.index_select(0, ix).contiguous()
        return self.comm.gather_rows(ids, local, self.local_index, self.owner)

    def _krum(self, krum_set, Un, hi, lo, norms, clip, slots_d=None, epoch=None) -> torch.Tensor:
        """Multi-Krum accept mask (uint8 [nk]) - stays on the device."""
        cfg = self.cfg
        nk = len(krum_set)
        if self.fused:
            _, acc = self.comm.krum_fused(krum_set, clip, None, rows=slots_d)
        elif getattr(self, "_compact", False):
            _, acc, _ = krum_ops.krum(hi[:nk], self.d, clip, hi[:nk], lo[:nk], norms[:nk], impl=cfg.gram_impl)
        else:
            ix = self._idx([self.local_index[p] for p in krum_set], torch.int64) if self.comm is None else None
            Uk = self._gather_rows(krum_set, Un, ix)
            hik = self._gather_rows(krum_set, hi, ix) if hi is not None else None
            lok = self._gather_rows(krum_set, lo, ix) if lo is not None else None
            nrm = self._gather_rows(krum_set, norms.unsqueeze(1), ix).squeeze(1) if norms is not None else None
            _, acc, _ = krum_ops.krum(Uk, self.d, clip, hik, lok, nrm, impl=cfg.gram_impl)
        return acc.to(torch.uint8)

    def _roni(self, krum_set, Un) -> torch.Tensor:
        """RONI (main.go:191-233, honest.go:598-629): reject when validation error rises by > threshold."""
        Uk = self._gather_rows(krum_set, Un)
        base_err = self.model.eval_errors(self.w32.unsqueeze(0), None, self.data.X_test, self.data.y_test)[0]
        errs = self.model.eval_errors(Uk, self.w32, self.data.X_test, self.data.y_test)
        return ((errs - base_err) <= self.cfg.roni_threshold).to(torch.uint8)

    def evaluate(self):
        w = self.w32.unsqueeze(0)
        te = self.model.eval_errors(w, None, self.data.X_test, self.data.y_test)
        ar = self.model.eval_errors(w, None, self.data.X_attack, self.data.y_attack)
        return te, ar

    def verify_last_aggregate(self, samples=((0, 0), (0, 13), (1, 20))):
        """What a miner can check about the last block's aggregation (kyber.go:244-287 + verifySecret): the sum of
        the contributors' witnesses opens the sum of their chunk commitments to the sum of their shares.  Waits
        for the off-chain stream.  Single-rank engines only (each rank aggregates its own contributors).  Returns
        None when there is nothing to check."""
        if self.last_aggregate is None or self.world != 1 or getattr(self, "_last_ysum", None) is None:
            return None
        if self._wevent is not None:
            self._wevent.synchronize()
        if getattr(self, "_agg_event", None) is not None:
            self._agg_event.synchronize()
        aggC, aggW = self.last_aggregate
        Cb, Wb = self.commit.to_bytes(self.commit.normalize(aggC) if aggC.shape[-1] == 24 else aggC), \
            self.commit.to_bytes(self.commit.normalize(aggW) if aggW.shape[-1] == 24 else aggW)
        Y = self._last_ysum.cpu()
        S = self.S
        ok = True
        for c, s in samples:
            if c >= self.cfg.nchunk or s >= S:
                continue
            off = 64 * (c * S + s)
            ok = ok and self.commit.verify_share_trapdoor(Cb[64 * c: 64 * c + 64], Wb[off: off + 64], c, s, int(Y[c, s]))
        return ok

    def _launch_eval(self, rep: BlockReport):
        """Queue the error-rate kernels of the new weights and their read-back; `rep` gets a resolver that waits for
        that read-back when somebody looks at the numbers (honest.go:141-162 logs them per block)."""
        it = rep.iteration
        log = self.log

        def publish(te, ar):
            rep._test_error, rep._attack_rate, rep._resolve = float(te), float(ar), None
            log.train_error(it, rep._test_error)
            log.attack_rate(it, rep._attack_rate)

        if self.device.type == "cuda" and self.model.kind in ("softmax", "logistic"):
            # both error counts into one int32[2] and one read-back; four slots, resolved at the latest two blocks on
            if not hasattr(self, "_ev_slots"):
                self._ev_slots = [(torch.zeros(2, dtype=torch.int32, device=self.device),
                                   torch.zeros(2, dtype=torch.int32).pin_memory(), torch.cuda.Event()) for _ in range(4)]
            cnt_d, cnt_h, ev = self._ev_slots[it & 3]
            cnt_d.zero_()
            w = self.w32.unsqueeze(0)
            d = self.data
            if self.model.kind == "softmax":
                C = self.cfg.spec.n_classes
                ml.eval_softmax_errors(w, None, d.X_test, d.y_test, C, wrong_out=cnt_d[0:1])
                ml.eval_softmax_errors(w, None, d.X_attack, d.y_attack, C, wrong_out=cnt_d[1:2])
            else:
                ml.eval_logistic_errors(w, None, d.X_test, d.y_test, wrong_out=cnt_d[0:1])
                ml.eval_logistic_errors(w, None, d.X_attack, d.y_attack, wrong_out=cnt_d[1:2])
            cnt_h.copy_(cnt_d, non_blocking=True)
            ev.record()
            nt, na = d.X_test.shape[0], d.X_attack.shape[0]

            def resolve():
                ev.synchronize()
                c = cnt_h.numpy()
                publish(c[0] / max(1, nt), c[1] / max(1, na))

            rep._resolve = resolve
            self._eval_waiting.append(rep)
            return
        te, ar = self.evaluate()
        if self.device.type != "cuda":
            publish(te[0], ar[0])
            return
        te_h, ar_h = te.to("cpu", non_blocking=False), ar.to("cpu", non_blocking=False)     # CNN models: library path
        publish(te_h[0], ar_h[0])

    def _collect(self, tensors: dict) -> dict:
        """Device tensors -> numpy, through persistent pinned buffers, with a single synchronisation - the one host
        wait of a block.  self._collect_ev marks the copies (i.e. the end of the block's chain) in the stream."""
        if self.device.type != "cuda":
            return {k: v.detach().cpu().numpy().copy() for k, v in tensors.items()}
        if not hasattr(self, "_pinned"):
            self._pinned = {}
            self._collect_ev = torch.cuda.Event()
        out = {}
        for k, v in tensors.items():
            key = (k, tuple(v.shape), v.dtype)
            buf = self._pinned.get(key)
            if buf is None:
                buf = torch.empty(v.shape, dtype=v.dtype).pin_memory()
                self._pinned[key] = buf
            buf.copy_(v, non_blocking=True)
            out[k] = buf
        # bytes this block reads back from the device (weights, counters, contributor list, commitment||signature
        # rows; + 8 for the two error counts, which _launch_eval reads the same way)
        self.last_d2h_bytes = sum(b.numel() * b.element_size() for b in out.values()) + 8
        self._collect_ev.record()
        self._collect_ev.synchronize()
        return {k: b.numpy() for k, b in out.items()}

    def _finish_block(self, rep: BlockReport, phases, t0, host=None, updates=(), packed=None, then=None) -> BlockReport:
        cfg = self.cfg
        it = self.iteration
        cuda = self.device.type == "cuda"
        with self._phase("block"):
            if host is None:
                host = self._collect({"w": self.w64})
            nonempty = bool(updates) or packed is not None
            ts = (1 + it if nonempty else 0) if cfg.deterministic_time else -1
            if packed is not None:
                t_a = time.perf_counter()
                rep.block_hash = self.chain.append_packed(it, host["w"], packed[0], packed[1], packed[2], self.stake,
                                                          STAKE_UNIT, timestamp=ts)
                phases["h_append"] = (time.perf_counter() - t_a) * 1e3
            else:
                rep.block_hash = self.chain.add_block(it, host["w"], list(updates), self.stake, timestamp=ts)
            rep.empty = not nonempty
            # The next block's noiser lottery hangs on this hash: start every peer's VRF now, before anything else
            # (the kernel is ~0.4 ms of pure latency at the head of the next block's chain).
            if cuda and cfg.mode != "fedsys" and cfg.noising and cfg.num_noisers > 0 and cfg.epsilon > 0:
                self._vrf_early = (rep.block_hash, self.vrf.launch_dev(rep.block_hash))
        t_after = time.perf_counter()
        if then is not None:
            then(getattr(self, "_collect_ev", None) if cuda else None, rep)
            phases["host_after"] = (time.perf_counter() - t_after) * 1e3
        elif cfg.eval_every and (it % cfg.eval_every == 0):
            self._launch_eval(rep)          # empty blocks and FedSys rounds: nothing else is queued behind the block
        # error rates of older blocks have long been read back: publish them (keeps the "Train Error" lines in order)
        while len(self._eval_waiting) > (2 if self.pipelined else 0):
            r = self._eval_waiting.pop(0)
            if r._resolve is not None:
                r._resolve()
        rep.phase_ms = dict(phases)
        rep.phase_ms["total_host"] = (time.perf_counter() - t0) * 1e3
        self.reports.append(rep)
        self.iteration += 1
        return rep

    def join_streams(self):
        """Order the caller's current stream after EVERYTHING this engine has queued (chain, off-chain, approval, witness and
        VRF streams) - what a benchmark records its closing event behind."""
        if self.device.type != "cuda" or self._main is None:
            return
        cur = torch.cuda.current_stream(self.device)
        for st in (self._main, self._side, self._wit, self._agg, self.signer.stream if self.signer is not None else None,
                   getattr(self.vrf, "_dside", None)):
            if st is not None:
                cur.wait_stream(st)

    def flush_reports(self):
        """Wait for every outstanding error-rate read-back (end of a run, before the logs are parsed)."""
        while self._eval_waiting:
            r = self._eval_waiting.pop(0)
            if r._resolve is not None:
                r._resolve()

    # ------------------------------------------------------------------------------------------
    def run(self, iterations: int | None = None):
        """Main loop (main.go:1059-1187): runs until iterationCount > MAX_ITERATIONS, i.e. max+1 blocks
        after genesis (102 logged blocks with the genesis one)."""
        total = (self.cfg.max_iterations + 1) if iterations is None else iterations
        for _ in range(total):
            self.step()
        self.flush_reports()
        return self.reports

    def print_chain(self) -> str:
        return self.chain.dump()

    def mint_empty_block(self) -> BlockReport:
        """Block deadline (startBlockDeadlineTimer, main.go:2326-2355): a peer that saw no block for this height mints
        the deterministic empty one (block.go:32-38) and moves on.  Weights and stake are unchanged."""
        rep = BlockReport(self.iteration, b"", True, [], [], 0, 0, 0, 0)
        self._phases = {}
        self._pre = None
        self._ctx_next = None
        return self._finish_block(rep, self._phases, time.perf_counter())

    def close(self):
        """Release what the engine holds outside the Python heap (peer-memory mappings of the fused path)."""
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)
        if self.comm is not None and hasattr(self.comm, "close"):
            self.comm.close()
        # the window tables are the bulk of the footprint (up to half the device): drop them now rather than when the
        # garbage collector gets to the engine, so that the next engine of a campaign finds the memory free
        self._pre = self._ctx_next = self._vrf_early = None
        self._ns = None
        self._ns_ok = False
        self.commit = None
        if self.device.type == "cuda":
            import gc

            gc.unfreeze()          # what _native() froze goes back under the collector, so this engine's cycles can be reclaimed
            gc.collect()
            torch.cuda.empty_cache()

    # rejoin: a restarted peer adopts the longest chain it is offered (main.go:1000-1013)
    def adopt_chain(self, serialized: bytes) -> bool:
        other = Blockchain.deserialize(serialized)
        if len(other) > len(self.chain) and other.verify():
            self._install_chain(other)
            return True
        return False

    def _install_chain(self, other: Blockchain):
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)
        self._pre = None
        self._ctx_next = None
        self._vrf_early = None
        self.chain = other
        self.iteration = other.latest_iteration + 1
        w = torch.from_numpy(other.weights()).to(self.device)
        self.w64.copy_(w)
        self.w32.copy_(w.float())
        info = other.block_info(-1)
        if info["stake"]:
            for k, v in info["stake"].items():
                self.stake[k] = v

    def offer_block(self, serialized_block: bytes) -> int:
        """RegisterBlock from a peer that was on the other side of a partition (main.go:1238-1330): a block for the
        height this engine already holds replaces it iff the held one is empty and the offered one is not and both
        extend the same parent (evaluateBlockQuality, honest.go:631-653).  Returns accept_block's code: 2 = replaced
        (weights, stake and iteration follow the new tip), 1 = appended, 0 = ignored, < 0 = rejected."""
        rc = self.chain.accept_block(serialized_block)
        if rc in (1, 2):
            self._install_chain(self.chain)
        return rc


def cp_order() -> int:
    from ..oracle.bn256 import ORDER

    return ORDER


class _Phase:
    """Phase timer: CUDA events when timing is enabled on a GPU engine, wall clock otherwise."""

    def __init__(self, eng: BiscottiEngine, name: str):
        self.eng, self.name = eng, name

    def __enter__(self):
        e = self.eng
        if e._timing and e.device.type == "cuda":
            self.ev0 = torch.cuda.Event(enable_timing=True)
            self.ev1 = torch.cuda.Event(enable_timing=True)
            self.ev0.record()
        else:
            self.t0 = time.perf_counter()
        return self

    def __exit__(self, *a):
        e = self.eng
        if e._timing and e.device.type == "cuda":
            self.ev1.record()
            self.ev1.synchronize()
            ms = self.ev0.elapsed_time(self.ev1)
        else:
            ms = (time.perf_counter() - self.t0) * 1e3
        e._phases[self.name] = e._phases.get(self.name, 0.0) + ms
        return False
