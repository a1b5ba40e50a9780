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

There are no per-peer processes, sockets or wall-clock timeouts: peers are rows of device tensors,
message arrival order is an explicit seeded permutation (so "first NUM_SAMPLES/2 shares win" and
Krum's "first KRUM_UPDATETHRESH updates" are reproducible), and every rank replays the same control
flow from the same block hash, so the ledger is replicated without gossip.
"""
from __future__ import annotations

import contextlib
import hashlib
import math
import random
import struct
import time
from dataclasses import dataclass, field

import numpy as np
import torch

from ..config import BiscottiConfig, DEFAULT_STAKE, STAKE_UNIT, POLY_SIZE, NOISE_TABLE_ITERS
from ..crypto import primitives as cp
from ..crypto.schnorr_batch import SchnorrBatchSigner
from ..data.synthetic import make_data, poisoner_ids
from ..ledger.chain import Blockchain, UpdateRecord
from ..models.registry import Model
from ..ops import ml, krum as krum_ops, secagg
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
    test_error: float | None = None
    attack_rate: float | None = None
    loss: float | None = None
    bad_shares: int = 0
    phase_ms: dict = field(default_factory=dict)
    unmasked_updates: int = 0


class BiscottiEngine:
    def __init__(self, cfg: BiscottiConfig, device="cpu", rank: int = 0, world: int = 1, comm=None, log_stream=None,
                 external_inputs: bool = False):
        self.cfg = cfg
        self.device = torch.device(device)
        self.rank, self.world, self.comm = rank, world, comm
        n = cfg.num_nodes
        per = (n + world - 1) // world
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
        self._pre = None          # prefetched (Q, chunkC, fullC) of the next iteration, see _prefetch_next
        self._next_inputs, self._fed_inputs = None, False
        self._eval_pending = None
        self._vrf_early = None
        self._main = self._wstream = self._wevent = None     # streams of step() / _witness_side
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
                                           window=cfg.msm_window)
        self.vrf = cp.BatchedVRF([k.vrf_noise for k in self.keys], self.device)
        self.signer = SchnorrBatchSigner(self.device) if (cfg.verification and cfg.sign_approvals) else None
        self.reports: list[BlockReport] = []
        self._rng = np.random.RandomState(cfg.seed * 65537 % (2**31))
        self._timing = False

    # ------------------------------------------------------------------------------------------
    def _random_tau(self) -> int:
        h = hashlib.sha256(f"bsc-tau-{self.cfg.seed}".encode()).digest()
        return int.from_bytes(h, "big") % cp_order()

    def _arrival_order(self, ids: list[int], tag: str) -> list[int]:
        """Deterministic stand-in for network arrival order, identical on every rank."""
        h = hashlib.sha256(f"{self.cfg.seed}-{self.iteration}-{tag}".encode()).digest()
        out = list(ids)
        random.Random(int.from_bytes(h[:8], "little")).shuffle(out)   # Mersenne twister: same order on every rank
        return out

    def _phase(self, name: str):
        return _Phase(self, name)

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
        """Several small host lists -> device tensors with ONE staging copy.  specs: (values, dtype) pairs; each
        tensor starts on a 16-byte boundary of the shared buffer."""
        if self.device.type != "cuda":
            return [torch.tensor(v, dtype=dt) for v, dt in specs]
        if not hasattr(self, "_pack_stage"):
            self._pack_stage = [torch.empty(1 << 16, dtype=torch.uint8).pin_memory() for _ in range(8)]
            self._pack_i = 0
        buf = self._pack_stage[self._pack_i % len(self._pack_stage)]
        self._pack_i += 1
        offs, off = [], 0
        for v, dt in specs:
            nb = len(v) * torch.empty(0, dtype=dt).element_size()
            offs.append((off, nb))
            off = (off + nb + 15) // 16 * 16
        if off > buf.numel():
            return [self._idx(v, dt) for v, dt in specs]
        for (o, nb), (v, dt) in zip(offs, specs):
            if nb:
                buf[o: o + nb].view(dt).copy_(torch.as_tensor(v, dtype=dt))
        dev_buf = buf[: max(off, 16)].to(self.device, non_blocking=True)
        return [dev_buf[o: o + nb].view(dt) for (o, nb), (v, dt) in zip(offs, specs)]

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
            from ..ops._call import lib, ptr, stream, check

            check(lib().bsc_sample_minibatch(self.cfg.seed & (2**64 - 1), it, ptr(self.data.peer_ids), self.P, Ns, B,
                                             ptr(self._mb_idx), stream(self.device)), "sample_minibatch")
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
    def _local_and_commit(self, inputs, it: int, timed: bool):
        """Local SGD step of every local peer from the current global weights, then quantisation and the
        polynomial commitments of the update (phases 2 and 3 of an iteration).  Returns (Q, chunkC, fullC)."""
        cfg, dev = self.cfg, self.device
        slot = it % NOISE_TABLE_ITERS
        phase = self._phase if timed else (lambda name: contextlib.nullcontext())
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
        Q = chunkC = fullC_dev = None
        if self.commit is not None and self.P:
            with phase("commit"):
                Q = secagg.quantize(self.U, self.d, cfg.precision)
                chunkC = self.commit.chunk_commitments(Q, affine=cfg.affine_wire)    # [P, nchunk, 16|24]
                # full-vector commitments stay Jacobian on the device; the rows that enter the block are
                # normalised on the host with one shared inversion when the block is assembled
                fullC_dev = self.commit.full_commitments(chunkC, jacobian=True).view(torch.uint8)   # [P, 96]
        return Q, chunkC, fullC_dev

    def _prefetch_next(self, it_next: int):
        """Queue the next iteration's local step and commitment MSM behind this iteration's aggregation kernel.
        They read only the new global weights (already updated in stream order) and the resident shards, so the
        device works on them while the host assembles, signs and hashes the current block.  The reference has no
        such overlap (phases are strictly sequential); the values are identical either way."""
        cfg = self.cfg
        if not cfg.prefetch or not cfg.secure_agg or cfg.mode == "fedsys":
            return   # the plain path still reads U when the block is assembled
        nxt = self._next_inputs
        if nxt is None and (self._fed_inputs or self.external_inputs):
            return   # minibatches come from the caller and the next one has not been handed over
        if nxt is not None and len(nxt) == 3:      # (X, y, event): the H2D copy was issued on another stream
            torch.cuda.current_stream(self.device).wait_event(nxt[2])
            nxt = (nxt[0], nxt[1])
        Q, chunkC, fullC = self._local_and_commit(nxt, it_next, timed=False)
        self._pre = {"it": it_next, "Q": Q, "chunkC": chunkC, "fullC": fullC,
                     "src": None if nxt is None else nxt[0].data_ptr()}

    def step(self, inputs=None, next_inputs=None) -> BlockReport:
        """Run one block iteration.  `inputs` = (X [P,B,F], y [P,B]) device tensors when the caller feeds
        minibatches (bench end-to-end path); otherwise minibatches are drawn from the resident shards.
        `next_inputs` = the following call's `inputs` (optionally with a CUDA event that marks their H2D copy
        complete), if the caller already has them: a double-buffered input pipeline lets the engine start the
        next local step while this block is being assembled.

        On a GPU the iteration runs on the engine's own high-priority stream (ordered after the caller's stream on
        entry, and the caller's stream after it on exit), so that the witness MSM — queued on a default-priority
        stream, see _witness_side — yields SM slots to the kernels the block is waiting for."""
        if self.device.type != "cuda" or self.cfg.mode == "fedsys" or not self.cfg.overlap_witness:
            return self._step(inputs, next_inputs)
        if self._main is None:
            self._main = torch.cuda.Stream(self.device, priority=-1)
        cur = torch.cuda.current_stream(self.device)
        self._main.wait_stream(cur)
        prev_w = self._wevent
        with torch.cuda.stream(self._main):
            rep = self._step(inputs, next_inputs)
        cur.wait_stream(self._main)
        if prev_w is not None:
            # the previous block's witnesses had this whole iteration to finish; waiting for them here keeps every
            # iteration's device timeline (events on the caller's stream) accountable for one block of witness work
            cur.wait_event(prev_w)
        return rep

    def _witness_side(self):
        """Stream for the KZG witness MSM and the miners' homomorphic aggregation.  Nothing that goes into the block
        depends on their result (the block carries the contributors' commitments, the reconstructed weights and the
        approvals), so they run beside share aggregation, reconstruction, block assembly and the next block's VRF
        instead of in front of them.  Instrumented runs keep them in line so that phase times stay attributable."""
        if self.device.type != "cuda" or self._main is None or self._timing:
            return None
        if self._wstream is None:
            self._wstream = torch.cuda.Stream(self.device, priority=0)
        return self._wstream

    def _step(self, inputs=None, next_inputs=None) -> BlockReport:
        cfg = self.cfg
        self._next_inputs, self._fed_inputs = next_inputs, inputs is not None
        if cfg.mode == "fedsys":
            from .fedsys import fedsys_step

            return fedsys_step(self, inputs)
        t0 = time.perf_counter()
        it = self.iteration
        n = cfg.num_nodes
        dev = self.device
        phases = {}
        self._phases = phases
        # ---- fault injection (main.go:1117-1120): a failing peer sits this block out and rejoins
        if cfg.fail_prob > 0:
            for p in range(1, n):
                self.alive[p] = self._rng.rand() >= cfg.fail_prob
        # ---- scheduled network partitions (DistSys/blockNode.sh): the peer is cut off for a range of blocks; it
        # holds no role and contributes nothing meanwhile, and is back (with the chain everybody shares) afterwards
        if cfg.partitions:
            cut = {int(p) for p, a, b in cfg.partitions if int(a) <= it <= int(b)}
            was = getattr(self, "_cut", set())
            for p in was - cut:
                self.alive[p] = True
            for p in cut:
                self.alive[p] = False
            self._cut = cut
        # ---- 1. roles
        verifiers, miners, prev_hash = self.select_roles()
        committee = set(verifiers) | set(miners)
        workers = [p for p in range(n) if p not in committee and self.alive[p]]
        slot = it % NOISE_TABLE_ITERS
        rep = BlockReport(it, b"", True, verifiers, miners, len(workers), 0, 0, 0)
        if not verifiers or not miners or not workers:
            return self._finish_block(rep, [], None, phases, t0)   # main.go:1696-1707, 1934-1944

        # VRF evaluation of every peer on the previous block hash runs on a side stream, overlapped with
        # the local step and the commitment MSM; the noise phase collects it.
        use_noise = cfg.noising and cfg.num_noisers > 0 and cfg.epsilon > 0
        early, self._vrf_early = self._vrf_early, None
        if use_noise and early is not None and early[0] == prev_hash:
            vrf_handle = early[1]          # launched the moment the previous block was hashed (_finish_block)
        else:
            vrf_handle = self.vrf.launch(prev_hash) if use_noise else None

        # ---- 2/3. local step + commitments of my peers.  Neither depends on the roles or on the block hash, so
        # the previous step may already have queued them behind its aggregation kernel (see _prefetch_next).
        pre, self._pre = self._pre, None
        if pre is not None and pre["it"] == it and pre["src"] == (None if inputs is None else inputs[0].data_ptr()):
            Q, chunkC, fullC_dev = pre["Q"], pre["chunkC"], pre["fullC"]
        else:
            Q, chunkC, fullC_dev = self._local_and_commit(inputs, it, timed=True)

        # host-side bookkeeping that does not need the VRF outputs goes first: the noise phase below waits for them
        arrival = self._arrival_order(workers, "verify")
        krum_set = sorted(arrival[: cfg.krum_thresh])          # krum.go:296-314: first thresh, sorted by id
        rep.n_krum = len(krum_set)
        # Approvals (krum.go:338-358): the signature halves T = v*G do not depend on the Krum verdict or on the message,
        # only on (block, update, verifier), so they are started for every candidate update now — the host would
        # otherwise sit idle waiting for the VRF — and only those of the updates that reach the block are completed.
        signers = [v for v in verifiers if self.alive[v]]
        sig_handle = None
        if self.signer is not None and self.commit is not None and signers and krum_set:
            nk, ns = len(krum_set), len(signers)
            if not hasattr(self, "_sk_mat"):     # all peers' signing keys as one [n, 32] array (pKeyG1.json holds them all)
                self._sk_mat = np.frombuffer(b"".join(k.schnorr_secret for k in self.keys), dtype=np.uint8).reshape(n, 32)
            tags = np.empty(nk * ns, dtype=[("it", "<i8"), ("p", "<i4"), ("v", "<i4")])     # = struct.pack("<qii", it, p, v)
            tags["it"], tags["p"], tags["v"] = it, np.repeat(krum_set, ns), np.tile(signers, nk)
            sig_handle = self.signer.launch(np.tile(self._sk_mat[signers], (nk, 1)), tags.view(np.uint8).reshape(nk * ns, 16))
        # ---- 4. noisers + noise
        with self._phase("noise"):
            nn_cols = max(1, cfg.num_noisers)
            tab_np = np.full((max(1, self.P), nn_cols), -1, dtype=np.int32)
            noisers_of = {}
            if use_noise:
                stake = [s if a else 0 for s, a in zip(self.stake, self.alive)]
                is_worker = np.zeros(n, dtype=np.uint8)
                is_worker[workers] = 1
                full_tab = self.vrf.noiser_table(vrf_handle, stake, is_worker, cfg.num_noisers)   # [n, nn]
                if self.P:
                    lo_id = self.local_ids[0]
                    tab_np = full_tab[lo_id: lo_id + self.P]
                if self.colluder_ids:
                    noisers_of = {p: [int(q) for q in full_tab[p] if q >= 0] for p in workers}
            scale = cfg.noise_scale if use_noise else 0.0
            # Single-rank tcgen05 Krum: the kernel writes the candidates' bf16 planes and norms compacted and in Krum
            # order (row = position in krum_set), so the Gram kernel needs no gather; the fp32 noised rows are only
            # materialised for the paths that read them (RONI, SIMT Gram, NCCL gathers).
            self._compact = (dev.type == "cuda" and self.comm is None and cfg.verification and cfg.defense != "RONI"
                             and cfg.gram_impl in ("auto", "tcgen05") and len(krum_set) >= 3)
            kpos = {p: i for i, p in enumerate(krum_set)}
            rows = [kpos.get(p, -1) for p in self.local_ids] if self._compact else []
            # noiser table and row map travel in one staging copy
            noiser_tab, dst_row = self._idx_pack((tab_np.reshape(-1), torch.int32), (rows, torch.int32))
            noiser_tab = noiser_tab.view(-1, nn_cols)
            if not self._compact:
                dst_row = None
            Un, hi, lo, norms = ml.noise_mix(self.U, noiser_tab, cfg.seed, slot, scale, self.d,
                                             ldh=self.ldh, out_hi=self.comm.hi if self.fused else None,
                                             out_lo=self.comm.lo if self.fused else None, dst_row=dst_row,
                                             want_un=not self._compact)

        # ---- 5. verification
        accepted: list[int] = []
        with self._phase("verify"):
            if not cfg.verification:
                accepted = list(krum_set)
            elif cfg.defense == "RONI":
                accepted = self._roni(krum_set, Un)
            else:
                accepted = self._krum(krum_set, Un, hi, lo, norms)
        rep.n_accepted = len(accepted)
        # collusion accounting (main.go:1026-1057, 1138-1144)
        if self.colluder_ids:
            vcol = any(v in self.colluder_ids for v in verifiers)
            rep.unmasked_updates = sum(1 for p in krum_set if vcol and noisers_of.get(p) and
                                       all(q in self.colluder_ids for q in noisers_of[p]))

        # ---- 6. approvals: every reachable verifier signs the commitment of each accepted update; a worker
        # proceeds when it holds >= floor(nv/2) signatures (main.go:1686).  All verifiers run the same Krum on
        # the same set here, so the signature count is the number of live verifiers.
        approved = accepted if (signers and len(signers) >= len(verifiers) // 2) else []

        # ---- 7/8. shares, witnesses, aggregation, reconstruction, SGD step
        share_order = [p for p in self._arrival_order(approved, "shares")]
        trigger = max(1, cfg.num_samples // 2) if cfg.share_trigger == "half" else len(share_order)
        contributors = sorted(share_order[:trigger])            # main.go:360 + leader intersection
        alive_miners = [m for m in miners if self.alive[m]]
        shares_avail = self._available_shares(miners, alive_miners)
        enough = len(contributors) >= 2 and sum(shares_avail) >= POLY_SIZE and len(alive_miners) > 0
        if not enough:
            return self._finish_block(rep, [], None, phases, t0)   # main.go:2099-2117
        rep.n_contributors = len(contributors)
        local_contrib = [self.local_index[p] for p in contributors if p in self.local_index]
        appr_rows = [self.local_index[p] for p in approved if p in self.local_index]
        pos = {r: i for i, r in enumerate(appr_rows)}          # positions of contributors inside the approved list
        xs = [s for s, a in enumerate(shares_avail) if a][:POLY_SIZE]
        sel, appr_local, csel, xs_t, av_t = self._idx_pack(
            (local_contrib, torch.int32), (appr_rows, torch.int32), ([pos[r] for r in local_contrib], torch.int32),
            (xs, torch.int32), (shares_avail, torch.uint8))
        agg_box = [None]
        launch_witnesses = None
        if cfg.secure_agg:
            if self.commit is not None and self.P:
                # Every approved worker generates shares + KZG witnesses for the miners (kyber.go:456-482) and the
                # miners add the contributors' witnesses and chunk commitments homomorphically (kyber.go:244-287).
                # Nothing that enters the block depends on them, so they are *queued after* the share aggregation,
                # the reconstruction and the block's device->host collection (see after_copies below): the device
                # works on them while the host assembles, signs and hashes the block.

                def launch_witnesses():
                    side = self._witness_side()
                    if side is None:
                        with self._phase("witness"):
                            W = self.commit.witnesses(Q, appr_local, affine=cfg.affine_wire) if appr_rows else None
                        with self._phase("miner_aggregate"):
                            if W is not None and local_contrib:
                                aggW = self.commit.aggregate(W, csel, affine=cfg.affine_wire)        # [nchunk, S, 16|24]
                                aggC = self.commit.aggregate(chunkC, sel, affine=cfg.affine_wire)    # [nchunk, 16|24]
                                agg_box[0] = (aggC, aggW)
                    elif appr_rows:
                        ready = torch.cuda.Event()
                        ready.record()                     # Q, chunkC and the index tensors are complete at this point
                        side.wait_event(ready)
                        with torch.cuda.stream(side):
                            W = self.commit.witnesses(Q, appr_local, affine=cfg.affine_wire)
                            if local_contrib:
                                aggW = self.commit.aggregate(W, csel, affine=cfg.affine_wire)
                                aggC = self.commit.aggregate(chunkC, sel, affine=cfg.affine_wire)
                                agg_box[0] = (aggC, aggW)
                            self._wevent = torch.cuda.Event()
                            self._wevent.record()
                        for t in (Q, chunkC, appr_local, csel, sel):  # allocated on the main stream, read on `side`
                            t.record_stream(side)
            with self._phase("secagg"):
                if self.fused:
                    # reduce-scatter + reconstruct + SGD step + all-gather in one kernel over peer memory
                    if sel.numel():
                        secagg.share_sum(self.U, sel, self.d, self.S, cfg.precision, out=self.comm.ypart)
                    else:
                        self.comm.ypart.zero_()
                    bad = self.comm.secagg_fused(xs, shares_avail, cfg.precision, it + 1)
                else:
                    Ysum = secagg.share_sum(self.U, sel, self.d, self.S, cfg.precision) if sel.numel() else \
                        torch.zeros(cfg.nchunk, self.S, dtype=torch.int64, device=dev)
                    if self.comm is not None:
                        Ysum = self.comm.all_reduce_i64(Ysum)
                    self._last_ysum = Ysum
                    _, bad = secagg.reconstruct_step(Ysum, xs_t, av_t, self.d, self.w64, self.w32, cfg.precision)
        else:
            with self._phase("secagg"):
                if self.comm is not None and self.world > 1:
                    rows = self.comm.gather_rows(contributors, self.U, self.local_index, self.owner)
                    rsel = torch.arange(len(contributors), dtype=torch.int32, device=dev)
                    secagg.plain_sum_step(rows, rsel, self.d, self.w64, self.w32)
                else:
                    secagg.plain_sum_step(self.U, sel, self.d, self.w64, self.w32)
                bad = torch.zeros(2, dtype=torch.int32, device=dev)

        # ---- 9. block assembly (host).  Evaluation of the new weights is launched first so that one
        # device->host collection (weights, commitments, share-check counters, error rates) and one stream
        # synchronisation serve the whole tail of the block.
        if self._timing and launch_witnesses is not None:
            launch_witnesses()            # instrumented runs: in line, so the phases do not nest
            launch_witnesses = None
        with self._phase("block"):
            updates = []
            commits = {}
            want = {"w": self.w64, "bad": bad}
            do_eval = bool(cfg.eval_every) and (it % cfg.eval_every == 0)
            gathered = False
            if self.commit is not None and self.P:
                if self.comm is not None and hasattr(self.comm, "all_gather_rows_u8"):
                    want["cb"] = self.comm.all_gather_rows_u8(fullC_dev)             # [world*per, 96], row = peer id
                    gathered = True
                else:
                    want["cb"] = fullC_dev
            if sig_handle is not None and sig_handle["T_dev"] is not None:
                torch.cuda.current_stream(dev).wait_event(sig_handle["event"])
                want["sigT"] = sig_handle["T_dev"]
            # Evaluation of the new weights is not part of the block: it is queued behind the collection (with the
            # next iteration's prefetch) and read back after the block has been assembled and hashed.
            def after_copies():
                if launch_witnesses is not None:
                    launch_witnesses()
                if do_eval:
                    self._launch_eval()
                self._prefetch_next(it + 1)

            host = self._collect(want, before_sync=after_copies)
            if "cb" in host:
                cb = host["cb"]
                cbj = cb.view(np.int32).reshape(cb.shape[0], 24)
                if gathered:
                    for p, b in zip(contributors, CommitmentEngine.jac_rows_to_bytes(cbj, list(contributors))):
                        commits[p] = b
                else:
                    mine = [p for p in contributors if p in self.local_index]
                    for p, b in zip(mine, CommitmentEngine.jac_rows_to_bytes(cbj, [self.local_index[p] for p in mine])):
                        commits[p] = b
                    if self.comm is not None:
                        commits = self.comm.merge_dicts(commits)
            all_sigs = None
            if sig_handle is not None and all(len(commits.get(p, b"")) == 64 for p in contributors):
                msgs = [commits[p] for p in contributors for _ in signers]
                kpos = {p: i for i, p in enumerate(krum_set)}
                rows = [kpos[p] * len(signers) + j for p in contributors for j in range(len(signers))]
                all_sigs = self.signer.finish(self.signer.select(sig_handle, rows, host.get("sigT")), msgs)
            nv = len(signers)
            fast = None
            if (cfg.secure_agg and contributors and all(len(commits.get(p, b"")) == 64 for p in contributors)
                    and (all_sigs is not None or sig_handle is None)):
                # the common block: (source, commitment, approvals) per update, handed to the ledger as three arrays
                fast = (contributors, b"".join(commits[p] for p in contributors),
                        b"".join(all_sigs) if all_sigs else b"", nv if all_sigs else 0)
                for p in contributors:
                    self.stake[p] += STAKE_UNIT
            for ci, p in enumerate(contributors if fast is None else ()):
                c = commits.get(p, b"")
                sigs = all_sigs[ci * nv: (ci + 1) * nv] if all_sigs is not None else []
                rec = UpdateRecord(source=p, iteration=it, commitment=c, accepted=True, signatures=sigs)
                if not cfg.secure_agg and p in self.local_index and self.world == 1:
                    rec.delta = self.U[self.local_index[p], : self.d].double().cpu().numpy()
                updates.append(rec)
                self.stake[p] += STAKE_UNIT
            if not cfg.secure_agg:
                for p in krum_set:
                    if p not in accepted:
                        self.stake[p] -= STAKE_UNIT                      # honest.go:362-370 (plain path only)
            rep.bad_shares = int(host["bad"].sum())
        return self._finish_block(rep, updates, agg_box[0], phases, t0, host, fast)

    # ------------------------------------------------------------------------------------------
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

    def _krum(self, krum_set, Un, hi, lo, norms):
        cfg = self.cfg
        nk = len(krum_set)
        if nk < 3:
            return list(krum_set)
        clip = int(math.floor(cfg.krum_adversary_frac * nk))
        if self.fused:
            _, acc = self.comm.krum_fused(krum_set, clip, self.iteration + 1)
        elif getattr(self, "_compact", False):
            _, acc, _ = krum_ops.krum(hi[:nk], self.d, clip, hi[:nk], lo[:nk], norms[:nk], impl=cfg.gram_impl)
        else:
            ix = self._idx([self.local_index[p] for p in krum_set], torch.int64) if self.comm is None else None
            Uk = self._gather_rows(krum_set, Un, ix)
            hik = self._gather_rows(krum_set, hi, ix) if hi is not None else None
            lok = self._gather_rows(krum_set, lo, ix) if lo is not None else None
            nrm = self._gather_rows(krum_set, norms.unsqueeze(1), ix).squeeze(1) if norms is not None else None
            _, acc, _ = krum_ops.krum(Uk, self.d, clip, hik, lok, nrm, impl=cfg.gram_impl)
        mask = acc.cpu().tolist()
        return [p for p, a in zip(krum_set, mask) if a]

    def _roni(self, krum_set, Un):
        """RONI (main.go:191-233, honest.go:598-629): reject when validation error rises by > threshold."""
        Uk = self._gather_rows(krum_set, Un)
        base_err = self.model.eval_errors(self.w32.unsqueeze(0), None, self.data.X_test, self.data.y_test)[0]
        errs = self.model.eval_errors(Uk, self.w32, self.data.X_test, self.data.y_test)
        ok = ((errs - base_err) <= self.cfg.roni_threshold).cpu().tolist()
        return [p for p, a in zip(krum_set, ok) if a]

    def evaluate(self):
        w = self.w32.unsqueeze(0)
        te = self.model.eval_errors(w, None, self.data.X_test, self.data.y_test)
        ar = self.model.eval_errors(w, None, self.data.X_attack, self.data.y_attack)
        return te, ar

    def verify_last_aggregate(self, samples=((0, 0), (0, 13), (1, 20))):
        """What a miner can check about the last block's aggregation (kyber.go:244-287 + verifySecret): the sum of
        the contributors' witnesses opens the sum of their chunk commitments to the sum of their shares.  Waits
        for the witness stream.  Single-rank engines only (each rank aggregates its own contributors).  Returns
        None when there is nothing to check."""
        if self.last_aggregate is None or self.world != 1 or getattr(self, "_last_ysum", None) is None:
            return None
        if self._wevent is not None:
            self._wevent.synchronize()
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

    def _launch_eval(self):
        """Queue the error-rate kernels and their read-back; _finish_block picks the numbers up."""
        if self.device.type == "cuda" and self.model.kind in ("softmax", "logistic"):
            # both error counts into one persistent int32[2], one read-back; the rates are formed on the host
            if not hasattr(self, "_ev_cnt"):
                self._ev_cnt = torch.zeros(2, dtype=torch.int32, device=self.device)
                self._ev_cnt_pinned = torch.zeros(2, dtype=torch.int32).pin_memory()
                self._eval_ev = torch.cuda.Event()
            self._ev_cnt.zero_()
            w = self.w32.unsqueeze(0)
            d = self.data
            if self.model.kind == "softmax":
                C = self.cfg.spec.n_classes
                ml.eval_softmax_errors(w, None, d.X_test, d.y_test, C, wrong_out=self._ev_cnt[0:1])
                ml.eval_softmax_errors(w, None, d.X_attack, d.y_attack, C, wrong_out=self._ev_cnt[1:2])
            else:
                ml.eval_logistic_errors(w, None, d.X_test, d.y_test, wrong_out=self._ev_cnt[0:1])
                ml.eval_logistic_errors(w, None, d.X_attack, d.y_attack, wrong_out=self._ev_cnt[1:2])
            self._ev_cnt_pinned.copy_(self._ev_cnt, non_blocking=True)
            self._eval_ev.record()
            self._eval_pending = ("counts", self._ev_cnt_pinned, self._eval_ev, d.X_test.shape[0], d.X_attack.shape[0])
            return
        te, ar = self.evaluate()
        if self.device.type != "cuda":
            self._eval_pending = (te.detach().cpu(), ar.detach().cpu(), None)
            return
        if not hasattr(self, "_eval_pinned"):
            self._eval_pinned = (torch.empty_like(te, device="cpu").pin_memory(), torch.empty_like(ar, device="cpu").pin_memory())
            self._eval_ev = torch.cuda.Event()
        self._eval_pinned[0].copy_(te, non_blocking=True)
        self._eval_pinned[1].copy_(ar, non_blocking=True)
        self._eval_ev.record()
        self._eval_pending = (self._eval_pinned[0], self._eval_pinned[1], self._eval_ev)

    def _collect(self, tensors: dict, before_sync=None) -> dict:
        """Device tensors -> numpy, through persistent pinned buffers, with a single synchronisation.  `before_sync`
        runs after the copies are queued and before the host waits for them: work it launches sits behind the
        copies in the stream and is not waited for."""
        if self.device.type != "cuda":
            out = {k: v.detach().cpu().numpy().copy() for k, v in tensors.items()}
            if before_sync is not None:
                before_sync()
            return out
        if not hasattr(self, "_pinned"):
            self._pinned = {}
        out = {}
        for k, v in tensors.items():
            key = (k, tuple(v.shape), v.dtype)
            buf = self._pinned.get(key)
            if buf is None:
                buf = torch.empty(v.shape, dtype=v.dtype).pin_memory()
                self._pinned[key] = buf
            buf.copy_(v, non_blocking=True)
            out[k] = buf
        # bytes this block reads back from the device (weights, commitments, share-check counters, signature
        # halves; + 8 for the two error rates, which _launch_eval reads the same way)
        self.last_d2h_bytes = sum(b.numel() * b.element_size() for b in out.values()) + 8
        if before_sync is None:
            torch.cuda.current_stream(self.device).synchronize()
        else:
            if not hasattr(self, "_collect_ev"):
                self._collect_ev = torch.cuda.Event()
            self._collect_ev.record()
            before_sync()
            self._collect_ev.synchronize()
        return {k: b.numpy() for k, b in out.items()}

    def _finish_block(self, rep: BlockReport, updates, agg_commit, phases, t0, host=None, fast=None) -> BlockReport:
        cfg = self.cfg
        it = self.iteration
        with self._phase("block"):
            if host is None:
                want = {"w": self.w64}
                if cfg.eval_every and (it % cfg.eval_every == 0):
                    want["te"], want["ar"] = self.evaluate()
                host = self._collect(want)
            nonempty = bool(updates) or fast is not None
            ts = (1 + it if nonempty else 0) if cfg.deterministic_time else -1
            if fast is not None:
                rep.block_hash = self.chain.add_commit_block(it, host["w"], fast[0], fast[1], fast[2], fast[3], self.stake,
                                                             timestamp=ts)
            else:
                rep.block_hash = self.chain.add_block(it, host["w"], updates, self.stake, timestamp=ts)
            rep.empty = not nonempty
            # The next block's noiser lottery hangs on this hash: start every peer's VRF now rather than after the
            # bookkeeping below and the caller's loop overhead (the kernel is ~0.4 ms of pure latency).
            if self.device.type == "cuda" and cfg.mode != "fedsys" and cfg.noising and cfg.num_noisers > 0 and cfg.epsilon > 0:
                self._vrf_early = (rep.block_hash, self.vrf.launch(rep.block_hash))
        pend, self._eval_pending = self._eval_pending, None
        if pend is not None and isinstance(pend[0], str):      # ("counts", pinned int32[2], event, n_test, n_attack)
            pend[2].synchronize()
            cnt = pend[1].numpy()
            host = dict(host, te=[cnt[0] / max(1, pend[3])], ar=[cnt[1] / max(1, pend[4])])
        elif pend is not None:
            if pend[2] is not None:
                pend[2].synchronize()
            host = dict(host, te=pend[0].numpy(), ar=pend[1].numpy())
        if "te" in host:
            rep.test_error, rep.attack_rate = float(host["te"][0]), float(host["ar"][0])
            self.log.train_error(it, rep.test_error)
            self.log.attack_rate(it, rep.attack_rate)
        rep.phase_ms = dict(phases)
        rep.phase_ms["total_host"] = (time.perf_counter() - t0) * 1e3
        self.last_aggregate = agg_commit
        self.reports.append(rep)
        self.iteration += 1
        return rep

    # ------------------------------------------------------------------------------------------
    def run(self, iterations: int | None = None):
        """Main loop (main.go:1059-1187): runs until iterationCount > MAX_ITERATIONS, i.e. max+1 blocks
        after genesis (102 logged blocks with the genesis one)."""
        total = (self.cfg.max_iterations + 1) if iterations is None else iterations
        for _ in range(total):
            self.step()
        return self.reports

    def print_chain(self) -> str:
        return self.chain.dump()

    # rejoin: a restarted peer adopts the longest chain it is offered (main.go:1000-1013)
    def adopt_chain(self, serialized: bytes) -> bool:
        other = Blockchain.deserialize(serialized)
        if len(other) > len(self.chain) and other.verify():
            self._pre = None
            self.chain = other
            self.iteration = other.latest_iteration + 1
            w = torch.from_numpy(other.weights()).to(self.device)
            self.w64.copy_(w)
            self.w32.copy_(w.float())
            info = other.block_info(-1)
            if info["stake"]:
                for k, v in info["stake"].items():
                    self.stake[k] = v
            return True
        return False


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
