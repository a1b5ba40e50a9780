"""The standard block as native command lists (ops/cmdlist.py -> csrc/exec.cu).

protocol/engine.py describes a block in Python: every kernel is a ctypes call, every dependency a stream or event call.
That is the reference implementation of the schedule - it handles every configuration - but ~45 Python-issued launches
cost the host about a millisecond per block, more than the device needs for the block's dependency chain.  For the
configuration the benchmarks and the campaigns run (softmax / logistic peers, Multi-Krum, secure aggregation with
commitments, approvals, DP noise; one GPU or the fused peer-memory path) the same launches, in the same order, on the same
three streams, are recorded ONCE here as command lists over persistent buffers and replayed with one native call per
segment; per block only a handful of argument words are patched (iteration, candidate count, epoch, ...).  The ledger is
byte-identical to the Python schedule's (tests/test_chain_kernels_gpu.py).

Segments, per block parity (buffers that two blocks in flight could both touch are double-buffered by parity):
    LOCAL   side stream : minibatch draw, local step -> U                                     records u_ready
    EVAL    side stream : error counts of the new weights, read-back                          records eval_done
    FRONT   chain stream: waits u_ready / side_done / VRF; noisers, noise, Gram, Krum, plan, share sums
    BACK    chain stream: waits eval_done; reconstruct + SGD step (or the fused reduce-scatter kernel); waits the
                          approvals; gathers the commitment||signature rows; the block's device->host copies
    COMMIT  side stream : quantise, commitment MSM, column sums, normalise, wire bytes        records c_ready
    SIGN    sign stream : nonces, T = v*G, normalise, bytes; waits c_ready; challenge + response -> pack rows
    WITNESS wit stream  : witness MSM, miners' homomorphic aggregation                        records side_done
"""
from __future__ import annotations

import ctypes

import numpy as np
import torch

from ..config import NOISE_TABLE_ITERS, POLY_SIZE
from ..ops import gram_tc as gram_ops
from ..ops.cmdlist import CmdList, NativeEvent
from ..ops.msm import AFF, JAC
from ..ops.secagg import recover_matrix, _RecoverMat

MAIN, SIDE, SIGN, WIT, AGG = 0, 1, 2, 3, 4


class NativeSchedule:
    @staticmethod
    def supported(eng) -> bool:
        cfg = eng.cfg
        return (eng.device.type == "cuda" and cfg.pipeline and cfg.native_schedule and cfg.mode == "biscotti"
                and eng.model.kind in ("softmax", "logistic") and cfg.secure_agg and eng.commit is not None
                and eng.commit.c_table is not None and not cfg.affine_wire and cfg.verification and cfg.defense == "KRUM"
                and cfg.gram_impl in ("auto", "tcgen05") and not eng.colluder_ids and not eng.external_inputs
                and (eng.comm is None or eng.fused) and eng.P > 0 and cfg.shard_size <= 1024
                and not (cfg.dp_in_model and eng.model.kind == "softmax" and False))

    def __init__(self, eng):
        self.eng = eng
        cfg, dev = eng.cfg, eng.device
        n, P, d = cfg.num_nodes, eng.P, eng.d
        self.n, self.P = n, P
        eng._ensure_bufs()
        i32 = dict(dtype=torch.int32, device=dev)
        streams = [eng._main, eng._side, eng.signer.stream if eng.signer is not None else eng._side, eng._wit, eng._agg]
        self.streams = streams
        # ---- events
        NB = self.NB = eng.NB                     # ring depth of the per-block buffers, events and command lists
        ev = lambda: [NativeEvent() for _ in range(NB)]   # noqa: E731
        self.u_ready, self.c_ready, self.plan_ready, self.sig_done, self.side_done = ev(), ev(), ev(), ev(), ev()
        self.w_done = NativeEvent()       # the block's SGD step has been applied (recorded right behind the reconstruction)
        self.agg_done = ev()
        self.eval_done, self.collect_ev = NativeEvent(), NativeEvent()
        self.eval_host = [NativeEvent() for _ in range(4)]
        self._vrf_placeholder = NativeEvent()
        # ---- the plan's fixed device layout (one pinned staging copy per block)
        self.nv = max(1, cfg.num_verifiers)
        off = 0

        def take(nbytes):
            nonlocal off
            o = off
            off = (off + nbytes + 15) & ~15
            return o

        self.o_stake, self.o_isw, self.o_dst = take(4 * n), take(n), take(4 * max(1, P))
        self.o_lrow, self.o_rank, self.o_ids = take(4 * n), take(4 * n), take(4 * n)
        self.o_sign, self.o_xs, self.o_av, self.o_slots, self.o_hash = take(4 * self.nv), take(40), take(64), take(4 * n), take(64)
        self.plan_bytes = off
        self.plan_pinned = [torch.zeros(off, dtype=torch.uint8).pin_memory() for _ in range(4)]
        self.plan_np = [b.numpy() for b in self.plan_pinned]
        self.plan_dev = [torch.zeros(off, dtype=torch.uint8, device=dev) for _ in range(NB)]
        self._plan_views = []
        for b in self.plan_np:
            def vw(o, cnt, dt, b=b):
                return b[o: o + cnt * np.dtype(dt).itemsize].view(dt)
            self._plan_views.append(dict(stake=vw(self.o_stake, n, np.int32), isw=vw(self.o_isw, n, np.uint8),
                                         dst=vw(self.o_dst, max(1, P), np.int32), lrow=vw(self.o_lrow, n, np.int32),
                                         rank=vw(self.o_rank, n, np.int32), ids=vw(self.o_ids, n, np.int32),
                                         sign=vw(self.o_sign, self.nv, np.int32), xs=vw(self.o_xs, 10, np.int32),
                                         av=vw(self.o_av, 64, np.uint8), slots=vw(self.o_slots, n, np.int32),
                                         hash=vw(self.o_hash, 64, np.uint8)))
        self._plan_ptrs = [{k_: v_.ctypes.data_as(ctypes.c_void_p) for k_, v_ in vw_.items()} for vw_ in self._plan_views]
        zi = lambda k_: np.zeros(k_, dtype=np.int32)   # noqa: E731
        sc = dict(stake=zi(n), alive=np.zeros(n, dtype=np.uint8), verifiers=zi(64), miners=zi(64), workers=zi(n), counts=zi(8))
        for k_ in list(sc):
            sc["p_" + k_] = sc[k_].ctypes.data_as(ctypes.c_void_p)
        self._scratch = sc
        self._pos = np.full(n, -1, dtype=np.int32)
        self._arange = np.arange(n, dtype=np.int32)
        self._local_ids = np.asarray(eng.local_ids, dtype=np.int64)
        self._local_row_of = np.full(n, -1, dtype=np.int32)
        self._local_row_of[self._local_ids] = np.arange(P, dtype=np.int32)
        self._slot_of = (np.asarray([eng.comm.slot_of(p) for p in range(n)], dtype=np.int32) if eng.fused
                         else np.zeros(n, dtype=np.int32))
        # ---- persistent intermediates
        ldh, ldu = eng.ldh, eng.ldu
        B = cfg.batch_size
        self.mb_idx = torch.zeros(P, B, **i32)
        self.loss = torch.zeros(P, dtype=torch.float32, device=dev)
        self.norms = torch.zeros(max(P, n), dtype=torch.float32, device=dev)
        self.scores = torch.zeros(n, dtype=torch.float32, device=dev)
        self.accept = torch.zeros(n, dtype=torch.uint8, device=dev)
        if not eng.fused:
            self.hi = torch.zeros(P, ldh, dtype=torch.bfloat16, device=dev)
            self.lo = torch.zeros(P, ldh, dtype=torch.bfloat16, device=dev)
            gmax = 0
            for nk in range(3, n + 1):
                ldg = (nk + 127) // 128 * 128
                gmax = max(gmax, gram_ops.ksplit_for(nk, ldh) * ldg * ldg)
            self.Gpart = torch.zeros(gmax, dtype=torch.float32, device=dev)
        C = eng.commit
        self.dpad, self.nchunk, self.S = C.dpad, C.nchunk, eng.S
        self.Q = [torch.zeros(P, C.dpad, dtype=torch.int64, device=dev) for _ in range(NB)]
        self.chunkC = [torch.zeros(P, C.nchunk, JAC, **i32) for _ in range(NB)]
        self.fullJ = torch.zeros(P, JAC, **i32)
        self.fullA = torch.zeros(P, AFF, **i32)
        self.wrows = min(P, cfg.krum_thresh)
        self.W = torch.zeros(self.wrows, C.nchunk * eng.S, JAC, **i32)
        self.aggQ = torch.zeros(1, C.dpad, dtype=torch.int64, device=dev)
        self.aggW = [torch.zeros(C.nchunk * eng.S, JAC, **i32) for _ in range(NB)]
        self.aggC = [torch.zeros(C.nchunk, JAC, **i32) for _ in range(NB)]
        if eng.signer is not None:
            eng.signer._ensure(n * self.nv)
        self.salt = (ctypes.c_ubyte * 64)()
        self.avail_host = (ctypes.c_ubyte * 64)()
        self.rm_fallback = _RecoverMat()
        # ---- the block's read-back
        self.h_w = torch.zeros(d, dtype=torch.float64).pin_memory()
        self.h_bad = torch.zeros(4, dtype=torch.int32).pin_memory()
        self.h_counts = torch.zeros(8, dtype=torch.int32).pin_memory()
        self.h_cpos = torch.zeros(n, dtype=torch.int32).pin_memory()
        rows = eng.world * eng.per
        self.h_pack = torch.zeros(rows, eng.pack_stride, dtype=torch.uint8).pin_memory()
        self.d2h_bytes = (self.h_w.numel() * 8 + 16 + 32 + 4 * n + rows * eng.pack_stride + 8)
        eng._ensure_eval_slots()
        self.cl = [self._build(par) for par in range(NB)]

    # ------------------------------------------------------------------------------------------------------------------
    def _view(self, par, off, n, dtype):
        return self.plan_dev[par][off: off + n * torch.empty(0, dtype=dtype).element_size()].view(dtype)

    def _build(self, par: int):
        eng = self.eng
        cfg, dev, P, n, d = eng.cfg, eng.device, self.P, self.n, eng.d
        C = eng.commit
        cl = CmdList(self.streams)
        ix = {}
        pout = eng._plan_out[par]
        counts = pout["counts"]
        n_con, n_lc, n_la = counts[1:2], counts[2:3], counts[3:4]
        pd = self.plan_dev[par]
        stake_d, isw_d = self._view(par, self.o_stake, n, torch.int32), self._view(par, self.o_isw, n, torch.uint8)
        dst_d, lrow_d = self._view(par, self.o_dst, max(1, P), torch.int32), self._view(par, self.o_lrow, n, torch.int32)
        rank_d, ids_d = self._view(par, self.o_rank, n, torch.int32), self._view(par, self.o_ids, n, torch.int32)
        sign_d, xs_d = self._view(par, self.o_sign, self.nv, torch.int32), self._view(par, self.o_xs, 10, torch.int32)
        av_d, slots_d = self._view(par, self.o_av, 64, torch.uint8), self._view(par, self.o_slots, n, torch.int32)
        hash_d = self._view(par, self.o_hash, 64, torch.uint8)
        spec = cfg.spec
        F, Cn, B, Ns = spec.n_features, spec.n_classes, cfg.batch_size, cfg.shard_size
        data = eng.data
        seed = cfg.seed & (2**64 - 1)
        scale4 = float(10 ** cfg.precision)

        # ---- LOCAL ------------------------------------------------------------------------------------------------
        ix["local0"] = cl.n
        # behind the previous block's SGD step (which is behind every reader of U on the chain stream) - not behind its
        # read-back: the step of block t+1 is queued as soon as block t's back is, before the host waits for block t
        cl.wait(SIDE, self.w_done)
        ix["mb"] = cl.add("sample_minibatch", SIDE, seed, 0, data.peer_ids, P, Ns, B, self.mb_idx)
        if eng.model.kind == "softmax":
            ix["step"] = cl.add("local_softmax_step", SIDE, eng.w32, data.X, data.y, self.mb_idx, eng.U, self.loss, P, Ns, B, F,
                                Cn, eng.ldu, cfg.clip_norm)
            if cfg.dp_in_model and cfg.noising and cfg.epsilon > 0:
                own = data.peer_ids.view(-1, 1).contiguous()
                ix["own"] = cl.add("noise_mix", SIDE, eng.U, own, 1, (cfg.seed ^ 0x6F776E6E) & (2**64 - 1), 0, cfg.noise_scale,
                                   eng.U, None, None, None, P, d, eng.ldu, eng.ldh)
        else:
            own = cfg.noise_scale if (cfg.noising and cfg.epsilon > 0) else 0.0
            ix["step"] = cl.add("local_logistic_step", SIDE, eng.w32, data.X, data.y, self.mb_idx, eng.U, P, Ns, B, F, eng.ldu,
                                cfg.logistic_alpha, cfg.logistic_lambda, own, seed, data.peer_ids, 0)
        cl.record(SIDE, self.u_ready[par])
        ix["local1"] = cl.n
        # ---- EVAL (of the weights the previous block produced) -------------------------------------------------------
        ix["eval0"] = cl.n
        cl.wait(AGG, self.collect_ev)
        cnt_d, cnt_h, _ = eng._ev_slots[0]
        ix["ev_set"] = cl.memset(AGG, cnt_d, 0)
        if eng.model.kind == "softmax":
            ix["ev_a"] = cl.add("eval_softmax", AGG, eng.w32, None, data.X_test, data.y_test, cnt_d[0:1], 1, data.X_test.shape[0],
                                F, Cn, eng.w32.numel())
            ix["ev_b"] = cl.add("eval_softmax", AGG, eng.w32, None, data.X_attack, data.y_attack, cnt_d[1:2], 1,
                                data.X_attack.shape[0], F, Cn, eng.w32.numel())
        else:
            ix["ev_a"] = cl.add("eval_logistic", AGG, eng.w32, None, data.X_test, data.y_test, cnt_d[0:1], 1, data.X_test.shape[0],
                                F, eng.w32.numel())
            ix["ev_b"] = cl.add("eval_logistic", AGG, eng.w32, None, data.X_attack, data.y_attack, cnt_d[1:2], 1,
                                data.X_attack.shape[0], F, eng.w32.numel())
        ix["ev_cp"] = cl.d2h(AGG, cnt_h, cnt_d)
        cl.record(AGG, self.eval_done)
        ix["ev_rec"] = cl.record(AGG, self.eval_host[0])
        ix["eval1"] = cl.n
        # ---- PLAN upload -------------------------------------------------------------------------------------------------
        ix["plan0"] = cl.n
        ix["plan_cp"] = cl.h2d(MAIN, pd, self.plan_pinned[0])
        cl.record(MAIN, self.plan_ready[par])
        ix["plan1"] = cl.n
        # ---- FRONT --------------------------------------------------------------------------------------------------------
        ix["front0"] = cl.n
        cl.wait(MAIN, self.u_ready[par])
        cl.wait(MAIN, self.side_done[par])
        cl.wait(MAIN, self.agg_done[par])
        ix["vrf_wait"] = cl.wait(MAIN, self._vrf_placeholder)
        nn = max(1, cfg.num_noisers)
        ix["noisers"] = cl.add("vrf_noisers", MAIN, None, hash_d, 32, stake_d, n, isw_d, data.peer_ids, P, cfg.num_noisers,
                               eng._noiser_tab, None)
        if eng.fused:
            comm = eng.comm
            ix["noise"] = cl.add("noise_mix", MAIN, eng.U, eng._noiser_tab, nn, seed, 0, 0.0, None, comm.hi, comm.lo, self.norms,
                                 P, d, eng.ldu, comm.hi.stride(0))
            ix["gram"] = cl.add("krum_gram_fused", MAIN, comm._tables["hi"], comm._tables["lo"], comm._tables["gpart"],
                                comm._tables["flags"], comm.world, comm.rank, comm.Sr, comm.ldh, comm.ksplit, comm.ldg,
                                comm.base + comm.o_lpart, comm.base + comm.o_cnt, 0, 0)
            ix["select"] = cl.add("krum_select_mc_wait", MAIN, comm.base + comm.o_gpart, comm.world, comm.ldg, slots_d, 0, 0,
                                  comm.base + comm.o_flags, comm.world, 0, self.scores, self.accept)
        else:
            ix["noise"] = cl.add("noise_mix_rows", MAIN, eng.U, eng._noiser_tab, nn, seed, 0, 0.0, None, self.hi, self.lo,
                                 self.norms, P, d, eng.ldu, eng.ldh, dst_d)
            ix["gram"] = cl.add("gram_tc", MAIN, self.hi, self.lo, 0, P, eng.ldh, 1, self.Gpart, 128)
            ix["select"] = cl.add("krum_select_mc", MAIN, self.Gpart, 1, 128, self.norms, None, 0, 0, self.scores, self.accept,
                                  None)
        ix["plan"] = cl.add("block_plan", MAIN, self.accept, rank_d, lrow_d, 0, 0, 0, counts, pout["contrib_pos"], pout["sel"],
                            pout["appr_local"], pout["csel"])
        ysum = eng.comm.ypart if eng.fused else eng._ysum[par]
        cl.add("share_sum_dc", MAIN, eng.U, pout["sel"], pout["sel"].numel(), n_lc, d, eng.ldu, 0, self.nchunk, self.S, scale4,
               ysum)
        ix["front1"] = cl.n
        # ---- BACK ---------------------------------------------------------------------------------------------------------
        ix["back0"] = cl.n
        cl.wait(MAIN, self.eval_done)
        if eng.fused:
            comm = eng.comm
            c0, nc = comm.chunk_range()
            ix["recon"] = cl.add("secagg_fused", MAIN, comm._tables["ypart"], comm._tables["w64"], comm._tables["w32"],
                                 comm._tables["flags"], comm.world, comm.rank, self.S, d, c0, nc, self.rm_fallback, 1,
                                 self.avail_host, scale4, comm.base + comm.o_cnt + 512, comm.base + comm.o_cnt + 768,
                                 comm.base + comm.o_bad, n_con, 2, 0)
        else:
            ix["recon"] = cl.add("reconstruct_mat", MAIN, ysum, self.rm_fallback, av_d, self.S, 0, self.nchunk, d, scale4,
                                 eng.w64, eng.w32, None, eng._bad, n_con, 2)
        ix["w_rec"] = cl.record(MAIN, self.w_done)
        ix["back_mid"] = cl.n
        cl.wait(MAIN, self.sig_done[par])
        pack = eng._pack[par]
        src_pack = pack
        if eng.fused and eng.world > 1:
            comm = eng.comm
            self._cb_tabs = getattr(self, "_cb_tabs", {})
            tabs = [comm._table(comm.o_cb + b * comm.cb_bytes) for b in (0, 1)]
            self._cb_tabs[par] = tabs
            ix["gather"] = cl.add("symm_allgather_rows", MAIN, tabs[0], comm._tables["flags"], comm.base + comm.o_flags,
                                  comm.world, comm.rank, pack, comm.per * comm.cb_row, 4, 0)
            ix["pack_cp"] = cl._push(-4, MAIN, [self.h_pack.data_ptr(), comm.base + comm.o_cb, comm.cb_bytes])
        else:
            ix["pack_cp"] = cl.d2h(MAIN, self.h_pack, src_pack, src_pack.numel())
        cl.d2h(MAIN, self.h_w, eng.w64)
        cl.d2h(MAIN, self.h_bad, eng._bad)
        cl.d2h(MAIN, self.h_counts, counts)
        cl.d2h(MAIN, self.h_cpos, pout["contrib_pos"], 4 * n)
        cl.record(MAIN, self.collect_ev)
        ix["back1"] = cl.n
        # ---- COMMIT -------------------------------------------------------------------------------------------------------
        ix["commit0"] = cl.n
        Q, chunkC = self.Q[par], self.chunkC[par]
        cl.wait(SIDE, self.side_done[par])      # the witnesses of the block two heights back still read Q / chunkC of this parity
        cl.wait(SIDE, self.agg_done[par])       # ... and so does the miners' aggregation
        cl.add("quantize", SIDE, eng.U, Q, P, d, eng.ldu, C.dpad, scale4)
        cl.add("msm_table_dc", SIDE, Q, Q.stride(0), None, P, None, C.c_sidx, C.c_bidx, None, C.nchunk, POLY_SIZE, C.c_table,
               C.w, C.nw, chunkC, C.nchunk)
        cl.add("g1_sum_cols_jac", SIDE, chunkC, P, C.nchunk, self.fullJ)
        cl.add("g1_normalize", SIDE, self.fullJ, self.fullA, P)
        cl.add("g1_affine_to_bytes", SIDE, self.fullA, eng._cbytes[par], P)
        cl.record(SIDE, self.c_ready[par])
        ix["commit1"] = cl.n
        # ---- SIGN ---------------------------------------------------------------------------------------------------------
        ix["sign0"] = cl.n
        sg = eng.signer
        if sg is not None:
            from ..crypto.schnorr_batch import W_BITS, N_WIN

            cl.wait(SIGN, self.plan_ready[par])
            ix["nonce"] = cl.add("schnorr_nonces_dev", SIGN, sg.secrets_dev, self.salt, 0, ids_d, 0, sign_d, 0, sg._k)
            ix["comb"] = cl.add("g1_comb_mul", SIGN, sg._k, 0, sg._comb._table, W_BITS, N_WIN, sg._jac)
            ix["tnorm"] = cl.add("g1_normalize", SIGN, sg._jac, sg._aff, 0)
            ix["tbytes"] = cl.add("g1_affine_to_bytes", SIGN, sg._aff, sg._T, 0)
            ix["sign_mid"] = cl.n
            cl.wait(SIGN, self.c_ready[par])
            ix["finish"] = cl.add("schnorr_finish_dev", SIGN, sg._T, sg._k, sg.secrets_dev, eng._cbytes[par], lrow_d, 0, sign_d, 0,
                                  pack, pack.stride(0))
            ix["packonly"] = cl.add("pack_commitments", SIGN, eng._cbytes[par], P, pack, pack.stride(0))
            cl.record(SIGN, self.sig_done[par])
        else:
            ix["sign_mid"] = cl.n
            cl.wait(SIDE, self.c_ready[par])
            ix["packonly"] = cl.add("pack_commitments", SIDE, eng._cbytes[par], P, pack, pack.stride(0))
            cl.record(SIDE, self.sig_done[par])
        ix["sign1"] = cl.n
        # ---- WITNESS ------------------------------------------------------------------------------------------------------
        # Their own lowest-priority stream: the witness MSM fills the GPU for most of a block, while the side stream's
        # kernels (local step, error rates, column sums, normalisation) are short and latency-bound - side by side they
        # overlap instead of queueing behind one another.
        ix["wit0"] = cl.n
        G = C.nchunk * self.S
        cl.wait(WIT, self.collect_ev)           # the plan's row lists and counts of this block
        cl.wait(WIT, self.c_ready[par])         # Q and chunkC of this block
        ix["wit"] = cl.add("msm_table_dc", WIT, Q, Q.stride(0), pout["appr_local"], self.wrows, n_la, C.w_sidx, C.w_bidx, None,
                           G, POLY_SIZE - 1, C.w_table, C.w, C.nw, self.W, G)
        cl.record(WIT, self.side_done[par])
        # the miners' aggregate = commitment / witnesses of the contributors' SUMMED polynomial (CommitmentEngine.
        # aggregate_by_scalars): one more MSM row each instead of a point addition per contributor and group.  Two small,
        # latency-bound launches: their own stream (shared with the error-rate kernels), beside the witness MSM
        cl.wait(AGG, self.collect_ev)
        cl.wait(AGG, self.c_ready[par])
        cl.add("sum_rows_i64_dc", AGG, Q, Q.stride(0), pout["sel"], pout["sel"].numel(), n_lc, C.dpad, self.aggQ)
        cl.add("msm_table_dc", AGG, self.aggQ, C.dpad, None, 1, None, C.c_sidx, C.c_bidx, None, C.nchunk, POLY_SIZE, C.c_table,
               C.w, C.nw, self.aggC[par], C.nchunk)
        cl.add("msm_table_dc", AGG, self.aggQ, C.dpad, None, 1, None, C.w_sidx, C.w_bidx, None, G, POLY_SIZE - 1, C.w_table,
               C.w, C.nw, self.aggW[par], G)
        cl.record(AGG, self.agg_done[par])
        ix["wit1"] = cl.n
        cl.ix = ix
        return cl

    # ------------------------------------------------------------------------------------------------------------------
    def run(self, par: int, seg: str):
        cl = self.cl[par]
        cl.run(cl.ix[seg + "0"], cl.ix[seg + "1"])

    def queue_local_and_eval(self, it_next: int, rep, do_eval: bool, inputs=None):
        """Behind block `it_next - 1`: the next block's local step (minibatches drawn from the resident shards, or the
        caller's `inputs` = (X [P,B,F], y [P,B])), then the error rates of the new weights."""
        eng = self.eng
        par = it_next % self.NB
        cl = self.cl[par]
        ix = cl.ix
        slot = it_next % NOISE_TABLE_ITERS
        cl.patch(ix["mb"], 1, it_next)
        soft = eng.model.kind == "softmax"
        cl.enable(ix["mb"], inputs is None)
        if inputs is None:
            X, y, idx, Ns = eng.data.X, eng.data.y, self.mb_idx.data_ptr(), eng.cfg.shard_size
        else:
            X, y, idx, Ns = inputs[0], inputs[1], 0, inputs[0].shape[1]
            self._fed_keep = inputs
        cl.patch(ix["step"], 1, X.data_ptr())
        cl.patch(ix["step"], 2, y.data_ptr())
        cl.patch(ix["step"], 3, idx)
        cl.patch(ix["step"], 7 if soft else 6, Ns)
        if eng.model.kind == "logistic":
            cl.patch(ix["step"], 15, slot)
        elif "own" in ix:
            cl.patch(ix["own"], 4, slot)
        self._last_src = None if inputs is None else inputs[0].data_ptr()
        self.run(par, "local")
        if do_eval:
            self.queue_eval(rep, par)

    def queue_eval(self, rep, par: int):
        """Error rates of the weights block `rep.iteration` produced (side stream, behind the block's read-back)."""
        eng = self.eng
        cl = self.cl[par]
        ix = cl.ix
        k = rep.iteration & 3
        cnt_d, cnt_h, _ = eng._ev_slots[k]
        cl.patch(ix["ev_set"], 0, cnt_d.data_ptr())
        cl.patch(ix["ev_a"], 4, cnt_d[0:1].data_ptr())
        cl.patch(ix["ev_b"], 4, cnt_d[1:2].data_ptr())
        cl.patch(ix["ev_cp"], 0, cnt_h.data_ptr())
        cl.patch(ix["ev_cp"], 1, cnt_d.data_ptr())
        cl.patch(ix["ev_rec"], 0, self.eval_host[k].handle)
        self.run(par, "eval")
        eng._attach_eval_resolver(rep, cnt_h, self.eval_host[k])

    def plan_block(self, plan_cls):
        """Host plan of the next block written STRAIGHT into the pinned staging buffer (ledger.cpp: bsc_plan_block fills the
        stake / worker / candidate / arrival-rank / signer / share arrays in place), plus the three row maps this rank adds.
        Returns the engine's _Plan; upload_plan() then only has to queue the copy."""
        eng = self.eng
        cfg, it, n = eng.cfg, eng.iteration, self.n
        from ..ops import chain as chain_ops
        from ..ops._call import lib

        k = self._plan_i = (getattr(self, "_plan_i", -1) + 1) % 4
        v, pp = self._plan_views[k], self._plan_ptrs[k]
        sc = self._scratch
        sc["stake"][:] = eng.stake
        sc["alive"][:] = eng.alive
        prev_hash = eng.chain.latest_hash()
        lib().bsc_plan_block(prev_hash, len(prev_hash), sc["p_stake"], sc["p_alive"], n, cfg.num_verifiers, cfg.num_miners,
                             cfg.krum_thresh, cfg.num_samples, int(cfg.random_sampling),
                             chain_ops.seed64(cfg.seed, it, "verify"), chain_ops.seed64(cfg.seed, it, "shares"),
                             chain_ops.seed64("rs", it), ctypes.c_double(cfg.krum_adversary_frac), cfg.shares_per_miner,
                             self.S, sc["p_verifiers"], sc["p_miners"], sc["p_workers"], pp["ids"], pp["rank"], pp["sign"],
                             pp["av"], pp["xs"], pp["isw"], pp["stake"], sc["p_counts"])
        nv, nm, nw, nk, ns, ok, enough, clip = (int(x) for x in sc["counts"])
        pl = plan_cls(sc["verifiers"][:nv].tolist(), sc["miners"][:nm].tolist(), None, prev_hash)
        pl.n_workers = nw
        pl.stake_m, pl.is_worker = v["stake"], v["isw"]
        if nv == 0 or nm == 0 or nw == 0:
            pl.workers = []
            return pl
        pl.workers = True                    # only its truth value and count are used on this path
        ks = v["ids"][:nk]
        pl.krum_set = ks.copy()
        pl.clip = clip
        pl.share_rank = v["rank"][:nk]
        pl.signers = v["sign"][:ns].tolist()
        pl.approved_ok = bool(ok)
        pl.trigger = max(1, cfg.num_samples // 2) if cfg.share_trigger == "half" else nk
        pl.shares_avail = v["av"][: self.S].tolist()
        pl.xs = v["xs"][:10].tolist()
        pl.enough_shares = bool(enough)
        pos = self._pos
        pos.fill(-1)
        pos[ks] = self._arange[:nk]
        v["dst"][: self.P] = pos[self._local_ids]
        v["lrow"][:nk] = self._local_row_of[ks]
        if eng.fused:
            v["slots"][:nk] = self._slot_of[ks]
        v["hash"][:32] = np.frombuffer(prev_hash, dtype=np.uint8)
        pl._staged = k
        return pl

    def upload_plan(self, ctx):
        """Queue the ONE pinned staging copy of everything the chain reads about the block (fixed layout, see __init__)."""
        par = ctx.par
        k = ctx.pl._staged
        cl = self.cl[par]
        cl.patch(cl.ix["plan_cp"], 1, self.plan_pinned[k].data_ptr())
        self.run(par, "plan")

    def queue_front(self, ctx):
        eng, cfg, pl = self.eng, self.eng.cfg, ctx.pl
        par, nk, it = ctx.par, ctx.nk, ctx.it
        cl = self.cl[par]
        ix = cl.ix
        slot = it % NOISE_TABLE_ITERS
        if ctx.use_noise:
            xyz, vev, _ = ctx.vrf_handle
            cl.patch(ix["vrf_wait"], 0, vev.handle)
            cl.patch(ix["noisers"], 0, xyz.data_ptr())
        cl.enable(ix["vrf_wait"], ctx.use_noise)
        cl.enable(ix["noisers"], ctx.use_noise)
        cl.patch(ix["noise"], 4, slot)
        cl.patch(ix["noise"], 5, cfg.noise_scale if ctx.use_noise else 0.0, "f")
        if eng.fused:
            comm = eng.comm
            comm._gram_seq = getattr(comm, "_gram_seq", 0) + 1
            epoch = comm.next_epoch("krum")      # per launch, not per block: a replanned block runs this front twice
            cl.patch(ix["gram"], 12, comm._gram_seq)
            cl.patch(ix["gram"], 13, epoch)
            cl.patch(ix["select"], 4, nk)
            cl.patch(ix["select"], 5, pl.clip)
            cl.patch(ix["select"], 8, epoch)
        else:
            ldg = (nk + 127) // 128 * 128
            ks = gram_ops.ksplit_for(nk, eng.ldh)
            cl.patch(ix["gram"], 2, nk)
            cl.patch(ix["gram"], 5, ks)
            cl.patch(ix["gram"], 7, ldg)
            cl.patch(ix["select"], 1, ks)
            cl.patch(ix["select"], 2, ldg)
            cl.patch(ix["select"], 5, nk)
            cl.patch(ix["select"], 6, pl.clip)
        cl.patch(ix["plan"], 3, nk)
        cl.patch(ix["plan"], 4, int(pl.approved_ok))
        cl.patch(ix["plan"], 5, pl.trigger)
        self.run(par, "front")
        ctx.front_queued = True

    def queue_sign_T(self, ctx):
        """The approvals' nonces and T = v*G need only the plan (chain position, candidates, signers): queued first thing
        behind the block's tail, they are the longest off-chain dependency of the NEXT tail."""
        eng = self.eng
        par, nk, it = ctx.par, ctx.nk, ctx.it
        cl = self.cl[par]
        ix = cl.ix
        if eng.signer is not None:
            ns = ctx.ns
            signing = ns > 0
            for name in ("nonce", "comb", "tnorm", "tbytes", "finish"):
                cl.enable(ix[name], signing)
            cl.enable(ix["packonly"], not signing)
            if signing:
                ctypes.memmove(self.salt, ctx.salt64, 64)
                N = nk * ns
                cl.patch(ix["nonce"], 2, it)
                cl.patch(ix["nonce"], 4, nk)
                cl.patch(ix["nonce"], 6, ns)
                cl.patch(ix["comb"], 1, N)
                cl.patch(ix["tnorm"], 2, N)
                cl.patch(ix["tbytes"], 2, N)
                cl.patch(ix["finish"], 5, nk)
                cl.patch(ix["finish"], 7, ns)
        cl.run(ix["sign0"], ix["sign_mid"])

    def queue_commit(self, it: int) -> dict:
        """Commitments of block `it`'s updates (side stream, right behind its local step).  Returns the block's `pre` record."""
        eng = self.eng
        par = it % self.NB
        self.run(par, "commit")
        return {"it": it, "Q": self.Q[par], "chunkC": self.chunkC[par], "cbytes": eng._cbytes[par], "u_ready": None,
                "c_ready": None, "src": self._last_src, "native": True}

    def queue_sign_finish(self, ctx):
        """The approvals' challenge + response: waits (on the approval stream) for T = v*G and for the commitments."""
        cl = self.cl[ctx.par]
        cl.run(cl.ix["sign_mid"], cl.ix["sign1"])

    def queue_commit_and_sign(self, ctx):
        ctx.pre = self.queue_commit(ctx.it)
        self.queue_sign_finish(ctx)

    def queue_back(self, ctx) -> bool:
        """Returns False when this block's abscissae have no safe recovery matrix (the Python schedule then queues the
        divided-difference kernel instead)."""
        eng, cfg, pl = self.eng, self.eng.cfg, ctx.pl
        par, it = ctx.par, ctx.it
        cl = self.cl[par]
        ix = cl.ix
        rm = recover_matrix(pl.xs)
        if eng.fused:
            comm = eng.comm
            use = rm is not None
            if not use:
                rm = self.rm_fallback
                for i, x in enumerate(pl.xs):
                    rm.xs[i] = int(x)
            cl.patch(ix["recon"], 10, ctypes.addressof(rm))
            cl.patch(ix["recon"], 11, int(use))
            for i in range(64):
                self.avail_host[i] = pl.shares_avail[i] if i < len(pl.shares_avail) else 0
            cl.patch(ix["recon"], 19, comm.next_epoch("secagg"))
            if "gather" in ix:
                comm._cb_seq = getattr(comm, "_cb_seq", 0) + 1
                buf = comm._cb_seq & 1
                cl.patch(ix["gather"], 0, ctypes.addressof(self._cb_tabs[par][buf]))
                cl.patch(ix["gather"], 8, comm._cb_seq)
                cl.patch(ix["pack_cp"], 1, comm.base + comm.o_cb + buf * comm.cb_bytes)
        else:
            if rm is None:
                # no safe matrix for these abscissae: the divided-difference kernel, then the rest of the segment
                from ..ops import secagg

                cl.run(ix["back0"], ix["recon"])
                with torch.cuda.stream(eng._main):
                    secagg.reconstruct_step_dev(eng._ysum[par], pl.xs, self._view(par, self.o_xs, 10, torch.int32),
                                                self._view(par, self.o_av, 64, torch.uint8)[: self.S], eng.d, eng.w64, eng.w32,
                                                cfg.precision, bad=eng._bad, gate=eng._plan_out[par]["counts"][1:2], gate_min=2)
                cl.run(ix["w_rec"], ix["back1"])
                return True
            cl.patch(ix["recon"], 1, ctypes.addressof(rm))
        self.run(par, "back")
        return True

    def collect(self) -> dict:
        self.collect_ev.synchronize()
        self.eng.last_d2h_bytes = self.d2h_bytes
        return {"w": self.h_w.numpy(), "bad": self.h_bad.numpy(), "counts": self.h_counts.numpy(), "cpos": self.h_cpos.numpy(),
                "pack": self.h_pack.numpy()}

    def queue_witness(self, ctx):
        par = ctx.par
        cl = self.cl[par]
        cl.patch(cl.ix["wit"], 3, min(self.wrows, ctx.nk))
        self.run(par, "wit")
        return (self.aggC[par], self.aggW[par])
