"""Batched Schnorr approvals: every verifier signs the commitment of every update Multi-Krum accepted
(DistSys/krum.go:338-358 -> SchnorrSign, kyber.go:873-896).

The expensive part of a signature is the commitment T = v*G.  All (verifier, update) pairs of a block
are done in one launch of the fixed-base comb kernel (g1_comb_mul_kernel: 33 mixed additions per
scalar over a table of the generator's window multiples); hashing c = H(T || msg) with BLAKE2Xb and the
scalar arithmetic r = v - x*c stay on the host (microseconds).  The launch is asynchronous, so it
overlaps with the witness MSM; `finish()` is called when the block is assembled.
"""
from __future__ import annotations

import ctypes

import numpy as np
import torch

from .. import _native
from ..ops.msm import AFF, JAC

W_BITS, N_WIN = 8, 33          # 33 * 8 = 264 >= 257: the recoding carry of a 256-bit scalar has a window


class SchnorrBatchSigner:
    def __init__(self, device="cpu"):
        self.device = torch.device(device)
        self.L = _native.load()
        self._table = None

    def _ensure_table(self):
        if self._table is not None or self.device.type != "cuda":
            return
        L = self.L
        gen = ctypes.create_string_buffer(64)
        L.bsc_g1_generator(gen)
        g_aff = torch.empty(1, AFF, dtype=torch.int32)
        L.bsc_g1_bytes_to_affine_host(gen.raw, ctypes.c_void_p(g_aff.data_ptr()), 1)
        H = 1 << (W_BITS - 1)
        jac = torch.empty(N_WIN * H, JAC, dtype=torch.int32, device=self.device)
        stream = ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        g_dev = g_aff.to(self.device)
        rc = L.bsc_msm_table_build(ctypes.c_void_p(g_dev.data_ptr()), 1, W_BITS, N_WIN, ctypes.c_void_p(jac.data_ptr()), stream)
        assert rc == 0
        self._table = torch.empty(N_WIN * H, AFF, dtype=torch.int32, device=self.device)
        rc = L.bsc_g1_normalize(ctypes.c_void_p(jac.data_ptr()), ctypes.c_void_p(self._table.data_ptr()), N_WIN * H, stream)
        assert rc == 0

    def launch(self, secrets: list[bytes], tags: list[bytes]):
        """Start signing: one (secret key, unique tag) pair per signature.  Returns a handle for finish()."""
        L = self.L
        if isinstance(secrets, np.ndarray):      # [n, 32] uint8 secrets and [n, taglen] uint8 tags, already packed
            n = secrets.shape[0]
            x = np.ascontiguousarray(secrets, dtype=np.uint8).reshape(-1)
            tg = np.ascontiguousarray(tags, dtype=np.uint8).reshape(-1)
            taglen = tags.shape[1] if n else 0
        else:
            n = len(secrets)
            x = np.frombuffer(b"".join(secrets), dtype=np.uint8).copy()
            tg = np.frombuffer(b"".join(tags), dtype=np.uint8).copy()
            taglen = len(tags[0]) if n else 0
        vw = np.empty((max(1, n), 8), dtype=np.uint32)
        vbe = np.empty(max(1, n) * 32, dtype=np.uint8)
        if n:
            L.bsc_schnorr_nonces(n, tg.ctypes.data_as(ctypes.c_void_p), taglen, x.ctypes.data_as(ctypes.c_void_p),
                                 vw.ctypes.data_as(ctypes.c_void_p), vbe.ctypes.data_as(ctypes.c_void_p))
        T_dev, event = None, None
        if self.device.type == "cuda" and n:
            self._ensure_table()
            if not hasattr(self, "_side"):
                torch.cuda.current_stream(self.device).synchronize()   # once: the comb table build has finished
                self._side = torch.cuda.Stream(self.device, priority=-1)   # needed at block assembly
                self._k_pinned = torch.empty(4096, 8, dtype=torch.int32).pin_memory()
            main = torch.cuda.current_stream(self.device)
            kp = self._k_pinned[:n] if n <= self._k_pinned.shape[0] else torch.empty(n, 8, dtype=torch.int32).pin_memory()
            kp.copy_(torch.from_numpy(vw.view(np.int32)[:n]))
            # side stream: the comb kernel is a single small CTA-wave and runs under the witness MSM.  It does not
            # wait for the compute stream (its inputs are the nonces uploaded here and the static comb table), so
            # it starts at once even though the witness / aggregation kernels were queued first.
            # persistent device buffers (one launch per block; the previous block's results were read back, with a
            # host-side wait, before this launch) instead of four allocations + record_stream calls per block
            if getattr(self, "_cap", 0) < n:
                self._cap = max(256, 2 * n)
                with torch.cuda.stream(self._side):
                    self._k_dev = torch.empty(self._cap, 8, dtype=torch.int32, device=self.device)
                    self._jac = torch.empty(self._cap, JAC, dtype=torch.int32, device=self.device)
                    self._aff = torch.empty(self._cap, AFF, dtype=torch.int32, device=self.device)
                    self._T = torch.empty(self._cap, 64, dtype=torch.uint8, device=self.device)
                self._event = torch.cuda.Event()
            with torch.cuda.stream(self._side):
                stream = ctypes.c_void_p(self._side.cuda_stream)
                k = self._k_dev[:n]
                k.copy_(kp, non_blocking=True)
                rc = L.bsc_g1_comb_mul(ctypes.c_void_p(k.data_ptr()), n, ctypes.c_void_p(self._table.data_ptr()), W_BITS,
                                       N_WIN, ctypes.c_void_p(self._jac.data_ptr()), stream)
                assert rc == 0, rc
                assert L.bsc_g1_normalize(ctypes.c_void_p(self._jac.data_ptr()), ctypes.c_void_p(self._aff.data_ptr()), n,
                                          stream) == 0
                T_dev = self._T[:n]
                assert L.bsc_g1_affine_to_bytes(ctypes.c_void_p(self._aff.data_ptr()), ctypes.c_void_p(T_dev.data_ptr()), n,
                                                stream) == 0
                event = self._event
                event.record(self._side)
        return dict(n=n, x=x, vw=vw, vbe=vbe, T_dev=T_dev, event=event)

    def select(self, handle, rows: list[int], T_host: np.ndarray | None = None):
        """Sub-handle for a subset of the launched signatures (the updates that reached the block).  T_host: the
        [n, 64] commitment bytes of the *full* launch if the caller already copied them to the host."""
        n = handle["n"]
        idx = np.asarray(rows, dtype=np.int64)
        if T_host is None and handle["T_dev"] is not None:
            if handle.get("event") is not None:
                torch.cuda.current_stream(self.device).wait_event(handle["event"])
            T_host = handle["T_dev"].cpu().numpy()
        sub = dict(n=len(rows), x=handle["x"].reshape(n, 32)[idx].reshape(-1).copy(),
                   vw=handle["vw"][:n][idx].copy(), vbe=handle["vbe"].reshape(-1, 32)[:n][idx].reshape(-1).copy(),
                   T_dev=None, event=None, T_host=None if T_host is None else np.ascontiguousarray(np.asarray(T_host)[idx]))
        return sub

    def finish(self, handle, msgs: list[bytes], T_host: np.ndarray | None = None) -> list[bytes]:
        """Complete the signatures over `msgs` (all the same length).  T_host: the [n, 64] commitment bytes if
        the caller already copied handle['T_dev'] to the host; otherwise they are fetched (or computed on the
        CPU when there is no GPU)."""
        n = handle["n"]
        if n == 0:
            return []
        L = self.L
        if T_host is None:
            T_host = handle.get("T_host")
        if T_host is None:
            if handle["T_dev"] is not None:
                if handle.get("event") is not None:
                    torch.cuda.current_stream(self.device).wait_event(handle["event"])
                T_host = handle["T_dev"].cpu().numpy()
            else:
                gen = ctypes.create_string_buffer(64)
                L.bsc_g1_generator(gen)
                g_aff = np.empty(AFF, dtype=np.int32)
                L.bsc_g1_bytes_to_affine_host(gen.raw, g_aff.ctypes.data_as(ctypes.c_void_p), 1)
                jac = np.empty((n, JAC), dtype=np.int32)
                L.bsc_g1_fixed_base_mul_host(handle["vw"].ctypes.data_as(ctypes.c_void_p), n,
                                             g_aff.ctypes.data_as(ctypes.c_void_p), jac.ctypes.data_as(ctypes.c_void_p))
                T_host = np.empty((n, 64), dtype=np.uint8)
                L.bsc_g1_jac_to_bytes_host(jac.ctypes.data_as(ctypes.c_void_p), T_host.ctypes.data_as(ctypes.c_void_p), n)
        T_host = np.ascontiguousarray(T_host, dtype=np.uint8)
        m = np.frombuffer(b"".join(msgs), dtype=np.uint8).copy()
        sigs = np.empty(n * 64, dtype=np.uint8)
        L.bsc_schnorr_finish_batch(n, m.ctypes.data_as(ctypes.c_void_p), len(msgs[0]), handle["x"].ctypes.data_as(ctypes.c_void_p),
                                   handle["vbe"].ctypes.data_as(ctypes.c_void_p), T_host.ctypes.data_as(ctypes.c_void_p),
                                   sigs.ctypes.data_as(ctypes.c_void_p))
        raw = sigs.tobytes()
        return [raw[64 * i: 64 * i + 64] for i in range(n)]


class ApprovalSigner:
    """The verifiers' approvals of one block, produced where the commitments live.

    GPU: every (Krum candidate, live verifier) pair of this rank's peers is signed before the Krum verdict is
    known - nonce derivation (schnorr_nonce_kernel), T = v*G (g1_comb_mul_kernel), c = H(T || commitment) and
    r = v - x*c (schnorr_finish_kernel) all run on a side stream while the VRF and Krum kernels run, and the
    result lands in the per-peer `pack` row (commitment || signatures) that is gathered across ranks and copied
    to the host once, at the block's tail.  The host never touches a nonce, a hash or a scalar.
    CPU: only the contributors' approvals are computed, at the tail, with the host restatements.

    Nonce tag = per-run salt || previous block hash || (iteration, update, verifier) - see ops.chain.approval_tag.
    """

    def __init__(self, device, secrets: list[bytes]):
        self.device = torch.device(device)
        self.L = _native.load()
        self.n = len(secrets)
        self._sk_np = np.frombuffer(b"".join(secrets), dtype=np.uint8).reshape(self.n, 32).copy()
        self._comb = SchnorrBatchSigner(device)
        self._cap = 0
        if self.device.type == "cuda":
            self.secrets_dev = torch.from_numpy(self._sk_np).to(self.device)
            self._comb._ensure_table()
            self.stream = torch.cuda.Stream(self.device, priority=-2)
            self.event = torch.cuda.Event()

    def _ensure(self, n):
        if self._cap >= n:
            return
        self._cap = max(256, 2 * n)
        dev = self.device
        self._k = torch.empty(self._cap, 8, dtype=torch.int32, device=dev)
        self._jac = torch.empty(self._cap, JAC, dtype=torch.int32, device=dev)
        self._aff = torch.empty(self._cap, AFF, dtype=torch.int32, device=dev)
        self._T = torch.empty(self._cap, 64, dtype=torch.uint8, device=dev)

    def launch(self, salt64: bytes, it: int, cand_ids: torch.Tensor, cand_rows: torch.Tensor, nc: int,
               signers: torch.Tensor, ns: int, cbytes: torch.Tensor, pack: torch.Tensor, after: "torch.cuda.Event | None",
               serial: bool = False, commit_ready: "torch.cuda.Event | None" = None):
        """Queue the approvals of `nc` candidates by `ns` signers.  cand_ids / cand_rows / signers: int32 device tensors
        (global ids; rows of `cbytes` and `pack`, -1 = not this rank's peer).  `after`: event that marks the index tensors
        uploaded; `commit_ready`: event that marks `cbytes` complete.  The nonces and T = v*G need only the chain position,
        so they start at once; only the last kernel (challenge + response) waits for the commitments.  Returns the event
        that marks `pack` complete (None when run on the current stream)."""
        L, dev = self.L, self.device
        N = nc * ns
        if N == 0:
            return None
        self._ensure(N)
        p = ctypes.c_void_p
        ctx = torch.cuda.stream(self.stream) if not serial else _null()
        with ctx:
            if not serial and after is not None:
                self.stream.wait_event(after)
            st = p(torch.cuda.current_stream(dev).cuda_stream)
            assert L.bsc_schnorr_nonces_dev(p(self.secrets_dev.data_ptr()), salt64, it, p(cand_ids.data_ptr()), nc,
                                            p(signers.data_ptr()), ns, p(self._k.data_ptr()), st) == 0
            tab = self._comb._table
            assert L.bsc_g1_comb_mul(p(self._k.data_ptr()), N, p(tab.data_ptr()), W_BITS, N_WIN, p(self._jac.data_ptr()), st) == 0
            assert L.bsc_g1_normalize(p(self._jac.data_ptr()), p(self._aff.data_ptr()), N, st) == 0
            assert L.bsc_g1_affine_to_bytes(p(self._aff.data_ptr()), p(self._T.data_ptr()), N, st) == 0
            if not serial and commit_ready is not None:
                self.stream.wait_event(commit_ready)
            assert L.bsc_schnorr_finish_dev(p(self._T.data_ptr()), p(self._k.data_ptr()), p(self.secrets_dev.data_ptr()),
                                            p(cbytes.data_ptr()), p(cand_rows.data_ptr()), nc, p(signers.data_ptr()), ns,
                                            p(pack.data_ptr()), pack.stride(0), st) == 0
            if serial:
                return None
            self.event.record(self.stream)
        return self.event

    def sign_host(self, salt64: bytes, it: int, peers: list[int], signers: list[int], commitments: list[bytes]) -> list[bytes]:
        """CPU path: signatures of `signers` over each peer's commitment, row-major (peer, signer)."""
        from ..ops.chain import approval_tag

        if not peers or not signers:
            return []
        secrets = [self._sk_np[v].tobytes() for _ in peers for v in signers]
        tags = [approval_tag(salt64, it, p_, v) for p_ in peers for v in signers]
        h = self._comb.launch(secrets, tags) if self.device.type != "cuda" else _host_handle(self.L, secrets, tags)
        msgs = [c for c in commitments for _ in signers]
        return SchnorrBatchSigner("cpu").finish(h, msgs)


def _host_handle(L, secrets, tags):
    return SchnorrBatchSigner("cpu").launch(secrets, tags)


class _null:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False
