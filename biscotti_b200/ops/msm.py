"""Polynomial commitments and KZG witnesses on BN256 G1, batched over peers.

Reference: DistSys/kyber.go:172-202 (per-chunk commitments with SRS slice [10c, 10c+10)), :533-562
(createCommitment), :579-646 (createShareAndWitness), :244-287 (aggregateSecret), :456-482
(generateMinerSecretShares); DistSys/publicKey.go:28-66 (SRS = powers of tau, tau = 2).

Points live on the device as ``uint32[..., 16]`` (affine, Montgomery form) or ``uint32[..., 24]``
(Jacobian); the 64-byte big-endian wire form (point.go:109-167) is produced only for what enters
the ledger.
"""
from __future__ import annotations

import ctypes
import os

import torch

from ._call import lib, ptr, stream, check
from .secagg import POLY_SIZE, nchunks

AFF = 16  # uint32 words per affine point
JAC = 24  # uint32 words per Jacobian point


def _i32(x, device):
    return torch.tensor(x, dtype=torch.int32, device=device)


class CommitmentEngine:
    """Holds the SRS and the witness base table for one model dimension and runs the MSM kernels."""

    def __init__(self, d: int, total_shares: int, tau: int = 2, device="cpu", srs_bytes: bytes | None = None,
                 window: int | None = None, max_scalar_bits: int = 20, table_budget_bytes: int | None = None,
                 window2: int | None = None):
        self.d, self.S, self.device = d, total_shares, torch.device(device)
        self.nchunk = nchunks(d)
        self.dpad = self.nchunk * POLY_SIZE
        self.tau = tau
        L = lib()
        # ---- SRS (PKG1[i] = tau^i * G), padded with the generator (padding scalars are always 0)
        if srs_bytes is None:
            srs_bytes = self._generate_srs(tau)
        self.srs_bytes = srs_bytes
        gen = ctypes.create_string_buffer(64)
        L.bsc_g1_generator(gen)
        full = srs_bytes + gen.raw * (self.dpad - d)
        host_aff = torch.empty(self.dpad, AFF, dtype=torch.int32)
        L.bsc_g1_bytes_to_affine_host(full, ptr(host_aff), self.dpad)
        self.bases = host_aff.to(self.device)
        # ---- index tables
        c = torch.arange(self.nchunk, dtype=torch.int32)
        self.c_sidx = (c * POLY_SIZE).to(self.device)
        self.c_bidx = (c * POLY_SIZE).to(self.device)
        S = self.S
        cs = torch.arange(self.nchunk * S, dtype=torch.int32)
        self.w_sidx = ((cs // S) * POLY_SIZE + 1).to(self.device)          # a_1..a_9 of the chunk
        self.w_bidx = (cs * (POLY_SIZE - 1)).to(self.device)               # T[c, s, 0..8]
        self.wtable = self._build_witness_table()
        # ---- fixed-base window tables (device path; window=0 disables them)
        # Layout per base: window 0 has 2^(w-1) entries, the nw-1 windows above it 2^(w1-1) each (see kernels_msm.cu,
        # "two-width tables").  self.nw carries (count | w1 << 8) for the kernels; self.nwin / self.w1 are the plain values.
        self.w = self.nw = self.nwin = self.w1 = 0
        self.c_table = self.w_table = None
        self.table_bytes = 0
        nbases = self.dpad + self.wtable.shape[0]
        need = max_scalar_bits + 1

        def layout(w0, w1=None):
            """(nwin, w1, entries per base) for a first window of w0 bits; w1=None: uniform windows."""
            if w1 is None:
                nwin = -(-need // w0)
                return nwin, 0, nwin * (1 << (w0 - 1))
            nwin = 1 + max(0, -(-(need - w0) // w1))
            return nwin, w1, (1 << (w0 - 1)) + (nwin - 1) * (1 << (w1 - 1))

        # HBM budget of the window tables: half of the device (90 GB of a B200's 180 GB) unless told otherwise
        if table_budget_bytes is None:
            table_budget_bytes = 48 << 30
            if self.device.type == "cuda":
                table_budget_bytes = torch.cuda.get_device_properties(self.device).total_memory // 2
        env_gb = os.environ.get("BISCOTTI_MSM_TABLE_GB")
        if env_gb:
            table_budget_bytes = int(float(env_gb) * (1 << 30))
        env_w = os.environ.get("BISCOTTI_MSM_WINDOW")
        if env_w and window is None:
            window = int(env_w)
        choice = None
        if window is None:
            # Widest first window that fits the budget, with narrow windows above it: ~99 % of a clipped update's fixed-point
            # coordinates are below 2^12, so a 13-bit first window makes a term ONE mixed addition; the rare larger values
            # take a second addition from an 8-bit window.  Fall back to uniform windows when even that does not fit.
            if self.device.type == "cuda" and table_budget_bytes > 0:     # budget 0: no tables at all (msm_groups_kernel)
                for w0, w1 in ((13, 8), (12, 9), (11, 10)):
                    lay = layout(w0, w1)
                    if nbases * lay[2] * 64 <= table_budget_bytes:
                        choice = (w0,) + lay
                        break
                if choice is None:
                    for w0 in (7, 6, 5, 4, 3):
                        lay = layout(w0)
                        if nbases * lay[2] * 64 <= table_budget_bytes or w0 == 3:
                            choice = (w0,) + lay
                            break
        elif window and window2:
            choice = (window,) + layout(window, window2)
        elif window:
            w0 = window
            while w0 > 2 and nbases * layout(w0)[2] * 64 > table_budget_bytes:
                w0 -= 1
            choice = (w0,) + layout(w0)
        if choice is not None:
            self.w, self.nwin, self.w1, self.per_base = choice
            self.nw = self.nwin | (self.w1 << 8)
            self.table_bytes = nbases * self.per_base * 64
            self.c_table = self._build_window_table(self.bases)
            self.w_table = self._build_window_table(self.wtable)

    # -- SRS ------------------------------------------------------------------------------------
    def _generate_srs(self, tau: int) -> bytes:
        L = lib()
        d = self.d
        tb = int(tau).to_bytes(32, "big")
        if tau == 2 or d <= 4096 or self.device.type != "cuda":
            out = ctypes.create_string_buffer(64 * d)
            L.bsc_srs_g1_host(tb, d, out)
            return out.raw
        words = torch.empty(d, 8, dtype=torch.int32)
        L.bsc_fr_powers(tb, d, ptr(words))
        gen = ctypes.create_string_buffer(64)
        L.bsc_g1_generator(gen)
        g_aff = torch.empty(AFF, dtype=torch.int32)
        L.bsc_g1_bytes_to_affine_host(gen.raw, ptr(g_aff), 1)
        k = words.to(self.device)
        jac = torch.empty(d, JAC, dtype=torch.int32, device=self.device)
        check(L.bsc_g1_fixed_base_mul(ptr(k), d, ptr(g_aff), ptr(jac), stream(self.device)), "g1_fixed_base_mul")
        return self.to_bytes(self.normalize(jac))

    # -- witness base table T[c, s, j-1] = sum_{k<j} x_s^{j-1-k} * PKG1[10c+k], j = 1..9 ------------
    def _build_witness_table(self):
        S, T = self.S, POLY_SIZE - 1
        # "peers" of this MSM are the share indices s; scalars depend only on (s, j, k)
        Q = torch.zeros(S, T * T, dtype=torch.int64)
        for s in range(S):
            x = s - 10
            for j in range(1, POLY_SIZE):
                for k in range(j):
                    Q[s, (j - 1) * T + k] = x ** (j - 1 - k)
        G = self.nchunk * T
        g = torch.arange(G, dtype=torch.int32)
        cc, jj = g // T, g % T
        sidx = (jj * T)
        bidx = (cc * POLY_SIZE)
        og = (cc * (S * T) + jj)
        out = self._msm(Q.to(self.device), None, S, sidx.to(self.device), bidx.to(self.device), og.to(self.device), G, T,
                        self.bases, ostride=T, nout=self.nchunk * S * T)
        return self.normalize(out)

    def _build_window_table(self, bases_aff: torch.Tensor) -> torch.Tensor:
        """table[b][i][k-1] = k * 2^(w*i) * B_b (affine), built in chunks to bound the Jacobian scratch."""
        L = lib()
        nb = bases_aff.shape[0]
        per = self.per_base
        table = torch.empty(nb * per, AFF, dtype=torch.int32, device=self.device)
        chunk = max(1, (1 << 23) // per)
        for b0 in range(0, nb, chunk):
            b1 = min(nb, b0 + chunk)
            jac = torch.empty((b1 - b0) * per, JAC, dtype=torch.int32, device=self.device)
            src = bases_aff[b0:b1].contiguous()
            if self.device.type == "cuda":
                check(L.bsc_msm_table_build(ptr(src), b1 - b0, self.w, self.nw, ptr(jac), stream(self.device)),
                      "msm_table_build")
            else:
                L.bsc_msm_table_build_host(ptr(src), b1 - b0, self.w, self.nw, ptr(jac))
            table[b0 * per: b1 * per] = self.normalize(jac)
        return table

    # -- raw kernels -----------------------------------------------------------------------------
    def _msm_table(self, Q, sel, P, sidx, bidx, og, G, T, table, ostride, nout, count=None):
        out = torch.empty(nout, JAC, dtype=torch.int32, device=self.device)
        L = lib()
        if self.device.type == "cuda":
            check(L.bsc_msm_table_dc(ptr(Q), Q.stride(0), ptr(sel), P, ptr(count), ptr(sidx), ptr(bidx), ptr(og), G, T,
                                     ptr(table), self.w, self.nw, ptr(out), ostride, stream(self.device)), "msm_table")
        else:
            L.bsc_msm_table_host(ptr(Q), Q.stride(0), ptr(sel), P, ptr(sidx), ptr(bidx), ptr(og), G, T, ptr(table),
                                 self.w, self.nw, ptr(out), ostride)
        return out

    def _msm(self, Q, sel, P, sidx, bidx, og, G, T, bases, ostride, nout):
        out = torch.empty(nout, JAC, dtype=torch.int32, device=self.device)
        L = lib()
        if self.device.type == "cuda":
            check(L.bsc_msm_groups(ptr(Q), Q.stride(0), ptr(sel), P, ptr(sidx), ptr(bidx), ptr(og), G, T, ptr(bases),
                                   ptr(out), ostride, stream(self.device)), "msm_groups")
        else:
            L.bsc_msm_groups_host(ptr(Q), Q.stride(0), ptr(sel), P, ptr(sidx), ptr(bidx), ptr(og), G, T, ptr(bases),
                                  ptr(out), ostride)
        return out

    def normalize(self, jac: torch.Tensor) -> torch.Tensor:
        shape = jac.shape[:-1]
        N = jac.numel() // JAC
        out = torch.empty(*shape, AFF, dtype=torch.int32, device=jac.device)
        if jac.is_cuda:
            check(lib().bsc_g1_normalize(ptr(jac), ptr(out), N, stream(jac.device)), "g1_normalize")
        else:
            lib().bsc_g1_normalize_host(ptr(jac), ptr(out), N)
        return out

    def to_bytes(self, aff: torch.Tensor) -> bytes:
        N = aff.numel() // AFF
        if aff.is_cuda:
            out = torch.empty(N, 64, dtype=torch.uint8, device=aff.device)
            check(lib().bsc_g1_affine_to_bytes(ptr(aff), ptr(out), N, stream(aff.device)), "g1_affine_to_bytes")
            return out.cpu().numpy().tobytes()
        buf = ctypes.create_string_buffer(64 * N)
        lib().bsc_g1_affine_to_bytes_host(ptr(aff.contiguous()), buf, N)
        return buf.raw

    def to_bytes_into(self, aff: torch.Tensor, out: torch.Tensor) -> torch.Tensor:
        """Wire bytes of N affine points into the first N rows of a persistent [>= N, 64] uint8 tensor (no host sync)."""
        N = aff.numel() // AFF
        if aff.is_cuda:
            check(lib().bsc_g1_affine_to_bytes(ptr(aff), ptr(out), N, stream(aff.device)), "g1_affine_to_bytes")
        else:
            buf = ctypes.create_string_buffer(64 * N)
            lib().bsc_g1_affine_to_bytes_host(ptr(aff.contiguous()), buf, N)
            out[:N] = torch.frombuffer(bytearray(buf.raw), dtype=torch.uint8).view(N, 64)
        return out

    def to_bytes_tensor(self, aff: torch.Tensor) -> torch.Tensor:
        """Device-resident wire bytes [N, 64] (no host sync)."""
        N = aff.numel() // AFF
        out = torch.empty(N, 64, dtype=torch.uint8, device=aff.device)
        if aff.is_cuda:
            check(lib().bsc_g1_affine_to_bytes(ptr(aff), ptr(out), N, stream(aff.device)), "g1_affine_to_bytes")
        else:
            buf = ctypes.create_string_buffer(64 * N)
            lib().bsc_g1_affine_to_bytes_host(ptr(aff.contiguous()), buf, N)
            out = torch.frombuffer(bytearray(buf.raw), dtype=torch.uint8).view(N, 64)
        return out

    # -- protocol-level operations -----------------------------------------------------------------
    # Point tensors are recognised by their last dimension: 16 words = affine, 24 words = Jacobian.  With
    # affine=False the kernels hand Jacobian points to each other and no field inversion is paid until
    # something is serialised (workers and miners share device memory here, so the 64-byte wire form of
    # kyber.go:88-168 is only needed for what enters the ledger).
    def chunk_commitments(self, Q: torch.Tensor, sel: torch.Tensor | None = None, affine: bool = True) -> torch.Tensor:
        """C[p, c] = sum_j Q[p, 10c+j] * PKG1[10c+j]  -> [P, nchunk, 16 | 24]."""
        P = Q.shape[0] if sel is None else sel.numel()
        if self.c_table is not None:
            jac = self._msm_table(Q, sel, P, self.c_sidx, self.c_bidx, None, self.nchunk, POLY_SIZE, self.c_table,
                                  ostride=self.nchunk, nout=P * self.nchunk)
        else:
            jac = self._msm(Q, sel, P, self.c_sidx, self.c_bidx, None, self.nchunk, POLY_SIZE, self.bases,
                            ostride=self.nchunk, nout=P * self.nchunk)
        if not affine:
            return jac.view(P, self.nchunk, JAC)
        return self.normalize(jac).view(P, self.nchunk, AFF)

    def full_commitments(self, chunk_pts: torch.Tensor, jacobian: bool = False) -> torch.Tensor:
        """Full-vector commitment = sum of the chunk commitments -> affine [P, 16]
        (or Jacobian [P, 24] when `jacobian`, to be normalised on the host for the rows that enter a block)."""
        P, G = chunk_pts.shape[0], chunk_pts.shape[1]
        isj = chunk_pts.shape[-1] == JAC
        jac = torch.empty(P, JAC, dtype=torch.int32, device=chunk_pts.device)
        L = lib()
        if chunk_pts.is_cuda:
            fn = L.bsc_g1_sum_cols_jac if isj else L.bsc_g1_sum_cols
            check(fn(ptr(chunk_pts), P, G, ptr(jac), stream(chunk_pts.device)), "g1_sum_cols")
        else:
            (L.bsc_g1_sum_cols_jac_host if isj else L.bsc_g1_sum_cols_host)(ptr(chunk_pts), P, G, ptr(jac))
        return jac if jacobian else self.normalize(jac)

    @staticmethod
    def jac_rows_to_bytes(jac_host, rows) -> list:
        """Wire bytes of the selected rows of a host [N, 24] int32 Jacobian array, one shared inversion."""
        import numpy as np

        jac_host = np.ascontiguousarray(jac_host, dtype=np.int32)
        idx = np.asarray(rows, dtype=np.int32)
        out = np.empty((len(rows), 64), dtype=np.uint8)
        lib().bsc_g1_jac_to_bytes_batch_host(jac_host.ctypes.data_as(ctypes.c_void_p), idx.ctypes.data_as(ctypes.c_void_p),
                                             len(rows), out.ctypes.data_as(ctypes.c_void_p))
        return [out[i].tobytes() for i in range(len(rows))]

    def witnesses(self, Q: torch.Tensor, sel: torch.Tensor | None = None, affine: bool = True, count=None,
                  max_rows: int | None = None) -> torch.Tensor:
        """W[p, c, s] = KZG opening proof of chunk polynomial c of peer p at x_s -> [P, nchunk, S, 16 | 24].
        count (int32 [1] on the device) / max_rows: `sel` holds `count` live rows, at most `max_rows`; the kernel reads
        the count itself, so the approved set never has to visit the host.  Rows >= count are left unwritten."""
        P = Q.shape[0] if sel is None else sel.numel()
        if max_rows is not None:
            P = min(P, max_rows)
        if count is not None and self.device.type != "cuda":
            P = int(count[0])
            sel, count = sel[:P], None
        G = self.nchunk * self.S
        if P == 0:
            return torch.empty(0, self.nchunk, self.S, AFF if affine else JAC, dtype=torch.int32, device=self.device)
        if self.w_table is not None:
            jac = self._msm_table(Q, sel, P, self.w_sidx, self.w_bidx, None, G, POLY_SIZE - 1, self.w_table, ostride=G,
                                  nout=P * G, count=count)
        else:
            jac = self._msm(Q, sel, P, self.w_sidx, self.w_bidx, None, G, POLY_SIZE - 1, self.wtable, ostride=G,
                            nout=P * G)
        if not affine:
            return jac.view(P, self.nchunk, self.S, JAC)
        return self.normalize(jac).view(P, self.nchunk, self.S, AFF)

    def verify_commitment(self, Q_row: torch.Tensor, commitment: bytes) -> bool:
        """verifyCommitment (kyber.go:564-577): recompute the full-vector commitment and compare bytes."""
        C = self.chunk_commitments(Q_row.view(1, -1), affine=False)
        return self.to_bytes(self.full_commitments(C)) == commitment

    def verify_share_trapdoor(self, chunk_commit: bytes, witness: bytes, chunk: int, s: int, y: int) -> bool:
        """Opening check C - y*tau^(10c)*G == (tau - x_s)*W.  verifySecret (kyber.go:650-673) does this with a
        pairing because tau is secret; with the reference's public tau = 2 (publicKey.go:36) — or any tau this
        engine was built with — the same relation can be checked in G1 directly.  Used by tests and the
        peer-scale 'Shamir verify' sweep; an unknown-tau deployment uses `verify_share_pairing`."""
        from ..oracle import bn256 as O

        x = s - 10
        tau = self.tau % O.ORDER
        lhs = O.add(O.from_bytes(chunk_commit), O.neg(O.mul(O.GEN, (y * pow(tau, 10 * chunk, O.ORDER)) % O.ORDER)))
        rhs = O.mul(O.from_bytes(witness), (tau - x) % O.ORDER)
        return lhs == rhs

    def verify_share_pairing(self, chunk_commit: bytes, witness: bytes, chunk: int, s: int, y: int) -> bool:
        """verifySecret (kyber.go:650-673) proper: e(C, g2) == e(W, tau*g2 - x_s*g2) * e(PKG1[10c], g2)^y, using only
        public SRS points (g2 and tau*g2), so it also holds for a secret tau.  Host pairing, ~10 ms per check."""
        from ..crypto import pairing

        if not hasattr(self, "_g2_srs"):
            g2 = pairing.g2_generator()
            self._g2_srs = (g2, pairing.g2_mul(g2, self.tau))
        base = self.srs_bytes[64 * POLY_SIZE * chunk: 64 * POLY_SIZE * chunk + 64]
        return pairing.verify_secret(chunk_commit, witness, s - 10, y, self._g2_srs[0], self._g2_srs[1], base=base)

    def aggregate_by_scalars(self, Q: torch.Tensor, sel: torch.Tensor, count=None, affine: bool = False):
        """The miners' homomorphic aggregate over the contributing peers (kyber.go:244-287) computed from the SUMMED
        polynomial: sum_p C[p] and sum_p W[p] are the chunk commitments / witnesses of sum_p Q[p] (same group elements),
        so the aggregation is one more MSM row instead of one point addition per contributor and group.
        Returns (aggC [nchunk, 16|24], aggW [nchunk, S, 16|24]); count: int32 [1] device-resident live length of sel."""
        if Q.is_cuda:
            aggQ = torch.empty(1, Q.shape[1], dtype=torch.int64, device=Q.device)
            check(lib().bsc_sum_rows_i64_dc(ptr(Q), Q.stride(0), ptr(sel), sel.numel(), ptr(count), Q.shape[1], ptr(aggQ),
                                            stream(Q.device)), "sum_rows_i64")
        else:
            rows = sel if count is None else sel[: int(count[0])]
            aggQ = Q[rows.long()].sum(dim=0, keepdim=True) if rows.numel() else torch.zeros(1, Q.shape[1], dtype=torch.int64)
        aggC = self.chunk_commitments(aggQ, affine=affine)[0]
        aggW = self.witnesses(aggQ, affine=affine)[0]
        return aggC, aggW

    def aggregate(self, pts: torch.Tensor, sel: torch.Tensor, affine: bool = True, count=None) -> torch.Tensor:
        """Miner-side homomorphic add over contributing peers: pts [P, G..., 16|24] -> [G..., 16|24].
        count (int32 [1] on the device): live entries of `sel`, read by the kernel."""
        inner = pts.shape[1:-1]
        words = pts.shape[-1]
        G = 1
        for k in inner:
            G *= k
        jac = torch.empty(G, JAC, dtype=torch.int32, device=pts.device)
        L = lib()
        if count is not None and not pts.is_cuda:
            sel, count = sel[: int(count[0])], None
        if pts.is_cuda:
            fn = L.bsc_g1_sum_rows_jac_dc if words == JAC else L.bsc_g1_sum_rows_dc
            check(fn(ptr(pts), ptr(sel), sel.numel(), ptr(count), G, ptr(jac), stream(pts.device)), "g1_sum_rows")
        else:
            (L.bsc_g1_sum_rows_jac_host if words == JAC else L.bsc_g1_sum_rows_host)(ptr(pts), ptr(sel), sel.numel(), G,
                                                                                   ptr(jac))
        if not affine:
            return jac.view(*inner, JAC)
        return self.normalize(jac).view(*inner, AFF)
