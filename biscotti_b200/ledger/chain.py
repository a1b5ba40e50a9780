"""Python face of the native ledger (csrc/ledger.cpp).

Field / rule parity: DistSys/block.go, blockData.go, update.go, blockchain.go, honest.go:631-653.
"""
from __future__ import annotations

import os

import ctypes
from dataclasses import dataclass, field

import numpy as np

from .. import _native

c_i, c_ll, c_p, c_sz = ctypes.c_int, ctypes.c_longlong, ctypes.c_void_p, ctypes.c_size_t


def _lib():
    L = _native.load()
    if not getattr(L, "_ledger_declared", False):
        L.bsc_chain_new.restype = c_p
        L.bsc_chain_new.argtypes = [c_i]
        L.bsc_chain_free.argtypes = [c_p]
        L.bsc_chain_len.argtypes = [c_p]
        L.bsc_chain_dim.argtypes = [c_p]
        L.bsc_chain_latest_iteration.restype = c_ll
        L.bsc_chain_latest_iteration.argtypes = [c_p]
        L.bsc_block_new.restype = c_p
        L.bsc_block_new.argtypes = [c_ll, c_p, c_i]
        L.bsc_block_free.argtypes = [c_p]
        L.bsc_block_add_update.argtypes = [c_p, c_i, c_i, c_i, c_p, c_i, c_p, c_i, c_p, c_i, c_p, c_i, c_p, c_i, c_i]
        L.bsc_block_set_stake.argtypes = [c_p, c_p, c_p, c_i]
        L.bsc_block_add_commit_updates.argtypes = [c_p, c_i, c_p, c_i, c_p, c_i, c_p, c_i, c_i]
        L.bsc_chain_append.argtypes = [c_p, c_p, c_ll]
        L.bsc_chain_block_index.argtypes = [c_p, c_ll]
        L.bsc_chain_hash.argtypes = [c_p, c_i, c_p]
        L.bsc_chain_prev_hash.argtypes = [c_p, c_i, c_p, c_p]
        L.bsc_chain_weights.argtypes = [c_p, c_i, c_p]
        L.bsc_chain_block_info.restype = c_ll
        L.bsc_chain_block_info.argtypes = [c_p, c_i, c_p, c_p, c_p]
        L.bsc_chain_block_stake.argtypes = [c_p, c_i, c_p, c_p]
        L.bsc_chain_update_info.argtypes = [c_p, c_i, c_i, c_p, c_p, c_p, c_p]
        L.bsc_chain_verify.argtypes = [c_p]
        L.bsc_chain_serialize.restype = c_sz
        L.bsc_chain_serialize.argtypes = [c_p, c_p, c_sz]
        L.bsc_chain_deserialize.restype = c_p
        L.bsc_chain_deserialize.argtypes = [ctypes.c_char_p, c_sz]
        L.bsc_chain_block_serialize.restype = c_sz
        L.bsc_chain_block_serialize.argtypes = [c_p, c_i, c_p, c_sz]
        L.bsc_chain_accept_block.argtypes = [c_p, ctypes.c_char_p, c_sz]
        L.bsc_chain_dump.restype = c_p
        L.bsc_chain_dump.argtypes = [c_p, c_p]
        L.bsc_chain_equal.argtypes = [c_p, c_p]
        L.bsc_go_float.restype = ctypes.c_char_p
        L.bsc_go_float.argtypes = [ctypes.c_double]
        L._ledger_declared = True
    return L


@dataclass
class UpdateRecord:
    """update.go:13-22.  In secure-aggregation blocks only source/iteration/commitment/signatures are set."""
    source: int
    iteration: int
    commitment: bytes = b""
    delta: np.ndarray | None = None
    noise: np.ndarray | None = None
    noised: np.ndarray | None = None
    accepted: bool = True
    signatures: list = field(default_factory=list)


def _f64(a):
    if a is None:
        return None, 0
    a = np.ascontiguousarray(a, dtype=np.float64)
    return a, a.size


class Blockchain:
    def __init__(self, dim: int | None = None, _handle=None):
        self._L = _lib()
        self._h = _handle if _handle is not None else self._L.bsc_chain_new(dim)

    def __del__(self):
        try:
            if self._h:
                self._L.bsc_chain_free(self._h)
                self._h = None
        except Exception:
            pass

    def __len__(self):
        return self._L.bsc_chain_len(self._h)

    @property
    def dim(self) -> int:
        return self._L.bsc_chain_dim(self._h)

    @property
    def latest_iteration(self) -> int:
        return int(self._L.bsc_chain_latest_iteration(self._h))

    def latest_hash(self) -> bytes:
        return self.block_hash(-1)

    def block_hash(self, index: int) -> bytes:
        buf = ctypes.create_string_buffer(32)
        self._L.bsc_chain_hash(self._h, index, buf)
        return buf.raw

    def prev_hash(self, index: int) -> bytes:
        buf = ctypes.create_string_buffer(32)
        n = c_i(0)
        self._L.bsc_chain_prev_hash(self._h, index, buf, ctypes.byref(n))
        return buf.raw[: n.value]

    def weights(self, index: int = -1) -> np.ndarray:
        """getLatestGradient / getLatestBlockModel (blockchain.go:31-41): a copy of GlobalW."""
        out = np.empty(self.dim, dtype=np.float64)
        self._L.bsc_chain_weights(self._h, index, out.ctypes.data_as(c_p))
        return out

    def get_block_index(self, iteration: int) -> int:
        """getBlock (blockchain.go:77-96): index of the block for `iteration`, -1 if absent; raises on the
        'multiple iterations appended' inconsistency the reference aborts on."""
        r = self._L.bsc_chain_block_index(self._h, iteration)
        if r == -2:
            raise RuntimeError("blocks for multiple iterations have been appended")
        return r

    def block_info(self, index: int = -1) -> dict:
        it, nu, ns = c_ll(0), c_i(0), c_i(0)
        ts = self._L.bsc_chain_block_info(self._h, index, ctypes.byref(it), ctypes.byref(nu), ctypes.byref(ns))
        ids = (c_i * ns.value)()
        vals = (c_i * ns.value)()
        self._L.bsc_chain_block_stake(self._h, index, ids, vals)
        ups = []
        for u in range(nu.value):
            src, acc, nsig = c_i(0), c_i(0), c_i(0)
            cbuf = ctypes.create_string_buffer(64)
            dl = self._L.bsc_chain_update_info(self._h, index, u, ctypes.byref(src), ctypes.byref(acc), cbuf,
                                               ctypes.byref(nsig))
            ups.append(dict(source=src.value, accepted=bool(acc.value), commitment=cbuf.raw, nsigs=nsig.value,
                            delta_len=dl))
        return dict(timestamp=int(ts), iteration=int(it.value), updates=ups, stake=dict(zip(ids, vals)),
                    hash=self.block_hash(index), prev_hash=self.prev_hash(index))

    def add_block(self, iteration: int, weights, updates=(), stake: dict | None = None, timestamp: int = -1) -> bytes:
        """NewBlock + AddBlock (block.go:30-44, blockchain.go:12-16).  Empty `updates` => Timestamp 0 so every
        peer mints the same empty block.  Returns the new block's hash."""
        w, d = _f64(weights)
        b = self._L.bsc_block_new(iteration, w.ctypes.data_as(c_p), d)
        simple = len(updates) > 0 and all(u.delta is None and u.noise is None and u.noised is None and u.accepted
                                          and len(u.commitment) == len(updates[0].commitment)
                                          and len(u.signatures) == len(updates[0].signatures) for u in updates)
        if simple:
            # commitment-only updates of a secure-aggregation block: one native call for all of them
            src = np.fromiter((u.source for u in updates), dtype=np.int32, count=len(updates))
            cm = b"".join(u.commitment for u in updates)
            ns = len(updates[0].signatures)
            sg = b"".join(s_ for u in updates for s_ in u.signatures)
            sl = len(updates[0].signatures[0]) if ns else 0
            self._L.bsc_block_add_commit_updates(b, len(updates), src.ctypes.data_as(c_p), updates[0].iteration,
                                                 cm or None, len(updates[0].commitment), sg or None, ns, sl)
            updates = ()
        for u in updates:
            dl, ndl = _f64(u.delta)
            nz, nnz = _f64(u.noise)
            nd, nnd = _f64(u.noised)
            sigs = b"".join(u.signatures)
            siglen = len(u.signatures[0]) if u.signatures else 0
            self._L.bsc_block_add_update(
                b, u.source, u.iteration, int(u.accepted), u.commitment or None, len(u.commitment or b""),
                dl.ctypes.data_as(c_p) if dl is not None else None, ndl,
                nz.ctypes.data_as(c_p) if nz is not None else None, nnz,
                nd.ctypes.data_as(c_p) if nd is not None else None, nnd,
                sigs or None, len(u.signatures), siglen)
        if stake is not None and len(stake):
            if isinstance(stake, dict):
                ids = np.fromiter(stake.keys(), dtype=np.int32, count=len(stake))
                vals = np.fromiter(stake.values(), dtype=np.int32, count=len(stake))
            else:                                   # dense list: node id == position
                vals = np.asarray(stake, dtype=np.int32)
                ids = np.arange(len(vals), dtype=np.int32)
            self._L.bsc_block_set_stake(b, ids.ctypes.data_as(c_p), vals.ctypes.data_as(c_p), len(vals))
        self._L.bsc_chain_append(self._h, b, timestamp)
        return self.latest_hash()

    def add_commit_block(self, iteration: int, weights, sources, commitments: bytes, signatures: bytes, n_sigs: int,
                         stake, timestamp: int = -1) -> bytes:
        """add_block for the common secure-aggregation block, without per-update Python objects: every update is
        (source, iteration, 64-byte commitment, n_sigs 64-byte signatures, accepted).  Byte-identical to add_block
        with the equivalent UpdateRecord list."""
        w, d = _f64(weights)
        b = self._L.bsc_block_new(iteration, w.ctypes.data_as(c_p), d)
        src = np.ascontiguousarray(sources, dtype=np.int32)
        n = len(src)
        if n:
            self._L.bsc_block_add_commit_updates(b, n, src.ctypes.data_as(c_p), iteration, commitments or None,
                                                 len(commitments) // n, signatures or None, n_sigs, 64 if n_sigs else 0)
        vals = np.asarray(stake, dtype=np.int32)
        ids = np.arange(len(vals), dtype=np.int32)
        self._L.bsc_block_set_stake(b, ids.ctypes.data_as(c_p), vals.ctypes.data_as(c_p), len(vals))
        self._L.bsc_chain_append(self._h, b, timestamp)
        return self.latest_hash()

    def append_packed(self, iteration: int, weights, sources, pack: np.ndarray, n_sigs: int, stake: list, stake_unit: int,
                      timestamp: int = -1) -> bytes:
        """The tail of a secure-aggregation block in one native call (ledger.cpp: bsc_chain_append_packed): `pack` is
        the [n_peers, 64*(1+k)] array of commitment || approvals rows (row = peer id), `sources` the contributors.
        Credits stake_unit to every contributor in `stake` (a Python list, updated in place), seals, appends and
        returns the block hash.  Byte-identical to add_commit_block with the same content."""
        w, d = _f64(weights)
        src = np.ascontiguousarray(sources, dtype=np.int32)
        pk = np.ascontiguousarray(pack, dtype=np.uint8)
        st = np.asarray(stake, dtype=np.int32)
        out = ctypes.create_string_buffer(32)
        self._L.bsc_chain_append_packed(self._h, iteration, w.ctypes.data_as(c_p), d, src.ctypes.data_as(c_p), len(src),
                                        pk.ctypes.data_as(c_p), pk.strides[0], n_sigs, st.ctypes.data_as(c_p), len(st),
                                        stake_unit, timestamp, out)
        stake[:] = st.tolist()
        return out.raw

    def verify(self) -> bool:
        return bool(self._L.bsc_chain_verify(self._h))

    def serialize(self) -> bytes:
        n = self._L.bsc_chain_serialize(self._h, None, 0)
        buf = ctypes.create_string_buffer(n)
        self._L.bsc_chain_serialize(self._h, buf, n)
        return buf.raw

    @classmethod
    def deserialize(cls, data: bytes) -> "Blockchain":
        h = _lib().bsc_chain_deserialize(data, len(data))
        if not h:
            raise ValueError("malformed chain bytes")
        return cls(_handle=h)

    def serialize_block(self, index: int = -1) -> bytes:
        n = self._L.bsc_chain_block_serialize(self._h, index, None, 0)
        buf = ctypes.create_string_buffer(n)
        self._L.bsc_chain_block_serialize(self._h, index, buf, n)
        return buf.raw

    def accept_block(self, data: bytes) -> int:
        """RegisterBlock/processBlock (main.go:1238-1330): 1 appended, 2 replaced an empty block, 0 ignored, <0 rejected."""
        return self._L.bsc_chain_accept_block(self._h, data, len(data))

    def dump(self) -> str:
        """PrintChain (blockchain.go:43-53) — the text the localTest equality check diffs."""
        n = c_sz(0)
        p = self._L.bsc_chain_dump(self._h, ctypes.byref(n))
        return ctypes.string_at(p, n.value).decode()

    def __eq__(self, other):
        return isinstance(other, Blockchain) and bool(self._L.bsc_chain_equal(self._h, other._h))

    def save(self, path: str):
        """Atomic: the bytes go to a temporary file that is renamed over `path`, so a process killed in the middle of a
        checkpoint leaves the previous checkpoint intact."""
        tmp = f"{path}.tmp.{os.getpid()}"
        with open(tmp, "wb") as f:
            f.write(self.serialize())
            f.flush()
            os.fsync(f.fileno())
        os.replace(tmp, path)

    @classmethod
    def load(cls, path: str) -> "Blockchain":
        with open(path, "rb") as f:
            return cls.deserialize(f.read())


def go_float(v: float) -> str:
    return _lib().bsc_go_float(float(v)).decode()
