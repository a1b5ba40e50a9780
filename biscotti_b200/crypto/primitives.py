"""Lottery, VRF, Schnorr and key generation — Python faces of the native runtime.

Reference: DistSys/vrf.go (lottery + VRF keys), DistSys/kyber.go:873-933 (Schnorr),
DistSys/publicKey.go + keyGeneration/generateBootstrapFile.go (SRS and per-peer keys, JSON-lines
bootstrap files), DistSys/main.go:486-565 (role encoding as a product of primes).
"""
from __future__ import annotations

import base64
import ctypes
import hashlib
import json
import os
from dataclasses import dataclass

import numpy as np
import torch

from .. import _native
from ..config import VERIFIER_PRIME, MINER_PRIME, NOISER_PRIME, BASE_PORT

c_i = ctypes.c_int


def _L():
    return _native.load()


def _buf(n):
    return ctypes.create_string_buffer(n)


# ---------------------------------------------------------------------------------------------
# hashing
# ---------------------------------------------------------------------------------------------
def sha256(data: bytes) -> bytes:
    out = _buf(32)
    _L().bsc_sha256(data, len(data), out)
    return out.raw


# ---------------------------------------------------------------------------------------------
# lottery (vrf.go:54-182) — bit-exact consensus rule
# ---------------------------------------------------------------------------------------------
def lottery_roles(block_hash: bytes, stake: list[int], n_verifiers: int, n_miners: int):
    """Verifiers then miners, drawn from the raw previous block hash (vrf.go:103-182)."""
    n = len(stake)
    st = (c_i * n)(*stake)
    v = (c_i * max(1, n_verifiers))()
    m = (c_i * max(1, n_miners))()
    nv, nm = c_i(0), c_i(0)
    rc = _L().bsc_lottery_roles(block_hash, len(block_hash), st, n, n_verifiers, n_miners, v, m, ctypes.byref(nv),
                                ctypes.byref(nm))
    if rc != 0:
        raise RuntimeError("lottery needs at least one ticket")
    return list(v[: nv.value]), list(m[: nm.value])


def lottery_noisers(vrf_output: bytes, stake: list[int], source: int, want: int):
    """Noisers for one worker from its private VRF output, itself excluded (vrf.go:54-100)."""
    n = len(stake)
    st = (c_i * n)(*stake)
    out = (c_i * max(1, want))()
    got = _L().bsc_lottery_noisers(vrf_output, len(vrf_output), st, n, source, want, out)
    return list(out[:got])


def role_ids(n: int, verifiers, miners, noisers=()) -> list[int]:
    """getRoles (main.go:486-527): product of role primes per node id."""
    r = [1] * n
    for v in verifiers:
        r[v] *= VERIFIER_PRIME
    for m in miners:
        r[m] *= MINER_PRIME
    for q in noisers:
        r[q] *= NOISER_PRIME
    return r


# ---------------------------------------------------------------------------------------------
# VRF
# ---------------------------------------------------------------------------------------------
class VRFKey:
    """One coniks-style VRF keypair (vrf.go:16-32 creates two per peer: roles + noise)."""

    def __init__(self, seed: bytes):
        assert len(seed) == 32
        self.seed = seed
        pk = _buf(32)
        _L().bsc_vrf_pubkey(seed, pk)
        self.public = pk.raw

    def prove(self, msg: bytes):
        vrf, proof = _buf(32), _buf(96)
        _L().bsc_vrf_prove(self.seed, msg, len(msg), vrf, proof)
        return vrf.raw, proof.raw

    def secret_scalar_words(self) -> np.ndarray:
        out = np.empty(8, dtype=np.uint32)
        _L().bsc_vrf_secret_scalar(self.seed, out.ctypes.data_as(ctypes.c_void_p))
        return out


def vrf_verify(public: bytes, msg: bytes, vrf: bytes, proof: bytes) -> bool:
    return bool(_L().bsc_vrf_verify(public, msg, len(msg), vrf, proof))


class BatchedVRF:
    """All peers' VRF outputs on one block hash in one launch: Gamma_p = x_p * H(m) on the device
    (ed25519_scalarmult_kernel), output_p = SHAKE256(Gamma_p || m) on the host."""

    def __init__(self, keys: list[VRFKey], device="cpu", kernel: str = "quad"):
        self.keys = keys
        self.kernel = kernel   # "quad": four lanes per scalar multiplication; "r25": one thread each
        self.device = torch.device(device)
        words = np.stack([k.secret_scalar_words() for k in keys]).astype(np.uint32)
        self.scalars = torch.from_numpy(words.view(np.int32)).to(self.device)

    def outputs(self, msg: bytes, sync: bool = True):
        L = _L()
        n = len(self.keys)
        hpt = torch.empty(32, dtype=torch.int32)
        L.bsc_vrf_hash_to_curve(msg, len(msg), ctypes.c_void_p(hpt.data_ptr()), None)
        gam = torch.empty(n, 32, dtype=torch.uint8, device=self.device)
        if self.device.type == "cuda":
            hd = hpt.to(self.device)
            rc = L.bsc_ed25519_scalarmult_batch(ctypes.c_void_p(self.scalars.data_ptr()), ctypes.c_void_p(hd.data_ptr()),
                                                1, n, ctypes.c_void_p(gam.data_ptr()),
                                                ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream))
            if rc:
                raise RuntimeError("ed25519_scalarmult_batch failed")
            self._pending = (gam, hd)
        else:
            L.bsc_ed25519_scalarmult_batch_host(ctypes.c_void_p(self.scalars.data_ptr()),
                                                ctypes.c_void_p(hpt.data_ptr()), 1, n, ctypes.c_void_p(gam.data_ptr()))
        if not sync:
            return gam
        return self.finish(gam, msg)

    # ---- asynchronous form used by the engine: launch on a side stream at the top of a block, collect
    # the noiser committees (one native call) when the noise phase needs them.
    def launch(self, msg: bytes):
        L = _L()
        n = len(self.keys)
        if self.device.type == "cuda" and hasattr(self, "_side"):
            hxy = self._hpt_pinned          # hash-to-curve writes straight into the pinned staging buffer
        else:
            hxy = torch.empty(64, dtype=torch.uint8)
        L.bsc_vrf_hash_to_curve_xy(msg, len(msg), ctypes.c_void_p(hxy.data_ptr()))
        if self.device.type != "cuda":
            gam = torch.empty(n, 32, dtype=torch.uint8)
            L.bsc_ed25519_scalarmult_r25_host(ctypes.c_void_p(self.scalars.data_ptr()), ctypes.c_void_p(hxy.data_ptr()),
                                              n, ctypes.c_void_p(gam.data_ptr()))
            return (gam, None, msg)
        if not hasattr(self, "_side"):
            torch.cuda.current_stream(self.device).synchronize()   # once: the scalars' upload has landed
            self._side = torch.cuda.Stream(self.device, priority=-1)   # on the block's critical path
            self._host_gam = torch.empty(n, 32, dtype=torch.uint8).pin_memory()
            self._hpt_pinned = torch.empty(64, dtype=torch.uint8).pin_memory()
            self._hd = torch.empty(64, dtype=torch.uint8, device=self.device)
            self._gam = torch.empty(n, 32, dtype=torch.uint8, device=self.device)
            self._xyz = torch.empty(n, 30, dtype=torch.int32, device=self.device)
            self._host_xyz = torch.empty(n, 30, dtype=torch.int32).pin_memory()
        if hxy is not self._hpt_pinned:
            self._hpt_pinned.copy_(hxy)
        # No wait on the compute stream: the kernel reads only the static secret scalars and the point copied here,
        # so it may run beside whatever the main stream still has queued (the prefetched local step / commitment MSM).
        with torch.cuda.stream(self._side):
            self._hd.copy_(self._hpt_pinned, non_blocking=True)
            if self.kernel == "quad":
                # projective results; the single shared inversion + compression happens on the host
                rc = L.bsc_ed25519_scalarmult_quad_xyz(ctypes.c_void_p(self.scalars.data_ptr()),
                                                       ctypes.c_void_p(self._hd.data_ptr()), n,
                                                       ctypes.c_void_p(self._xyz.data_ptr()),
                                                       ctypes.c_void_p(self._side.cuda_stream))
                self._host_xyz.copy_(self._xyz, non_blocking=True)
            else:
                rc = L.bsc_ed25519_scalarmult_r25(ctypes.c_void_p(self.scalars.data_ptr()),
                                                  ctypes.c_void_p(self._hd.data_ptr()), n,
                                                  ctypes.c_void_p(self._gam.data_ptr()),
                                                  ctypes.c_void_p(self._side.cuda_stream))
                self._host_gam.copy_(self._gam, non_blocking=True)
            if rc:
                raise RuntimeError("ed25519 scalar multiplication kernel failed")
            if not hasattr(self, "_ev"):
                self._ev = torch.cuda.Event()    # one launch in flight at a time: the event is reused
            ev = self._ev
            ev.record(self._side)
        return (self._host_gam, ev, msg)

    def launch_dev(self, msg: bytes):
        """GPU only: Gamma_p = x_p * H(msg) for every peer, left projective ON THE DEVICE: (xyz int32 [n, 30], event,
        msg).  The consumer is ops.chain.vrf_noisers (inversion, SHAKE256, stake lottery in one kernel), so the VRF
        outputs never visit the host.  Runs on this object's high-priority stream and waits for nothing."""
        L = _L()
        n = len(self.keys)
        if not hasattr(self, "_dside"):
            torch.cuda.current_stream(self.device).synchronize()   # once: the scalars' upload has landed
            self._dside = torch.cuda.Stream(self.device, priority=-3)
            self._dh_pinned = [torch.empty(64, dtype=torch.uint8).pin_memory() for _ in range(4)]
            self._dhd = [torch.empty(64, dtype=torch.uint8, device=self.device) for _ in range(4)]
            self._dxyz = [torch.empty(n, 30, dtype=torch.int32, device=self.device) for _ in range(2)]
            from ..ops.cmdlist import NativeEvent

            self._dev = [NativeEvent() for _ in range(2)]
            self._di = 0
        i = self._di
        self._di += 1
        hp, hd, xyz, ev = self._dh_pinned[i % 4], self._dhd[i % 4], self._dxyz[i % 2], self._dev[i % 2]
        # one native call: hash-to-curve on the host, 64-byte upload, the kernel, the event record (this sits between the
        # block hash and the next block's chain, every microsecond of interpreter here is on the critical path)
        rc = L.bsc_vrf_launch(msg, len(msg), ctypes.c_void_p(hp.data_ptr()), ctypes.c_void_p(hd.data_ptr()),
                              ctypes.c_void_p(self.scalars.data_ptr()), n, ctypes.c_void_p(xyz.data_ptr()),
                              0 if self.kernel == "quad4" else 1, ctypes.c_void_p(self._dside.cuda_stream),
                              ctypes.c_void_p(ev.handle))
        if rc:
            raise RuntimeError(f"VRF launch failed with code {rc}")
        return (xyz, ev, msg)

    def noiser_table(self, handle, stake: list[int], is_worker: np.ndarray, want: int) -> np.ndarray:
        """[n, want] int32 table of noiser ids (-1 = none) for every worker (vrf.go:54-100)."""
        gam, ev, msg = handle
        n = len(self.keys)
        if ev is not None:
            ev.synchronize()
            if self.kernel == "quad":
                _L().bsc_ed25519_compress_batch(ctypes.c_void_p(self._host_xyz.data_ptr()), n,
                                                ctypes.c_void_p(gam.data_ptr()))
        out = np.full((n, max(1, want)), -1, dtype=np.int32)
        st = np.asarray(stake, dtype=np.int32)
        iw = np.ascontiguousarray(is_worker, dtype=np.uint8)
        _L().bsc_noisers_batch(ctypes.c_void_p(gam.data_ptr()), msg, len(msg), st.ctypes.data_as(ctypes.c_void_p), n,
                               iw.ctypes.data_as(ctypes.c_void_p), want, out.ctypes.data_as(ctypes.c_void_p), None)
        return out

    def finish(self, gam: torch.Tensor, msg: bytes):
        L = _L()
        raw = gam.cpu().numpy().tobytes()
        outs = []
        o = _buf(32)
        for i in range(len(self.keys)):
            L.bsc_vrf_output_from_gamma(raw[32 * i: 32 * i + 32], msg, len(msg), o)
            outs.append(o.raw)
        return outs


# ---------------------------------------------------------------------------------------------
# Schnorr (kyber.go:873-933)
# ---------------------------------------------------------------------------------------------
def schnorr_pubkey(secret: bytes) -> bytes:
    out = _buf(64)
    _L().bsc_schnorr_pubkey(secret, out)
    return out.raw


def schnorr_sign(msg: bytes, secret: bytes, nonce: bytes | None = None) -> bytes:
    """sig = c || r (64 bytes).  nonce=None derives v deterministically from (secret, msg)."""
    if nonce is None:
        nonce = hashlib.sha256(b"bsc-schnorr-nonce" + secret + msg).digest()
    out = _buf(64)
    _L().bsc_schnorr_sign(msg, len(msg), secret, nonce, out)
    return out.raw


def schnorr_verify(msg: bytes, public: bytes, sig: bytes) -> bool:
    return bool(_L().bsc_schnorr_verify(msg, len(msg), public, sig))


# ---------------------------------------------------------------------------------------------
# key generation / bootstrap files (keyGeneration/generateBootstrapFile.go:47-191)
# ---------------------------------------------------------------------------------------------
@dataclass
class PeerKeys:
    schnorr_secret: bytes
    schnorr_public: bytes
    vrf_roles: VRFKey
    vrf_noise: VRFKey


def derive_peer_keys(n: int, seed: int) -> list[PeerKeys]:
    out = []
    for i in range(n):
        base = hashlib.sha256(f"bsc-peer-{seed}-{i}".encode()).digest()
        sk = hashlib.sha256(base + b"schnorr").digest()
        out.append(PeerKeys(sk, schnorr_pubkey(sk), VRFKey(hashlib.sha256(base + b"vrf-roles").digest()),
                            VRFKey(hashlib.sha256(base + b"vrf-noise").digest())))
    return out


def generate_srs_bytes(d: int, tau: int = 2) -> bytes:
    """PKG1[i] = tau^i * G as 64-byte wire points (publicKey.go:28-66, tau = 2 is the reference's constant)."""
    out = _buf(64 * d)
    _L().bsc_srs_g1_host(int(tau).to_bytes(32, "big"), d, out)
    return out.raw


def write_bootstrap_files(outdir: str, n_nodes: int, dims: int, hosts: list[str] | None = None, seed: int = 1,
                          tau: int = 2, nodes_per_host: int | None = None):
    """Write peersfile.txt, commitKey.json, pKeyG1.json in the reference's JSON-lines layout
    (one {Id,Pkey,Skey} object per line, base64 byte fields like Go's encoding/json): commitKey.json carries
    PKG1[i] (64-byte G1) as Pkey and PKG2[i] (129-byte G2) as Skey, generateBootstrapFile.go:102-133."""
    from . import pairing

    os.makedirs(outdir, exist_ok=True)
    srs = generate_srs_bytes(dims, tau)
    srs2 = pairing.srs_g2(tau, dims)
    with open(os.path.join(outdir, "commitKey.json"), "w") as f:
        for i in range(dims):
            pt = base64.b64encode(srs[64 * i: 64 * i + 64]).decode()
            f.write(json.dumps({"Id": i, "Pkey": pt, "Skey": base64.b64encode(srs2[i]).decode()}) + "\n")
    keys = derive_peer_keys(n_nodes, seed)
    with open(os.path.join(outdir, "pKeyG1.json"), "w") as f:
        for i, k in enumerate(keys):
            f.write(json.dumps({"Id": i, "Pkey": base64.b64encode(k.schnorr_public).decode(),
                                "Skey": base64.b64encode(k.schnorr_secret).decode()}) + "\n")
    hosts = hosts or ["127.0.0.1"]
    per = nodes_per_host or (n_nodes + len(hosts) - 1) // len(hosts)
    with open(os.path.join(outdir, "peersfile.txt"), "w") as f:
        i = 0
        for h in hosts:
            for k in range(per):
                if i >= n_nodes:
                    break
                f.write(f"{h}:{BASE_PORT + k}\n")
                i += 1
    return srs, keys


def load_commit_key(path: str) -> bytes:
    pts = []
    with open(path) as f:
        for line in f:
            if line.strip():
                pts.append(base64.b64decode(json.loads(line)["Pkey"]))
    return b"".join(pts)


def load_commit_key_g2(path: str) -> list[bytes]:
    """The G2 half of commitKey.json (PKG2[i] = tau^i * G2, honest.go:848-864)."""
    pts = []
    with open(path) as f:
        for line in f:
            if line.strip():
                pts.append(base64.b64decode(json.loads(line)["Skey"]))
    return pts


def g1_generator() -> bytes:
    out = _buf(64)
    _L().bsc_g1_generator(out)
    return out.raw
