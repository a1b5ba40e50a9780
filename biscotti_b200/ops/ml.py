"""Local training steps, DP-noise mixing, evaluation: CUDA kernels (kernels_ml.cu) with PyTorch CPU
restatements that double as the numerics oracle.

Reference semantics: ML/Pytorch/client.py:38-65 (softmax step), ML/code/logistic_model.py:92-140
(logistic step), ML/Pytorch/client_obj.py:97-98 + DistSys/main.go:1592-1660 (noise),
ML/Pytorch/client.py:136-172 (error / attack rate), client_obj.py:100-112 (RONI).
"""
from __future__ import annotations

import ctypes
import os

import torch

from ._call import lib, ptr, stream, check


def pad_to(x: int, m: int) -> int:
    return (x + m - 1) // m * m


# ---------------------------------------------------------------------------------------------
# softmax step
# ---------------------------------------------------------------------------------------------
def softmax_step_reference(w, X, y, n_classes: int, clip: float = 100.0):
    """fp32 PyTorch restatement: per peer one minibatch, CE(mean), global-norm clip, U = -grad.
    w: [C*F+C], X: [P,B,F], y: [P,B] -> U [P, C*F+C], loss [P]."""
    P, B, F = X.shape
    C = n_classes
    W = w[: C * F].view(C, F).float()
    b = w[C * F: C * F + C].float()
    logits = torch.einsum("pbf,cf->pbc", X.float(), W) + b
    lse = torch.logsumexp(logits, dim=-1)
    loss = (lse - logits.gather(-1, y.long().unsqueeze(-1)).squeeze(-1)).mean(dim=1)
    prob = torch.softmax(logits, dim=-1)
    onehot = torch.nn.functional.one_hot(y.long(), C).float()
    dl = (prob - onehot) / B
    gW = torch.einsum("pbc,pbf->pcf", dl, X.float())
    gb = dl.sum(dim=1)
    g = torch.cat([gW.reshape(P, -1), gb], dim=1)
    total = g.norm(dim=1, keepdim=True)
    coef = torch.clamp(clip / (total + 1e-6), max=1.0)
    return -(g * coef), loss


def softmax_step(w, X, y, idx, n_classes: int, clip: float = 100.0, out=None, ldu=None):
    """X: [P,Ns,F] shard (idx [P,B] selects the minibatch) or [P,B,F] with idx=None."""
    P, Ns, F = X.shape
    C = n_classes
    d = C * F + C
    ldu = ldu or d
    B = idx.shape[1] if idx is not None else Ns
    if not X.is_cuda:
        Xb = X if idx is None else torch.gather(X, 1, idx.long().unsqueeze(-1).expand(P, B, F))
        yb = y if idx is None else torch.gather(y, 1, idx.long())
        U, loss = softmax_step_reference(w, Xb, yb, C, clip)
        if out is None:
            out = torch.zeros(P, ldu, dtype=torch.float32)
        out[:, :d] = U
        return out, loss
    if out is None:
        out = torch.empty(P, ldu, dtype=torch.float32, device=X.device)
    loss = torch.empty(P, dtype=torch.float32, device=X.device)
    check(lib().bsc_local_softmax_step(ptr(w), ptr(X), ptr(y), ptr(idx), ptr(out), ptr(loss), P, Ns, B, F, C,
                                       out.stride(0), ctypes.c_float(clip), stream(X.device)), "local_softmax_step")
    return out, loss


# ---------------------------------------------------------------------------------------------
# logistic step
# ---------------------------------------------------------------------------------------------
def logistic_step_reference(w, X, y, alpha: float, lam: float):
    P, B, F = X.shape
    yXw = y.float() * torch.einsum("pbf,f->pb", X.float(), w.float())
    res = -y.float() / (1.0 + torch.exp(yXw))
    g = torch.einsum("pbf,pb->pf", X.float(), res) / B + lam * w.float()
    return -alpha * g


def logistic_step(w, X, y, idx, alpha, lam, own_noise_scale=0.0, seed=0, peer_ids=None, slot=0, out=None, ldu=None):
    P, Ns, F = X.shape
    ldu = ldu or F
    B = idx.shape[1] if idx is not None else Ns
    if not X.is_cuda:
        Xb = X if idx is None else torch.gather(X, 1, idx.long().unsqueeze(-1).expand(P, B, F))
        yb = y if idx is None else torch.gather(y, 1, idx.long())
        U = logistic_step_reference(w, Xb, yb, alpha, lam)
        if own_noise_scale != 0.0:
            ids = peer_ids.tolist() if peer_ids is not None else list(range(P))
            nz = torch.stack([noise_row(F, seed ^ 0x6F776E6E, i, slot, own_noise_scale) for i in ids])
            U = U + nz
        if out is None:
            out = torch.zeros(P, ldu, dtype=torch.float32)
        out[:, :F] = U
        return out
    if out is None:
        out = torch.empty(P, ldu, dtype=torch.float32, device=X.device)
    check(lib().bsc_local_logistic_step(ptr(w), ptr(X), ptr(y), ptr(idx), ptr(out), P, Ns, B, F, out.stride(0),
                                        ctypes.c_float(alpha), ctypes.c_float(lam), ctypes.c_float(own_noise_scale),
                                        ctypes.c_ulonglong(seed & (2**64 - 1)), ptr(peer_ids), slot, stream(X.device)),
          "local_logistic_step")
    return out


# ---------------------------------------------------------------------------------------------
# noise
# ---------------------------------------------------------------------------------------------
def noise_row(d: int, seed: int, peer: int, slot: int, scale: float, device="cpu"):
    """One row of a peer's (virtual) presampled noise table: scale * N(0,1), Philox counter RNG."""
    out = torch.empty(d, dtype=torch.float32, device=device)
    if out.is_cuda:
        check(lib().bsc_noise_row(ptr(out), d, ctypes.c_ulonglong(seed & (2**64 - 1)), peer, slot,
                                  ctypes.c_float(scale), stream(out.device)), "noise_row")
    else:
        lib().bsc_noise_row_host(ptr(out), d, ctypes.c_ulonglong(seed & (2**64 - 1)), peer, slot, ctypes.c_float(scale))
    return out


def noise_mix(U, noisers, seed: int, slot: int, scale: float, d: int, want_planes: bool = True, ldh=None,
              out_hi=None, out_lo=None, dst_row=None, want_un: bool = True, in_place: bool = False):
    """Un = U + mean_q noise(q); returns (Un fp32 [P,ldu], hi bf16 [P,ldh], lo bf16 [P,ldh], norms [P]).
    noisers: int32 [P, nn] global ids (-1 = unused slot)."""
    P, ldu = U.shape
    ldh = ldh or pad_to(d, 64)
    nn = noisers.shape[1] if noisers is not None else 0
    if not U.is_cuda:
        Un = U if in_place else U.clone()
        if nn:
            for p in range(P):
                ids = [int(q) for q in noisers[p].tolist() if q >= 0]
                if ids:
                    acc = sum(noise_row(d, seed, q, slot, 1.0) for q in ids)
                    Un[p, :d] += acc * (scale / len(ids))
        hi = Un[:, :d].to(torch.bfloat16)
        lo = (Un[:, :d] - hi.float()).to(torch.bfloat16)
        pad = ldh - d
        if pad:
            hi = torch.nn.functional.pad(hi, (0, pad))
            lo = torch.nn.functional.pad(lo, (0, pad))
        return Un, hi.contiguous(), lo.contiguous(), (Un[:, :d] ** 2).sum(dim=1)
    Un = U if in_place else (torch.empty_like(U) if want_un else None)   # in place: each element is read and written once
    if out_hi is not None:      # caller-provided planes (rows >= P), e.g. views into symmetric peer memory
        hi, lo = out_hi, out_lo
        ldh = hi.stride(0)
    else:
        hi = torch.empty(P, ldh, dtype=torch.bfloat16, device=U.device) if want_planes else None
        lo = torch.empty(P, ldh, dtype=torch.bfloat16, device=U.device) if want_planes else None
    norms = torch.empty(P, dtype=torch.float32, device=U.device)
    if dst_row is not None:
        # (CUDA only) planes / norms of peer p go to row dst_row[p] (< 0: none): compact, Krum-ordered candidates
        check(lib().bsc_noise_mix_rows(ptr(U), ptr(noisers), nn, ctypes.c_ulonglong(seed & (2**64 - 1)), slot,
                                       ctypes.c_float(scale), ptr(Un), ptr(hi), ptr(lo), ptr(norms), P, d, ldu, ldh,
                                       ptr(dst_row), stream(U.device)), "noise_mix")
        return Un, hi, lo, norms
    check(lib().bsc_noise_mix(ptr(U), ptr(noisers), nn, ctypes.c_ulonglong(seed & (2**64 - 1)), slot,
                              ctypes.c_float(scale), ptr(Un), ptr(hi), ptr(lo), ptr(norms), P, d, ldu, ldh,
                              stream(U.device)), "noise_mix")
    return Un, hi, lo, norms


# ---------------------------------------------------------------------------------------------
# evaluation
# ---------------------------------------------------------------------------------------------
def eval_softmax_errors(Wc, base, X, y, n_classes: int, wrong_out=None):
    """Misclassification rate of each candidate row of Wc (+ base) on (X, y). Wc: [M, >=d].
    wrong_out (CUDA): zero-filled int32 [M] to accumulate into; the raw counts are returned instead of rates."""
    M = Wc.shape[0]
    N, F = X.shape
    C = n_classes
    if not X.is_cuda:
        W = Wc[:, : C * F + C].float()
        if base is not None:
            W = W + base[: C * F + C].float()
        logits = torch.einsum("nf,mcf->mnc", X.float(), W[:, : C * F].reshape(M, C, F)) + W[:, None, C * F:]
        return (logits.argmax(-1) != y.long()[None]).float().mean(dim=1)
    wrong = torch.zeros(M, dtype=torch.int32, device=X.device) if wrong_out is None else wrong_out
    # tcgen05 path for batches of candidates (RONI, sweeps): 10x at M = 70, 28x at M = 280 (124 TFLOP/s useful).  A
    # single candidate is 25 MFLOP and three launches on that path vs one on the SIMT kernel (54 vs 43 us), so the
    # per-block test-error / attack-rate evaluation stays on the SIMT kernel.  BISCOTTI_EVAL_IMPL=tc|simt forces one.
    impl = os.environ.get("BISCOTTI_EVAL_IMPL", "auto")
    if C <= 16 and hasattr(lib(), "bsc_eval_tc") and (impl == "tc" or (impl == "auto" and M >= 4)):
        _eval_softmax_tc(Wc, base, X, y, wrong, M, N, F, C)
    else:
        check(lib().bsc_eval_softmax(ptr(Wc), ptr(base), ptr(X), ptr(y), ptr(wrong), M, N, F, C, Wc.stride(0),
                                     stream(X.device)), "eval_softmax")
    return wrong if wrong_out is not None else wrong.float() / N


_x_planes: dict = {}     # sample matrices are static (test / attack sets): their bf16x3 planes are built once
_w_scratch: dict = {}


def _eval_softmax_tc(Wc, base, X, y, wrong, M, N, F, C):
    """tcgen05 path (csrc/gram_tc.cu eval_tc_kernel): logits GEMM + bias + arg-max + error count in one kernel."""
    L = lib()
    dev = X.device
    ldk = pad_to(F, 64)
    key = (X.data_ptr(), N, F, X._version)
    planes = _x_planes.get(key)
    if planes is None:
        if len(_x_planes) > 16:
            _x_planes.clear()
        rows_pad = pad_to(N, 128)
        hi = torch.empty(rows_pad, ldk, dtype=torch.bfloat16, device=dev)
        lo = torch.empty_like(hi)
        check(L.bsc_split_planes(ptr(X), None, N, 1, 1, F, 0, F, ptr(hi), ptr(lo), ldk, rows_pad, None, 0, stream(dev)),
              "split_planes")
        planes = (hi, lo, rows_pad, X)     # keep X alive so the pointer key stays valid
        _x_planes[key] = planes
    b_rows = pad_to(16 * M, 128)
    wk = (dev, b_rows, ldk)
    scr = _w_scratch.get(wk)
    if scr is None:
        scr = (torch.empty(b_rows, ldk, dtype=torch.bfloat16, device=dev), torch.empty(b_rows, ldk, dtype=torch.bfloat16, device=dev),
               torch.empty(b_rows, dtype=torch.float32, device=dev))
        _w_scratch[wk] = scr
    whi, wlo, bias = scr
    check(L.bsc_split_planes(ptr(Wc), ptr(base), M, C, 16, Wc.stride(0), F, F, ptr(whi), ptr(wlo), ldk, b_rows, ptr(bias),
                             C * F, stream(dev)), "split_planes")
    check(L.bsc_eval_tc(ptr(planes[0]), ptr(planes[1]), planes[2], ptr(whi), ptr(wlo), b_rows, ldk, ptr(bias), ptr(y),
                        ptr(wrong), N, M, stream(dev)), "eval_tc")


def eval_logistic_errors(Wc, base, X, y, wrong_out=None):
    M = Wc.shape[0]
    N, F = X.shape
    if not X.is_cuda:
        W = Wc[:, :F].float()
        if base is not None:
            W = W + base[:F].float()
        pred = torch.where(X.float() @ W.t() >= 0, 1.0, -1.0)  # [N, M]
        return (pred != y.float()[:, None]).float().mean(dim=0)
    wrong = torch.zeros(M, dtype=torch.int32, device=X.device) if wrong_out is None else wrong_out
    check(lib().bsc_eval_logistic(ptr(Wc), ptr(base), ptr(X), ptr(y), ptr(wrong), M, N, F, Wc.stride(0),
                                  stream(X.device)), "eval_logistic")
    return wrong if wrong_out is not None else wrong.float() / N
