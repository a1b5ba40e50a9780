"""ctypes wrapper of the CNN local-step kernels (csrc/kernels_cnn.cu): one-convolution models (mnist_cnn,
cifar_cnn) and the two-convolution + max-pool model (lfw_cnn)."""
from __future__ import annotations

import ctypes

import torch

from .. import _native
from ._call import ptr, stream, check

_scratch = {}


def supports(kind: str) -> bool:
    """True when the native kernels for this CNN are in the loaded library (models.registry falls back to the PyTorch
    functional path otherwise - CPU runs and the CPU oracle)."""
    if kind not in ("mnist_cnn", "cifar_cnn", "lfw_cnn"):
        return False
    try:
        return hasattr(_native.load(), "bsc_cnn2_local_step" if kind == "lfw_cnn" else "bsc_cnn1_local_step")
    except Exception:
        return False


def local_step(kind: str, w32: torch.Tensor, Xb: torch.Tensor, yb: torch.Tensor, clip: float, out: torch.Tensor):
    """Xb: [P, B, IC*H*W] minibatch, yb: [P, B] int32, out: [P, ldu] <- -clip(grad)."""
    from ..models.registry import CNN_ARCH

    arch = CNN_ARCH[kind]
    if len(arch["convs"]) == 2:
        return _local_step2(arch, w32, Xb, yb, clip, out)
    c = arch["convs"][0]
    IC, H, W = arch["image"]
    P, B = Xb.shape[0], Xb.shape[1]
    L = _native.load()
    L.bsc_cnn1_scratch_floats.restype = ctypes.c_longlong
    n = L.bsc_cnn1_scratch_floats(P, B, IC, c.cout, c.k, c.pad, H, W, arch["n_classes"])
    key = (Xb.device, n)
    buf = _scratch.get(key)
    if buf is None:
        _scratch.clear()
        buf = torch.empty(n, dtype=torch.float32, device=Xb.device)
        _scratch[key] = buf
    Xc = Xb.contiguous().float()
    yc = yb.contiguous().int()
    check(L.bsc_cnn1_local_step(ptr(w32), ptr(Xc), ptr(yc), ptr(out), None, ptr(buf), P, B, IC, c.cout, c.k, c.pad, H, W,
                                arch["n_classes"], out.stride(0), ctypes.c_float(clip), stream(Xb.device)),
          "cnn1_local_step")
    return out


_eval_scratch = {}


def eval_errors(kind: str, w32: torch.Tensor, X: torch.Tensor, y: torch.Tensor):
    """Error rate of one flat weight vector over X [N, IC*H*W] with the one-conv model, on the local-step kernels
    (convolution + classifier partial sums) and an arg-max count; None when the shape has no native path."""
    from ..models.registry import CNN_ARCH

    arch = CNN_ARCH[kind]
    if len(arch["convs"]) != 1 or not X.is_cuda:
        return None
    c = arch["convs"][0]
    IC, H, W = arch["image"]
    N = X.shape[0]
    L = _native.load()
    L.bsc_cnn1_eval_scratch_floats.restype = ctypes.c_longlong
    n = L.bsc_cnn1_eval_scratch_floats(N, IC, c.cout, c.k, c.pad, H, W, arch["n_classes"])
    key = (X.device, n)
    buf = _eval_scratch.get(key)
    if buf is None:
        _eval_scratch.clear()
        buf = (torch.empty(n, dtype=torch.float32, device=X.device), torch.zeros(1, dtype=torch.int32, device=X.device))
        _eval_scratch[key] = buf
    Xc = X.contiguous().float()
    yc = y.contiguous().int()
    wc = w32.contiguous().float()
    rc = L.bsc_cnn1_eval_errors(ptr(wc), ptr(Xc), ptr(yc), ptr(buf[1]), ptr(buf[0]), N, IC, c.cout, c.k, c.pad, H, W,
                                arch["n_classes"], stream(X.device))
    if rc == -1:
        return None
    check(rc, "cnn1_eval_errors")
    return buf[1].float()[0] / N


def _scratch_buf(device, n):
    key = (device, n)
    buf = _scratch.get(key)
    if buf is None:
        _scratch.clear()
        buf = torch.empty(n, dtype=torch.float32, device=device)
        _scratch[key] = buf
    return buf


def _local_step2(arch, w32, Xb, yb, clip, out):
    """conv+ReLU+MaxPool2 twice, then the linear head (lfw_cnn_model.py:11-29)."""
    c1, c2 = arch["convs"]
    assert c1.pool == 2 and c2.pool == 2
    IC, H, W = arch["image"]
    P, B = Xb.shape[0], Xb.shape[1]
    L = _native.load()
    n = L.bsc_cnn2_scratch_floats(P, B, IC, H, W, c1.cout, c1.k, c1.pad, c2.cout, c2.k, c2.pad, arch["n_classes"])
    buf = _scratch_buf(Xb.device, n)
    Xc = Xb.contiguous().float()
    yc = yb.contiguous().int()
    check(L.bsc_cnn2_local_step(ptr(w32), ptr(Xc), ptr(yc), ptr(out), None, ptr(buf), P, B, IC, H, W, c1.cout, c1.k, c1.pad,
                                c2.cout, c2.k, c2.pad, arch["n_classes"], out.stride(0), ctypes.c_float(clip),
                                stream(Xb.device)), "cnn2_local_step")
    return out
