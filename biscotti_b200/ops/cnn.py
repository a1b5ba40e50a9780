"""Hand-written sm_100a CNN local-step kernels (K2).  `supports()` gates on the kernels being built
into libbiscotti_b200.so; models.registry falls back to the PyTorch functional path otherwise."""
from __future__ import annotations

from .. import _native


def supports(kind: str) -> bool:
    if kind not in ("mnist_cnn", "cifar_cnn", "lfw_cnn"):
        return False
    try:
        return hasattr(_native.load(), "bsc_cnn2_local_step" if kind == "lfw_cnn" else "bsc_cnn1_local_step")
    except Exception:
        return False


def local_step(kind, w32, Xb, yb, clip, out):
    from . import cnn_native

    return cnn_native.local_step(kind, w32, Xb, yb, clip, out)
