"""Thin helpers shared by the op wrappers: pointer extraction, current stream, return-code checks."""
from __future__ import annotations

import ctypes

import torch

from .. import _native


def lib():
    return _native.load()


def ptr(t):
    if t is None:
        return None
    assert t.is_contiguous(), "native kernels take contiguous tensors"
    return ctypes.c_void_p(t.data_ptr())


def stream(device=None):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def check(rc: int, what: str):
    if rc != 0:
        raise RuntimeError(f"biscotti_b200 native call {what} failed with code {rc}")


def require_cuda_lib():
    """On a GPU box the CUDA path is the only path: never fall back silently."""
    if not _native.available():
        raise RuntimeError("libbiscotti_b200.so is not built; run `python -m biscotti_b200.build`")
