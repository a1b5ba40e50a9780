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


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def stream(device=None):
    """Raw handle of the current CUDA stream of `device` (the C entry points take it as void*)."""
    if _raw_stream is not None:
        if isinstance(device, str):
            device = torch.device(device)
        idx = device.index if isinstance(device, torch.device) else device
        if idx is None:
            idx = torch.cuda.current_device()
        return ctypes.c_void_p(_raw_stream(idx))     # one C call, no Stream object
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def check(rc: int, what: str):
    if rc != 0:
        raise RuntimeError(f"biscotti_b200 native call {what} failed with code {rc}")


def require_cuda_lib():
    """On a GPU box the CUDA path is the only path: never fall back silently."""
    if not _native.available():
        raise RuntimeError("libbiscotti_b200.so is not built; run `python -m biscotti_b200.build`")
