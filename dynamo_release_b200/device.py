"""Device plumbing: PyTorch is used only as allocator, stream owner and (for multi-GPU) rendezvous.

Every numeric step is a call into libdvf_b200 (see _lib.py); nothing here computes on the CPU.
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib


def require_cuda():
    if not torch.cuda.is_available():
        raise RuntimeError("dynamo_release_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")


def stream_ptr():
    return torch.cuda.current_stream().cuda_stream


def dev(a, dtype=torch.float64):
    """numpy array / torch tensor -> contiguous CUDA tensor of `dtype` (exact upcast, like scipy's cdist does)."""
    require_cuda()
    if isinstance(a, torch.Tensor):
        t = a
        if t.device.type != "cuda":
            t = t.to("cuda", non_blocking=True)
        if t.dtype != dtype:
            t = t.to(dtype)
        return t.contiguous()
    arr = np.asarray(a)
    if arr.dtype == np.float64 and dtype == torch.float64 and arr.flags.c_contiguous and arr.flags.writeable:
        return torch.from_numpy(arr).to("cuda", non_blocking=True)  # pinned sources copy asynchronously
    np_dtype = {torch.float64: np.float64, torch.float32: np.float32, torch.int32: np.int32,
                torch.int64: np.int64}[dtype]
    return torch.from_numpy(np.ascontiguousarray(arr, dtype=np_dtype)).to("cuda", non_blocking=True)


def ptr(t):
    return None if t is None else t.data_ptr()


def empty(shape, dtype=torch.float64):
    return torch.empty(shape, dtype=dtype, device="cuda")


def host(t, pinned=True):
    """CUDA tensor -> numpy array (through torch's cached pinned allocator, so repeated calls do not re-pin)."""
    if not pinned or t.numel() == 0:
        return t.cpu().numpy()
    buf = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
    buf.copy_(t, non_blocking=True)
    torch.cuda.current_stream().synchronize()
    return buf.numpy()


def pinned_like(a):
    """Copy a numpy array into pinned host memory (used by bench.py / callers that want async H2D)."""
    t = torch.empty(a.shape, dtype=torch.from_numpy(np.empty(0, a.dtype)).dtype, pin_memory=True)
    t.numpy()[...] = a
    return t.numpy()


def lib():
    require_cuda()
    return _lib.load()
