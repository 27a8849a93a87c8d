"""Shared helpers of the per-stage classes (CifHr, CifSeeds, CafScored, ...)."""
import ctypes as C

import numpy as np
import torch

from .. import _lib


def cuda_device():
    if not torch.cuda.is_available():
        raise RuntimeError('openpifpaf_b200 needs a CUDA device (B200); there is no CPU path')
    return torch.device('cuda', torch.cuda.current_device())


def to_device_f32(array, device=None):
    """numpy array / torch tensor -> contiguous float32 CUDA tensor (no copy if already one)."""
    device = device or cuda_device()
    if isinstance(array, torch.Tensor):
        t = array
    else:
        a = np.asarray(array)
        if a.dtype != np.float32:
            raise ValueError("Buffer dtype mismatch, expected 'float' but got %r" % a.dtype)
        t = torch.from_numpy(np.ascontiguousarray(a))
    if t.dtype != torch.float32:
        raise ValueError("Buffer dtype mismatch, expected 'float' but got %r" % t.dtype)
    return t.to(device, non_blocking=True).contiguous()


def ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def stream_ptr():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def shape_caps(B, F, L, H, W, stride, ann_cap=256):
    shape = _lib.Shape(B, F, L, H, W, stride)
    caps = _lib.Caps()
    _lib.check(_lib.lib().pp_default_caps(C.byref(shape), C.byref(caps)))
    caps.ann_cap = ann_cap
    return shape, caps


def workspace(shape, caps, device):
    n = _lib.lib().pp_workspace_bytes(C.byref(shape), C.byref(caps))
    if n == 0:
        raise _lib.PifPafError(_lib.PP_ERR_ARG, 'bad shape for workspace')
    return torch.empty((n,), dtype=torch.uint8, device=device), n


def raise_on_status(status):
    words = status.cpu().numpy() & _lib.PP_STATUS_OVERFLOW_MASK
    if words.any():
        from .engine import describe_status
        raise _lib.PifPafError(_lib.PP_ERR_CAPACITY, describe_status(words))
