"""Inference epilogue of the fused head networks on the GPU.

The reference computes the decoder's input fields from the raw head convolution in
``CompositeFieldFused.forward`` (``network/heads.py:406-448``: PixelShuffle(2) +
crop, sigmoid, exp, four channel-group views) followed by
``CifCafCollector.forward`` (``network/heads.py:65-88``: concatenation, index add,
CAF channel reorder) -- roughly ten eager passes over the field tensors.  Here the
same result is one kernel (``pp_head_epilogue``, csrc/head.cu), writing the
``(B, n_fields, 5 | 9, H, W)`` layout ``CifCaf.decode_batch`` reads.

Names follow the reference so a model's tail can be swapped by import.
"""
import torch

from .. import _lib
from ..decoder import _stage


def index_field_torch(shape, *, device=None):
    """(2, H, W) cell index, channel 0 = x (column), 1 = y (row); heads.py:15-26."""
    ys, xs = torch.meshgrid(torch.arange(shape[0], dtype=torch.float32, device=device),
                            torch.arange(shape[1], dtype=torch.float32, device=device),
                            indexing='ij')
    return torch.stack((xs, ys), dim=0)


class FusedHeadEpilogue:
    """``FusedHeadEpilogue(n_fields, n_vectors, quad)(x)``: raw convolution output
    ``(B, n_fields * (1 + 4 n_vectors) * 4**quad, h, w)`` of one
    ``CompositeFieldFused`` (eval mode) to its collected field
    ``(B, n_fields, 1 + 4 n_vectors, H, W)``."""

    def __init__(self, n_fields, n_vectors, quad=1):
        if quad not in (0, 1):
            raise NotImplementedError('quad must be 0 or 1 (heads.py:367 uses 1)')
        self.n_fields, self.n_vectors, self.quad = int(n_fields), int(n_vectors), int(quad)

    @classmethod
    def from_meta(cls, meta, quad=1):
        """``meta``: an IntensityMeta / AssociationMeta-like object (heads.py:217-251)."""
        if meta.n_confidences != 1 or meta.n_scales != meta.n_vectors:
            raise NotImplementedError('only cif / caf shaped heads (1 confidence, one scale per vector)')
        return cls(meta.n_fields, meta.n_vectors, quad)

    def out_shape(self, x):
        B, C, h, w = x.shape
        nch = 1 + 4 * self.n_vectors
        if C != self.n_fields * nch * (4 ** self.quad):
            raise ValueError('expected %d channels, got %d' % (self.n_fields * nch * 4 ** self.quad, C))
        H, W = (2 * h - 1, 2 * w - 1) if self.quad else (h, w)
        return (B, self.n_fields, nch, H, W)

    def __call__(self, x, out=None):
        if not x.is_cuda:
            raise _lib.PifPafError(_lib.PP_ERR_ARG, 'FusedHeadEpilogue runs on CUDA tensors only (no CPU path)')
        x = x.detach()
        if x.dtype != torch.float32 or not x.is_contiguous():
            x = x.float().contiguous()
        shape = self.out_shape(x)
        if out is None:
            out = torch.empty(shape, dtype=torch.float32, device=x.device)
        elif tuple(out.shape) != shape or out.dtype != torch.float32 or not out.is_contiguous():
            raise ValueError('out must be contiguous float32 of shape %s' % (shape,))
        if shape[0] == 0:
            return out
        _lib.check(_lib.lib().pp_head_epilogue(
            _stage.ptr(x), _stage.ptr(out), shape[0], self.n_fields, self.n_vectors,
            x.shape[2], x.shape[3], self.quad, _stage.stream_ptr()))
        return out


class CifCafCollector:
    """Raw head outputs -> (cif, caf) decoder fields (heads.py:28-88).

    ``CifCafCollector(cif_meta, caf_meta)(x_cif, x_caf)`` with the two raw
    convolution outputs.  Only the single-head form is built (the reference can
    concatenate several heads along the field axis)."""

    def __init__(self, cif_meta, caf_meta, quad=1):
        self.cif = FusedHeadEpilogue.from_meta(cif_meta, quad)
        self.caf = FusedHeadEpilogue.from_meta(caf_meta, quad)

    def __call__(self, x_cif, x_caf):
        return self.cif(x_cif), self.caf(x_caf)
