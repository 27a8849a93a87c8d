"""CifSeeds: seed candidates from CIF fields re-scored against CifHr.

Mirror of the reference's ``decoder/cif_seeds.py:13-64``; the work is
``pp_cif_seeds`` (csrc/seeds.cu).  ``get()`` returns the reference's list of
``(v, field, x, y, scale)`` tuples in ``sorted(..., reverse=True)`` order.
"""
import ctypes as C
import logging
import time

import numpy as np
import torch

from .. import _lib
from . import _stage
from .engine import DecoderParams
from .field_config import FieldConfig

LOG = logging.getLogger(__name__)


class CifSeeds:
    threshold = None      # must be configured (decoder.configure), as in the reference
    score_scale = 1.0
    debug_visualizer = None

    def __init__(self, cifhr, config: FieldConfig):
        # cifhr: (F,Hh,Wh) numpy array or CUDA tensor (CifHr.accumulated[_device])
        self.cifhr = cifhr
        self.config = config
        self.seeds = []
        self._sorted_blocks = []

    def fill_cif(self, cif, stride, *, min_scale=0.0, seed_mask=None):
        start = time.perf_counter()
        if self.threshold is None:
            raise TypeError("'>' not supported between instances of 'float' and 'NoneType' "
                            "(CifSeeds.threshold is not configured)")
        cif = _stage.to_device_f32(cif)
        hr = _stage.to_device_f32(self.cifhr, cif.device)
        F, five, H, W = cif.shape
        assert five == 5
        shape, caps = _stage.shape_caps(1, F, 0, H, W, int(stride))
        ws, ws_bytes = _stage.workspace(shape, caps, cif.device)
        seeds = torch.empty((caps.seed_cap, 5), dtype=torch.float32, device=cif.device)
        n_seeds = torch.zeros((1,), dtype=torch.int32, device=cif.device)
        status = torch.zeros((1,), dtype=torch.int32, device=cif.device)
        params = DecoderParams(seed_threshold=self.threshold, seed_score_scale=self.score_scale,
                               cif_min_scale=min_scale, seed_mask=seed_mask).to_c()
        _lib.check(_lib.lib().pp_cif_seeds(
            _stage.ptr(cif), _stage.ptr(hr), _stage.ptr(seeds), _stage.ptr(n_seeds), _stage.ptr(ws),
            ws_bytes, C.byref(shape), C.byref(caps), C.byref(params), _stage.ptr(status),
            _stage.stream_ptr()))
        _stage.raise_on_status(status)
        n = int(n_seeds.item())
        block = seeds[:n].cpu().numpy()
        self._sorted_blocks.append(block)
        self.seeds.extend(_as_tuples(block))
        LOG.debug('seeds %d, %.3fs', len(self.seeds), time.perf_counter() - start)
        return self

    def get(self):
        if self.debug_visualizer is not None:
            self.debug_visualizer.predicted(self.seeds)
        if len(self._sorted_blocks) == 1:
            return list(self.seeds)            # already in the reference's order
        return sorted(self.seeds, reverse=True)

    def fill(self, fields):
        for cif_i, stride, min_scale in zip(self.config.cif_indices, self.config.cif_strides,
                                            self.config.cif_min_scales):
            self.fill_cif(fields[cif_i], stride, min_scale=min_scale,
                          seed_mask=self.config.seed_mask)
        return self


def _as_tuples(block):
    f = block[:, 1].astype(np.int64)
    return [(block[i, 0], int(f[i]), block[i, 2], block[i, 3], block[i, 4])
            for i in range(block.shape[0])]
