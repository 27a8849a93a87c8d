"""Occupancy map with the reference's interface (decoder/occupancy.py:10-47).

The batch decoder never rasterises occupancy (csrc/assemble.cu counts covering
boxes instead).  This class keeps the reference's object for callers and tests
that drive it directly: a uint8 map on the device, updated and queried through
the parity-shim primitives ``pp_scalar_square_add_single`` (decoder/utils.py:61-66)
and ``pp_scalar_nonzero_clipped_with_reduction`` (functional.pyx:281-286).
"""
import logging

import torch

from .. import _lib
from . import _stage

LOG = logging.getLogger(__name__)


class Occupancy:
    def __init__(self, shape, reduction, *, min_scale=None):
        assert len(shape) == 3
        if min_scale is None:
            min_scale = reduction
        assert min_scale >= reduction
        self.reduction = reduction
        self.min_scale = min_scale
        self.min_scale_reduced = min_scale / reduction
        self.device = _stage.cuda_device()
        self.occupancy_device = torch.zeros(
            (shape[0], int(shape[1] / reduction), int(shape[2] / reduction)),
            dtype=torch.uint8, device=self.device)
        self._probe = torch.zeros((1,), dtype=torch.uint8, device=self.device)
        LOG.debug('shape = %s, min_scale = %d', self.occupancy_device.shape, self.min_scale_reduced)

    @property
    def occupancy(self):
        return self.occupancy_device.cpu().numpy()

    def __len__(self):
        return self.occupancy_device.shape[0]

    def set(self, f, x, y, sigma):
        """Mark a square centred at the rounded (x, y); half-to-even like Python's round."""
        if f >= len(self):
            return
        xi = round(x / self.reduction)
        yi = round(y / self.reduction)
        si = round(max(self.min_scale_reduced, sigma / self.reduction))
        field = self.occupancy_device[f]
        _lib.check(_lib.lib().pp_scalar_square_add_single(
            _stage.ptr(field), field.shape[0], field.shape[1], int(xi), int(yi), int(si), 1,
            _stage.stream_ptr()))

    def get(self, f, x, y):
        """Occupancy at the floor of (x, y) / reduction, clipped into the map."""
        if f >= len(self):
            return 1.0
        field = self.occupancy_device[f]
        _lib.check(_lib.lib().pp_scalar_nonzero_clipped_with_reduction(
            _stage.ptr(field), field.shape[0], field.shape[1], float(x), float(y),
            float(self.reduction), _stage.ptr(self._probe), _stage.stream_ptr()))
        return int(self._probe.item())
