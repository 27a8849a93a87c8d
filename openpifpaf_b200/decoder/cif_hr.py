"""CifHr: high-resolution confidence map accumulated from CIF fields.

Mirror of the reference's ``decoder/cif_hr.py:14-81`` (same class attributes,
``fill`` / ``fill_cif`` / ``fill_multiple`` entry points and ``accumulated``
result); the work is ``pp_cifhr_accumulate`` (csrc/cifhr.cu).
"""
import ctypes as C
import logging
import time

import torch

from .. import _lib
from . import _stage
from .engine import DecoderParams
from .field_config import FieldConfig

LOG = logging.getLogger(__name__)


class CifHr:
    neighbors = 16       # fixed in the CUDA kernel (v / 16), cif_hr.py:15
    v_threshold = 0.1
    debug_visualizer = None

    def __init__(self, config: FieldConfig):
        self.config = config
        self.accumulated_device = None     # (F, Hh, Wh) CUDA tensor
        self._host = None

    @property
    def accumulated(self):
        """(F, Hh, Wh) float32 numpy array, like the reference's attribute."""
        if self.accumulated_device is None:
            return None
        if self._host is None:
            self._host = self.accumulated_device.cpu().numpy()
        return self._host

    def fill_cif(self, cif, stride, min_scale=0.0):
        return self.fill_multiple([cif], stride, min_scale)

    def fill_multiple(self, cifs, stride, min_scale=0.0):
        start = time.perf_counter()
        if len(cifs) != 1:
            raise NotImplementedError('multi-head accumulation (hflip pairs) is not built yet')
        if self.neighbors != 16:
            raise NotImplementedError('the CUDA kernel fixes CifHr.neighbors at 16')
        cif = _stage.to_device_f32(cifs[0])
        F, five, H, W = cif.shape
        assert five == 5
        shape, caps = _stage.shape_caps(1, F, 0, H, W, int(stride))
        ws, ws_bytes = _stage.workspace(shape, caps, cif.device)
        hr = torch.empty((F, (H - 1) * int(stride) + 1, (W - 1) * int(stride) + 1),
                         dtype=torch.float32, device=cif.device)
        status = torch.zeros((1,), dtype=torch.int32, device=cif.device)
        params = DecoderParams(cif_threshold=self.v_threshold, cif_min_scale=min_scale).to_c()
        _lib.check(_lib.lib().pp_cifhr_accumulate(
            _stage.ptr(cif), _stage.ptr(hr), _stage.ptr(ws), ws_bytes, C.byref(shape),
            C.byref(caps), C.byref(params), _stage.ptr(status), _stage.stream_ptr()))
        _stage.raise_on_status(status)
        if self.accumulated_device is None:
            self.accumulated_device = hr
        else:   # multi-scale merge, cif_hr.py:62
            self.accumulated_device = torch.maximum(hr, self.accumulated_device)
        self._host = None
        LOG.debug('target_intensities %.3fs', time.perf_counter() - start)
        return self

    def fill(self, fields):
        if len(self.config.cif_indices) == 10:
            raise NotImplementedError('multi-scale + hflip CifHr is not built yet')
        for cif_i, stride, min_scale in zip(self.config.cif_indices, self.config.cif_strides,
                                            self.config.cif_min_scales):
            self.fill_cif(fields[cif_i], stride, min_scale=min_scale)
        if self.debug_visualizer is not None:
            self.debug_visualizer.predicted(self.accumulated)
        return self
