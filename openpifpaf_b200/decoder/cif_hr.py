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
    neighbors = 16       # v / neighbors / len(cifs), cif_hr.py:15,40
    v_threshold = 0.1
    debug_visualizer = None
    _planes = 5          # channels of a cif field; CifDetHr: 7
    _det = 0

    def __init__(self, config: FieldConfig):
        self.config = config
        self.accumulated_device = None     # (F, Hh, Wh) CUDA tensor
        self._host = None
        self._hr_shape = None              # (H, W, stride) of the head that defined the map

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
        """cif_hr.py:42-64: the heads of ``cifs`` (one, or the two of an hflip pair) are accumulated
        into one map with ``v / neighbors / len(cifs)``; the first call stores it, later calls
        (other scales) merge theirs with ``np.maximum`` -- both inside the splat kernel."""
        start = time.perf_counter()
        if not 1 <= len(cifs) <= 2:
            raise NotImplementedError('one head or one hflip pair per call (what cif_hr.py:66-78 passes)')
        heads = [_stage.to_device_f32(c) for c in cifs]
        F, planes, H, W = heads[0].shape
        assert planes == self._planes, heads[0].shape
        assert all(h.shape == heads[0].shape for h in heads)
        stride = int(stride)
        device = heads[0].device
        merge = self.accumulated_device is not None
        if not merge:
            self._hr_shape = (H, W, stride)
            self.accumulated_device = torch.empty((F, (H - 1) * stride + 1, (W - 1) * stride + 1),
                                                  dtype=torch.float32, device=device)
        shape = _lib.Shape(1, F, 0, H, W, stride)
        shape_hr = _lib.Shape(1, F, 0, *self._hr_shape)
        caps = _lib.Caps(F * H * W, H * W, 256, len(heads) * H * W)
        lib = _lib.lib()
        # scratch: the CifHr part of the decode workspace of the map-defining shape
        big = _lib.Shape(1, F, 0, max(H, self._hr_shape[0]), max(W, self._hr_shape[1]),
                         max(stride, self._hr_shape[2]))
        ws_bytes = (lib.pp_cifdet_workspace_bytes if self._det else lib.pp_workspace_bytes)(
            C.byref(big), C.byref(caps))
        if ws_bytes == 0:
            raise _lib.PifPafError(_lib.PP_ERR_ARG, 'bad shape for the CifHr workspace')
        ws = torch.empty((ws_bytes,), dtype=torch.uint8, device=device)
        status = torch.zeros((1,), dtype=torch.int32, device=device)
        params = DecoderParams(cif_threshold=self.v_threshold, cif_min_scale=min_scale,
                               cif_neighbors=self.neighbors).to_c()
        _lib.check(lib.pp_cifhr_accumulate_ex(
            _stage.ptr(heads[0]), _stage.ptr(heads[1]) if len(heads) == 2 else None,
            _stage.ptr(self.accumulated_device), int(merge), self._det, _stage.ptr(ws), ws_bytes,
            C.byref(shape), C.byref(shape_hr), C.byref(caps), C.byref(params), _stage.ptr(status),
            _stage.stream_ptr()))
        _stage.raise_on_status(status)
        self._host = None
        LOG.debug('target_intensities %.3fs', time.perf_counter() - start)
        return self

    def fill(self, fields):
        cfg = self.config
        if len(cfg.cif_indices) == 10:                                     # cif_hr.py:66-72
            for i1, i2, stride, min_scale in zip(cfg.cif_indices[:5], cfg.cif_indices[5:],
                                                 cfg.cif_strides[:5], cfg.cif_min_scales[:5]):
                self.fill_multiple([fields[i1], fields[i2]], stride, min_scale=min_scale)
        else:
            for cif_i, stride, min_scale in zip(cfg.cif_indices, cfg.cif_strides, cfg.cif_min_scales):
                self.fill_cif(fields[cif_i], stride, min_scale=min_scale)
        if self.debug_visualizer is not None:
            self.debug_visualizer.predicted(self.accumulated)
        return self
