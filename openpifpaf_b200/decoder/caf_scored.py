"""CafScored: CAF candidates re-scored against CifHr, per link and direction.

Mirror of the reference's ``decoder/caf_scored.py:13-98``; the work is
``pp_caf_scored`` (csrc/caf_scored.cu).  ``forward[l]`` / ``backward[l]`` are
``(9, K)`` float32 arrays with the reference's row order.
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


class CafScored:
    default_score_th = 0.1

    def __init__(self, cifhr, config: FieldConfig, skeleton, *, score_th=None, cif_floor=0.1):
        self.cifhr = cifhr
        self.config = config
        self.skeleton = skeleton
        self.score_th = score_th or self.default_score_th
        self.cif_floor = cif_floor
        self.forward = None
        self.backward = None

    def directed(self, caf_i, forward):
        if forward:
            return self.forward[caf_i], self.backward[caf_i]
        return self.backward[caf_i], self.forward[caf_i]

    def fill_caf(self, caf, stride, min_distance=0.0, max_distance=None):
        start = time.perf_counter()
        caf = _stage.to_device_f32(caf)
        hr = _stage.to_device_f32(self.cifhr, caf.device)
        L, nine, H, W = caf.shape
        assert nine == 9
        if self.forward is None:
            self.forward = [np.empty((9, 0), dtype=np.float32) for _ in range(L)]
            self.backward = [np.empty((9, 0), dtype=np.float32) for _ in range(L)]
        shape, caps = _stage.shape_caps(1, hr.shape[0], L, H, W, int(stride))
        fwd = torch.empty((L, 9, caps.caf_cap), dtype=torch.float32, device=caf.device)
        bwd = torch.empty((L, 9, caps.caf_cap), dtype=torch.float32, device=caf.device)
        nf = torch.zeros((L,), dtype=torch.int32, device=caf.device)
        nb = torch.zeros((L,), dtype=torch.int32, device=caf.device)
        status = torch.zeros((1,), dtype=torch.int32, device=caf.device)
        params = DecoderParams(caf_threshold=self.score_th, cif_floor=self.cif_floor,
                               caf_min_distance=min_distance, caf_max_distance=max_distance).to_c()
        sk = _lib.Skeleton.from_list(self.skeleton)
        _lib.check(_lib.lib().pp_caf_scored(
            _stage.ptr(caf), _stage.ptr(hr), C.byref(sk), C.c_float(self.score_th), _stage.ptr(fwd),
            _stage.ptr(nf), _stage.ptr(bwd), _stage.ptr(nb), C.byref(shape), C.byref(caps),
            C.byref(params), _stage.ptr(status), _stage.stream_ptr()))
        _stage.raise_on_status(status)
        nf_h, nb_h = nf.cpu().numpy(), nb.cpu().numpy()
        kmax = int(max(nf_h.max(initial=0), nb_h.max(initial=0)))
        fwd_h, bwd_h = fwd[:, :, :kmax].cpu().numpy(), bwd[:, :, :kmax].cpu().numpy()
        for l in range(L):
            self.forward[l] = np.concatenate((self.forward[l], fwd_h[l, :, :nf_h[l]]), axis=1)
            self.backward[l] = np.concatenate((self.backward[l], bwd_h[l, :, :nb_h[l]]), axis=1)
        LOG.debug('scored caf (%d, %d) in %.3fs', sum(f.shape[1] for f in self.forward),
                  sum(b.shape[1] for b in self.backward), time.perf_counter() - start)
        return self

    def fill(self, fields):
        for caf_i, stride, min_distance, max_distance in zip(
                self.config.caf_indices, self.config.caf_strides,
                self.config.caf_min_distances, self.config.caf_max_distances):
            self.fill_caf(fields[caf_i], stride, min_distance=min_distance,
                          max_distance=max_distance)
        return self
