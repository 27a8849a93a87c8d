"""Non-maximum suppression (reference: decoder/nms.py).

``Keypoints`` keeps the reference's class-attribute knobs and its
``annotations(anns)`` entry point; the work is ``pp_keypoint_nms``
(csrc/nms.cu).  Inside ``CifCaf`` the same kernel runs fused after assembly,
without going through Python objects.
"""
import ctypes as C
import logging
import time

import numpy as np
import torch

from .. import _lib
from . import _stage
from .engine import DecoderParams

LOG = logging.getLogger(__name__)


class Keypoints:
    suppression = 0.0
    instance_threshold = 0.0
    keypoint_threshold = 0.0
    occupancy_visualizer = None

    def annotations(self, anns):
        start = time.perf_counter()
        if not anns:
            return anns
        device = _stage.cuda_device()
        n, F = len(anns), len(anns[0].data)
        for ann in anns:
            assert ann.joint_scales is not None
            assert len(ann.data) == F
        shape, caps = _stage.shape_caps(1, F, 0, 1, 1, 8, ann_cap=n)
        data = torch.from_numpy(np.stack([ann.data for ann in anns]).astype(np.float32)).to(device)
        scales = torch.from_numpy(
            np.stack([ann.joint_scales for ann in anns]).astype(np.float32)).to(device)
        count = torch.full((1,), n, dtype=torch.int32, device=device)
        source = torch.zeros((n,), dtype=torch.int32, device=device)
        score = torch.zeros((n,), dtype=torch.float64, device=device)
        ws, ws_bytes = _stage.workspace(shape, caps, device)
        params = DecoderParams(nms_keypoint_threshold=self.keypoint_threshold,
                               nms_instance_threshold=self.instance_threshold,
                               nms_suppression=self.suppression).to_c()
        _lib.check(_lib.lib().pp_keypoint_nms(
            _stage.ptr(data), _stage.ptr(scales), _stage.ptr(count), _stage.ptr(score),
            _stage.ptr(source), _stage.ptr(ws), ws_bytes, C.byref(shape), C.byref(caps),
            C.byref(params), _stage.stream_ptr()))
        m = int(count.item())
        data_h, src_h = data[:m].cpu().numpy(), source[:m].cpu().numpy()
        out = []
        for r in range(m):          # the reference mutates and returns the same objects
            ann = anns[int(src_h[r])]
            ann.data[:] = data_h[r]
            out.append(ann)
        LOG.debug('nms = %.3fs', time.perf_counter() - start)
        return out


class Detection:
    """Box NMS of the CifDet decoder (decoder/nms.py:60-102): not on the CifCaf path."""
    suppression = 0.1
    suppression_soft = 0.3
    instance_threshold = 0.1
    iou_threshold = 0.7
    iou_threshold_soft = 0.5

    def annotations(self, anns):
        raise NotImplementedError('CifDet decoding is outside the CifCaf hot path (SURVEY.md 8f)')
