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
    """Soft NMS by IoU for detections (reference: decoder/nms.py:60-102).  Class attributes are
    the reference's knobs; inside ``CifDet`` the same routine runs fused after the seed loop."""
    suppression = 0.1
    suppression_soft = 0.3
    instance_threshold = 0.1
    iou_threshold = 0.7
    iou_threshold_soft = 0.5

    def annotations(self, anns):
        """List[AnnotationDet] -> the surviving objects (scores updated in place), score order."""
        start = time.perf_counter()
        if not anns:
            return anns
        from .generator.cifdet import det_params
        device = _stage.cuda_device()
        n = len(anns)
        field = torch.tensor([[a.field_i for a in anns]], dtype=torch.int32, device=device)
        score = torch.from_numpy(np.array([[a.score for a in anns]], np.float32)).to(device)
        box = torch.from_numpy(np.stack([np.asarray(a.bbox, np.float32) for a in anns])[None]).to(device)
        count = torch.full((1,), n, dtype=torch.int32, device=device)
        out_field, out_score, out_box = torch.zeros_like(field), torch.zeros_like(score), torch.zeros_like(box)
        out_src = torch.zeros_like(field)
        out_n = torch.zeros((1,), dtype=torch.int32, device=device)
        ws_bytes = _lib.lib().pp_detection_nms_workspace_bytes(1, n)
        ws = torch.empty((ws_bytes,), dtype=torch.uint8, device=device)
        with CifSeedsThresholdGuard():
            params = det_params()
        params.nms_suppression, params.nms_suppression_soft = self.suppression, self.suppression_soft
        params.nms_instance_threshold = self.instance_threshold
        params.nms_iou_threshold, params.nms_iou_threshold_soft = self.iou_threshold, self.iou_threshold_soft
        _lib.check(_lib.lib().pp_detection_nms(
            _stage.ptr(field), _stage.ptr(score), _stage.ptr(box), _stage.ptr(count),
            _stage.ptr(out_field), _stage.ptr(out_score), _stage.ptr(out_box), _stage.ptr(out_src),
            _stage.ptr(out_n), _stage.ptr(ws), ws_bytes, 1, n, C.byref(params), _stage.stream_ptr()))
        m = int(out_n.item())
        new_score = out_score[0, :m].cpu().numpy()
        src = out_src[0, :m].cpu().numpy()
        out = []
        for r in range(m):          # the reference mutates and returns the same objects
            ann = anns[int(src[r])]
            ann.score = new_score[r]
            out.append(ann)
        LOG.debug('nms = %.3fs', time.perf_counter() - start)
        return out


class CifSeedsThresholdGuard:
    """det_params() reads CifSeeds.threshold, which the standalone NMS does not need."""

    def __enter__(self):
        from .cif_seeds import CifSeeds
        self._cls, self._saved = CifSeeds, CifSeeds.threshold
        if CifSeeds.threshold is None:
            CifSeeds.threshold = 0.0
        return self

    def __exit__(self, *exc):
        self._cls.threshold = self._saved
        return False
