"""CifDet: detection decoder on the GPU.

Mirror of the reference's ``decoder/generator/cifdet.py:17-52`` (same constructor,
``__call__(fields)`` and ``batch``); ``CifDetHr``, ``CifDetSeeds``, the seed loop
with its occupancy map and ``nms.Detection`` run as one device pipeline
(``pp_cifdet_decode_batch``: csrc/cifhr.cu + seeds.cu with the detection sigma /
record layout, csrc/det.cu).  Knobs are read from the same class attributes the
reference's ``configure`` sets (``CifHr.v_threshold``, ``CifSeeds.threshold`` /
``score_scale``, ``nms.Detection.*``).
"""
import ctypes as C
import logging
import time

import numpy as np
import torch

from ... import _lib
from ...annotation import AnnotationDet
from .. import _stage, nms
from ..cif_hr import CifHr
from ..cif_seeds import CifSeeds
from ..field_config import FieldConfig
from .generator import Generator

LOG = logging.getLogger(__name__)


def det_params(*, min_scale=0.0, with_nms=True):
    """pp_det_params from the reference's class attributes."""
    p = _lib.DetParams()
    _lib.lib().pp_det_default_params(C.byref(p))
    if CifSeeds.threshold is None:
        raise TypeError("'>' not supported between instances of 'float' and 'NoneType' "
                        "(CifSeeds.threshold is not configured)")
    p.cif_threshold = CifHr.v_threshold
    p.seed_threshold = CifSeeds.threshold
    p.seed_score_scale = CifSeeds.score_scale
    p.cif_min_scale = min_scale or 0.0
    d = nms.Detection
    p.nms_suppression, p.nms_suppression_soft = d.suppression, d.suppression_soft
    p.nms_instance_threshold = d.instance_threshold
    p.nms_iou_threshold, p.nms_iou_threshold_soft = d.iou_threshold, d.iou_threshold_soft
    p.nms = int(bool(with_nms))
    return p


def run_cifdet(field, stride, *, min_scale=0.0, with_nms=True, ann_cap=1024, want=()):
    """``field (B, C, 7, H, W)`` float32 (CUDA tensor, numpy or CPU tensor) through
    pp_cifdet_decode_batch.  Returns a dict of CUDA tensors: det_field / det_score / det_box /
    n_det, plus whichever of hr, seeds, n_seeds, pre_* are named in ``want``."""
    field = _stage.to_device_f32(field)
    B, Cn, seven, H, W = field.shape
    if seven != 7:
        raise ValueError('CifDet fields have 7 channels [v, x, y, _, w, h, _], got %d' % seven)
    dev = field.device
    shape, caps = _stage.shape_caps(B, Cn, 0, H, W, int(stride), ann_cap=ann_cap)
    n = _lib.lib().pp_cifdet_workspace_bytes(C.byref(shape), C.byref(caps))
    if n == 0:
        raise _lib.PifPafError(_lib.PP_ERR_ARG, 'bad shape for the CifDet workspace')
    ws = torch.empty((n,), dtype=torch.uint8, device=dev)
    hh, hw = (H - 1) * int(stride) + 1, (W - 1) * int(stride) + 1
    t = {
        'det_field': torch.zeros((B, ann_cap), dtype=torch.int32, device=dev),
        'det_score': torch.zeros((B, ann_cap), dtype=torch.float32, device=dev),
        'det_box': torch.zeros((B, ann_cap, 4), dtype=torch.float32, device=dev),
        'n_det': torch.zeros((B,), dtype=torch.int32, device=dev),
        'status': torch.zeros((B,), dtype=torch.int32, device=dev),
    }
    optional = {
        'hr': lambda: torch.empty((B, Cn, hh, hw), dtype=torch.float32, device=dev),
        'seeds': lambda: torch.zeros((B, caps.seed_cap, 6), dtype=torch.float32, device=dev),
        'n_seeds': lambda: torch.zeros((B,), dtype=torch.int32, device=dev),
        'pre_field': lambda: torch.zeros((B, ann_cap), dtype=torch.int32, device=dev),
        'pre_score': lambda: torch.zeros((B, ann_cap), dtype=torch.float32, device=dev),
        'pre_box': lambda: torch.zeros((B, ann_cap, 4), dtype=torch.float32, device=dev),
        'n_pre': lambda: torch.zeros((B,), dtype=torch.int32, device=dev),
    }
    for name in want:
        t[name] = optional[name]()
    out = _lib.DetOutputs(**{name: tensor.data_ptr() for name, tensor in t.items()})
    params = det_params(min_scale=min_scale, with_nms=with_nms)
    _lib.check(_lib.lib().pp_cifdet_decode_batch(
        _stage.ptr(field), C.byref(out), _stage.ptr(ws), n, C.byref(shape), C.byref(caps),
        C.byref(params), _stage.stream_ptr()))
    _stage.raise_on_status(t['status'])
    return t


class CifDetHr(CifHr):
    """``CifDetHr(config).fill(fields).accumulated`` (cif_hr.py:84-100): CifHr.fill_multiple with the
    detection sigma, ``max(1, 0.1 min(w, h) stride)``, and the min-scale filter on both sizes."""
    _planes = 7
    _det = 1


class CifDetSeeds(CifSeeds):
    """``CifDetSeeds(hr, config).fill(fields).get()`` -> [(v, f, x, y, w, h)] (cif_seeds.py:67-90).
    The hr map passed in is recomputed on the device (same kernels, same bits)."""

    def fill_cif(self, cif, stride, *, min_scale=0.0, seed_mask=None):
        field = _stage.to_device_f32(cif)
        t = run_cifdet(field[None], stride, min_scale=min_scale, with_nms=False,
                       want=('seeds', 'n_seeds'))
        block = t['seeds'][0, :int(t['n_seeds'][0].item())].cpu().numpy()
        if seed_mask is not None:       # cif_seeds.py:72-73: masked categories contribute no seeds
            keep = np.asarray(seed_mask, dtype=bool)[block[:, 1].astype(np.int64)]
            block = block[keep]
        self._sorted_blocks.append(block)
        f = block[:, 1].astype(np.int64)
        self.seeds.extend((block[i, 0], int(f[i]), block[i, 2], block[i, 3], block[i, 4], block[i, 5])
                          for i in range(block.shape[0]))
        return self


class CifDet(Generator):
    occupancy_visualizer = None
    max_detections = 1024          # result capacity per image (PifPafError beyond it)

    def __init__(self, field_config: FieldConfig, categories, *, worker_pool=None):
        super().__init__(worker_pool)
        self.field_config = field_config
        self.categories = categories

    def decode_batch(self, field, metas=None):
        """``field (B, C, 7, H, W)`` -> List[List[AnnotationDet]].  ``metas``: optional per-image
        preprocessing meta dicts; the boxes are then mapped back to the original image on the device
        (``Preprocess.anndet_inverse``, transforms/preprocess.py:84-94)."""
        cfg = self.field_config
        if len(cfg.cif_indices) != 1:
            raise NotImplementedError('CifDet decodes one head (cifdet.py:31-35)')
        t = run_cifdet(field, cfg.cif_strides[0], min_scale=cfg.cif_min_scales[0],
                       ann_cap=self.max_detections)
        if metas is not None:
            from ...transforms.preprocess import pack_metas
            if len(metas) != t['det_box'].shape[0]:
                raise ValueError('one meta dict per image')
            d_meta = pack_metas(metas, t['det_box'].device)
            inv = _lib.Inverse(d_meta.data_ptr(), None)
            _lib.check(_lib.lib().pp_detections_inverse(
                _stage.ptr(t['det_box']), _stage.ptr(t['n_det']), C.byref(inv), t['det_box'].shape[0],
                t['det_box'].shape[1], _stage.stream_ptr()))
        n_det = t['n_det'].cpu().numpy()
        nmax = int(n_det.max(initial=0))
        fields = t['det_field'][:, :nmax].cpu().numpy()
        scores = t['det_score'][:, :nmax].cpu().numpy()
        boxes = t['det_box'][:, :nmax].cpu().numpy()
        return [[AnnotationDet(self.categories).set(int(fields[b, i]), scores[b, i], boxes[b, i].copy())
                 for i in range(int(n))] for b, n in enumerate(n_det)]

    def decode_heads(self, heads):
        return self.decode_batch(heads[self.field_config.cif_indices[0]])

    def __call__(self, fields):
        start = time.perf_counter()
        cfg = self.field_config
        if cfg.cif_visualizers:
            for vis, cif_i in zip(cfg.cif_visualizers, cfg.cif_indices):
                vis.predicted(fields[cif_i])
        field = fields[cfg.cif_indices[0]]
        if isinstance(field, torch.Tensor):
            field = field.unsqueeze(0)
        else:
            field = np.ascontiguousarray(field, dtype=np.float32)[None]
        annotations = self.decode_batch(field)[0]
        LOG.info('annotations %d, decoder = %.3fs', len(annotations), time.perf_counter() - start)
        return annotations
