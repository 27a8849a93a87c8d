"""CPU restatement (numpy) of Preprocess.annotations_inverse.  TEST INFRASTRUCTURE ONLY.

Follows the reference's transforms/preprocess.py:35-82 on packed arrays instead of
Annotation objects, with numpy's casting spelled out: the rotation is float32
arithmetic (Python scalars against float32 arrays), offset / scale / flip are
computed in float64 and rounded to float32 after every statement (in-place
operators and item assignment on float32 arrays).  Pinned against the running
reference by tests/test_oracle_vs_reference.py::test_annotations_inverse.
"""
import math

import numpy as np


def annotations_inverse(data, joint_scales, decoding_xy, meta, hflip_target=None):
    """data (N, F, 3) float32, joint_scales (N, F) float32, decoding_xy (M, 2) float32 (the
    coordinates stored in decoding_order entries).  Returns transformed copies."""
    f32, f64 = np.float32, np.float64
    data = np.array(data, f32)
    js = np.array(joint_scales, f32)
    dxy = np.array(decoding_xy, f32).reshape(-1, 2)
    rot = meta.get('rotation') or {'angle': 0.0}
    angle = -rot['angle']
    if angle != 0.0:                                           # preprocess.py:51-56
        c, s = f32(math.cos(angle / 180.0 * math.pi)), f32(math.sin(angle / 180.0 * math.pi))
        cx, cy = f32((rot['width'] - 1) / 2), f32((rot['height'] - 1) / 2)
        xo, yo = data[..., 0] - cx, data[..., 1] - cy
        data[..., 0] = (cx + c * xo) + s * yo
        data[..., 1] = (cy - s * xo) + c * yo
    off = np.asarray(meta['offset'], f64)
    scale = np.asarray(meta['scale'], f64)
    for i in (0, 1):                                           # preprocess.py:59-64
        data[..., i] = (data[..., i].astype(f64) + off[i]).astype(f32)
        data[..., i] = (data[..., i].astype(f64) / scale[i]).astype(f32)
    js = (js.astype(f64) / scale[0]).astype(f32)               # preprocess.py:65
    if meta['hflip']:                                          # preprocess.py:69-73
        w1 = f64(meta['width_height'][0] - 1)
        data[..., 0] = ((-data[..., 0]).astype(f64) + w1).astype(f32)
        if hflip_target is not None:
            swapped = np.zeros_like(data)
            swapped[:, np.asarray(hflip_target)] = data
            data = swapped
    for i in (0, 1):                                           # preprocess.py:75-80
        dxy[:, i] = (dxy[:, i].astype(f64) + off[i]).astype(f32)
        dxy[:, i] = (dxy[:, i].astype(f64) / scale[i]).astype(f32)
    return data, js, dxy


def detections_inverse(boxes, meta):
    """Preprocess.anndet_inverse (transforms/preprocess.py:84-94 with utils.rotate_box,
    transforms/utils.py:5-28) on boxes (N, 4) x, y, w, h float32.  Returns a transformed copy."""
    f32, f64 = np.float32, np.float64
    boxes = np.array(boxes, f32).reshape(-1, 4)
    rot = meta.get('rotation') or {'angle': 0.0}
    angle = -rot['angle']
    if angle != 0.0:
        c, s = f32(math.cos(angle / 180.0 * math.pi)), f32(math.sin(angle / 180.0 * math.pi))
        hw, hh = f32((rot['width'] - 1) / 2), f32((rot['height'] - 1) / 2)      # width/2 of rotate_box
        x0, y0 = boxes[:, 0], boxes[:, 1]
        x1, y1 = x0 + boxes[:, 2], y0 + boxes[:, 3]
        cx = np.stack([x0, x1, x0, x1], axis=1)
        cy = np.stack([y0, y0, y1, y1], axis=1)
        xo, yo = cx - hw, cy - hh
        rx = (hw + c * xo) + s * yo
        ry = (hh - s * xo) + c * yo
        x, y = rx.min(axis=1), ry.min(axis=1)
        boxes = np.stack([x, y, rx.max(axis=1) - x, ry.max(axis=1) - y], axis=1).astype(f32)
    off = np.asarray(meta['offset'], f64)
    scale = np.asarray(meta['scale'], f64)
    boxes[:, :2] = (boxes[:, :2].astype(f64) + off).astype(f32)
    boxes[:, :2] = (boxes[:, :2].astype(f64) / scale).astype(f32)
    boxes[:, 2:] = (boxes[:, 2:].astype(f64) / scale).astype(f32)
    return boxes
