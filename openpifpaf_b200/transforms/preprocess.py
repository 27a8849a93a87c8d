"""``Preprocess.annotations_inverse`` on the device.

Mirror of the reference's ``transforms/preprocess.py:35-82``: the same meta dict
(``offset``, ``scale``, ``hflip``, ``width_height``, ``rotation``,
``horizontal_swap``) maps decoded coordinates back to the original image.  The
work is ``pp_annotations_inverse`` (csrc/inverse.cu); in a batched decode it runs
on the result buffers before they leave the GPU
(``CifCaf.decode_batch(..., metas=[...])``), so the host never loops over joints.
"""
import ctypes as C
import math

import numpy as np
import torch

from .. import _lib


def _meta_record(meta):
    """One reference meta dict -> pp_image_meta, with the scalar conversions numpy would apply."""
    rot = meta.get('rotation') or {'angle': 0.0, 'width': None, 'height': None}
    angle = -rot['angle']
    m = _lib.ImageMeta()
    m.offset_x, m.offset_y = float(meta['offset'][0]), float(meta['offset'][1])
    m.scale_x, m.scale_y = float(meta['scale'][0]), float(meta['scale'][1])
    m.hflip = int(bool(meta['hflip']))
    m.flip_w1 = float(meta['width_height'][0] - 1) if m.hflip else 0.0
    m.rotate = int(angle != 0.0)
    if m.rotate:
        m.rot_cos = math.cos(angle / 180.0 * math.pi)
        m.rot_sin = math.sin(angle / 180.0 * math.pi)
        m.rot_cx, m.rot_cy = (rot['width'] - 1) / 2, (rot['height'] - 1) / 2
    return m


def pack_metas(metas, device):
    """List of meta dicts -> (B, sizeof(pp_image_meta)) uint8 CUDA tensor."""
    arr = (_lib.ImageMeta * len(metas))(*[_meta_record(m) for m in metas])
    raw = np.frombuffer(arr, dtype=np.uint8).reshape(len(metas), C.sizeof(_lib.ImageMeta)).copy()
    return torch.from_numpy(raw).to(device)


def hflip_target(metas, n_keypoints, device):
    """Row permutation of ``meta['horizontal_swap']`` (hflip.py:13-29): row j moves to
    ``target[j]``.  Probed by swapping an index array; None when no image is flipped."""
    swap = None
    for m in metas:
        if m['hflip'] and m.get('horizontal_swap'):
            if swap is not None and m['horizontal_swap'] is not swap:
                probe_a = np.asarray(swap(np.arange(n_keypoints, dtype=np.float64)[:, None] + 1.0))
                probe_b = np.asarray(m['horizontal_swap'](
                    np.arange(n_keypoints, dtype=np.float64)[:, None] + 1.0))
                if not np.array_equal(probe_a, probe_b):
                    raise NotImplementedError('one horizontal_swap per batch')
            swap = m['horizontal_swap']
    if swap is None:
        return None
    moved = np.asarray(swap(np.arange(n_keypoints, dtype=np.float64)[:, None] + 1.0))[:, 0]
    target = np.empty(n_keypoints, np.int32)
    if sorted(moved.tolist()) != list(range(1, n_keypoints + 1)):
        raise NotImplementedError('horizontal_swap must be a permutation of the keypoints')
    target[(moved - 1).astype(np.int64)] = np.arange(n_keypoints, dtype=np.int32)
    return torch.from_numpy(target).to(device)


class Preprocess:
    """Only the inverse mapping of the reference's abstract base class."""

    @staticmethod
    def _detections_inverse(annotations, meta, device):
        """anndet_inverse (transforms/preprocess.py:84-94) through pp_detections_inverse."""
        from ..annotation import AnnotationDet
        n = len(annotations)
        box = torch.from_numpy(np.stack([np.asarray(a.bbox, np.float32) for a in annotations])[None]).to(device)
        count = torch.tensor([n], dtype=torch.int32, device=device)
        d_meta = pack_metas([meta], device)
        inv = _lib.Inverse(d_meta.data_ptr(), None)
        _lib.check(_lib.lib().pp_detections_inverse(
            C.c_void_p(box.data_ptr()), C.c_void_p(count.data_ptr()), C.byref(inv), 1, n,
            C.c_void_p(torch.cuda.current_stream(device).cuda_stream)))
        out = box[0].cpu().numpy()
        return [AnnotationDet(a.categories).set(a.field_i, a.score, out[i].copy())
                for i, a in enumerate(annotations)]

    @staticmethod
    def annotations_inverse(annotations, meta):
        """Reference signature: a list of Annotation and one meta dict; returns new
        annotations in original-image coordinates (the inputs are left untouched).

        Runs ``pp_annotations_inverse`` on a packed copy; requires a CUDA device."""
        from ..annotation import Annotation, AnnotationDet
        if not annotations:
            return []
        if not torch.cuda.is_available():
            raise _lib.PifPafError(_lib.PP_ERR_CUDA, 'annotations_inverse runs on the GPU (no CPU path)')
        device = torch.device('cuda', torch.cuda.current_device())
        if any(isinstance(a, AnnotationDet) for a in annotations):
            if not all(isinstance(a, AnnotationDet) for a in annotations):
                raise NotImplementedError('mixed pose / detection lists')
            return Preprocess._detections_inverse(annotations, meta, device)
        n, F = len(annotations), len(annotations[0].data)
        data = np.stack([a.data for a in annotations]).astype(np.float32)[None]
        scales = np.stack([a.joint_scales for a in annotations]).astype(np.float32)[None]
        orders = [a.decoding_order for a in annotations]
        # decoding_order coordinates: one (source, target) pair of rows per step
        flat = [(ai, k, side) for ai, o in enumerate(orders) for k in range(len(o)) for side in (2, 3)]
        rows = np.zeros((max(1, len(flat)), 3), np.float32)
        for r, (ai, k, side) in enumerate(flat):
            rows[r] = orders[ai][k][side]
        n_rows = -(-len(rows) // F)
        packed = np.zeros((n_rows * F, 3), np.float32)
        packed[:len(rows)] = rows
        cap = max(n, n_rows)
        d_data = torch.zeros((1, cap, F, 3), dtype=torch.float32, device=device)
        d_scales = torch.zeros((1, cap, F), dtype=torch.float32, device=device)
        d_pre = torch.zeros((1, cap, F, 3), dtype=torch.float32, device=device)
        d_data[0, :n] = torch.from_numpy(data[0]).to(device)
        d_scales[0, :n] = torch.from_numpy(scales[0]).to(device)
        d_pre[0, :n_rows] = torch.from_numpy(packed.reshape(n_rows, F, 3)).to(device)
        n_ann = torch.tensor([n], dtype=torch.int32, device=device)
        n_pre = torch.tensor([n_rows], dtype=torch.int32, device=device)
        d_meta = pack_metas([meta], device)
        d_target = hflip_target([meta], F, device)
        inv = _lib.Inverse(d_meta.data_ptr(), d_target.data_ptr() if d_target is not None else None)
        stream = C.c_void_p(torch.cuda.current_stream(device).cuda_stream)
        _lib.check(_lib.lib().pp_annotations_inverse(
            C.c_void_p(d_data.data_ptr()), C.c_void_p(d_scales.data_ptr()), C.c_void_p(n_ann.data_ptr()),
            C.c_void_p(d_pre.data_ptr()), C.c_void_p(n_pre.data_ptr()), C.byref(inv), 1, F, cap, stream))
        out_data = d_data[0, :n].cpu().numpy()
        out_scales = d_scales[0, :n].cpu().numpy()
        out_rows = d_pre[0, :n_rows].cpu().numpy().reshape(-1, 3)
        if np.any(np.isnan(out_data)):
            raise AssertionError('NaN coordinates (preprocess.py:67)')
        result = []
        for ai, ann in enumerate(annotations):
            new = Annotation(ann.keypoints, ann.skeleton, category_id=ann.category_id,
                             suppress_score_index=ann.suppress_score_index)
            new.set(out_data[ai].copy(), out_scales[ai].copy(), fixed_score=ann.fixed_score)
            new.frontier_order = list(ann.frontier_order)
            new.decoding_order = []
            result.append(new)
        for r, (ai, k, side) in enumerate(flat):
            if side == 2:
                s, e = orders[ai][k][0], orders[ai][k][1]
                result[ai].decoding_order.append((s, e, out_rows[r].copy(), out_rows[r + 1].copy()))
        return result
