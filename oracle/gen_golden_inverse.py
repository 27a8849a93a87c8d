#!/usr/bin/env python3
"""Generate tests/golden/inverse_mid.npz: Preprocess.annotations_inverse of the UNMODIFIED
reference (oracle/_ref) applied to the annotations of the golden case mid_321x481_8p under four
meta dicts.  TEST INFRASTRUCTURE; run in the build container:

    python oracle/gen_golden_inverse.py

Inputs are tests/golden/mid_321x481_8p.npz's ann_data / ann_scales / ann_order; outputs are the
reference's data, joint_scales and decoding_order coordinates per meta.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, 'tests', 'golden')

META_KINDS = ('plain', 'scaled', 'flipped', 'rotated')


def make_meta(kind, swap=None):
    """The four meta dicts (shape as built by the reference's transforms)."""
    meta = {'offset': np.array((0.0, 0.0)), 'scale': np.array((1.0, 1.0)), 'hflip': False,
            'width_height': np.array((640, 480)),
            'rotation': {'angle': 0.0, 'width': None, 'height': None}}
    if kind in ('scaled', 'flipped', 'rotated'):
        meta['offset'] = np.array((-12.0, 31.5))
        meta['scale'] = np.array((0.8013, 0.7977))
    if kind in ('flipped', 'rotated'):
        meta['hflip'] = True
        meta['horizontal_swap'] = swap
    if kind == 'rotated':
        meta['rotation'] = {'angle': 17.5, 'width': 513, 'height': 385}
    return meta


def detection_boxes():
    """Seeded detection boxes (x, y, w, h) for the anndet_inverse fixture."""
    rng = np.random.RandomState(5)
    return np.concatenate([rng.uniform(0, 400, (12, 2)), rng.uniform(5, 200, (12, 2))], axis=1).astype(np.float32)


def main():
    from oracle import build_ref
    ref = build_ref.import_reference()
    import openpifpaf.transforms.hflip as ref_hflip
    from openpifpaf_b200.datasets.constants import COCO_KEYPOINTS, COCO_PERSON_SKELETON, HFLIP
    g = np.load(os.path.join(GOLDEN, 'mid_321x481_8p.npz'))
    swap = ref_hflip._HorizontalSwap(COCO_KEYPOINTS, HFLIP)    # pylint: disable=protected-access
    out = {}
    for kind in META_KINDS:
        anns = []
        for data, js, order in zip(g['ann_data'], g['ann_scales'], g['ann_order']):
            a = ref.Annotation(COCO_KEYPOINTS, COCO_PERSON_SKELETON).set(data.copy(), js.copy())
            a.decoding_order = [(int(s), int(e), data[s].copy(), data[e].copy())
                                for s, e in order if s >= 0]
            anns.append(a)
        res = ref.transforms.Preprocess.annotations_inverse(anns, make_meta(kind, swap))
        out[kind + '_data'] = np.stack([a.data for a in res]).astype(np.float32)
        out[kind + '_scales'] = np.stack([a.joint_scales for a in res]).astype(np.float32)
        out[kind + '_dxy'] = np.array(
            [c[:2] for a in res for _, __, c1, c2 in a.decoding_order for c in (c1, c2)], np.float32)
        dets = [ref.annotation.AnnotationDet(['a', 'b']).set(i % 2, np.float32(0.5), b.copy())
                for i, b in enumerate(detection_boxes())]
        dets = ref.transforms.Preprocess.annotations_inverse(dets, make_meta(kind, swap))
        out[kind + '_detbox'] = np.stack([np.asarray(a.bbox, np.float32) for a in dets])
    np.savez_compressed(os.path.join(GOLDEN, 'inverse_mid.npz'), **out)
    print('inverse_mid', {k: v.shape for k, v in out.items() if k.startswith('rotated')})


if __name__ == '__main__':
    main()
