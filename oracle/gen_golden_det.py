#!/usr/bin/env python3
"""Generate tests/golden/det_*.npz by running the UNMODIFIED reference CifDet decoder
(oracle/_ref) on fields made by ``openpifpaf_b200.encoder.synthetic_det_fields``.
TEST INFRASTRUCTURE; run in the build container:   python oracle/gen_golden_det.py

Stored per case: the input recipe + sha256 of the regenerated field; sha256 of
CifDetHr(...).accumulated; CifDetSeeds(...).get() as (N, 6) float32; the detections of
CifDet(...)(fields): field_i, score (float32), bbox.
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, 'tests', 'golden')

# name: (image_hw, n_boxes, box_seed, jitter_seed, n_categories, cif_min_scale)
CASES = {
    'det_321x481_8b': ((321, 481), 8, 0, 1, 5, 0.0),
    'det_641_20b': ((641, 641), 20, 3, 4, 12, 0.0),
    'det_161_3b': ((161, 161), 3, 5, 6, 3, 0.0),
    'det_641_minscale': ((641, 641), 12, 7, 8, 40, 60.0),
    'det_empty_97': ((97, 97), 0, 9, 10, 4, 0.0),
}


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def configure(dec):
    dec.CifHr.v_threshold = 0.1
    dec.CifSeeds.threshold = 0.2
    dec.CifSeeds.score_scale = 1.0
    det = dec.nms.Detection
    det.suppression, det.suppression_soft, det.instance_threshold = 0.1, 0.3, 0.1
    det.iou_threshold, det.iou_threshold_soft = 0.7, 0.5


def run_reference(ref, field, min_scale):
    dec = ref.decoder
    configure(dec)
    cfg = dec.FieldConfig()
    cfg.cif_min_scales = [min_scale]
    hr = dec.cif_hr.CifDetHr(cfg).fill([field.copy()]).accumulated
    seeds = dec.cif_seeds.CifDetSeeds(hr, cfg).fill([field.copy()]).get()
    seeds = np.array([list(s) for s in seeds], np.float32).reshape(-1, 6)
    cats = ['c%d' % i for i in range(field.shape[0])]
    anns = dec.generator.cifdet.CifDet(cfg, cats)([field.copy()])
    return hr, seeds, anns


def main():
    from oracle import build_ref
    from openpifpaf_b200.encoder import synthetic_det_fields
    ref = build_ref.import_reference()
    os.makedirs(GOLDEN, exist_ok=True)
    for name, (hw, nb, bs, js, ncat, min_scale) in CASES.items():
        field, _ = synthetic_det_fields(hw, nb, box_seed=bs, jitter_seed=js, n_categories=ncat)
        hr, seeds, anns = run_reference(ref, field, min_scale)
        np.savez_compressed(
            os.path.join(GOLDEN, name + '.npz'),
            image_hw=np.array(hw), n_boxes=nb, box_seed=bs, jitter_seed=js, n_categories=ncat,
            cif_min_scale=min_scale, field_sha256=np.array(sha(field)), hr_sha256=np.array(sha(hr)),
            hr_shape=np.array(hr.shape), seeds=seeds,
            det_field=np.array([a.field_i for a in anns], np.int32),
            det_score=np.array([a.score for a in anns], np.float32),
            det_box=np.array([a.bbox for a in anns], np.float32).reshape(-1, 4),
            versions=np.array('openpifpaf %s, numpy %s' % (ref.__version__, np.__version__)))
        print('%-20s seeds %5d  detections %3d' % (name, len(seeds), len(anns)))


if __name__ == '__main__':
    main()
