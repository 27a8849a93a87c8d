#!/usr/bin/env python3
"""Generate tests/golden/ms_*.npz: the UNMODIFIED reference decoder (oracle/_ref) under a
multi-scale FieldConfig (several head sets with their own stride, min-scale and distance filters;
hr maps merged with np.maximum; ten sets = hflip pairs).  TEST INFRASTRUCTURE; run in the build
container:   python oracle/gen_golden_ms.py
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, 'tests', 'golden')

MIN_SCALES = [0.0, 12.0, 16.0, 24.0, 40.0]                 # decoder/factory.py:163-165
MAX_DISTANCES = [160.0, 240.0, 320.0, 480.0, None]

# name: (image_hw, n_poses, strides, pose_seed, jitter_seeds per half)
CASES = {
    'ms_3scales_321x481': ((321, 481), 5, (8, 16, 24), 3, (4,)),
    'ms_10heads_321x481': ((321, 481), 5, (8, 12, 16, 24, 40), 3, (4, 90)),
}


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def make_sets(hw, n_poses, strides, pose_seed, jitter_seeds):
    """[(cif, caf, stride, cif_min_scale, caf_min_distance, caf_max_distance)] as the factory
    would configure them."""
    from openpifpaf_b200.encoder import synthetic_multiscale_fields
    sets = []
    for js in jitter_seeds:
        for k, (cif, caf, stride) in enumerate(synthetic_multiscale_fields(
                hw, n_poses, strides=strides, pose_seed=pose_seed, jitter_seed=js)):
            sets.append((cif, caf, stride, MIN_SCALES[k], 3.0 * MIN_SCALES[k], MAX_DISTANCES[k]))
    return sets


def field_config(dec, sets):
    cfg = dec.FieldConfig()
    n = len(sets)
    cfg.cif_indices = [2 * k for k in range(n)]
    cfg.caf_indices = [2 * k + 1 for k in range(n)]
    cfg.cif_strides = [s[2] for s in sets]
    cfg.caf_strides = [s[2] for s in sets]
    cfg.cif_min_scales = [s[3] for s in sets]
    cfg.caf_min_distances = [s[4] for s in sets]
    cfg.caf_max_distances = [s[5] for s in sets]
    return cfg


def fields_of(sets):
    out = []
    for s in sets:
        out += [s[0].copy(), s[1].copy()]
    return out


def main():
    from oracle import build_ref
    from oracle.gen_golden import configure, pack_annotations
    from openpifpaf_b200.datasets.constants import COCO_KEYPOINTS, COCO_PERSON_SKELETON
    ref = build_ref.import_reference()
    dec = ref.decoder
    skeleton = [list(l) for l in COCO_PERSON_SKELETON]
    for name, (hw, n_poses, strides, pose_seed, jitter_seeds) in CASES.items():
        configure(dec, {})
        sets = make_sets(hw, n_poses, strides, pose_seed, jitter_seeds)
        cfg = field_config(dec, sets)
        hr = dec.CifHr(cfg).fill(fields_of(sets)).accumulated
        seeds = dec.CifSeeds(hr, cfg).fill(fields_of(sets)).get()
        seeds = np.array([list(s) for s in seeds], np.float32).reshape(-1, 5)
        scored = dec.CafScored(hr, cfg, skeleton).fill(fields_of(sets))
        anns = dec.CifCaf(cfg, keypoints=COCO_KEYPOINTS, skeleton=skeleton)(fields_of(sets))
        data, js, order = pack_annotations(anns, len(COCO_KEYPOINTS))
        np.savez_compressed(
            os.path.join(GOLDEN, name + '.npz'),
            inputs_sha256=np.array(sha(np.concatenate([np.concatenate([s[0].ravel(), s[1].ravel()])
                                                       for s in sets]))),
            hr_sha256=np.array(sha(hr)), hr_shape=np.array(hr.shape), seeds=seeds,
            caf_nf=np.array([f.shape[1] for f in scored.forward], np.int32),
            caf_nb=np.array([b.shape[1] for b in scored.backward], np.int32),
            caf_f_sha256=np.array([sha(f.astype(np.float32)) for f in scored.forward]),
            caf_b_sha256=np.array([sha(b.astype(np.float32)) for b in scored.backward]),
            ann_data=data, ann_scales=js, ann_order=order,
            ann_score=np.array([a.score() for a in anns], np.float64))
        print('%-22s sets %2d seeds %5d caf %5d anns %d' % (
            name, len(sets), len(seeds), int(sum(f.shape[1] for f in scored.forward)), len(anns)))


if __name__ == '__main__':
    main()
