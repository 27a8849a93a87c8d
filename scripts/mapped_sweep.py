#!/usr/bin/env python3
"""Sweep sub-batch size / lanes of the host-mapped public API call."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

import openpifpaf_b200.decoder as decoder  # noqa: E402
from openpifpaf_b200.datasets.constants import COCO_KEYPOINTS, COCO_PERSON_SKELETON  # noqa: E402
from openpifpaf_b200.encoder import synthetic_fields  # noqa: E402

B = 64
decoder.CifHr.v_threshold, decoder.CifSeeds.threshold = 0.1, 0.2
decoder.CafScored.default_score_th = 0.1
decoder.CifCaf.keypoint_threshold = 0.001
decoder.nms.Keypoints.keypoint_threshold = 0.001
fields = [synthetic_fields((641, 641), 5, pose_seed=b, jitter_seed=1000 + b)[:2] for b in range(B)]
cif = torch.from_numpy(np.stack([f[0] for f in fields])).pin_memory()
caf = torch.from_numpy(np.stack([f[1] for f in fields])).pin_memory()
dec = decoder.CifCaf(decoder.FieldConfig(), keypoints=COCO_KEYPOINTS, skeleton=[list(l) for l in COCO_PERSON_SKELETON])
dec.host_transfer = 'mapped'
for sub, lanes in ((4, 8), (8, 4), (8, 8), (16, 2), (16, 4), (32, 2), (64, 1)):
    dec.mapped_chunk, dec.host_lanes = sub, lanes
    for _ in range(3):
        dec.decode_batch(cif, caf)
    t0 = time.perf_counter()
    for _ in range(10):
        dec.decode_batch(cif, caf)
    ms = (time.perf_counter() - t0) / 10 * 1e3
    print('sub %2d lanes %d: %.2f ms  %.0f img/s' % (sub, lanes, ms, B / ms * 1e3))
