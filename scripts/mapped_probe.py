#!/usr/bin/env python3
"""Per-stage device times of the host entries: full copy against host_transfer='mapped'."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

import openpifpaf_b200.decoder as decoder  # noqa: E402
from openpifpaf_b200.datasets.constants import COCO_KEYPOINTS, COCO_PERSON_SKELETON  # noqa: E402
from openpifpaf_b200.decoder.engine import StageProfile  # noqa: E402
from openpifpaf_b200.encoder import synthetic_fields  # noqa: E402

B, CHUNK = int(os.environ.get('B', 64)), int(os.environ.get('CHUNK', 64))
decoder.CifHr.v_threshold, decoder.CifSeeds.threshold = 0.1, 0.2
decoder.CafScored.default_score_th = 0.1
decoder.CifCaf.keypoint_threshold = 0.001
decoder.nms.Keypoints.keypoint_threshold = 0.001
fields = [synthetic_fields((641, 641), 5, pose_seed=b, jitter_seed=1000 + b)[:2] for b in range(B)]
cif = torch.from_numpy(np.stack([f[0] for f in fields])).pin_memory()
caf = torch.from_numpy(np.stack([f[1] for f in fields])).pin_memory()
dec = decoder.CifCaf(decoder.FieldConfig(), keypoints=COCO_KEYPOINTS, skeleton=[list(l) for l in COCO_PERSON_SKELETON])
dec.host_chunk = CHUNK
out = {}
for mode in ('copy', 'mapped'):
    dec.host_transfer = mode
    eng = dec.engine()
    for _ in range(3):
        dec.decode_batch(cif, caf)
    n_chunks = (B + CHUNK - 1) // CHUNK
    prof = StageProfile(8 * n_chunks)
    params = dec.params()
    params.profile = prof.handle
    t0 = time.perf_counter()
    for _ in range(8):
        for _buf, _b0, _b1 in eng.run_host_chunked(cif, caf, params, CHUNK):
            pass
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) / 8
    ms = prof.read().mean(axis=0)
    out[mode] = {'wall_ms_no_python_objects': wall * 1e3, 'stages_ms_per_chunk': dict(zip(prof.names, [float(v) for v in ms])),
                 'kernels_ms_per_chunk': float(ms.sum())}
    t0 = time.perf_counter()
    for _ in range(8):
        dec.decode_batch(cif, caf)
    out[mode]['wall_ms_public_api'] = (time.perf_counter() - t0) / 8 * 1e3
print(json.dumps(out, indent=1))
