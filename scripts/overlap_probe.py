#!/usr/bin/env python3
"""Probe: decode steps issued round-robin on S streams (one engine + workspace per stream) against
the single-stream loop.  Same work per step; shows how much of the latency-bound stages (assembly,
sort, NMS: one CTA per image) hides behind another batch's HBM-bound stages."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

from openpifpaf_b200.datasets.constants import COCO_PERSON_SKELETON  # noqa: E402
from openpifpaf_b200.decoder.engine import DecoderParams, DeviceDecoder  # noqa: E402
from openpifpaf_b200.encoder import synthetic_fields  # noqa: E402

B, STEPS = int(os.environ.get('B', 64)), int(os.environ.get('STEPS', 40))
fields = [synthetic_fields((641, 641), 5, pose_seed=b, jitter_seed=1000 + b)[:2] for b in range(B)]
cif = torch.from_numpy(np.stack([f[0] for f in fields])).cuda()
caf = torch.from_numpy(np.stack([f[1] for f in fields])).cuda()
params = DecoderParams()
out = {}
for S in (1, 2, 3):
    engines = [DeviceDecoder(COCO_PERSON_SKELETON) for _ in range(S)]
    streams = [torch.cuda.Stream() for _ in range(S)]
    for k in range(2 * S):
        with torch.cuda.stream(streams[k % S]):
            engines[k % S].run_device(cif, caf, params)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for s in streams:
        s.wait_event(a)
    for k in range(STEPS):
        with torch.cuda.stream(streams[k % S]):
            engines[k % S].run_device(cif, caf, params)
    for s in streams:
        torch.cuda.current_stream().wait_stream(s)
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / STEPS
    out['streams_%d' % S] = {'ms_per_step': ms, 'images_per_s': B / ms * 1e3}
print(json.dumps(out))
