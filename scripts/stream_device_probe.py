#!/usr/bin/env python3
"""decode_stream on CUDA fields: ms per batch against the number of lanes and the length of the stream."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from openpifpaf_b200 import decoder
from openpifpaf_b200.datasets.constants import COCO_KEYPOINTS
bench.configure_profile(decoder)
cif, caf, skel = bench.make_batch('coco17_641_5p', 64)
dec = decoder.CifCaf(decoder.FieldConfig(), keypoints=COCO_KEYPOINTS, skeleton=skel)
d = (torch.from_numpy(cif).cuda(), torch.from_numpy(caf).cuda())
for lanes, n in ((8, 20), (8, 200), (12, 200), (16, 200), (4, 200)):
    for rep in range(2):
        t0 = time.perf_counter()
        cnt = sum(len(a) for out in dec.decode_stream((d for _ in range(n)), n_lanes=lanes) for a in out)
        dt = time.perf_counter() - t0
    st = dec.stream_stats
    print('lanes %2d, %3d batches: %.3f ms/batch (%.0f img/s)  submit %.3f wait %.3f build %.3f' % (
        lanes, n, 1e3 * dt / n, 64 * n / dt, 1e3 * st['submit_s'] / n, 1e3 * st['wait_s'] / n, 1e3 * st['build_s'] / n), flush=True)
