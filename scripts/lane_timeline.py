#!/usr/bin/env python3
"""Timeline of batches in flight: the library's stage events (pp_profile_read_abs) of K steps issued
round-robin over the lanes, drawn as one text row per lane (one character per RES microseconds):
  c field scan   S splat   e seed segments   o seed merge   a CafScored + bins   A assembly + NMS   p pack
An interval runs from the event before a stage to the event before the next one on that lane's stream, i.e.
it contains the time the stage's kernel waited for resources.  Usage: lane_timeline.py [workload] [lanes] [steps]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault('CUDA_DEVICE_MAX_CONNECTIONS', '32')
import torch  # noqa: E402
import bench  # noqa: E402
from openpifpaf_b200 import decoder  # noqa: E402
from openpifpaf_b200.datasets.constants import COCO_KEYPOINTS  # noqa: E402
from openpifpaf_b200.decoder.engine import StageProfile  # noqa: E402

WL = sys.argv[1] if len(sys.argv) > 1 else 'coco17_641_5p'
LANES = int(sys.argv[2]) if len(sys.argv) > 2 else 8
K = int(sys.argv[3]) if len(sys.argv) > 3 else 3 * LANES
COLS = int(os.environ.get('COLS', '150'))
bench.configure_profile(decoder)
cif, caf, skel = bench.make_batch(WL, bench.WORKLOADS[WL][2])
dec = decoder.CifCaf(decoder.FieldConfig(), keypoints=COCO_KEYPOINTS, skeleton=skel)
eng, params = dec.engine(), dec.params()
eng.use_splat_gate = bool(int(os.environ.get('GATE', '1')))
tuning = os.environ.get('TUNING', '')
if tuning:
    params.tuning = {k: int(v) for k, v in (kv.split('=') for kv in tuning.split(','))}
d_cif, d_caf = torch.from_numpy(cif).cuda(), torch.from_numpy(caf).cuda()
for _ in range(2):
    ts = [eng.submit(d_cif, d_caf, params, n_lanes=LANES) for _ in range(LANES)]
    for t in ts:
        t.wait()
prof = StageProfile(K)
params.profile = prof.handle
torch.cuda.synchronize()
tickets = []
for k in range(K):
    if len(tickets) == LANES:
        tickets.pop(0).wait()
    tickets.append(eng.submit(d_cif, d_caf, params, n_lanes=LANES))
for t in tickets:
    t.wait()
torch.cuda.synchronize()
ms = prof.read_abs().astype(np.float64) * 1e3        # us
sym = 'cSeoaAp'
end = ms.max()
RES = max(1.0, end / COLS)
print('%s, %d lanes, %d steps, splat gate %s: %.1f us per step overall; one char = %.0f us' % (
    WL, LANES, K, 'on' if eng.use_splat_gate else 'off', end / K, RES))
rows = [[' '] * (int(end / RES) + 2) for _ in range(LANES)]
for s in range(K):
    lane = s % LANES
    for k in range(len(sym)):
        a, b = int(ms[s, k] / RES), int(ms[s, k + 1] / RES)
        for x in range(a, max(b, a + 1)):
            rows[lane][x] = sym[k] if rows[lane][x] == ' ' else rows[lane][x]
for lane in range(LANES):
    print('lane %2d |%s' % (lane, ''.join(rows[lane])))
# how much of the time is some splat interval open, and how many at once
grid = np.zeros(int(end) + 2, np.int32)
for s in range(K):
    grid[int(ms[s, 1]):int(ms[s, 2])] += 1
steady = grid[int(end * 0.25):int(end * 0.75)]
print('steady state (middle half): splat intervals open: 0 -> %.0f%%, 1 -> %.0f%%, 2 -> %.0f%%, >=3 -> %.0f%% of the time' % (
    100 * (steady == 0).mean(), 100 * (steady == 1).mean(), 100 * (steady == 2).mean(), 100 * (steady >= 3).mean()))
dur = ms[:, 1:] - ms[:, :-1]
print('mean interval per stage (us): ' + ', '.join('%s %.0f' % (n, v) for n, v in zip(prof.names, dur[LANES:].mean(axis=0))))
