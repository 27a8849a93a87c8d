#!/usr/bin/env python3
"""Soak: decodes in flight for a given time in ONE process; reports how many completed and the first CUDA error.
usage: soak.py WORKLOAD SECONDS [lanes=8] [compress=1] [knob=value ...]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import bench  # noqa: E402
from openpifpaf_b200 import decoder  # noqa: E402
from openpifpaf_b200.datasets.constants import COCO_KEYPOINTS  # noqa: E402
from openpifpaf_b200.decoder import engine  # noqa: E402

wl, seconds = sys.argv[1], float(sys.argv[2])
opts = dict(kv.split('=') for kv in sys.argv[3:])
lanes = int(opts.pop('lanes', 8))
skip = opts.pop('skip', None)      # library built with -DPP_DEBUG_SKIP: stages left out after the first full round
engine.BatchBuffers.compress_hr = bool(int(opts.pop('compress', 1)))
bench.configure_profile(decoder)
cif, caf, skel = bench.make_batch(wl, bench.WORKLOADS[wl][2])
dec = decoder.CifCaf(decoder.FieldConfig(), keypoints=COCO_KEYPOINTS, skeleton=skel)
eng, params = dec.engine(), dec.params()
params.tuning = {k: int(v) for k, v in opts.items()}
d_cif, d_caf = torch.from_numpy(cif).cuda(), torch.from_numpy(caf).cuda()
done = 0
tickets = [eng.submit(d_cif, d_caf, params, n_lanes=lanes) for _ in range(lanes)]     # one full decode per lane
for t in tickets:
    t.wait()
if skip is not None:
    os.environ['PP_DEBUG_SKIP'] = skip
    opts['skip'] = skip
t0 = time.time()
tickets = []
try:
    while time.time() - t0 < seconds:
        for _ in range(64):
            if len(tickets) == lanes:
                tickets.pop(0).wait()
                done += 1
            tickets.append(eng.submit(d_cif, d_caf, params, n_lanes=lanes))
    for t in tickets:
        t.wait()
        done += 1
    torch.cuda.synchronize()
    print('%s lanes %d compress %d %s: %d decodes in %.0f s, no error' % (
        wl, lanes, engine.BatchBuffers.compress_hr, opts, done, time.time() - t0), flush=True)
except Exception as e:  # pylint: disable=broad-except
    print('%s lanes %d compress %d %s: ERROR after %d decodes (%.0f s): %s' % (
        wl, lanes, engine.BatchBuffers.compress_hr, opts, done, time.time() - t0, str(e).splitlines()[0]), flush=True)
    try:
        import ctypes as C
        from openpifpaf_b200 import _lib
        out = (C.c_int * 64)()
        n = _lib.lib().pp_debug_crumbs(out, 64)
        print('crumbs (stage each stream had reached: 10 scan, 11 strips, 30 wide splat, 12 seeds, 13 seed merge, '
              '14 CafScored, 15 assembly, 16 nms mark, 50 pack, 99 done):', list(out[:n]), flush=True)
    except Exception as e2:  # pylint: disable=broad-except
        print('no crumbs:', e2)
    os._exit(3)
