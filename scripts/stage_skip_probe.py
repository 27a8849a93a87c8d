"""Which stage costs what with batches in flight?  Library built with PP_EXTRA_NVCC=-DPP_DEBUG_SKIP; stages are
left out by mask (1 seeds + sort, 2 CafScored + binning, 4 assembly, 8 NMS, 16 CifHr) and the lanes throughput
is measured.  Timing experiment only: skipped stages leave the previous results in the buffers."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault('CUDA_DEVICE_MAX_CONNECTIONS', '32')
import torch
import bench
from openpifpaf_b200 import decoder
from openpifpaf_b200.datasets.constants import COCO_KEYPOINTS
bench.configure_profile(decoder)
cif, caf, skel = bench.make_batch('coco17_641_5p', 64)
dec = decoder.CifCaf(decoder.FieldConfig(), keypoints=COCO_KEYPOINTS, skeleton=skel)
eng, params = dec.engine(), dec.params()
d_cif, d_caf = torch.from_numpy(cif).cuda(), torch.from_numpy(caf).cuda()
LANES, K = 8, 64
os.environ['PP_DEBUG_SKIP'] = '0'
for _ in range(2):
    ts = [eng.submit(d_cif, d_caf, params, n_lanes=LANES) for _ in range(LANES)]
    for t in ts: t.wait()
names = {0: 'everything', 8: 'no NMS', 12: 'no assembly, NMS', 14: 'CifHr + seeds only', 15: 'CifHr only',
         16: 'all but CifHr', 31 - 4: 'assembly only', 31 - 2: 'CafScored + binning only', 4: 'no assembly',
         1: 'no seeds', 2: 'no CafScored'}
for mask in (0, 8, 4, 12, 2, 1, 14, 15, 16, 27, 29):
    os.environ['PP_DEBUG_SKIP'] = str(mask)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(K):
        t = eng.submit(d_cif, d_caf, params, n_lanes=LANES)
        t.done = True
    main = torch.cuda.current_stream()
    for lane in eng.lanes(LANES):
        main.wait_stream(lane.stream)
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / K
    print('mask %2d %-26s %.3f ms/step  %.0f img/s' % (mask, names.get(mask, ''), ms, 64 / ms * 1e3), flush=True)
