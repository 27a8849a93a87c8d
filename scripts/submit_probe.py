"""Is `value` bound by the host submitting batches?  Wall time of the submit loop against the device time."""
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
ts = [eng.submit(d_cif, d_caf, params, n_lanes=LANES) for _ in range(LANES)]
for t in ts: t.wait()
torch.cuda.synchronize()
# (a) submit only: never wait inside the loop (tickets marked done by hand; buffers are simply reused)
t0 = time.perf_counter()
for _ in range(K):
    t = eng.submit(d_cif, d_caf, params, n_lanes=LANES)
    t.done = True
t_submit = time.perf_counter() - t0
torch.cuda.synchronize()
t_total = time.perf_counter() - t0
print('submit loop %.3f ms/step, until the GPU is idle %.3f ms/step' % (t_submit / K * 1e3, t_total / K * 1e3))
# (b) one pp_decode_batch call alone
t0 = time.perf_counter()
for _ in range(K):
    eng.run_device(d_cif, d_caf, params)
t_call = time.perf_counter() - t0
torch.cuda.synchronize()
print('run_device (one stream) host time %.3f ms/step' % (t_call / K * 1e3))
