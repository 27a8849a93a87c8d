"""Host streaming (CifCaf.decode_stream on pinned batches): images/s against sub-batch size and lanes."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from openpifpaf_b200 import decoder
from openpifpaf_b200.datasets.constants import COCO_KEYPOINTS
bench.configure_profile(decoder)
cif, caf, skel = bench.make_batch('coco17_641_5p', 64)
dec = decoder.CifCaf(decoder.FieldConfig(), keypoints=COCO_KEYPOINTS, skeleton=skel)
h_cif, h_caf = torch.from_numpy(cif).pin_memory(), torch.from_numpy(caf).pin_memory()
K = 60
SETS = ((16, 4), (16, 8), (16, 12), (32, 3), (32, 4), (32, 6), (32, 8), (64, 2), (64, 3), (64, 4), (64, 6))
if len(sys.argv) > 1:
    SETS = tuple(tuple(int(v) for v in a.split('x')) for a in sys.argv[1:])
for chunk, lanes in SETS:
    dec.stream_chunk, dec.stream_lanes = chunk, lanes
    for _ in dec.decode_stream((h_cif, h_caf) for _ in range(4)):
        pass
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in dec.decode_stream((h_cif, h_caf) for _ in range(K)):
        pass
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) / K * 1e3
    st = dec.stream_stats
    print('chunk %2d lanes %2d: %.3f ms/step  %.0f img/s   host: submit %.2f wait %.2f build %.2f ms/step' % (
        chunk, lanes, ms, 64 / ms * 1e3, 1e3 * st['submit_s'] / K, 1e3 * st['wait_s'] / K, 1e3 * st['build_s'] / K), flush=True)
