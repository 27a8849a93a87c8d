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
K = 30
for chunk, lanes in ((8, 8), (8, 16), (16, 4), (16, 8), (16, 12), (32, 4), (32, 8), (64, 4)):
    dec.stream_chunk, dec.stream_lanes = chunk, lanes
    for _ in dec.decode_stream((h_cif, h_caf) for _ in range(4)):
        pass
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in dec.decode_stream((h_cif, h_caf) for _ in range(K)):
        pass
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) / K * 1e3
    print('chunk %2d lanes %2d: %.2f ms/step  %.0f img/s' % (chunk, lanes, ms, 64 / ms * 1e3), flush=True)
