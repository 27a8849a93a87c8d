"""cProfile of CifCaf.decode_batch on pinned host fields (where does the host time of e2e go)."""
import cProfile, pstats, sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from openpifpaf_b200 import decoder
from openpifpaf_b200.datasets.constants import COCO_KEYPOINTS
bench.configure_profile(decoder)
mode = sys.argv[1] if len(sys.argv) > 1 else 'mapped'
cif, caf, skel = bench.make_batch('coco17_641_5p', 64)
dec = decoder.CifCaf(decoder.FieldConfig(), keypoints=COCO_KEYPOINTS, skeleton=skel)
dec.host_transfer = mode
h_cif, h_caf = torch.from_numpy(cif).pin_memory(), torch.from_numpy(caf).pin_memory()
for _ in range(3):
    dec.decode_batch(h_cif, h_caf)
t0 = time.perf_counter()
for _ in range(20):
    dec.decode_batch(h_cif, h_caf)
print('%s: %.3f ms/step' % (mode, (time.perf_counter() - t0) / 20 * 1e3))
pr = cProfile.Profile()
pr.enable()
for _ in range(20):
    dec.decode_batch(h_cif, h_caf)
pr.disable()
pstats.Stats(pr).sort_stats('tottime').print_stats(18)
