"""Per-phase cycles of assemble_kernel (library built with PP_EXTRA_NVCC=-DPP_ASM_TIMING)."""
import ctypes as C, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from openpifpaf_b200 import _lib, decoder
from openpifpaf_b200.datasets.constants import COCO_KEYPOINTS
bench.configure_profile(decoder)
WL = sys.argv[1] if len(sys.argv) > 1 else 'coco17_641_5p'
NB = bench.WORKLOADS[WL][2]
if WL in bench.RANDINIT_MAX_SCALE:
    decoder.CifCaf.max_annotations = bench.RANDINIT_ANN_CAP
    NB = 8
cif, caf, skel = bench.make_batch(WL, NB)
dec = decoder.CifCaf(decoder.FieldConfig(), keypoints=COCO_KEYPOINTS, skeleton=skel)
eng, params = dec.engine(), dec.params()
d_cif, d_caf = torch.from_numpy(cif).cuda(), torch.from_numpy(caf).cuda()
for _ in range(3): eng.run_device(d_cif, d_caf, params)
torch.cuda.synchronize()
out = (C.c_ulonglong * 12)()
_lib.lib().pp_debug_asm_cycles(out, 1)
N = 3 if WL in bench.RANDINIT_MAX_SCALE else 10
for _ in range(N): eng.run_device(d_cif, d_caf, params)
torch.cuda.synchronize()
_lib.lib().pp_debug_asm_cycles(out, 1)
names = ['setup', 'seed_search', 'wave_pick', 'wave_grow', '-', 'store_boxes', 'occ_update', 'nms_tail']
tot = sum(out[:8])
print('%s: per image %.2f waves, %.2f grown poses, %.2f committed' % (
    WL, out[8] / N / NB, out[9] / N / NB, out[10] / N / NB))
for n, v in zip(names, out[:8]):
    print('%-12s %8.1f us/image  %5.1f%%' % (n, v / N / NB / 1965.0, 100.0 * v / tot))
print('total %.1f us/image' % (tot / N / NB / 1965.0))
