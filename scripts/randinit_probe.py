"""GPU probe: decode fields of a randomly initialised resnet50 head."""
import sys, time, json
import numpy as np
import torch
sys.path.insert(0, '.')
from openpifpaf_b200 import decoder, workloads
from openpifpaf_b200.datasets.constants import COCO_KEYPOINTS, COCO_PERSON_SKELETON
from openpifpaf_b200.decoder.engine import StageProfile
import bench

B = int(sys.argv[1]) if len(sys.argv) > 1 else 4
max_scale = float(sys.argv[2]) if len(sys.argv) > 2 and sys.argv[2] != 'none' else None
bench.configure_profile(decoder)
decoder.CifCaf.max_annotations = 1024
t0 = time.time()
cif, caf = workloads.random_init_head_fields(B, seed=0, max_scale=max_scale)
torch.cuda.synchronize()
print('fields in %.1fs' % (time.time() - t0), cif.shape, caf.shape, flush=True)
print(json.dumps(workloads.field_statistics(cif)), flush=True)
dec = decoder.CifCaf(decoder.FieldConfig(), keypoints=COCO_KEYPOINTS, skeleton=COCO_PERSON_SKELETON)
eng, params = dec.engine(), dec.params()
for nb in (1, B):
    c, a = cif[:nb].contiguous(), caf[:nb].contiguous()
    t0 = time.time()
    buf = eng.run_device(c, a, params)
    torch.cuda.synchronize()
    print('B=%d first decode %.3fs status %s n_ann %s' % (nb, time.time() - t0, buf.dev['status'].cpu().tolist(),
          buf.dev['n_ann'].cpu().tolist()), flush=True)
    prof = StageProfile(3)
    params.profile = prof.handle
    t0 = time.time()
    for _ in range(3):
        eng.run_device(c, a, params)
    torch.cuda.synchronize()
    dt = (time.time() - t0) / 3
    params.profile = None
    ms = prof.read().mean(axis=0)
    print('B=%d %.2f ms/step %.1f img/s' % (nb, dt * 1e3, nb / dt), {n: round(float(v), 3) for n, v in zip(prof.names, ms)}, flush=True)
