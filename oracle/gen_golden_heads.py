#!/usr/bin/env python3
"""Generate tests/golden/head_*.npz from the UNMODIFIED reference head modules
(oracle/_ref): CompositeFieldFused in eval mode + CifCafCollector, on seeded raw
convolution outputs.  TEST INFRASTRUCTURE; run in the build container:

    python oracle/gen_golden_heads.py

Stored: the recipe of the raw tensors (seed, shapes; regenerated at test time by
``raw_head_outputs`` below and checked by sha256) and the reference's (cif, caf).
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, 'tests', 'golden')

# name: (batch, h, w, seed)
CASES = {'head_2x5x7': (2, 5, 7, 21), 'head_1x3x3': (1, 3, 3, 22)}


def raw_head_outputs(batch, h, w, seed, n_cif=17, n_caf=19):
    """Seeded stand-ins for the two head convolutions' outputs (before the pixel shuffle)."""
    rng = np.random.RandomState(seed)
    x_cif = rng.normal(0.0, 2.0, (batch, n_cif * 5 * 4, h, w)).astype(np.float32)
    x_caf = rng.normal(0.0, 2.0, (batch, n_caf * 9 * 4, h, w)).astype(np.float32)
    return x_cif, x_caf


def reference_fields(ref, x_cif, x_caf):
    """Run the reference's own modules downstream of the convolution."""
    import torch
    from openpifpaf_b200.datasets.constants import (COCO_KEYPOINTS, COCO_PERSON_SIGMAS,
                                                    COCO_PERSON_SKELETON, COCO_UPRIGHT_POSE)
    heads = ref.network.heads
    metas = [heads.IntensityMeta('cif', COCO_KEYPOINTS, COCO_PERSON_SIGMAS, COCO_UPRIGHT_POSE,
                                 COCO_PERSON_SKELETON),
             heads.AssociationMeta('caf', COCO_KEYPOINTS, COCO_PERSON_SIGMAS, COCO_UPRIGHT_POSE,
                                   COCO_PERSON_SKELETON)]
    outs = []
    for meta, x in zip(metas, (x_cif, x_caf)):
        head = heads.CompositeFieldFused(meta, 8).eval()
        head.conv = torch.nn.Identity()          # feed the convolution's output directly
        with torch.no_grad():
            outs.append(head(torch.from_numpy(x.copy())))
    with torch.no_grad():
        cif, caf = heads.CifCafCollector([0], [1])(outs)
    return cif.numpy(), caf.numpy()


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    from oracle import build_ref
    ref = build_ref.import_reference()
    import openpifpaf.network.heads  # noqa: F401  (staged reference)
    os.makedirs(GOLDEN, exist_ok=True)
    for name, (batch, h, w, seed) in CASES.items():
        x_cif, x_caf = raw_head_outputs(batch, h, w, seed)
        cif, caf = reference_fields(ref, x_cif, x_caf)
        np.savez_compressed(os.path.join(GOLDEN, name + '.npz'), batch=batch, h=h, w=w, seed=seed,
                            x_cif_sha256=np.array(sha(x_cif)), x_caf_sha256=np.array(sha(x_caf)),
                            cif=cif, caf=caf)
        print(name, cif.shape, caf.shape)


if __name__ == '__main__':
    main()
