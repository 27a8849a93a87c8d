#!/usr/bin/env python3
"""Repeat-run determinism under load: the substitute for compute-sanitizer's racecheck, which is closed on
this pool.  A shared-memory or global race in a kernel of the path shows up as a result that differs from
one run to the next once several batches are in flight and the SMs are shared; every decode here must give
the same packed bytes on every repetition, and the first image of each workload must equal the CPU oracle.

    python scripts/determinism_stress.py [reps]
"""
import hashlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    reps = int(sys.argv[1]) if len(sys.argv) > 1 else 40
    from openpifpaf_b200 import decoder
    from openpifpaf_b200.datasets.constants import COCO_PERSON_SKELETON, DENSER_COCO_PERSON_CONNECTIONS
    from openpifpaf_b200.decoder.engine import DecoderParams, DeviceDecoder
    from openpifpaf_b200.encoder import synthetic_fields
    from oracle import oracle as orc
    import bench
    bench.configure_profile(decoder)
    skel = list(COCO_PERSON_SKELETON)
    dense = skel + list(DENSER_COCO_PERSON_CONNECTIONS)
    cases = [
        ('c1_641_5p x16', (641, 641), 5, False, 16, skel, 64),
        ('crowd_1281_50p x4', (1281, 1281), 50, False, 4, skel, 256),
        ('dense_641_5p x16', (641, 641), 5, True, 16, dense, 64),
    ]
    lanes = 8
    for label, hw, n_poses, is_dense, B, skeleton, ann_cap in cases:
        items = [synthetic_fields(hw, n_poses, pose_seed=b, jitter_seed=b + 1, dense=is_dense)[:2] for b in range(B)]
        cif = torch.from_numpy(np.stack([i[0] for i in items])).cuda()
        caf = torch.from_numpy(np.stack([i[1] for i in items])).cuda()
        eng = DeviceDecoder(skeleton, ann_cap=ann_cap)
        params = DecoderParams()
        digests = set()
        first = None
        for rep in range(reps):
            tickets = [eng.submit(cif, caf, params, n_lanes=lanes, fetch=True) for _ in range(lanes)]
            for t in tickets:
                res = eng.collect(t.wait())
                blob = b''.join(np.ascontiguousarray(a).tobytes() for a in (
                    res.n_ann, res.status, res.n_pre, res.score, res.source, res.data, res.scales, res.pre,
                    res.order))
                digests.add(hashlib.sha256(blob).hexdigest())
                if first is None:
                    first = res.data[:int(res.n_ann[0])].copy()
        o = orc.decode(items[0][0], items[0][1], skeleton, orc.Config(exp_mode=1))
        assert len(first) == len(o.data) and np.array_equal(first, o.data), label
        assert len(digests) == 1, (label, len(digests))
        print('%-20s %4d decodes in flight on %d lanes: one digest %s, image 0 equals the oracle' % (
            label, reps * lanes, lanes, next(iter(digests))[:16]), flush=True)


if __name__ == '__main__':
    main()
