#!/usr/bin/env python3
"""Small decodes for compute-sanitizer (memcheck / racecheck): every kernel of the path runs at least once
on inputs small enough for the tool's slow-down, and each result is checked against the CPU oracle.

    compute-sanitizer --tool memcheck  python scripts/sanitize_case.py
    compute-sanitizer --tool racecheck python scripts/sanitize_case.py
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))


def main():
    from openpifpaf_b200 import decoder
    from openpifpaf_b200.datasets.constants import (COCO_KEYPOINTS, COCO_PERSON_SKELETON,
                                                    DENSER_COCO_PERSON_CONNECTIONS)
    from openpifpaf_b200.decoder.engine import DecoderParams, DeviceDecoder
    from openpifpaf_b200.encoder import random_init_like_fields, synthetic_fields
    from oracle import oracle as orc
    import bench
    bench.configure_profile(decoder)
    skel = list(COCO_PERSON_SKELETON)

    def check(cif, caf, skeleton, label, ann_cap=256, **kw):
        eng = DeviceDecoder(skeleton, ann_cap=ann_cap)
        buf = eng.run_device(torch.from_numpy(cif[None]).cuda(), torch.from_numpy(caf[None]).cuda(),
                             DecoderParams(**kw), keep=('ann', 'hr', 'seeds', 'caf'))
        torch.cuda.synchronize()
        cfg = {'force_complete': 1, 'keypoint_th': 0.0, 'nms_keypoint_th': 0.0} if kw.get('force_complete') else {}
        if kw.get('connection_method') == 'max':
            cfg['connection_method'] = 1
        if kw.get('greedy'):
            cfg['greedy'] = 1
        o = orc.decode(cif, caf, skeleton, orc.Config(exp_mode=1, **cfg))
        n = int(buf.dev['n_ann'][0])
        assert n == len(o.data), (label, n, len(o.data))
        assert np.array_equal(buf.dev['ann_data'][0, :n].cpu().numpy(), o.data), label
        assert np.array_equal(buf.dev['hr'][0].cpu().numpy(), o.hr), label
        print('%-28s %3d annotations, bit-identical to the oracle' % (label, n), flush=True)

    cif, caf, _ = synthetic_fields((161, 161), 2, pose_seed=5, jitter_seed=6)
    check(cif, caf, skel, 'small 161 / 2 persons')
    check(cif, caf, skel, 'small, max + greedy', connection_method='max', greedy=True)
    check(cif, caf, skel, 'small, force_complete', force_complete=True, keypoint_threshold=0.0,
          nms_keypoint_threshold=0.0)
    check(cif, caf, skel, 'small, seed index mode', tuning={'asm_heavy_seeds': -1})
    cif, caf, _ = synthetic_fields((321, 481), 8, pose_seed=3, jitter_seed=4)
    check(cif, caf, skel, 'mid 321x481 / 8 persons')
    cif, caf, _ = synthetic_fields((161, 241), 3, pose_seed=11, jitter_seed=12, dense=True)
    check(cif, caf, skel + list(DENSER_COCO_PERSON_CONNECTIONS), 'dense 44 links')
    cif, caf = random_init_like_fields((13, 17), 1)
    check(cif, caf, skel, 'random-init-like (wide boxes)')
    # host entries: pinned fields, mapped transfer, packed results, streams
    items = [synthetic_fields((161, 161), 1 + b % 2, pose_seed=60 + b, jitter_seed=61 + b)[:2] for b in range(4)]
    h_cif = torch.from_numpy(np.stack([i[0] for i in items])).pin_memory()
    h_caf = torch.from_numpy(np.stack([i[1] for i in items])).pin_memory()
    dec = decoder.CifCaf(decoder.FieldConfig(), keypoints=COCO_KEYPOINTS, skeleton=skel)
    dec.stream_chunk, dec.stream_lanes = 2, 2
    want = dec.decode_batch(h_cif.cuda(), h_caf.cuda())
    for name, got in (('decode_batch (host, mapped)', dec.decode_batch(h_cif, h_caf)),
                      ('decode_stream (host)', next(iter(dec.decode_stream([(h_cif, h_caf)]))))):
        assert [len(a) for a in got] == [len(a) for a in want], name
        for ga, wa in zip(got, want):
            for g, w in zip(ga, wa):
                assert np.array_equal(g.data, w.data), name
        print('%-28s equals the device-field decode' % name, flush=True)
    print('sanitize_case: ok')


if __name__ == '__main__':
    main()
