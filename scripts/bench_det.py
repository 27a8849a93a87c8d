#!/usr/bin/env python3
"""Time the CifDet path (pp_cifdet_decode_batch) on batch 64 x 641 px, 20 boxes, 80 categories,
next to the reference's CifDet on the host cores (oracle/_ref, one process per core).
Prints one JSON line.  Not the headline bench (bench.py); a measurement of SURVEY 8f-4."""
import json
import multiprocessing
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

B = int(os.environ.get('B', 64))
NCAT = int(os.environ.get('NCAT', 80))
NBOX = int(os.environ.get('NBOX', 20))
STEPS = int(os.environ.get('STEPS', 20))


def _ref_worker(field):
    sys.path.insert(0, ROOT)
    from oracle import build_ref
    from oracle.gen_golden_det import run_reference
    ref = build_ref.import_reference()
    return len(run_reference(ref, field, 0.0)[2])


def main():
    import openpifpaf_b200.decoder as decoder
    from openpifpaf_b200.decoder.generator.cifdet import run_cifdet
    from openpifpaf_b200.encoder import synthetic_det_fields
    decoder.CifHr.v_threshold, decoder.CifSeeds.threshold = 0.1, 0.2
    fields = np.stack([synthetic_det_fields((641, 641), NBOX, box_seed=b, jitter_seed=100 + b,
                                            n_categories=NCAT)[0] for b in range(B)])
    dev = torch.from_numpy(fields).cuda()
    for _ in range(3):
        t = run_cifdet(dev, 8)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(STEPS):
        t = run_cifdet(dev, 8)
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / STEPS
    n_det = int(t['n_det'].sum())
    out = {'workload': 'cifdet: batch %d x 641x641, %d categories, %d boxes/image' % (B, NCAT, NBOX),
           'value': B / ms * 1e3, 'unit': 'images/s', 'ms_per_step': ms, 'detections_per_step': n_det,
           'note': 'includes per-call workspace/result allocation (run_cifdet)'}
    from oracle import build_ref
    if build_ref.staged():
        cores = os.cpu_count() or 1
        sample = [fields[i] for i in range(min(B, 2 * cores))]
        with multiprocessing.get_context('spawn').Pool(cores) as pool:
            pool.map(_ref_worker, sample[:cores])          # warm up / import
            t0 = time.perf_counter()
            counts = pool.map(_ref_worker, sample)
            dt = time.perf_counter() - t0
        out['cpu_baseline'] = {'value': len(sample) / dt, 'unit': 'images/s', 'cores': cores,
                               'kind': 'reference', 'sample': '%d images, pool of %d' % (len(sample), cores),
                               'detections': int(sum(counts))}
    print(json.dumps(out))


if __name__ == '__main__':
    main()
