#!/usr/bin/env python3
"""bench.py -- decoded images/s of the CifCaf decode path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
                    [--workload NAME] [--batch B]

A "step" is one pass of the hot path (CifHr -> seeds -> CafScored -> assembly
+ keypoint NMS -> packed results) over one batch of synthetic CIF/CAF fields.
One JSON line is printed by rank 0 (contract in DESIGN.md section "Measurement"):

  value         images/s, fields already resident in HBM, results left in HBM:
                K steps (one pp_decode_batch = 8 kernels each) issued round-robin
                over --lanes streams / workspaces and over as many DISTINCT input
                batches, CUDA events around a region of R such K-step passes issued
                back to back (the lanes fill and drain once per region); three
                regions (>= 0.5 s in all) are timed and the median ms per K steps
                is reported (`passes` = 3 R);
  value_one_lane, stages_ms
                the same K steps strictly one at a time on one stream, with the
                library's per-stage CUDA-event profile;
  e2e           images/s through the public API CifCaf.decode_stream(batches) with
                PINNED HOST fields and default settings: host -> device transfer,
                kernels, D2H of the packed results and the construction of the
                Annotation objects are all inside the timed region (median of 5
                passes of K steps; e2e.one_call_per_step: one synchronous
                CifCaf.decode_batch per step).  The default host_transfer='auto'
                moves what the decode reads: the confidence planes by DMA, the
                other planes of the cells above threshold by zero-copy reads;
  e2e_device_fields
                the reference's real call path after the network
                (Generator.batch, generator.py:84-101): CUDA field tensors ->
                CifCaf.decode_stream -> List[List[Annotation]] on the host;
  e2e_full_copy the e2e call with host_transfer='copy': every plane of every
                field is copied (PCIe-bound: 430 MB per 64 images);
  roofline      the CifHr strips kernel (the kernel that writes the hr maps, into
                compressible memory: pp_hr_alloc) against the measured HBM copy peak
                of MEASURED_PEAKS.json, timed in the one-lane pass; a memset of the
                same size into plain and into compressible memory beside it;
                in_flight = while lanes overlap;
  cpu_baseline  the UNMODIFIED reference decoder (oracle/_ref, its Cython
                functional module) on this box's host cores, bounded sample;
  --impl reference   the same reference decoder as its own arm.

--workload shard_4096 is BASELINE configs[3]: 4096 distinct images, contiguous
shards of 4096 / N per rank, chunks of 64, wall time from the first launch to
the last Annotation on the host (strong scaling).

Under torchrun every rank decodes its own images (image sharding, no
collective on the data path); torch.distributed is used for the barrier and the
max-over-ranks of the times only.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
# more than 8 streams only run side by side with more hardware work queues (read when the CUDA context is
# created); the default workload is indifferent to it, the 16-lane workloads gain ~20 %
os.environ.setdefault('CUDA_DEVICE_MAX_CONNECTIONS', '32')

METRIC = 'decoded images/s (641px COCO-17 CIF/CAF fields, CifCaf decode)'
UNIT = 'images/s'

# name -> (image_hw, persons per image, default batch, dense skeleton)
WORKLOADS = {
    'coco17_641_5p': ((641, 641), 5, 64, False),
    'crowd_1281_50p': ((1281, 1281), 50, 16, False),
    'dense_641_5p': ((641, 641), 5, 64, True),
    'dense_641_5p_b256': ((641, 641), 5, 256, True),          # BASELINE configs[4] at its stated batch
    # BASELINE configs[3]: 4096 distinct coco17_641_5p images sharded over the ranks, chunks of 64
    'shard_4096': ((641, 641), 5, 64, False),
    # CifDet: 80 categories, 20 synthetic boxes per image; a secondary line (value + reference)
    'cifdet_641_20b': ((641, 641), 20, 64, False),
    # BASELINE configs[1] literally: fields of a randomly initialised resnet50 + CifCaf heads on torch.randn
    # images (degenerate, exp() scales up to 1e18, CifHr compute-bound, ~100 s/image for
    # the reference), and the documented tamed variant (scale channels clamped to 20 cells)
    'randinit_641': ((641, 641), 0, 64, False),
    'randinit_641_tamed': ((641, 641), 0, 64, False),
}
RANDINIT_MAX_SCALE = {'randinit_641': None, 'randinit_641_tamed': 20.0}
RANDINIT_ANN_CAP = 4096          # these fields give ~950 annotations per image (CifCaf.max_annotations)
DET_CATEGORIES = 80
DEFAULT_WORKLOAD = 'coco17_641_5p'
SHARD_IMAGES = 4096
MIN_TIMED_S = 0.5                # the K-step pass is repeated until this much has been timed
E2E_PASSES = 5


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='b200', choices=('b200', 'reference'))
    ap.add_argument('--workload', default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument('--batch', type=int, default=0, help='images per GPU per step (0 = workload default)')
    ap.add_argument('--lanes', type=int, default=0,
                    help='value: batches in flight (streams + workspaces), 1 = strictly one at a time; '
                         '0 = 8 (16 for the crowd / dense workloads, whose assembly CTAs are few and long)')
    ap.add_argument('--distinct', type=int, default=0,
                    help='distinct input batches the steps cycle through (0 = as many as lanes, at most 8)')
    ap.add_argument('--images', type=int, default=SHARD_IMAGES, help='shard_4096: images of the whole job')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--host-chunk', type=int, default=0, help='e2e: images per pipeline chunk (0 = decoder default)')
    ap.add_argument('--cpu-workers', type=int, default=0, help='reference arm: pool size (0 = all cores)')
    ap.add_argument('--tuning', default='', help='experiments: schedule knobs of pp_params, e.g. cifhr_ctas_per_sm=2')
    ap.add_argument('--value-only', action='store_true', help='experiments: print value / one-lane stages and stop')
    return ap.parse_args()


# ----------------------------------------------------------------------------
# workload

def _skeleton(dense):
    from openpifpaf_b200.datasets.constants import COCO_PERSON_SKELETON, DENSER_COCO_PERSON_CONNECTIONS
    return list(COCO_PERSON_SKELETON) + (list(DENSER_COCO_PERSON_CONNECTIONS) if dense else [])


def _gen_image(job):
    from openpifpaf_b200.encoder import synthetic_fields
    hw, n_poses, seed, dense = job
    cif, caf, _ = synthetic_fields(hw, n_poses, pose_seed=seed, jitter_seed=seed + 1, dense=dense)
    return cif, caf


_FILL = {}


def _fill_image(job):
    """Pool worker of fill_images: writes image `i` into the fork-inherited shared arrays."""
    i, spec = job
    cif, caf = _gen_image(spec)
    _FILL['cif'][i] = cif
    _FILL['caf'][i] = caf
    return i


def shared_array(shape):
    """float32 array in anonymous shared memory (inherited by forked workers; no /dev/shm needed)."""
    import mmap
    n = int(np.prod(shape)) * 4
    buf = mmap.mmap(-1, max(n, mmap.PAGESIZE))
    return np.frombuffer(buf, dtype=np.float32, count=int(np.prod(shape))).reshape(shape)


def fill_images(workload, seeds, procs=0):
    """Fields of the images with pose seeds `seeds`: (cif (n,F,5,H,W), caf (n,L,9,H,W)) float32 in shared
    anonymous memory, generated by a pool of forked workers.  Call BEFORE CUDA is initialised."""
    import multiprocessing
    hw, n_poses, _, dense = WORKLOADS[workload]
    H, W = (hw[0] - 1) // 8 + 1, (hw[1] - 1) // 8 + 1
    L = len(_skeleton(dense))
    n = len(seeds)
    cif, caf = shared_array((n, 17, 5, H, W)), shared_array((n, L, 9, H, W))
    _FILL['cif'], _FILL['caf'] = cif, caf
    jobs = [(i, (hw, n_poses, int(s), dense)) for i, s in enumerate(seeds)]
    procs = procs or min(len(jobs), os.cpu_count() or 1)
    if procs > 1:
        with multiprocessing.get_context('fork').Pool(procs) as pool:
            for _ in pool.imap_unordered(_fill_image, jobs, chunksize=4):
                pass
    else:
        for j in jobs:
            _fill_image(j)
    _FILL.clear()
    return cif, caf


def make_batch(workload, batch, first_seed=0):
    """`batch` distinct synthetic images of the workload: (cif, caf, skeleton)."""
    hw, n_poses, _, dense = WORKLOADS[workload]
    skeleton = _skeleton(dense)
    if workload in RANDINIT_MAX_SCALE:
        import torch
        from openpifpaf_b200 import workloads
        device = 'cuda' if torch.cuda.is_available() else 'cpu'   # the network is workload tooling only
        cif, caf = workloads.random_init_head_fields(batch, hw, seed=0, image_seed=first_seed + 1, device=device,
                                                     max_scale=RANDINIT_MAX_SCALE[workload])
        return cif.cpu().numpy(), caf.cpu().numpy(), skeleton
    cache = '/tmp/pifpaf_bench_%s_%d_%d.npz' % (workload, batch, first_seed)
    if os.path.exists(cache):
        z = np.load(cache)
        return z['cif'], z['caf'], skeleton
    items = [_gen_image((hw, n_poses, first_seed + i, dense)) for i in range(batch)]
    cif, caf = np.stack([c for c, _ in items]), np.stack([a for _, a in items])
    try:
        np.savez(cache + '.tmp.npz', cif=cif, caf=caf)
        os.replace(cache + '.tmp.npz', cache)
    except OSError:
        pass
    return cif, caf, skeleton


def make_det_batch(workload, batch, first_seed=0):
    """`batch` distinct synthetic CifDet fields (B, C, 7, H, W)."""
    from openpifpaf_b200.encoder import synthetic_det_fields
    hw, n_boxes, _, _ = WORKLOADS[workload]
    return np.stack([synthetic_det_fields(hw, n_boxes, box_seed=first_seed + i, jitter_seed=first_seed + i + 100,
                                          n_categories=DET_CATEGORIES)[0] for i in range(batch)])


def _ref_det_decode(field):
    """One image through the reference's CifDet (worker process of the reference arm)."""
    from oracle import build_ref
    ref = build_ref.import_reference()
    dec = ref.decoder
    if dec.CifSeeds.threshold is None:
        dec.CifHr.v_threshold, dec.CifSeeds.threshold, dec.CifSeeds.score_scale = 0.1, 0.2, 1.0
    cats = ['c%d' % i for i in range(field.shape[0])]
    return len(dec.generator.cifdet.CifDet(dec.FieldConfig(), cats)([field]))


def configure_profile(decoder):
    """SURVEY.md 8d decoder settings (what decoder.cli + configure yield)."""
    decoder.CifHr.v_threshold = 0.1
    decoder.CifSeeds.threshold = 0.2
    decoder.CifSeeds.score_scale = 1.0
    decoder.CafScored.default_score_th = 0.1
    decoder.CifCaf.connection_method = 'blend'
    decoder.CifCaf.greedy = False
    decoder.CifCaf.force_complete = False
    decoder.CifCaf.keypoint_threshold = 0.001
    decoder.nms.Keypoints.keypoint_threshold = 0.001
    decoder.nms.Keypoints.instance_threshold = 0.0
    decoder.nms.Keypoints.suppression = 0.0


# ----------------------------------------------------------------------------
# reference arm: the reference's own CPU decoder on the host cores

def _port_decode(item):
    from oracle import oracle as orc
    orc.lib().orc_set_threads(1)          # the pool already uses every core
    cif, caf, skeleton = item
    return len(orc.decode(cif, caf, skeleton, orc.Config(exp_mode=0)).data)


def _ref_decode_guarded(dec, fields):
    """One image through the reference decoder; the reference raises TypeError when a frontier bound ties an
    evaluated score exactly (cifcaf.py:261,285: tuple vs None in its PriorityQueue), which degenerate
    random-init fields do trigger.  Returns (annotation count or -1, seconds)."""
    t0 = time.perf_counter()
    try:
        n = len(dec._mappable_annotations(fields, None))
    except TypeError:
        n = -1
    return n, time.perf_counter() - t0


class ReferenceRunner:
    """Times Generator.batch's decode region (reference generator.py:93-98):
    worker_pool.starmap(decoder, fields) with a multiprocessing.Pool."""

    def __init__(self, workload, batch, workers):
        from oracle import build_ref
        self.cores = workers or os.cpu_count() or 1
        self.batch = batch
        # raw random-init fields: per-image guard against the reference's own frontier-tie TypeError
        self.guarded = RANDINIT_MAX_SCALE.get(workload, 0) is None and workload in RANDINIT_MAX_SCALE
        self.crashed, self.ok_images, self.ok_seconds = 0, 0, 0.0
        self.kind = 'reference' if build_ref.staged() else 'port'
        if workload.startswith('cifdet'):
            import multiprocessing
            if self.kind != 'reference':
                raise SystemExit('the CifDet reference arm needs the staged reference (oracle/_ref)')
            self.kind = 'reference-det'
            self.det_fields = list(make_det_batch(workload, batch))
            self.pool = multiprocessing.Pool(self.cores)
            self.version = 'openpifpaf CifDet (oracle/_ref)'
            return
        self.cif, self.caf, self.skeleton = make_batch(workload, batch)
        if self.kind == 'reference':
            ref = build_ref.import_reference()
            configure_profile(ref.decoder)
            from openpifpaf_b200.datasets.constants import COCO_KEYPOINTS
            self.dec = ref.decoder.CifCaf(ref.decoder.FieldConfig(), keypoints=COCO_KEYPOINTS,
                                          skeleton=[list(l) for l in self.skeleton],
                                          worker_pool=self.cores if self.cores > 1 else None)
            self.fields = [[self.cif[i], self.caf[i]] for i in range(batch)]
            self.version = 'openpifpaf %s (oracle/_ref, Cython functional)' % ref.__version__
        else:
            import multiprocessing
            from oracle import oracle as orc
            orc.build()
            self.pool = multiprocessing.Pool(self.cores) if self.cores > 1 else None
            self.items = [(self.cif[i], self.caf[i], self.skeleton) for i in range(batch)]
            self.version = 'oracle/cifcaf_oracle.c (C restatement)'

    def step(self):
        if self.kind == 'reference-det':
            return sum(self.pool.map(_ref_det_decode, self.det_fields))
        if self.kind == 'reference' and self.guarded:
            out = self.dec.worker_pool.starmap(_ref_decode_guarded, [(self.dec, f) for f in self.fields])
            self.crashed += sum(1 for n, _ in out if n < 0)
            self.ok_seconds += sum(t for n, t in out if n >= 0)
            self.ok_images += sum(1 for n, _ in out if n >= 0)
            return sum(n for n, _ in out if n >= 0)
        if self.kind == 'reference':
            out = self.dec.worker_pool.starmap(self.dec._mappable_annotations,
                                               zip(self.fields, [None] * self.batch))
            return sum(len(a) for a in out)
        if self.pool is not None:
            return sum(self.pool.map(_port_decode, self.items))
        return sum(_port_decode(i) for i in self.items)

    def single_process(self, repeats=5, max_images=64):
        """the same decoder in ONE process (the reference's DummyPool, generator.py:14-17):
        1 warm-up, best of `repeats` over the first `max_images` images.  Returns images/s or None."""
        if self.kind != 'reference' or self.guarded:
            return None
        fields = self.fields[:max_images]
        best = None
        for r in range(repeats + 1):
            t0 = time.perf_counter()
            for f in fields:
                self.dec._mappable_annotations(f, None)
            dt = time.perf_counter() - t0
            if r > 0:
                best = dt if best is None else min(best, dt)
        return len(fields) / best

    def pickle_seconds(self):
        """What the parent process spends serialising one step's fields for the pool (starmap pickles
        every image's 6.7 MB of fields): the part of a step no number of cores shortens."""
        if self.kind != 'reference':
            return None
        import pickle
        t0 = time.perf_counter()
        for f in self.fields:
            pickle.dumps((f, None), protocol=pickle.HIGHEST_PROTOCOL)
        return time.perf_counter() - t0

    def close(self):
        pool = getattr(self, 'pool', None) or getattr(getattr(self, 'dec', None), 'worker_pool', None)
        if pool is not None and hasattr(pool, 'terminate'):
            pool.terminate()


def run_reference(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return 0
    _, _, default_batch, _ = WORKLOADS[args.workload]
    batch = args.batch or default_batch
    sample_workload = 'coco17_641_5p' if args.workload == 'shard_4096' else args.workload
    r = ReferenceRunner(sample_workload, batch, args.cpu_workers)
    n_ann = 0
    slow = args.workload in RANDINIT_MAX_SCALE     # ~100 s (raw) / ~10 s (tamed) per image and core
    for _ in range(args.warmup if slow else max(args.warmup, 1)):
        n_ann = r.step()
    step_s = []
    t0 = time.perf_counter()
    for _ in range(args.steps):
        t1 = time.perf_counter()
        n_ann = r.step()
        step_s.append(time.perf_counter() - t1)
    dt = time.perf_counter() - t0
    single = None if slow or args.workload.startswith('cifdet') else r.single_process()
    pickle_s = None if slow else r.pickle_seconds()
    r.close()
    value = batch * args.steps / dt
    sample = '%d steps x %d images of %s, pool of %d processes, %s' % (
        args.steps, batch, sample_workload, r.cores, r.version)
    if r.guarded:
        # images the reference could not decode are left out; the rate is per process-second of the others
        value = (r.ok_images / r.ok_seconds * min(r.cores, batch)) if r.ok_images else None
        sample += ('; the reference raised TypeError (frontier tie, cifcaf.py:261,285) on %d of %d images, '
                   'value = decoded images / their process-seconds x %d processes' % (
                       r.crashed, r.crashed + r.ok_images, min(r.cores, batch)))
    det = args.workload.startswith('cifdet')
    cpu = {'value': value, 'unit': UNIT, 'cores': r.cores, 'kind': r.kind.split('-')[0], 'sample': sample,
           # best of >= 5 repeats next to the mean over the K steps (`value`)
           'best_step': batch / min(step_s), 'median_step': batch / float(np.median(step_s)),
           'worst_step': batch / max(step_s), 'warmup_steps': args.warmup if slow else max(args.warmup, 1),
           'single_process': None if single is None else {
               'value': single, 'unit': UNIT, 'cores': 1,
               'how': 'the same decoder in one process (the reference DummyPool), 1 warm-up, best of 5'},
           'parent_pickle_s_per_step': pickle_s,
           'host_cpus': os.cpu_count(), 'affinity_cpus': len(os.sched_getaffinity(0))}
    line = {
        'impl': 'reference',
        'metric': 'decoded images/s (641px CifDet fields, %d categories)' % DET_CATEGORIES if det else METRIC,
        'value': value, 'unit': UNIT,
        'n_gpus': args.gpus, 'steps': args.steps, 'warmup': args.warmup,
        'ms_per_step': 1e3 * dt / args.steps, 'higher_is_better': True,
        'scaling': 'strong' if args.workload == 'shard_4096' else 'weak',
        'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
        'config': {'workload': '%s: batch %d x 641x641, %d categories' % (args.workload, batch, DET_CATEGORIES)}
        if det else workload_config(args.workload, batch, args.gpus),
        'cpu_baseline': cpu,
        'e2e': {'value': value, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0, 'annotations_per_step': n_ann,
    }
    print(json.dumps(line), flush=True)
    return 0


def workload_config(workload, batch, n_gpus):
    hw, n_poses, _, dense = WORKLOADS[workload]
    if workload in RANDINIT_MAX_SCALE:
        text = ('%s: batch %d x %dx%d per GPU, COCO 17-kp/19-link skeleton, fields of a randomly initialised '
                'resnet50 + cif/caf heads on torch.randn images (BASELINE configs[1], SURVEY 8d config 2)%s' % (
                    workload, batch, hw[0], hw[1],
                    '' if RANDINIT_MAX_SCALE[workload] is None else
                    ', scale channels clamped to %g cells (tamed variant)' % RANDINIT_MAX_SCALE[workload]))
    elif workload == 'shard_4096':
        text = ('shard_4096: 4096 distinct 641x641 images (pose seeds 0..4095, 5 synthetic persons each, COCO '
                '17-kp/19-link skeleton), contiguous shards of 4096 / N images per GPU, chunks of %d '
                '(BASELINE configs[3], SURVEY 8d config 4)' % batch)
    else:
        text = ('%s: batch %d x %dx%d per GPU, %s skeleton, %d synthetic persons/image, fields '
                "from the repo's encoder + jitter (SURVEY 8d config 1 generator)" % (
                    workload, batch, hw[0], hw[1], 'dense 44-link' if dense else 'COCO 17-kp/19-link', n_poses))
    return {
        'workload': text,
        'batch_per_gpu': batch, 'global_batch': batch * n_gpus, 'image_hw': list(hw),
        'persons_per_image': n_poses,
        'decoder': 'cif_th 0.1, seed_th 0.2, caf_th 0.1, blend, keypoint_th 0.001, nms on, '
                   'no force_complete',
        'l2': 'no flush needed: one step streams the fields (%.0f MB) and writes the hr maps '
              '(%.2f GB) per GPU, both larger than the 126 MB L2' % (
                  batch * (17 * 5 + (44 if dense else 19) * 9) * ((hw[0] - 1) // 8 + 1) *
                  ((hw[1] - 1) // 8 + 1) * 4 / 1e6, batch * 17 * hw[0] * hw[1] * 4 / 1e9),
        'parallelism': 'image-sharded x%d, no collective' % n_gpus,
    }


# ----------------------------------------------------------------------------
# clocks

CLOCK_QUERY = ('index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,'
               'clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
               'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')


class ClockSampler:
    """nvidia-smi sampled every 50 ms while the timed regions run."""

    def __init__(self, gpu_index):
        self.rows = []
        self.proc = None
        try:
            self.proc = subprocess.Popen(
                ['nvidia-smi', '-i', str(gpu_index), '--query-gpu=' + CLOCK_QUERY,
                 '--format=csv,noheader,nounits', '-lms', '50'],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(',')])

    def stop(self):
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, smax, reasons, power = [], [], set(), []
        names = ('hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap')
        for r in self.rows:
            if len(r) < 9:
                continue
            try:
                sm.append(float(r[1]))
                smax.append(float(r[2]))
                power.append(float(r[3]))
            except ValueError:
                continue
            for name, cell in zip(names, r[5:9]):
                if cell.lower().startswith('active'):
                    reasons.add(name)
        return {'sm_mhz': float(np.median(sm)) if sm else None,
                'sm_max_mhz': max(smax) if smax else None,
                'power_w_max': max(power) if power else None,
                'samples': len(sm), 'reasons': sorted(reasons)}


def measured_peak():
    path = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    try:
        with open(path) as f:
            return float(json.load(f)['hbm_gbs']), 'measured (MEASURED_PEAKS.json hbm_gbs)'
    except (OSError, KeyError, ValueError):
        return 6650.0, 'fallback (B200_PROFILING.md 6.65 TB/s)'


# ----------------------------------------------------------------------------
# B200 arm

def cpu_baseline_subprocess(args, batch, steps=20, warmup=3, workers=0):
    """Runs the reference arm in a child process (before CUDA is initialised in
    this one, so the pool's fork is clean and the cores are idle)."""
    cmd = [sys.executable, os.path.abspath(__file__), '--impl', 'reference', '--steps', str(steps),
           '--warmup', str(warmup), '--workload', args.workload, '--batch', str(batch),
           '--cpu-workers', str(workers)]
    env = dict(os.environ)
    for k in ('RANK', 'LOCAL_RANK', 'WORLD_SIZE'):
        env.pop(k, None)
    try:
        out = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True,
                             timeout=900, env=env, cwd=ROOT)
        for line in reversed(out.stdout.strip().splitlines()):
            if line.startswith('{'):
                return json.loads(line)['cpu_baseline']
        return {'value': None, 'unit': UNIT, 'cores': 0, 'kind': 'reference',
                'sample': 'failed: ' + out.stderr.strip()[-300:]}
    except (subprocess.TimeoutExpired, OSError, ValueError, KeyError) as e:
        return {'value': None, 'unit': UNIT, 'cores': 0, 'kind': 'reference', 'sample': 'failed: %r' % e}


def bind_to_gpu_numa(device_index):
    """Multi-GPU runs: keep this rank (and the pinned buffers it first touches) on the CPUs NVML
    reports as local to its GPU.  Best effort; returns the CPU list or None."""
    try:
        import pynvml
        import torch
        pynvml.nvmlInit()
        uuid = str(torch.cuda.get_device_properties(device_index).uuid)
        handle = None
        for i in range(pynvml.nvmlDeviceGetCount()):
            h = pynvml.nvmlDeviceGetHandleByIndex(i)
            u = pynvml.nvmlDeviceGetUUID(h)
            u = u.decode() if isinstance(u, bytes) else u
            if uuid in u:
                handle = h
                break
        if handle is None:
            return None
        ncpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(handle, (ncpu + 63) // 64)
        local = {c for c in range(ncpu) if (words[c // 64] >> (c % 64)) & 1}
        allowed = local & set(os.sched_getaffinity(0))
        if len(allowed) >= 2:
            os.sched_setaffinity(0, allowed)
            return sorted(allowed)
    except Exception:  # pylint: disable=broad-except
        pass
    return None


def run_b200_det(args):
    """Secondary line: the CifDet path (pp_cifdet_decode_batch), device-resident fields, one GPU."""
    import torch
    _, n_boxes, default_batch, _ = WORKLOADS[args.workload]
    batch = args.batch or default_batch
    cpu_base = None if args.no_cpu_baseline else cpu_baseline_subprocess(args, batch, steps=2)
    if not torch.cuda.is_available():
        raise SystemExit('bench.py: no CUDA device; the B200 arm has no CPU path')
    import openpifpaf_b200.decoder as decoder
    from openpifpaf_b200.decoder.generator.cifdet import run_cifdet
    decoder.CifHr.v_threshold, decoder.CifSeeds.threshold, decoder.CifSeeds.score_scale = 0.1, 0.2, 1.0
    dev = torch.from_numpy(make_det_batch(args.workload, batch)).cuda()
    for _ in range(max(args.warmup, 3)):
        t = run_cifdet(dev, 8)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(args.steps):
        t = run_cifdet(dev, 8)
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / args.steps
    hr_bytes = batch * DET_CATEGORIES * 641 * 641 * 4
    line = {'metric': 'decoded images/s (641px CifDet fields, %d categories)' % DET_CATEGORIES,
            'value': batch / ms * 1e3, 'unit': UNIT, 'n_gpus': 1, 'steps': args.steps,
            'warmup': max(args.warmup, 3), 'ms_per_step': ms, 'higher_is_better': True, 'dtype': 'f32',
            'data': 'synthetic',
            'config': {'workload': '%s: batch %d x 641x641, %d categories, %d boxes/image' % (
                args.workload, batch, DET_CATEGORIES, n_boxes)},
            'detections_per_step': int(t['n_det'].sum()),
            'hr_bytes_per_step': hr_bytes, 'hr_gbs_whole_path': hr_bytes / ms / 1e6,
            'note': 'includes per-call workspace / result allocation (run_cifdet)'}
    if cpu_base is not None:
        line['cpu_baseline'] = cpu_base
    print(json.dumps(line), flush=True)
    return 0



# ----------------------------------------------------------------------------
# B200 arm

def _median(xs):
    return float(np.median(np.asarray(xs, dtype=np.float64)))


def run_b200(args):
    if args.workload.startswith('cifdet'):
        return run_b200_det(args)
    if args.workload == 'shard_4096':
        return run_shard(args)
    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    hw, _, default_batch, dense = WORKLOADS[args.workload]
    batch = args.batch or default_batch
    randinit = args.workload in RANDINIT_MAX_SCALE
    # batches in flight: 8 (the optimum moved from 8-16 to 6-8 with the thin CifHr kernel); crowds have few
    # images per batch and long latency-bound stages: 16; batch 256 holds 7 GB of hr maps per lane: 4
    lanes = args.lanes if args.lanes > 0 else {'crowd_1281_50p': 16, 'dense_641_5p_b256': 4}.get(args.workload, 8)
    n_distinct = args.distinct if args.distinct > 0 else min(lanes, 8)
    if args.workload == 'dense_641_5p_b256':
        n_distinct = min(n_distinct, 2)          # 3.2 GB of fields per batch

    cpu_base = None
    if world == 1 and not args.no_cpu_baseline:
        if not randinit:
            cpu_base = cpu_baseline_subprocess(args, min(batch, 64), steps=args.steps, warmup=max(args.warmup, 3))
        elif RANDINIT_MAX_SCALE[args.workload] is None:
            # ~100 s per image and core => the reference is timed on 4 images (4 processes)
            cpu_base = cpu_baseline_subprocess(args, 4, steps=1, warmup=0, workers=4)
            if cpu_base.get('value'):
                cpu_base['extrapolated_all_cores'] = cpu_base['value'] / 4 * (os.cpu_count() or 1)
                cpu_base['sample'] += '; extrapolated_all_cores = value / 4 x os.cpu_count()'
        else:
            cpu_base = cpu_baseline_subprocess(args, os.cpu_count() or 1, steps=1, warmup=0)

    # distinct input batches, generated by forked workers before CUDA exists in this process
    shared = None
    if not randinit:
        seeds = [rank * n_distinct * batch + i for i in range(n_distinct * batch)]
        shared = fill_images(args.workload, seeds, procs=max(1, (os.cpu_count() or 1) // max(1, world)))

    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit('bench.py: no CUDA device; the B200 arm has no CPU path '
                         '(use --impl reference for the host-core baseline)')
    torch.cuda.set_device(local_rank)
    device = torch.device('cuda', local_rank)
    if world > 1:
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        dist.init_process_group('nccl', device_id=device)
        bind_to_gpu_numa(local_rank)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return float(x)
        t = torch.tensor([float(x)], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    from openpifpaf_b200 import _lib, decoder
    from openpifpaf_b200.datasets.constants import COCO_KEYPOINTS
    from openpifpaf_b200.decoder.engine import StageProfile
    configure_profile(decoder)
    if randinit:
        decoder.CifCaf.max_annotations = RANDINIT_ANN_CAP
    skeleton = _skeleton(dense)
    if randinit:
        parts = [make_batch(args.workload, batch, first_seed=(rank * n_distinct + k) * batch)[:2]
                 for k in range(n_distinct)]
        cifs, cafs = [p[0] for p in parts], [p[1] for p in parts]
    else:
        cifs = [shared[0][k * batch:(k + 1) * batch] for k in range(n_distinct)]
        cafs = [shared[1][k * batch:(k + 1) * batch] for k in range(n_distinct)]
    cif, caf = cifs[0], cafs[0]
    dec = decoder.CifCaf(decoder.FieldConfig(), keypoints=COCO_KEYPOINTS, skeleton=skeleton)
    if args.host_chunk:
        decoder.CifCaf.host_chunk = args.host_chunk
    eng = dec.engine()
    params = dec.params()
    if args.tuning:
        params.tuning = {k: int(v) for k, v in (kv.split('=') for kv in args.tuning.split(','))}
        if 'compress' in params.tuning:                 # experiments: hr maps in plain device memory
            from openpifpaf_b200.decoder.engine import BatchBuffers
            BatchBuffers.compress_hr = bool(params.tuning.pop('compress'))
        if 'gate' in params.tuning:                     # experiments: splat gate on / off
            eng.use_splat_gate = bool(params.tuning.pop('gate'))
    d_batches = [(torch.from_numpy(np.ascontiguousarray(c)).to(device),
                  torch.from_numpy(np.ascontiguousarray(a)).to(device)) for c, a in zip(cifs, cafs)]
    h_batches = [(torch.from_numpy(np.ascontiguousarray(c)).pin_memory(),
                  torch.from_numpy(np.ascontiguousarray(a)).pin_memory()) for c, a in zip(cifs, cafs)]
    d_cif, d_caf = d_batches[0]
    h_cif, h_caf = h_batches[0]
    steps = args.steps

    # ---- device-resident throughput ("value") --------------------------------
    n_ann_batches = []
    for k in range(max(args.warmup, 3, n_distinct)):
        buf = eng.run_device(*d_batches[k % n_distinct], params)
        if k < n_distinct:
            torch.cuda.synchronize()
            assert not buf.dev['status'].any().item(), 'capacity overflow in warm-up'
            n_ann_batches.append(int(buf.dev['n_ann'].sum().item()))
    barrier()
    n_ann = n_ann_batches[0]
    ann_per_step = float(np.mean(n_ann_batches))
    prof = StageProfile(steps)
    params.profile = prof.handle
    try:
        uuid = 'GPU-%s' % torch.cuda.get_device_properties(local_rank).uuid
    except Exception:  # pylint: disable=broad-except
        uuid = str(local_rank)
    clocks = ClockSampler(uuid) if rank == 0 else None     # one poller per job, on the printing rank's GPU
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    # (1) one batch at a time on one stream, with the per-stage profile: explains the step
    barrier()
    ev0.record()
    for k in range(steps):
        eng.run_device(*d_batches[k % n_distinct], params)
    ev1.record()
    barrier()
    single_ms = max_over_ranks(ev0.elapsed_time(ev1))
    stage_ms = prof.read().astype(np.float64)
    # (2) the throughput configuration: K steps round-robin over `lanes` streams / workspaces, so the
    # latency-bound stages of one batch (assembly, sort) overlap the HBM-bound stages of the next.
    # The K-step pass is repeated until MIN_TIMED_S have been timed; the median pass is reported.
    main = torch.cuda.current_stream()
    prof.reset()
    tickets = []
    for k in range(lanes):                       # allocate + warm the lanes
        tickets.append(eng.submit(*d_batches[k % n_distinct], params, n_lanes=lanes))
    for t in tickets:
        t.wait()

    def value_region(n_passes):
        """n_passes x K steps issued back to back (no synchronisation between the passes, so the lanes fill
        and drain once per region); returns the device time of the region in ms, max over ranks."""
        prof.reset()
        barrier()
        ev0.record()
        tickets = []
        for k in range(n_passes * steps):
            if len(tickets) == lanes:
                tickets.pop(0).wait()
            tickets.append(eng.submit(*d_batches[k % n_distinct], params, n_lanes=lanes))
        for lane in eng.lanes(lanes):
            main.wait_stream(lane.stream)
        ev1.record()
        barrier()
        for t in tickets:
            t.wait()
        return max_over_ranks(ev0.elapsed_time(ev1))

    est = value_region(1)
    n_passes = int(min(400, max(1, np.ceil(1e3 * MIN_TIMED_S / 3 / est))))
    region_ms = [value_region(n_passes) for _ in range(3)]
    pass_ms = [r / n_passes for r in region_ms]
    dev_ms = _median(pass_ms)
    lane_stage_ms = prof.read().astype(np.float64)
    params.profile = None
    if args.value_only:
        if rank == 0:
            print(json.dumps({'value': world * batch * steps / (dev_ms * 1e-3), 'ms_per_step': dev_ms / steps,
                              'one_lane_ms_per_step': single_ms / steps, 'tuning': args.tuning, 'lanes': lanes,
                              'stages_ms': {n: round(float(v), 4) for n, v in zip(prof.names, stage_ms.mean(axis=0))},
                              'stages_ms_in_flight': {n: round(float(v), 4) for n, v in
                                                      zip(prof.names, lane_stage_ms.mean(axis=0))}}), flush=True)
            clocks.stop()
        if world > 1:
            dist.destroy_process_group()
        return 0

    # ---- end to end through the public API ---------------------------------------------------------
    # Every leg: E2E_PASSES passes of K steps, wall clock around a pass (perf_counter between a barrier and
    # the last Annotation list on the host), max over ranks per pass, median pass reported.
    def timed_passes(run_pass, n_passes=E2E_PASSES):
        run_pass()                                   # warm-up pass (allocations, lanes, guesses)
        out = []
        for _ in range(n_passes):
            barrier()
            t0 = time.perf_counter()
            run_pass()
            torch.cuda.synchronize()
            out.append(max_over_ranks(time.perf_counter() - t0))
        return out

    def host_batches(n):
        for k in range(n):
            yield h_batches[k % n_distinct]

    def device_batches(n):
        for k in range(n):
            yield d_batches[k % n_distinct]

    counted = {}

    def stream_pass(batches_fn, key):
        def run():
            total = 0
            for anns_s in dec.decode_stream(batches_fn(steps), n_lanes=min(lanes, 8)):
                total += sum(len(a) for a in anns_s)
            counted[key] = total
            counted[key + '_stats'] = dict(dec.stream_stats)
        return run

    def sync_pass():
        for k in range(steps):
            counted['sync'] = dec.decode_batch(*h_batches[k % n_distinct])

    expect = sum(n_ann_batches[k % n_distinct] for k in range(steps))
    # (1) the call a user makes with host fields: CifCaf.decode_stream / decode_batch on pinned host tensors
    # with the default host_transfer='auto', which takes the 'mapped' route here (confidence planes by DMA,
    # the other planes of the cells above threshold read by the kernels from the pinned buffers over PCIe)
    assert dec.host_transfer == 'auto'
    e2e_mapped = dec._use_mapped(h_cif, h_caf, dec.params())
    e2e_pass_s = timed_passes(stream_pass(host_batches, 'host'))
    assert counted['host'] == expect, 'host stream and device path disagree on the annotation count'
    e2e_sync_pass_s = timed_passes(sync_pass, 3)
    anns = dec.decode_batch(h_cif, h_caf)
    # (1b) the reference's real call path after the network: fields already on the GPU
    e2e_dev_pass_s = timed_passes(stream_pass(device_batches, 'device'))
    assert counted['device'] == expect, 'device stream disagrees on the annotation count'
    # (2) host fields with host_transfer='copy': every plane of every field is copied
    dec.host_transfer = 'copy'
    e2e_copy_pass_s = timed_passes(stream_pass(host_batches, 'copy'), 3)
    e2e_copy_sync_pass_s = timed_passes(sync_pass, 2)
    anns_c = dec.decode_batch(h_cif, h_caf)
    dec.host_transfer = 'auto'
    assert counted['copy'] == expect
    assert [len(a) for a in anns_c] == [len(a) for a in anns], 'mapped and copied host paths disagree'
    for la, lc in zip(anns, anns_c):
        for x, y in zip(la, lc):
            assert np.array_equal(x.data, y.data), 'mapped and copied host paths disagree'
    assert sum(len(a) for a in anns) == n_ann, 'host and device paths disagree on the annotation count'
    e2e_s, e2e_sync_s, e2e_dev_s = _median(e2e_pass_s), _median(e2e_sync_pass_s), _median(e2e_dev_pass_s)
    e2e_copy_s, e2e_copy_sync_s = _median(e2e_copy_pass_s), _median(e2e_copy_sync_pass_s)

    # latency of one image through the reference-shaped call CifCaf.__call__([cif, caf]) (video.py's use)
    lat = {}
    if rank == 0 and not randinit:
        for name, fields in (('cuda_tensors', [d_cif[0], d_caf[0]]), ('numpy_arrays', [cif[0], caf[0]])):
            for _ in range(3):
                dec(fields)
            ts = []
            for _ in range(30):
                t0 = time.perf_counter()
                dec(fields)
                ts.append(time.perf_counter() - t0)
            lat[name + '_ms'] = 1e3 * float(np.median(ts))
    clock_info = clocks.stop() if clocks is not None else None
    # result bytes that cross PCIe per step: header + the records copied (live rows, at least the guess)
    rec_bytes = 4 * _lib.lib().pp_packed_record_words(cif.shape[1])
    from openpifpaf_b200.decoder.engine import DeviceDecoder
    guess = DeviceDecoder.fetch_guess           # records per image copied with the header (6, grown by the batches seen)
    d2h = 4 * int(_lib.lib().pp_packed_header_words(batch)) + rec_bytes * max(int(ann_per_step), guess * batch)
    h2d = cif.nbytes + caf.nbytes
    h2d_dma = cif.nbytes // 5 + caf.nbytes // 9
    h2d_zero = int((cif[:, :, 0] > params.cif_threshold).sum()) * 4 * 4 + \
        int((caf[:, :, 0] > params.caf_threshold).sum()) * 8 * 4
    h2d_mapped = h2d_dma + h2d_zero

    # pure host->device copy of one step's fields: the PCIe floor under e2e_full_copy
    for _ in range(2):
        d_cif.copy_(h_cif, non_blocking=True)
        d_caf.copy_(h_caf, non_blocking=True)
    torch.cuda.synchronize()
    ev0.record()
    for _ in range(5):
        d_cif.copy_(h_cif, non_blocking=True)
        d_caf.copy_(h_caf, non_blocking=True)
    ev1.record()
    torch.cuda.synchronize()
    h2d_ms = ev0.elapsed_time(ev1) / 5

    # pure write stream of the size of one step's hr maps (cudaMemset through torch): the ceiling a kernel
    # that only writes can reach on this part, next to the driver's copy peak (half reads, half writes)
    fill = torch.empty((batch * cif.shape[1] * hw[0] * hw[1],), dtype=torch.float32, device=device)
    for _ in range(2):
        fill.zero_()
    torch.cuda.synchronize()
    ev0.record()
    for _ in range(5):
        fill.zero_()
    ev1.record()
    torch.cuda.synchronize()
    write_peak_gbs = fill.numel() * 4 / (ev0.elapsed_time(ev1) / 5 * 1e-3) / 1e9
    fill_shape = tuple(fill.shape)
    del fill
    # the same into compressible memory (pp_hr_alloc), where the hr maps of this run live
    from openpifpaf_b200.decoder.engine import BatchBuffers, hr_tensor
    fill, hr_compressed = hr_tensor(fill_shape, device)
    fill.zero_()
    torch.cuda.synchronize()
    ev0.record()
    for _ in range(5):
        fill.zero_()
    ev1.record()
    torch.cuda.synchronize()
    write_peak_comp_gbs = fill.numel() * 4 / (ev0.elapsed_time(ev1) / 5 * 1e-3) / 1e9
    del fill

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    # ---- roofline of the CifHr splat kernel -----------------------------------
    peak, peak_src = measured_peak()
    names = prof.names
    mean_ms = stage_ms.mean(axis=0)
    i_splat, i_compact = names.index('cifhr_splat'), names.index('cif_compact')
    hr_bytes = batch * cif.shape[1] * hw[0] * hw[1] * 4
    cif_bytes = cif.nbytes
    splat_gbs = hr_bytes / (mean_ms[i_splat] * 1e-3) / 1e9
    stage_gbs = (hr_bytes + cif_bytes) / ((mean_ms[i_splat] + mean_ms[i_compact]) * 1e-3) / 1e9
    lane_mean_ms = lane_stage_ms.mean(axis=0)
    lane_splat_gbs = hr_bytes / (lane_mean_ms[i_splat] * 1e-3) / 1e9
    traffic = None
    try:
        with open(os.path.join(ROOT, 'profiles', 'traffic.json')) as f:
            traffic = json.load(f).get('cifhr_strips_kernel', {}).get(args.workload)
        if batch != default_batch:
            traffic = None       # the captured launch was at the workload's default batch
    except (OSError, ValueError):
        pass
    images = world * batch * steps

    def stats_ms(key):
        st = counted.get(key + '_stats') or {}
        return {k.replace('_s', '_ms_per_step'): 1e3 * v / steps for k, v in st.items()}

    # bytes a step has to move through HBM: the hr maps written once + the planes the decode reads (every
    # confidence plane, the other planes of the active cells only)
    whole_step_bytes = hr_bytes + h2d_mapped
    line = {
        'metric': METRIC,
        'value': images / (dev_ms * 1e-3),
        'unit': UNIT, 'n_gpus': world, 'steps': steps, 'warmup': max(args.warmup, 3, n_distinct),
        'ms_per_step': dev_ms / steps, 'higher_is_better': True, 'scaling': 'weak',
        'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
        'config': workload_config(args.workload, batch, world),
        # how `value` was issued (not part of the workload): streams / workspaces in flight and the distinct
        # input batches the steps cycle through
        'batches_in_flight': lanes, 'distinct_batches': n_distinct,
        'passes': 3 * n_passes, 'timed_region_s': sum(region_ms) / 1e3,
        'pass_ms': {'median': dev_ms, 'min': min(pass_ms), 'max': max(pass_ms),
                    'how': '3 regions of %d K-step passes issued back to back (lanes fill and drain once per '
                           'region); ms per K steps of each region, median reported' % n_passes},
        'value_note': 'device-resident: fields in HBM, results left in HBM (no copy, no Annotation objects); '
                      'the figure that compares like for like with the CPU reference is e2e / e2e_device_fields',
        'hbm_step': {'bytes_per_step': whole_step_bytes,
                     'achieved_gbs': whole_step_bytes / (dev_ms / steps * 1e-3) / 1e9,
                     'frac_of_peak': whole_step_bytes / (dev_ms / steps * 1e-3) / 1e9 / peak,
                     'what': 'hr maps written + field planes read (confidence planes, other planes of the active '
                             'cells) per step, over the in-flight step time'},
        'e2e': {'value': images / e2e_s, 'unit': UNIT,
                'h2d_bytes_per_step': h2d_mapped if e2e_mapped else h2d, 'd2h_bytes_per_step': d2h,
                'api': 'CifCaf.decode_stream(K pinned host (cif, caf) batches) -> List[List[Annotation]] per '
                       'batch, defaults',
                'ms_per_step': 1e3 * e2e_s / steps,
                'passes': len(e2e_pass_s), 'pass_ms_per_step': [1e3 * t / steps for t in e2e_pass_s],
                'breakdown': dict(stats_ms('host'),
                                  what='host thread of rank 0, last pass: enqueueing, blocked in '
                                       'pp_pipeline_wait, building Annotation objects'),
                'one_call_per_step': {'value': images / e2e_sync_s, 'unit': UNIT,
                                      'ms_per_step': 1e3 * e2e_sync_s / steps,
                                      'api': 'CifCaf.decode_batch(pinned host cif, caf), one synchronous call '
                                             'per step (no overlap between steps)'},
                'host_transfer': "auto -> %s" % ('mapped' if e2e_mapped else 'copy'),
                'h2d_dma_bytes_per_step': h2d_dma if e2e_mapped else h2d,
                'h2d_zero_copy_bytes_per_step': h2d_zero if e2e_mapped else 0,
                'field_bytes_per_step': h2d,
                'note': 'the fields stay in pinned host memory; per step the confidence planes are copied '
                        '(strided DMA) and the kernels read the other planes of the cells above threshold '
                        'from the pinned buffers over PCIe (zero-copy bytes = active cells x planes x 4 B, '
                        'sector granularity not counted); results identical to the full copy, which is '
                        'e2e_full_copy.  d2h = header + live records of the packed result block'},
        'e2e_device_fields': {'value': images / e2e_dev_s, 'unit': UNIT, 'ms_per_step': 1e3 * e2e_dev_s / steps,
                              'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': d2h,
                              'api': 'CifCaf.decode_stream(K CUDA (cif, caf) batches) -> List[List[Annotation]]: '
                                     'what Generator.batch does after the network (generator.py:84-101)',
                              'passes': len(e2e_dev_pass_s),
                              'pass_ms_per_step': [1e3 * t / steps for t in e2e_dev_pass_s],
                              'breakdown': stats_ms('device'),
                              'gap_to_value': 'value leaves the results in HBM; this leg adds one packed D2H copy '
                                              'per batch and the Python Annotation objects (breakdown.build)'},
        'e2e_full_copy': {'value': images / e2e_copy_s, 'unit': UNIT,
                          'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': d2h,
                          'api': "CifCaf.decode_stream with host_transfer='copy' (every plane copied)",
                          'ms_per_step': 1e3 * e2e_copy_s / steps,
                          'one_call_per_step': {'value': images / e2e_copy_sync_s,
                                                'unit': UNIT, 'ms_per_step': 1e3 * e2e_copy_sync_s / steps,
                                                'host_chunk': decoder.CifCaf.host_chunk},
                          'h2d_only_ms_per_step': h2d_ms, 'h2d_gbs': h2d / h2d_ms / 1e6,
                          'pcie_floor_frac': h2d_ms / (1e3 * e2e_copy_s / steps)},
        # kernels of this library launched in the reported K-step pass: per decode field scan, strips kernel, CTA-per-strip
        # kernel (wide boxes; no work here), seeds sorted per field, seed merge, CafScored scoring + binning,
        # assembly + NMS, result gather (8 kernels, 2 memsets)
        'gpu_launches': (_lib.lib().pp_decode_launch_count(None) - 2) * steps,
        'gpu_launches_per_step': _lib.lib().pp_decode_launch_count(None) - 2,
        'gpu_launches_all_passes': (_lib.lib().pp_decode_launch_count(None) - 2) * steps * 3 * n_passes,
        'gpu_memsets_per_step': 2,
        'roofline': {'bound': 'hbm', 'kernel': 'cifhr_strips_kernel', 'achieved': splat_gbs,
                     'peak': peak, 'unit': 'GB/s', 'frac': splat_gbs / peak, 'traffic': traffic,
                     'peak_source': peak_src, 'algorithmic_bytes_per_launch': hr_bytes,
                     'write_only_peak': {'gbs': write_peak_gbs, 'frac': splat_gbs / write_peak_gbs,
                                         'how': 'memset of a buffer of the same size, timed here (5 runs): the '
                                                'kernel only writes, the copy peak moves half its bytes as reads'},
                     'hr_memory': {'compressible': bool(hr_compressed and BatchBuffers.compress_hr),
                                   'memset_gbs': write_peak_comp_gbs,
                                   'what': 'the hr maps live in an allocation made with compute data compression '
                                           '(pp_hr_alloc: cuMemCreate, CU_MEM_ALLOCATION_COMP_GENERIC); the map is '
                                           '~88 % zeros and B200 writes such data faster than into cudaMalloc '
                                           'memory, which is how the kernel gets past the copy peak (a cudaMalloc '
                                           'figure).  memset_gbs: a memset of the same size into such an '
                                           'allocation'},
                     'ms_per_launch': float(mean_ms[i_splat]),
                     'timed': 'one batch at a time on one stream (%d steps, value_one_lane); with %d '
                              'batches in flight the event-bracketed duration also contains the time the '
                              'kernel shares the GPU with other lanes: see in_flight' % (steps, lanes),
                     'in_flight': {'ms_per_launch': float(lane_mean_ms[i_splat]), 'achieved': lane_splat_gbs,
                                   'frac': lane_splat_gbs / peak},
                     'cifhr_stage': {'kernels': 'cif_compact + cifhr_splat',
                                     'algorithmic_bytes': hr_bytes + cif_bytes,
                                     'achieved': stage_gbs, 'frac': stage_gbs / peak}},
        'value_one_lane': {'value': images / (single_ms * 1e-3), 'unit': UNIT,
                           'ms_per_step': single_ms / steps},
        'stages_ms': {n: float(v) for n, v in zip(names, mean_ms)},
        'stages_ms_in_flight': {n: float(v) for n, v in zip(names, lane_mean_ms)},
        'annotations_per_step': ann_per_step * world,
        'latency_one_image': dict(lat, api='CifCaf.__call__([cif, caf]) -> List[Annotation], median of 30 calls, '
                                           'synchronous (includes the result copy and the Annotation objects)'),
        'clocks': clock_info,
    }
    if randinit:
        from openpifpaf_b200 import workloads
        stats = workloads.field_statistics(d_cif)
        line['field_statistics'] = stats
        line['cifhr_splat_pixel_updates_per_s'] = (
            stats['splat_pixel_updates_per_image'] * batch / (mean_ms[i_splat] * 1e-3))
        line['roofline']['note'] = ('degenerate fields: the splat is compute-bound here (px-updates/s above), '
                                    'the HBM fraction is not the figure of merit for this workload')
    if cpu_base is not None:
        line['cpu_baseline'] = cpu_base
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


# ----------------------------------------------------------------------------
# BASELINE configs[3]: 4096 distinct images sharded over the ranks (strong scaling)

def run_shard(args):
    """4096 distinct config-1 images (pose seeds 0..4095), contiguous shards of 4096 / N per rank through
    ShardedDecoder in chunks of 64; timed: wall clock from the first launch to the last Annotation list on
    the host, max over ranks.  `value`: fields resident in HBM; `e2e`: fields in pinned host memory.  A sample
    of every shard is checked against the CPU oracle outside the timed regions."""
    from openpifpaf_b200.decoder.sharding import ShardedDecoder, shard_bounds
    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    hw, _, default_chunk, _ = WORKLOADS['shard_4096']
    chunk = args.batch or default_chunk
    n_images = args.images
    lo, hi = shard_bounds(n_images, world, rank)
    n_local = hi - lo

    cpu_base = None
    if world == 1 and not args.no_cpu_baseline:
        cpu_base = cpu_baseline_subprocess(args, 64, steps=20, warmup=3)
    t0 = time.perf_counter()
    cif, caf = fill_images('coco17_641_5p', range(lo, hi), procs=max(1, (os.cpu_count() or 1) // max(1, world)))
    gen_s = time.perf_counter() - t0

    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit('bench.py: no CUDA device; the B200 arm has no CPU path')
    torch.cuda.set_device(local_rank)
    device = torch.device('cuda', local_rank)
    if world > 1:
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        dist.init_process_group('nccl', device_id=device)
        bind_to_gpu_numa(local_rank)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return float(x)
        t = torch.tensor([float(x)], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    from openpifpaf_b200 import _lib, decoder
    from openpifpaf_b200.datasets.constants import COCO_KEYPOINTS
    configure_profile(decoder)
    skeleton = _skeleton(False)
    dec = decoder.CifCaf(decoder.FieldConfig(), keypoints=COCO_KEYPOINTS, skeleton=skeleton)
    # the shard's fields: in HBM (value) and in page-locked host memory (e2e); the shared arrays the
    # workers filled are registered with CUDA instead of being copied into a second pinned allocation
    h_cif, h_caf = torch.from_numpy(cif), torch.from_numpy(caf)
    rt = torch.cuda.cudart()
    for t in (h_cif, h_caf):
        rc = rt.cudaHostRegister(t.data_ptr(), t.numel() * 4, 3)       # portable | mapped
        assert int(rc) == 0, 'cudaHostRegister failed: %s' % rc
    assert h_cif.is_pinned() and h_caf.is_pinned()
    d_cif, d_caf = h_cif.to(device), h_caf.to(device)
    sd = ShardedDecoder(dec.decode_batch, chunk=chunk, decode_stream=dec.decode_stream)
    clocks = None
    if rank == 0:
        try:
            clocks = ClockSampler('GPU-%s' % torch.cuda.get_device_properties(local_rank).uuid)
        except Exception:  # pylint: disable=broad-except
            clocks = ClockSampler(str(local_rank))

    def job(cif_t, caf_t):
        def get_fields(a, b):
            return cif_t[a - lo:b - lo], caf_t[a - lo:b - lo]
        first, results = sd.decode_shard(n_images, get_fields)
        assert first == lo and len(results) == n_local
        return results

    import gc

    def timed(cif_t, caf_t, passes):
        out, results = [], None
        for _ in range(passes):
            results = None          # the previous pass's objects are freed and collected outside the timed region
            gc.collect()
            barrier()
            t0 = time.perf_counter()
            results = job(cif_t, caf_t)
            out.append(max_over_ranks(time.perf_counter() - t0))
        return out, results

    # warm-up: two chunks through each path (allocations, lanes, module load)
    for src_c, src_a in ((d_cif, d_caf), (h_cif, h_caf)):
        for _ in dec.decode_stream((src_c[k:k + chunk], src_a[k:k + chunk])
                                   for k in range(0, min(n_local, 3 * chunk), chunk)):
            pass
    passes = 7

    def flat(results):
        return [np.stack([a.data for a in anns]) if anns else np.zeros((0, 17, 3), np.float32) for anns in results]

    dev_s, res_dev = timed(d_cif, d_caf, passes)
    stats_dev = dict(dec.stream_stats)
    n_ann = sum(len(a) for a in res_dev)
    sample = sorted({0, n_local // 3, (2 * n_local) // 3, n_local - 1})
    sample_dev = {i: res_dev[i] for i in sample}
    flat_dev = flat(res_dev)
    res_dev = None
    host_s, res_host = timed(h_cif, h_caf, passes)
    stats_host = dict(dec.stream_stats)
    clock_info = clocks.stop() if clocks is not None else None

    # ---- checks outside the timed regions: both paths agree; a sample of the shard against the oracle
    for a, b in zip(flat_dev, flat(res_host)):
        assert np.array_equal(a, b), 'device-field and host-field paths disagree'
    res_dev = sample_dev
    from oracle import oracle as orc
    for i in sample:
        o = orc.decode(cif[i], caf[i], skeleton, orc.Config(exp_mode=1))
        got = res_dev[i]
        assert len(got) == len(o.data), 'image %d: %d annotations, oracle %d' % (lo + i, len(got), len(o.data))
        for ann, data, scales in zip(got, o.data, o.joint_scales):
            assert np.array_equal(ann.data, data) and np.array_equal(ann.joint_scales, scales), \
                'image %d differs from the oracle' % (lo + i)
    checked = [lo + i for i in sample]
    if world > 1:
        gathered = [None] * world if rank == 0 else None
        dist.gather_object((checked, n_ann), gathered, dst=0)
        if rank == 0:
            checked = sorted(i for c, _ in gathered for i in c)
            n_ann = sum(n for _, n in gathered)
    if rank != 0:
        dist.destroy_process_group()
        return 0

    dev_t, host_t = _median(dev_s), _median(host_s)
    n_chunks = (n_local + chunk - 1) // chunk
    rec_bytes = 4 * _lib.lib().pp_packed_record_words(17)
    from openpifpaf_b200.decoder.engine import DeviceDecoder
    d2h = 4 * int(_lib.lib().pp_packed_header_words(chunk)) + rec_bytes * DeviceDecoder.fetch_guess * chunk
    field_bytes = (cif[:chunk].nbytes + caf[:chunk].nbytes)
    h2d_mapped = (cif[:chunk].nbytes // 5 + caf[:chunk].nbytes // 9 +
                  int((cif[:chunk, :, 0] > 0.1).sum()) * 16 + int((caf[:chunk, :, 0] > 0.1).sum()) * 32)
    peak, peak_src = measured_peak()
    hr_bytes_chunk = chunk * 17 * hw[0] * hw[1] * 4
    line = {
        'metric': METRIC, 'value': n_images / dev_t, 'unit': UNIT, 'n_gpus': world,
        'steps': n_chunks, 'warmup': 3, 'ms_per_step': 1e3 * dev_t / n_chunks,
        'higher_is_better': True, 'scaling': 'strong', 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
        'config': workload_config('shard_4096', chunk, world),
        'images': n_images, 'images_per_rank': n_local, 'chunks_per_rank': n_chunks,
        'passes': passes, 'job_s': {'median': dev_t, 'all': dev_s},
        'value_note': 'whole job: wall clock from the first launch to the last List[Annotation] on the host, max '
                      'over ranks, median of %d passes; fields resident in HBM (ShardedDecoder -> '
                      'CifCaf.decode_stream, chunks of %d, 8 lanes)' % (passes, chunk),
        'e2e': {'value': n_images / host_t, 'unit': UNIT, 'job_s': {'median': host_t, 'all': host_s},
                'h2d_bytes_per_step': int(h2d_mapped), 'd2h_bytes_per_step': d2h,
                'field_bytes_per_step': field_bytes,
                'api': 'the same job with the fields in page-locked HOST memory (host_transfer auto -> mapped)',
                'breakdown': {k.replace('_s', '_s_per_job'): v for k, v in stats_host.items()}},
        'breakdown': {k.replace('_s', '_s_per_job'): v for k, v in stats_dev.items()},
        'hbm_step': {'bytes_per_step': hr_bytes_chunk + field_bytes,
                     'achieved_gbs': world * (hr_bytes_chunk + field_bytes) * n_chunks / dev_t / 1e9,
                     'frac_of_peak': (hr_bytes_chunk + field_bytes) * n_chunks / dev_t / 1e9 / peak,
                     'peak': peak, 'peak_source': peak_src},
        'gpu_launches': (_lib.lib().pp_decode_launch_count(None) - 2) * n_chunks,
        'annotations': n_ann, 'oracle_checked_images': checked,
        'field_generation_s': gen_s,
        'clocks': clock_info,
    }
    if cpu_base is not None:
        line['cpu_baseline'] = cpu_base
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    args = parse_args()
    # exactly one line on stdout: Python's prints keep the real stdout, whatever libraries write to file
    # descriptor 1 (NCCL's version banner under torchrun) goes to stderr
    sys.stdout.flush()
    sys.stdout = os.fdopen(os.dup(1), 'w', buffering=1)
    os.dup2(2, 1)
    if args.impl == 'reference':
        return run_reference(args)
    return run_b200(args)


if __name__ == '__main__':
    sys.exit(main())
