#!/usr/bin/env python3
"""bench.py -- decoded images/s of the CifCaf decode path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
                    [--workload NAME] [--batch B]

A "step" is one pass of the hot path (CifHr -> seeds -> CafScored -> assembly
-> keypoint NMS) over one batch of synthetic CIF/CAF fields.  One JSON line is
printed by rank 0 (contract in DESIGN.md section "Measurement"):

  value         images/s, fields already resident in HBM: K steps (one
                pp_decode_batch = 8 kernels each) issued round-robin over
                --lanes streams / workspaces, CUDA events around the K steps;
  value_one_lane, stages_ms
                the same K steps strictly one at a time on one stream, with the
                library's per-stage CUDA-event profile;
  e2e           images/s through the public API CifCaf.decode_stream(batches) with
                PINNED HOST fields and default settings (e2e.one_call_per_step:
                one synchronous CifCaf.decode_batch per step): host -> device
                transfer, kernels, D2H of the results and the construction of
                the Annotation objects are all inside the timed region.  The
                default host_transfer='auto' moves what the decode reads: the
                confidence planes by DMA, the other planes of the cells above
                threshold by zero-copy reads from the pinned buffers;
  e2e_full_copy the same call with host_transfer='copy': every plane of every
                field is copied (PCIe-bound: 430 MB per 64 images);
  roofline      the CifHr splat kernel (the kernel that writes the hr maps)
                against the measured HBM copy peak of MEASURED_PEAKS.json,
                timed in the one-lane pass; in_flight = while lanes overlap;
  cpu_baseline  the UNMODIFIED reference decoder (oracle/_ref, its Cython
                functional module) on this box's host cores, bounded sample;
  --impl reference   the same reference decoder as its own arm.

Under torchrun every rank decodes its own batch (image sharding, no
collective on the data path); torch.distributed is used for the barrier and the
max-over-ranks of the device time only.
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
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--host-chunk', type=int, default=0, help='e2e: images per pipeline chunk (0 = decoder default)')
    ap.add_argument('--cpu-workers', type=int, default=0, help='reference arm: pool size (0 = all cores)')
    return ap.parse_args()


# ----------------------------------------------------------------------------
# workload

def make_batch(workload, batch, first_seed=0):
    """`batch` distinct synthetic images of the workload: (cif, caf, skeleton)."""
    from openpifpaf_b200.encoder import synthetic_fields
    from openpifpaf_b200.datasets.constants import (COCO_PERSON_SKELETON,
                                                    DENSER_COCO_PERSON_CONNECTIONS)
    hw, n_poses, _, dense = WORKLOADS[workload]
    cache = '/tmp/pifpaf_bench_%s_%d_%d.npz' % (workload, batch, first_seed)
    skeleton = list(COCO_PERSON_SKELETON)
    if workload in RANDINIT_MAX_SCALE:
        import torch
        from openpifpaf_b200 import workloads
        device = 'cuda' if torch.cuda.is_available() else 'cpu'   # the network is workload tooling only
        cif, caf = workloads.random_init_head_fields(batch, hw, seed=0, image_seed=first_seed + 1, device=device,
                                                     max_scale=RANDINIT_MAX_SCALE[workload])
        return cif.cpu().numpy(), caf.cpu().numpy(), skeleton
    if dense:
        skeleton = skeleton + list(DENSER_COCO_PERSON_CONNECTIONS)
    if os.path.exists(cache):
        z = np.load(cache)
        return z['cif'], z['caf'], skeleton
    cifs, cafs = [], []
    for i in range(batch):
        cif, caf, _ = synthetic_fields(hw, n_poses, pose_seed=first_seed + i,
                                       jitter_seed=first_seed + i + 1, dense=dense)
        cifs.append(cif)
        cafs.append(caf)
    cif, caf = np.stack(cifs), np.stack(cafs)
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
    r = ReferenceRunner(args.workload, batch, args.cpu_workers)
    n_ann = 0
    slow = args.workload in RANDINIT_MAX_SCALE     # ~100 s (raw) / ~10 s (tamed) per image and core
    for _ in range(args.warmup if slow else max(args.warmup, 1)):
        n_ann = r.step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        n_ann = r.step()
    dt = time.perf_counter() - t0
    r.close()
    value = batch * args.steps / dt
    sample = '%d steps x %d images of %s, pool of %d processes, %s' % (
        args.steps, batch, args.workload, r.cores, r.version)
    if r.guarded:
        # images the reference could not decode are left out; the rate is per process-second of the others
        value = (r.ok_images / r.ok_seconds * min(r.cores, batch)) if r.ok_images else None
        sample += ('; the reference raised TypeError (frontier tie, cifcaf.py:261,285) on %d of %d images, '
                   'value = decoded images / their process-seconds x %d processes' % (
                       r.crashed, r.crashed + r.ok_images, min(r.cores, batch)))
    det = args.workload.startswith('cifdet')
    line = {
        'impl': 'reference',
        'metric': 'decoded images/s (641px CifDet fields, %d categories)' % DET_CATEGORIES if det else METRIC,
        'value': value, 'unit': UNIT,
        'n_gpus': args.gpus, 'steps': args.steps, 'warmup': args.warmup,
        'ms_per_step': 1e3 * dt / args.steps, 'higher_is_better': True, 'scaling': 'weak',
        'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
        'config': {'workload': '%s: batch %d x 641x641, %d categories' % (args.workload, batch, DET_CATEGORIES)}
        if det else workload_config(args.workload, batch, args.gpus),
        'cpu_baseline': {'value': value, 'unit': UNIT, 'cores': r.cores, 'kind': r.kind.split('-')[0],
                         'sample': sample},
        'e2e': {'value': value, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0, 'annotations_per_step': n_ann,
    }
    print(json.dumps(line), flush=True)
    return 0


def workload_config(workload, batch, n_gpus, lanes=1):
    hw, n_poses, _, dense = WORKLOADS[workload]
    if workload in RANDINIT_MAX_SCALE:
        text = ('%s: batch %d x %dx%d per GPU, COCO 17-kp/19-link skeleton, fields of a randomly initialised '
                'resnet50 + cif/caf heads on torch.randn images (BASELINE configs[1], SURVEY 8d config 2)%s' % (
                    workload, batch, hw[0], hw[1],
                    '' if RANDINIT_MAX_SCALE[workload] is None else
                    ', scale channels clamped to %g cells (tamed variant)' % RANDINIT_MAX_SCALE[workload]))
    else:
        text = ('%s: batch %d x %dx%d per GPU, %s skeleton, %d synthetic persons/image, fields '
                "from the repo's encoder + jitter (SURVEY 8d config 1 generator)" % (
                    workload, batch, hw[0], hw[1], 'dense 44-link' if dense else 'COCO 17-kp/19-link', n_poses))
    return {
        'batches_in_flight': lanes,
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
    """nvidia-smi sampled every 200 ms while the timed region runs."""

    def __init__(self, gpu_index):
        self.rows = []
        self.proc = None
        try:
            self.proc = subprocess.Popen(
                ['nvidia-smi', '-i', str(gpu_index), '--query-gpu=' + CLOCK_QUERY,
                 '--format=csv,noheader,nounits', '-lms', '200'],
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

def cpu_baseline_subprocess(args, batch, steps=20, warmup=1, workers=0):
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
                             timeout=600, env=env, cwd=ROOT)
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


def run_b200(args):
    if args.workload.startswith('cifdet'):
        return run_b200_det(args)
    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    hw, _, default_batch, dense = WORKLOADS[args.workload]
    batch = args.batch or default_batch

    cpu_base = None
    randinit = args.workload in RANDINIT_MAX_SCALE
    if world == 1 and not args.no_cpu_baseline:
        if not randinit:
            cpu_base = cpu_baseline_subprocess(args, batch)
        elif RANDINIT_MAX_SCALE[args.workload] is None:
            # ~100 s per image and core => the reference is timed on 4 images (4 processes)
            cpu_base = cpu_baseline_subprocess(args, 4, steps=1, warmup=0, workers=4)
            if cpu_base.get('value'):
                cpu_base['extrapolated_all_cores'] = cpu_base['value'] / 4 * (os.cpu_count() or 1)
                cpu_base['sample'] += '; extrapolated_all_cores = value / 4 x os.cpu_count()'
        else:
            cpu_base = cpu_baseline_subprocess(args, os.cpu_count() or 1, steps=1, warmup=0)

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
    cif, caf, skeleton = make_batch(args.workload, batch, first_seed=rank * batch)
    dec = decoder.CifCaf(decoder.FieldConfig(), keypoints=COCO_KEYPOINTS, skeleton=skeleton)
    if args.host_chunk:
        decoder.CifCaf.host_chunk = args.host_chunk
    eng = dec.engine()
    params = dec.params()
    d_cif = torch.from_numpy(cif).to(device)
    d_caf = torch.from_numpy(caf).to(device)
    h_cif = torch.from_numpy(cif).pin_memory()
    h_caf = torch.from_numpy(caf).pin_memory()

    # ---- device-resident throughput ("value") --------------------------------
    for _ in range(max(args.warmup, 3)):
        buf = eng.run_device(d_cif, d_caf, params)
    barrier()
    status = buf.dev['status'].cpu().numpy()
    assert not status.any(), 'capacity overflow in warm-up: %s' % status
    n_ann = int(buf.dev['n_ann'].sum().item())
    prof = StageProfile(args.steps)
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
    for _ in range(args.steps):
        eng.run_device(d_cif, d_caf, params)
    ev1.record()
    barrier()
    single_ms = max_over_ranks(ev0.elapsed_time(ev1))
    stage_ms = prof.read().astype(np.float64)
    # (2) the throughput configuration: K steps round-robin over `lanes` streams / workspaces, so the
    # latency-bound stages of one batch (assembly, sort, NMS) overlap the HBM-bound stages of the next
    lanes = args.lanes if args.lanes > 0 else (16 if args.workload in ('crowd_1281_50p', 'dense_641_5p') else 8)
    main = torch.cuda.current_stream()
    prof.reset()
    tickets = []
    for _ in range(lanes):                       # allocate + warm the lanes
        tickets.append(eng.submit(d_cif, d_caf, params, n_lanes=lanes))
    for t in tickets:
        t.wait()
    prof.reset()
    barrier()
    ev0.record()
    tickets = []
    for _ in range(args.steps):
        if len(tickets) == lanes:
            tickets.pop(0).wait()
        tickets.append(eng.submit(d_cif, d_caf, params, n_lanes=lanes))
    for lane in eng.lanes(lanes):
        main.wait_stream(lane.stream)
    ev1.record()
    barrier()
    for t in tickets:
        t.wait()
    dev_ms = max_over_ranks(ev0.elapsed_time(ev1))
    lane_stage_ms = prof.read().astype(np.float64)
    params.profile = None

    # ---- end to end through the public API, host fields ------------------------
    # (1) the call a user makes: CifCaf.decode_batch on pinned host tensors with the default
    # host_transfer='auto', which takes the 'mapped' route here (confidence planes by DMA, the other
    # planes of the cells above threshold read by the kernels from the pinned buffers over PCIe)
    assert dec.host_transfer == 'auto'
    for _ in range(2):
        anns = dec.decode_batch(h_cif, h_caf)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        anns = dec.decode_batch(h_cif, h_caf)
    torch.cuda.synchronize()
    e2e_sync_s = max_over_ranks(time.perf_counter() - t0)
    e2e_mapped = dec._use_mapped(h_cif, h_caf, dec.params())
    # (1b) the same K batches through CifCaf.decode_stream: every batch still crosses PCIe and comes back as
    # Annotation lists inside the timed region, but the transfer of batch k + 1 overlaps the tail of batch k
    # (this is how `value` is issued as well: K steps over the lanes)

    def host_batches(n):
        for _ in range(n):
            yield h_cif, h_caf

    for _ in dec.decode_stream(host_batches(2)):
        pass
    barrier()
    t0 = time.perf_counter()
    n_stream = 0
    for anns_s in dec.decode_stream(host_batches(args.steps)):
        n_stream += sum(len(a) for a in anns_s)
    torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    assert n_stream == args.steps * sum(len(a) for a in anns), 'streamed and single-batch host paths disagree'
    # (2) the same call with host_transfer='copy': every plane of every field is copied
    dec.host_transfer = 'copy'
    for _ in range(2):
        anns_c = dec.decode_batch(h_cif, h_caf)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        anns_c = dec.decode_batch(h_cif, h_caf)
    torch.cuda.synchronize()
    e2e_copy_sync_s = max_over_ranks(time.perf_counter() - t0)
    for _ in dec.decode_stream(host_batches(2)):
        pass
    barrier()
    t0 = time.perf_counter()
    for _ in dec.decode_stream(host_batches(args.steps)):
        pass
    torch.cuda.synchronize()
    e2e_copy_s = max_over_ranks(time.perf_counter() - t0)
    dec.host_transfer = 'auto'
    assert [len(a) for a in anns_c] == [len(a) for a in anns], 'mapped and copied host paths disagree'
    for la, lc in zip(anns, anns_c):
        for x, y in zip(la, lc):
            assert np.array_equal(x.data, y.data), 'mapped and copied host paths disagree'
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
    assert sum(len(a) for a in anns) == n_ann, 'host and device paths disagree on the annotation count'
    hbuf = eng.buffers(batch, cif.shape[1], cif.shape[3], cif.shape[4], host=True)
    d2h = sum(t.numel() * t.element_size() for t in hbuf.host.values())
    h2d = cif.nbytes + caf.nbytes
    h2d_dma = cif.nbytes // 5 + caf.nbytes // 9
    h2d_zero = int((cif[:, :, 0] > params.cif_threshold).sum()) * 4 * 4 + \
        int((caf[:, :, 0] > params.caf_threshold).sum()) * 8 * 4
    h2d_mapped = h2d_dma + h2d_zero

    # pure host->device copy of one step's fields: the PCIe floor under e2e
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
            traffic = json.load(f).get('cifhr_splat_kernel', {}).get(args.workload)
        if batch != default_batch:
            traffic = None       # the captured launch was at the workload's default batch
    except (OSError, ValueError):
        pass
    line = {
        'metric': METRIC,
        'value': world * batch * args.steps / (dev_ms * 1e-3),
        'unit': UNIT, 'n_gpus': world, 'steps': args.steps, 'warmup': max(args.warmup, 3),
        'ms_per_step': dev_ms / args.steps, 'higher_is_better': True, 'scaling': 'weak',
        'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
        'config': workload_config(args.workload, batch, world, lanes),
        'e2e': {'value': world * batch * args.steps / e2e_s, 'unit': UNIT,
                'h2d_bytes_per_step': h2d_mapped if e2e_mapped else h2d, 'd2h_bytes_per_step': d2h,
                'api': 'CifCaf.decode_stream(K pinned host (cif, caf) batches) -> List[List[Annotation]] per '
                       'batch, defaults',
                'ms_per_step': 1e3 * e2e_s / args.steps,
                'one_call_per_step': {'value': world * batch * args.steps / e2e_sync_s, 'unit': UNIT,
                                      'ms_per_step': 1e3 * e2e_sync_s / args.steps,
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
                        'e2e_full_copy'},
        'e2e_full_copy': {'value': world * batch * args.steps / e2e_copy_s, 'unit': UNIT,
                          'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': d2h,
                          'api': "CifCaf.decode_stream with host_transfer='copy' (every plane copied)",
                          'ms_per_step': 1e3 * e2e_copy_s / args.steps,
                          'one_call_per_step': {'value': world * batch * args.steps / e2e_copy_sync_s,
                                                'unit': UNIT, 'ms_per_step': 1e3 * e2e_copy_sync_s / args.steps,
                                                'host_chunk': decoder.CifCaf.host_chunk},
                          'h2d_only_ms_per_step': h2d_ms, 'h2d_gbs': h2d / h2d_ms / 1e6,
                          'pcie_floor_frac': h2d_ms / (1e3 * e2e_copy_s / args.steps)},
        # kernels of this library launched in the timed `value` region (8 per decode: compact, splat, seeds
        # from sources, seed sort, CafScored, candidate binning, assembly, NMS); 3 memsets per decode beside
        'gpu_launches': (_lib.lib().pp_decode_launch_count(None) - 3) * args.steps,
        'gpu_memsets': 3 * args.steps,
        'roofline': {'bound': 'hbm', 'kernel': 'cifhr_splat_kernel', 'achieved': splat_gbs,
                     'peak': peak, 'unit': 'GB/s', 'frac': splat_gbs / peak, 'traffic': traffic,
                     'peak_source': peak_src, 'algorithmic_bytes_per_launch': hr_bytes,
                     'write_only_peak': {'gbs': write_peak_gbs, 'frac': splat_gbs / write_peak_gbs,
                                         'how': 'memset of a buffer of the same size, timed here (5 runs): the '
                                                'kernel only writes, the copy peak moves half its bytes as reads'},
                     'ms_per_launch': float(mean_ms[i_splat]),
                     'timed': 'one batch at a time on one stream (%d steps, value_one_lane); with %d '
                              'batches in flight the event-bracketed duration also contains the time the '
                              'kernel shares the GPU with other lanes: see in_flight' % (args.steps, lanes),
                     'in_flight': {'ms_per_launch': float(lane_mean_ms[i_splat]), 'achieved': lane_splat_gbs,
                                   'frac': lane_splat_gbs / peak},
                     'cifhr_stage': {'kernels': 'cif_compact + cifhr_splat',
                                     'algorithmic_bytes': hr_bytes + cif_bytes,
                                     'achieved': stage_gbs, 'frac': stage_gbs / peak}},
        'value_one_lane': {'value': world * batch * args.steps / (single_ms * 1e-3), 'unit': UNIT,
                           'ms_per_step': single_ms / args.steps},
        'stages_ms': {n: float(v) for n, v in zip(names, mean_ms)},
        'stages_ms_in_flight': {n: float(v) for n, v in zip(names, lane_mean_ms)},
        'annotations_per_step': n_ann * world,
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
