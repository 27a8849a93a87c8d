"""CifCaf: composite intensity / association fields -> person annotations.

Drop-in for the reference's ``decoder/generator/cifcaf.py:24-351``: same
constructor, class-attribute knobs (read at call time), ``__call__(fields)``
and ``batch(model, images)``.  The reference runs CifHr, CifSeeds, CafScored,
the seed loop with ``_grow`` + ``Occupancy`` and keypoint NMS on the CPU, one
image per worker process; here one ``pp_decode_batch`` call does all of it for
the whole batch on the GPU (csrc/*.cu) and only the finished annotations come
back to the host.
"""
from collections import defaultdict
import dataclasses
import logging
import time

import numpy as np
import torch

from ...annotation import Annotation
from .. import nms as nms_module
from ..caf_scored import CafScored
from ..cif_hr import CifHr
from ..cif_seeds import CifSeeds
from ..engine import DecoderParams, DeviceDecoder
from ..field_config import FieldConfig
from ... import _lib
from .generator import Generator

LOG = logging.getLogger(__name__)


class CifCaf(Generator):
    """Generate CifCaf poses from fields.

    :param nms: set to None to switch off non-maximum suppression.
    """
    connection_method = 'blend'
    occupancy_visualizer = None
    force_complete = False
    greedy = False
    keypoint_threshold = 0.0
    deterministic_exp = True    # exp() of include/pifpaf_detmath.h (bit-reproducible)
    max_annotations = 256       # per-image capacity of the result buffers
    host_chunk = 16               # images per pipeline chunk when the fields come from the host
    head_collector = None         # network.CifCafCollector when the model emits raw head outputs
    # host fields: 'copy' moves every plane to the device; 'mapped' copies the confidence planes only
    # and lets the kernels fetch the other planes of the active cells from the pinned host buffers
    # (needs pinned inputs, seed_threshold >= CifHr.v_threshold, no force_complete; else it copies);
    # 'auto' (default) = 'mapped' when those conditions hold, else 'copy'.  Like the reference, which
    # reads every confidence value but the other planes only under the threshold mask
    # (cif_hr.py:27-31, caf_scored.py:42-45), the mapped route moves just the bytes the decode reads.
    host_transfer = 'auto'
    host_lanes = 8                # sub-batches in flight in 'mapped' mode ...
    mapped_chunk = 8              # ... of this many images (641 px, batch 64: 8 x 8 -> 2.4 ms, 64 x 1 -> 3.3 ms)
    # decode_stream on host batches: successive batches overlap, so larger sub-batches (fewer submissions,
    # fewer Annotation-building rounds) win: 641 px, batch 64: 32 x 4 lanes -> 1.59 ms per batch (the PCIe
    # time of the bytes that move), 8 x 8 -> 1.77 ms
    stream_chunk = 32
    stream_lanes = 4
    stream_stats = None           # host-time breakdown of the last decode_stream call
    stage_profile = None          # engine.StageProfile while decoder.profiler.Profiler wraps a call

    def __init__(self, field_config: FieldConfig, *,
                 keypoints,
                 skeleton,
                 out_skeleton=None,
                 confidence_scales=None,
                 worker_pool=None,
                 nms=True):
        super().__init__(worker_pool)
        if nms is True:
            nms = nms_module.Keypoints()
        self.field_config = field_config
        self.keypoints = keypoints
        self.skeleton = skeleton
        self.skeleton_m1 = np.asarray(skeleton) - 1
        self.out_skeleton = out_skeleton or skeleton
        self.confidence_scales = confidence_scales
        self.nms = nms
        self.timers = defaultdict(float)

        # neighbour tables in the reference's dict order (cifcaf.py:57-65)
        self.by_target = defaultdict(dict)
        self.by_source = defaultdict(dict)
        for caf_i, (j1, j2) in enumerate(self.skeleton_m1.tolist()):
            self.by_target[j2][j1] = (caf_i, True)
            self.by_target[j1][j2] = (caf_i, False)
            self.by_source[j1][j2] = (caf_i, True)
            self.by_source[j2][j1] = (caf_i, False)
        self._neighbours = {j: list(self.by_source[j].keys()) for j in range(len(keypoints))}
        self._engine = None

    def __getstate__(self):
        state = super().__getstate__()
        state.pop('_engine', None)
        return state

    # ------------------------------------------------------------ plumbing
    def engine(self):
        if getattr(self, '_engine', None) is None or self._engine.ann_cap != self.max_annotations:
            self._engine = DeviceDecoder(
                [tuple(link) for link in self.skeleton],
                confidence_scales=self.confidence_scales,
                stride=self._single_stride(),
                ann_cap=self.max_annotations)
        return self._engine

    def _single_stride(self):
        cfg = self.field_config      # multi-scale: the stride of set 0 (which defines the hr map)
        if cfg.cif_strides[0] != cfg.caf_strides[0]:
            raise NotImplementedError('cif and caf heads must share one stride')
        return int(cfg.cif_strides[0])

    def params(self):
        """Snapshot of the class-attribute configuration (decoder/factory.py:74-92)."""
        cfg = self.field_config
        if CifSeeds.threshold is None:
            raise TypeError("'>' not supported between instances of 'float' and 'NoneType' "
                            "(CifSeeds.threshold is not configured)")
        kp = self.nms
        return DecoderParams(
            cif_threshold=CifHr.v_threshold,
            cif_neighbors=CifHr.neighbors,
            seed_threshold=CifSeeds.threshold,
            seed_score_scale=CifSeeds.score_scale,
            caf_threshold=CafScored.default_score_th,
            keypoint_threshold=self.keypoint_threshold,
            nms_keypoint_threshold=kp.keypoint_threshold if kp is not None else 0.0,
            nms_instance_threshold=kp.instance_threshold if kp is not None else 0.0,
            nms_suppression=kp.suppression if kp is not None else 0.0,
            cif_min_scale=cfg.cif_min_scales[0],
            caf_min_distance=cfg.caf_min_distances[0],
            caf_max_distance=cfg.caf_max_distances[0],
            connection_method=self.connection_method,
            greedy=self.greedy,
            force_complete=self.force_complete,
            nms=kp is not None,
            deterministic_exp=self.deterministic_exp,
            seed_mask=cfg.seed_mask,
            host_mapped=self.host_transfer in ('mapped', 'auto'),
            profile=self.stage_profile.handle if self.stage_profile is not None else None,
        )

    def _use_mapped(self, cif, caf, params):
        """True when the host fields go the 'mapped' route (see ``host_transfer``)."""
        if self.host_transfer not in ('mapped', 'auto'):
            return False
        if self.host_transfer == 'mapped':
            return True       # as asked; the library itself falls back to copying when it must
        return (isinstance(cif, torch.Tensor) and isinstance(caf, torch.Tensor)
                and cif.is_pinned() and caf.is_pinned()
                and not params.force_complete and params.seed_threshold >= params.cif_threshold)

    def _annotations_from(self, res):
        """``engine.PackedResult`` of some images -> List[List[Annotation]].

        The records were copied out of the (reused, pinned) result block once; each Annotation owns
        row views of that private copy."""
        if res.total == 0:
            return [[] for _ in range(len(res.n_ann))]
        anns = Annotation.wrap_batch(self.keypoints, self.out_skeleton, res.data, res.scales, res.pre,
                                     res.order, self._neighbours)
        out, k = [], 0
        for n in res.n_ann.tolist():
            out.append(anns[k:k + n])
            k += n
        return out

    # -------------------------------------------------------------- decode
    def _initial_tensors(self, initial, device):
        """List (per image) of lists of Annotation -> padded CUDA tensors for pp_initial."""
        F = len(self.keypoints)
        cap = max(1, max(len(anns) for anns in initial))
        data = np.zeros((len(initial), cap, F, 3), np.float32)
        scales = np.zeros((len(initial), cap, F), np.float32)
        n = np.zeros((len(initial),), np.int32)
        for b, anns in enumerate(initial):
            n[b] = len(anns)
            for i, ann in enumerate(anns):
                data[b, i] = ann.data
                scales[b, i] = ann.joint_scales
        return (torch.from_numpy(data).to(device), torch.from_numpy(scales).to(device),
                torch.from_numpy(n).to(device))

    def decode_batch(self, cif, caf, initial_annotations=None, metas=None):
        """Decode a batch. ``cif (B,F,5,H,W)``, ``caf (B,L,9,H,W)`` float32.
        ``initial_annotations``: optional list (per image) of lists of Annotation that are
        grown and marked occupied before the seeds (cifcaf.py:95-98); like the reference,
        those objects are updated in place and returned in the result lists.
        ``metas``: optional list (per image) of the reference's preprocessing meta dicts;
        when given, ``Preprocess.annotations_inverse`` (transforms/preprocess.py:35-82) is
        applied on the device and the annotations come back in original-image coordinates.

        CUDA tensors are decoded in place on their device; host arrays (numpy or
        CPU tensors, ideally pinned) go through ``pp_decode_batch_host``, which
        includes the host->device and device->host copies.
        """
        start = time.perf_counter()
        params = self.params()
        eng = self.engine()
        has_initial = bool(initial_annotations) and any(len(a) for a in initial_annotations)
        if has_initial and not (isinstance(cif, torch.Tensor) and cif.is_cuda):
            device = eng._device()           # the pipelined host path carries no initial annotations
            cif = torch.as_tensor(np.ascontiguousarray(cif) if not isinstance(cif, torch.Tensor)
                                  else cif).to(device)
            caf = torch.as_tensor(np.ascontiguousarray(caf) if not isinstance(caf, torch.Tensor)
                                  else caf).to(device)
        if has_initial:
            params.initial = self._initial_tensors(initial_annotations, cif.device)
        if metas is not None:
            from ...transforms.preprocess import hflip_target, pack_metas
            if len(metas) != len(cif):
                raise ValueError('one meta dict per image')
            device = cif.device if isinstance(cif, torch.Tensor) and cif.is_cuda else eng._device()
            params.inverse = (pack_metas(metas, device),
                              hflip_target(metas, len(self.keypoints), device))
        if isinstance(cif, torch.Tensor) and cif.is_cuda:
            # kernels + one small copy (header and live rows of the packed result block), one sync
            buf = eng.run_device(cif, caf, params, fetch=True)
            torch.cuda.current_stream(cif.device).synchronize()
            res = eng.collect(buf)
        else:
            # pipelined: annotations of chunk k are built while chunk k+1.. are still
            # being copied / decoded (engine.run_host_chunked)
            result = []
            if self._use_mapped(cif, caf, params):
                # short copies: the GPU is the bottleneck, so sub-batches go to lanes of their own
                chunks = eng.run_host_lanes(cif, caf, params, self.mapped_chunk, self.host_lanes)
            else:
                chunks = eng.run_host_chunked(cif, caf, params, self.host_chunk)
            for buf, b0, b1 in chunks:
                result.extend(self._annotations_from(buf.host_result(b0, b1 - b0).check()))
            self.last_decoder_time = time.perf_counter() - start
            return result
        result = self._annotations_from(res)
        if has_initial:
            self._adopt_initial(result, res.source, initial_annotations)
        self.last_decoder_time = time.perf_counter() - start
        return result

    @staticmethod
    def _adopt_initial(result, source, initial):
        """The reference mutates and returns the caller's initial Annotation objects.
        ``source``: flat over the images, like the annotations of ``result``."""
        k = 0
        for b, anns in enumerate(result):
            for r, ann in enumerate(anns):
                src = int(source[k])
                k += 1
                if src < len(initial[b]):
                    orig = initial[b][src]
                    orig.data = ann.data
                    orig.joint_scales = ann.joint_scales
                    orig.decoding_order = list(orig.decoding_order) + list(ann.decoding_order)
                    orig.frontier_order = list(orig.frontier_order) + list(ann.frontier_order)
                    anns[r] = orig

    def decode_stream(self, batches, *, n_lanes=8, metas=None):
        """Generator over batches ``(cif, caf)``: yields each batch's List[List[Annotation]] in order,
        with the same results as ``decode_batch`` per batch.

        CUDA tensors: ``n_lanes`` batches are in flight on their own streams / workspaces.  Host
        tensors / arrays (ideally pinned, see ``host_transfer``): the sub-batches of successive batches
        follow each other through ``host_lanes`` lanes, so the transfer of the next batch overlaps the
        kernels, the result copies and the construction of the ``Annotation`` objects of this one
        (``stream_chunk`` images per sub-batch on ``stream_lanes`` lanes)."""
        eng = self.engine()
        params = self.params()
        if metas is not None:
            # ``metas``: one list of the reference's preprocessing meta dicts per batch; the annotations
            # then come back in original-image coordinates (Preprocess.annotations_inverse on the device)
            from ...transforms.preprocess import hflip_target, pack_metas
            device = eng._device()

            def with_inverse(pairs, meta_lists):
                for (cif, caf), m in zip(pairs, meta_lists):
                    if len(m) != len(cif):
                        raise ValueError('one meta dict per image')
                    yield cif, caf, (pack_metas(m, device), hflip_target(m, len(self.keypoints), device))
            batches = with_inverse(batches, metas)
        batches = iter(batches)
        first = next(batches, None)
        if first is None:
            return
        # where the host thread's time went (seconds, this call): enqueueing work, blocked waiting for
        # results, building the Annotation objects; the rest is the caller's own time between items
        stats = self.stream_stats = {'submit_s': 0.0, 'wait_s': 0.0, 'build_s': 0.0}
        clock = time.perf_counter
        if not (isinstance(first[0], torch.Tensor) and first[0].is_cuda):
            import itertools
            result = []
            for _, buf, nb, last in eng.run_host_stream(itertools.chain([first], batches), params,
                                                        self.stream_chunk, self.stream_lanes, stats=stats):
                t0 = clock()
                result.extend(self._annotations_from(buf.host_result(0, nb).check()))
                stats['build_s'] += clock() - t0
                if last:
                    yield result
                    result = []
            return
        import itertools
        batches = itertools.chain([first], batches)
        tickets = []

        def collect(ticket):
            t0 = clock()
            buf = ticket.wait()
            t1 = clock()
            out = self._annotations_from(eng.collect(buf))
            stats['wait_s'] += t1 - t0
            stats['build_s'] += clock() - t1
            return out

        for batch in batches:
            if len(tickets) == n_lanes:
                yield collect(tickets.pop(0))
            t0 = clock()
            p_b = params if len(batch) == 2 else dataclasses.replace(params, inverse=batch[2])
            tickets.append(eng.submit(batch[0], batch[1], p_b, n_lanes=n_lanes, fetch=True))
            stats['submit_s'] += clock() - t0
        while tickets:
            yield collect(tickets.pop(0))

    def decode_batch_multiscale(self, fields, initial_annotations=None):
        """Multi-scale FieldConfig (decoder/factory.py:153-180): ``fields[i]`` are the batched head
        tensors addressed by ``field_config.cif_indices`` / ``caf_indices`` (each with its own
        resolution and stride).  The reference pairs heads (k, k + 5) in CifHr when there are exactly
        ten (cif_hr.py:66-72).  Everything runs on the device; returns List[List[Annotation]]."""
        start = time.perf_counter()
        cfg = self.field_config
        cfg.verify()
        if cfg.cif_strides != cfg.caf_strides or len(cfg.cif_indices) != len(cfg.caf_indices):
            raise NotImplementedError('cif and caf heads of one scale share a stride')
        eng = self.engine()
        device = eng._device()

        def dev(t):
            if not isinstance(t, torch.Tensor):
                t = torch.from_numpy(np.ascontiguousarray(t, dtype=np.float32))
            return t.to(device)

        sets = [(dev(fields[ci]), dev(fields[ai]), stride, min_scale, min_d, max_d)
                for ci, ai, stride, min_scale, min_d, max_d in zip(
                    cfg.cif_indices, cfg.caf_indices, cfg.cif_strides, cfg.cif_min_scales,
                    cfg.caf_min_distances, cfg.caf_max_distances)]
        params = self.params()
        has_initial = bool(initial_annotations) and any(len(a) for a in initial_annotations)
        if has_initial:
            params.initial = self._initial_tensors(initial_annotations, device)
        buf = eng.run_device_multiscale(sets, params, hflip_pairs=len(sets) == 10, fetch=True)
        torch.cuda.current_stream(device).synchronize()
        res = eng.collect(buf)
        result = self._annotations_from(res)
        if has_initial:
            self._adopt_initial(result, res.source, initial_annotations)
        self.last_decoder_time = time.perf_counter() - start
        return result

    def decode_heads(self, heads):
        """Head tensors of a batch -> annotations.  With ``head_collector`` set (a
        ``network.CifCafCollector``) the heads are the raw convolution outputs of the fused head
        networks and the epilogue (heads.py:65-88,406-448) runs on the device first."""
        cfg = self.field_config
        if len(cfg.cif_indices) > 1:
            return self.decode_batch_multiscale(heads)
        cif, caf = heads[cfg.cif_indices[0]], heads[cfg.caf_indices[0]]
        if self.head_collector is not None:
            cif, caf = self.head_collector(cif, caf)
        return self.decode_batch(cif, caf)

    def __call__(self, fields, initial_annotations=None):
        start = time.perf_counter()
        cfg = self.field_config
        if cfg.cif_visualizers:
            for vis, cif_i in zip(cfg.cif_visualizers, cfg.cif_indices):
                vis.predicted(fields[cif_i])
        if cfg.caf_visualizers:
            for vis, caf_i in zip(cfg.caf_visualizers, cfg.caf_indices):
                vis.predicted(fields[caf_i])
        if len(cfg.cif_indices) > 1:
            batched = [f.unsqueeze(0) if isinstance(f, torch.Tensor)
                       else np.ascontiguousarray(f, dtype=np.float32)[None] for f in fields]
            return self.decode_batch_multiscale(
                batched, [list(initial_annotations)] if initial_annotations else None)[0]
        cif, caf = fields[cfg.cif_indices[0]], fields[cfg.caf_indices[0]]
        if isinstance(cif, torch.Tensor):
            cif, caf = cif.unsqueeze(0), caf.unsqueeze(0)
        else:
            cif = np.ascontiguousarray(cif, dtype=np.float32)[None]
            caf = np.ascontiguousarray(caf, dtype=np.float32)[None]
        annotations = self.decode_batch(
            cif, caf, [list(initial_annotations)] if initial_annotations else None)[0]
        LOG.debug('annotations %d, %.3fs', len(annotations), time.perf_counter() - start)
        if LOG.isEnabledFor(logging.INFO):
            LOG.info('%d annotations: %s', len(annotations),
                     [np.sum(ann.data[:, 2] > 0.1) for ann in annotations])
        return annotations
