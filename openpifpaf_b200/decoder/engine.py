"""Device-side plumbing of the decoder: buffers, parameter marshalling, calls
into libpifpaf_b200.so.  PyTorch is used for device memory, pinned host memory
and streams only; every computation happens in the CUDA library.
"""
import ctypes as C
import dataclasses
import logging
import time
from typing import Optional

import numpy as np
import torch

from .. import _lib

LOG = logging.getLogger(__name__)


@dataclasses.dataclass
class DecoderParams:
    """Python mirror of ``pp_params``; field meaning as in include/pifpaf_b200.h."""
    cif_threshold: float = 0.1
    seed_threshold: float = 0.2
    seed_score_scale: float = 1.0
    caf_threshold: float = 0.1
    cif_floor: float = 0.1
    keypoint_threshold: float = 0.001
    nms_keypoint_threshold: float = 0.001
    nms_instance_threshold: float = 0.0
    nms_suppression: float = 0.0
    cif_min_scale: float = 0.0
    cif_neighbors: float = 16.0
    caf_min_distance: float = 0.0
    caf_max_distance: Optional[float] = None
    connection_method: str = 'blend'
    greedy: bool = False
    force_complete: bool = False
    nms: bool = True
    deterministic_exp: bool = True
    seed_mask: Optional[list] = None
    profile: Optional[int] = None      # pp_profile* from StageProfile (bench/profiling only)
    # initial annotations: (data (B, cap, F, 3), scales (B, cap, F), n (B,) int32) CUDA tensors
    initial: Optional[tuple] = None
    # annotations_inverse: (meta (B, sizeof pp_image_meta) uint8 CUDA tensor, hflip_target (F,) int32
    # CUDA tensor or None), see transforms.preprocess.pack_metas
    inverse: Optional[tuple] = None
    # host entries: copy only the confidence planes, read the rest of the active cells from the pinned
    # (device-mapped) host buffers over PCIe (pp_params.host_mapped)
    host_mapped: bool = False
    # batches kept in flight on other streams (pp_params.in_flight); set by submit / run_host_lanes
    in_flight: int = 1
    # schedule knobs of pp_params (0 = the library's choice; results never change): asm_heavy_seeds,
    # asm_stage_kb, cifhr_group, cifhr_ctas_per_sm, cifhr_rshift, cifhr_warps
    tuning: Optional[dict] = None
    splat_gate: Optional[int] = None    # pp_gate* shared by the lanes of an engine (set by submit)

    def to_c(self):
        p = _lib.Params()
        _lib.lib().pp_default_params(C.byref(p))
        p.nms_instance_threshold = float(self.nms_instance_threshold)
        p.cif_threshold = self.cif_threshold
        p.seed_threshold = self.seed_threshold
        p.seed_score_scale = self.seed_score_scale
        p.caf_threshold = self.caf_threshold
        p.cif_floor = self.cif_floor
        # (1.0 - cif_floor) is a Python float in the reference (caf_scored.py:64)
        p.cif_floor_compl = float(np.float32(1.0 - self.cif_floor))
        p.keypoint_threshold = self.keypoint_threshold
        p.nms_keypoint_threshold = self.nms_keypoint_threshold
        p.nms_suppression = self.nms_suppression
        p.cif_min_scale = self.cif_min_scale or 0.0
        p.cif_neighbors = float(self.cif_neighbors)
        p.caf_min_distance = self.caf_min_distance or 0.0
        p.caf_max_distance = self.caf_max_distance or 0.0
        if self.connection_method not in ('blend', 'max'):
            raise Exception('connection method not known')   # cifcaf.py:145
        p.connection_method = 1 if self.connection_method == 'max' else 0
        p.greedy = int(bool(self.greedy))
        p.force_complete = int(bool(self.force_complete))
        p.nms = int(bool(self.nms))
        p.exp_mode = 1 if self.deterministic_exp else 0
        mask = 0
        if self.seed_mask is not None:
            if len(self.seed_mask) > _lib.PP_MAX_F:
                raise ValueError('seed_mask has %d entries, the library masks at most %d fields' % (
                    len(self.seed_mask), _lib.PP_MAX_F))
            for i, m in enumerate(self.seed_mask):
                if m:
                    mask |= 1 << i
            p.seed_mask_enabled = 1        # an all-False mask means "no field seeds"
        p.seed_mask = mask
        p.profile = self.profile
        p.host_mapped = int(bool(self.host_mapped))
        p.in_flight = int(self.in_flight)
        p.splat_gate = self.splat_gate
        for name, value in (self.tuning or {}).items():
            if name not in ('asm_heavy_seeds', 'asm_stage_kb', 'cifhr_group', 'cifhr_ctas_per_sm',
                            'cifhr_rshift', 'cifhr_warps', 'cifhr_depth'):
                raise KeyError('unknown tuning knob %r' % name)
            setattr(p, name, int(value))
        if self.initial is not None:
            data, scales, n = self.initial
            assert data.is_cuda and data.dtype == torch.float32 and data.is_contiguous()
            assert scales.is_cuda and scales.dtype == torch.float32 and scales.is_contiguous()
            assert n.is_cuda and n.dtype == torch.int32
            ini = _lib.Initial(data.data_ptr(), scales.data_ptr(), n.data_ptr(), data.shape[1])
            p._initial_struct = ini          # keep the struct alive as long as the params
            p.initial = C.pointer(ini)
        if self.inverse is not None:
            meta, target = self.inverse
            assert meta.is_cuda and meta.dtype == torch.uint8 and meta.is_contiguous()
            assert meta.shape[1] == C.sizeof(_lib.ImageMeta)
            assert target is None or (target.is_cuda and target.dtype == torch.int32)
            inv = _lib.Inverse(meta.data_ptr(), target.data_ptr() if target is not None else None)
            p._inverse_struct = inv
            p.inverse = C.pointer(inv)
        return p


class StageProfile:
    """Per-stage CUDA-event timer of the library (pp_profile_*)."""

    def __init__(self, max_steps):
        self.max_steps = max_steps
        self.handle = _lib.lib().pp_profile_create(max_steps)
        if not self.handle:
            raise _lib.PifPafError(_lib.PP_ERR_CUDA, 'pp_profile_create failed')
        self.names = [_lib.lib().pp_stage_name(i).decode() for i in range(_lib.PP_N_STAGES)]

    def reset(self):
        _lib.lib().pp_profile_reset(self.handle)

    def read(self):
        """Synchronises and returns ms per stage, shape (steps, PP_N_STAGES)."""
        ms = np.zeros((self.max_steps, _lib.PP_N_STAGES), np.float32)
        n = C.c_int32(0)
        _lib.check(_lib.lib().pp_profile_read(self.handle, ms.ctypes.data_as(C.c_void_p),
                                              C.byref(n)))
        return ms[:n.value]

    def read_abs(self):
        """Synchronises and returns ms from the first recorded event to every stage start and step end,
        shape (steps, PP_N_STAGES + 1): a timeline across streams."""
        ms = np.zeros((self.max_steps, _lib.PP_N_STAGES + 1), np.float32)
        n = C.c_int32(0)
        _lib.check(_lib.lib().pp_profile_read_abs(self.handle, ms.ctypes.data_as(C.c_void_p), C.byref(n)))
        return ms[:n.value]

    def close(self):
        if self.handle:
            _lib.lib().pp_profile_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:  # pylint: disable=broad-except
            pass


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def _stream_ptr():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


class PackedResult:
    """The live result rows of ``n`` images, parsed from one packed block (include/pifpaf_b200.h,
    "Packed results").  All arrays are views of ONE private copy of the records, flat over the images:
    ``n_ann (n,)`` rows per image; ``data (T, F, 3)``, ``scales (T, F)``, ``score (T,)`` float64,
    ``source (T,)``, ``pre (T, F, 3)`` / ``order (T, F, 2)`` = the as-grown rows / decoding order of the
    pre-NMS annotation each final annotation came from."""

    def __init__(self, words, nb, F):
        hdr = (4 + 3 * nb + 1) & ~1
        total, rw = int(words[2]), int(words[3])
        assert int(words[0]) == nb and int(words[1]) == F, 'packed block header does not match the call'
        self.n_ann = np.array(words[4:4 + nb])
        self.status = np.array(words[4 + nb:4 + 2 * nb])
        self.n_pre = np.array(words[4 + 2 * nb:4 + 3 * nb])
        rec = np.array(words[hdr:hdr + total * rw]).reshape(total, rw)     # the private copy
        self.total = total
        self.score = rec[:, 0:2].view(np.float64)[:, 0]
        self.source = rec[:, 2]
        self.data = rec[:, 4:4 + 3 * F].view(np.float32).reshape(total, F, 3)
        self.scales = rec[:, 4 + 3 * F:4 + 4 * F].view(np.float32)
        self.pre = rec[:, 4 + 4 * F:4 + 7 * F].view(np.float32).reshape(total, F, 3)
        self.order = rec[:, 4 + 7 * F:4 + 9 * F].reshape(total, F, 2)

    def check(self):
        if (self.status & _lib.PP_STATUS_OVERFLOW_MASK).any():
            raise _lib.PifPafError(_lib.PP_ERR_CAPACITY, describe_status(self.status))
        if (self.status & _lib.PP_STATUS_SCORE_ABOVE_ONE).any():
            LOG.warning('connection scores above 1 in images %s (CAF confidences above 1?): the decoding '
                        'order is not proven to be the reference\'s', np.nonzero(
                            self.status & _lib.PP_STATUS_SCORE_ABOVE_ONE)[0].tolist())
        return self


class HrMemory:
    """An hr map in a compressible device allocation (``pp_hr_alloc``: cuMemCreate with compute data
    compression, which torch's allocator cannot ask for).  A tensor made from it (``__cuda_array_interface__``)
    keeps this object alive; the memory goes back when the last such tensor dies."""

    def __init__(self, shape, device):
        self.shape = tuple(int(v) for v in shape)
        self.device = device
        self.ptr = C.c_void_p()
        comp = C.c_int32(0)
        with torch.cuda.device(device):
            _lib.check(_lib.lib().pp_hr_alloc(4 * int(np.prod(self.shape)), C.byref(self.ptr), C.byref(comp)))
        self.compressed = bool(comp.value)
        self.__cuda_array_interface__ = {'shape': self.shape, 'typestr': '<f4', 'data': (self.ptr.value, False),
                                         'version': 2, 'strides': None}

    def __del__(self):
        ptr, self.ptr = getattr(self, 'ptr', None), None
        if ptr is not None and ptr.value:
            try:
                torch.cuda.synchronize(self.device)
                _lib.lib().pp_hr_free(ptr)
            except Exception:  # pylint: disable=broad-except
                pass


def hr_tensor(shape, device):
    """(tensor, compressed): the hr map in compressible memory, or a plain torch tensor when the driver
    refuses (``BatchBuffers.compress_hr = False`` forces the plain one)."""
    if BatchBuffers.compress_hr:
        try:
            mem = HrMemory(shape, device)
            return torch.as_tensor(mem, device=device), mem.compressed     # the tensor keeps `mem` alive
        except Exception as e:  # pylint: disable=broad-except
            # any failure (driver without VMM, out of address space, ...): the plain allocation always works
            LOG.warning('pp_hr_alloc failed (%s): hr maps in ordinary device memory', e)
    return torch.empty(shape, dtype=torch.float32, device=device), False


class BatchBuffers:
    """Caller-owned device (and optional pinned host) buffers for one batch shape."""
    compress_hr = True       # hr maps in compressible memory (pp_hr_alloc): ~1.2 x the HBM write rate

    def __init__(self, shape, caps, device, *, keep=('ann',), host=False):
        B, F, L = shape.B, shape.F, shape.L
        hh, hw = (shape.H - 1) * shape.stride + 1, (shape.W - 1) * shape.stride + 1
        self.shape, self.caps, self.device = shape, caps, device
        dev = dict(device=device)
        f32, i32, f64 = torch.float32, torch.int32, torch.float64
        t = {}
        t['ann_data'] = torch.empty((B, caps.ann_cap, F, 3), dtype=f32, **dev)
        t['ann_scales'] = torch.empty((B, caps.ann_cap, F), dtype=f32, **dev)
        t['ann_score'] = torch.empty((B, caps.ann_cap), dtype=f64, **dev)
        t['ann_source'] = torch.empty((B, caps.ann_cap), dtype=i32, **dev)
        t['n_ann'] = torch.zeros((B,), dtype=i32, **dev)
        t['status'] = torch.zeros((B,), dtype=i32, **dev)
        t['pre_data'] = torch.empty((B, caps.ann_cap, F, 3), dtype=f32, **dev)
        t['pre_scales'] = torch.empty((B, caps.ann_cap, F), dtype=f32, **dev)
        t['pre_order'] = torch.empty((B, caps.ann_cap, F, 2), dtype=i32, **dev)
        t['n_pre'] = torch.zeros((B,), dtype=i32, **dev)
        # the hr map is always the caller's (never carved from the workspace): it lives in compressible memory
        t['hr'], self.hr_compressed = hr_tensor((B, F, hh, hw), device)
        if 'seeds' in keep:
            t['seeds'] = torch.empty((B, caps.seed_cap, 5), dtype=f32, **dev)
            t['n_seeds'] = torch.zeros((B,), dtype=i32, **dev)
        if 'caf' in keep:
            t['caf_fwd'] = torch.empty((B, L, 9, caps.caf_cap), dtype=f32, **dev)
            t['caf_bwd'] = torch.empty((B, L, 9, caps.caf_cap), dtype=f32, **dev)
            t['n_caf_fwd'] = torch.zeros((B, L), dtype=i32, **dev)
            t['n_caf_bwd'] = torch.zeros((B, L), dtype=i32, **dev)
        lib = _lib.lib()
        self.packed_bytes = lib.pp_packed_bytes(C.byref(shape), C.byref(caps))
        self.rec_words = lib.pp_packed_record_words(F)
        self.image_stride = self.packed_bytes // B          # block of a chunk starts at b0 * image_stride
        t['packed'] = torch.empty((self.packed_bytes,), dtype=torch.uint8, **dev)
        self.h_packed = None
        self.h_words = None
        self.fetched_records = 0
        self.dev = t
        self.d_out = _lib.Outputs()
        for name, tensor in t.items():
            setattr(self.d_out, name, tensor.data_ptr())
        ws_bytes = _lib.lib().pp_workspace_bytes_ex(C.byref(shape), C.byref(caps), 1)
        if ws_bytes == 0:
            raise _lib.PifPafError(_lib.PP_ERR_ARG, 'bad shape/caps for workspace')
        self.ws = torch.empty((ws_bytes,), dtype=torch.uint8, **dev)
        self.ws_bytes = ws_bytes
        self.host = None
        self.h_out = None
        self.d_cif = self.d_caf = None
        if host:
            self.enable_host()

    def pin(self):
        """Pinned host mirror of the packed result block (the only result bytes that cross PCIe)."""
        if self.h_packed is None:
            self.h_packed = torch.empty((self.packed_bytes,), dtype=torch.uint8, pin_memory=True)
            self.h_words = self.h_packed.numpy().view(np.int32)
        return self.h_packed

    def enable_host(self):
        """Pinned host mirror + device staging of the fields for pp_decode_batch_host."""
        if self.host is not None:
            return
        s = self.shape
        self.host = {'packed': self.pin()}
        self.h_out = _lib.Outputs()
        self.h_out.packed = self.h_packed.data_ptr()
        self.d_cif = torch.empty((s.B, s.F, 5, s.H, s.W), dtype=torch.float32, device=self.device)
        self.d_caf = torch.empty((s.B, s.L, 9, s.H, s.W), dtype=torch.float32, device=self.device)

    def prefix_bytes(self, nb, records):
        return 4 * (((4 + 3 * nb + 1) & ~1) + records * self.rec_words)

    def enqueue_fetch(self, records_per_image):
        """Device path: copy header + the first ``records_per_image * B`` records to the pinned mirror on
        the current stream (after the kernels already enqueued there)."""
        self.pin()
        B = self.shape.B
        self.fetched_records = min(records_per_image, self.caps.ann_cap) * B
        n = self.prefix_bytes(B, self.fetched_records)
        self.h_packed[:n].copy_(self.dev['packed'][:n], non_blocking=True)

    def finish_fetch(self):
        """After the stream (or the event behind ``enqueue_fetch``) has been waited for: fetch the rare
        remainder and parse.  Returns ``(PackedResult, records per image worth guessing next time)``."""
        B = self.shape.B
        total = int(self.h_words[2])
        if total > self.fetched_records:
            a, b = self.prefix_bytes(B, self.fetched_records), self.prefix_bytes(B, total)
            self.h_packed[a:b].copy_(self.dev['packed'][a:b])          # synchronous
            self.fetched_records = total
        return PackedResult(self.h_words, B, self.shape.F), (total + B - 1) // B + 2

    def host_result(self, b0=0, nb=None):
        """Host entries: the parsed block of the chunk that starts at image ``b0`` (``nb`` images)."""
        nb = self.shape.B if nb is None else nb
        return PackedResult(self.h_words[b0 * self.image_stride // 4:], nb, self.shape.F)


class DeviceDecoder:
    """Runs the CUDA decode for batches of one skeleton; caches buffers per shape."""

    def __init__(self, skeleton, *, confidence_scales=None, stride=8, ann_cap=256, device=None):
        self.skeleton = [tuple(int(j) for j in link) for link in skeleton]
        self.c_skeleton = _lib.Skeleton.from_list(self.skeleton, confidence_scales)
        self._confidence_scales = confidence_scales
        self.stride = stride
        self.ann_cap = ann_cap
        self.device = device
        self._buffers = {}
        self._pipeline = None
        _lib.lib()   # fail loudly right here if the CUDA library is missing

    def _device(self):
        if self.device is None:
            if not torch.cuda.is_available():
                raise RuntimeError('openpifpaf_b200 needs a CUDA device (B200); there is no CPU path')
            self.device = torch.device('cuda', torch.cuda.current_device())
        return self.device

    def buffers(self, B, F, H, W, *, keep=('ann',), host=False):
        key = (B, F, H, W, tuple(sorted(keep)))
        buf = self._buffers.get(key)
        if buf is None:
            shape = _lib.Shape(B, F, len(self.skeleton), H, W, self.stride)
            caps = _lib.Caps()
            _lib.check(_lib.lib().pp_default_caps(C.byref(shape), C.byref(caps)))
            caps.ann_cap = self.ann_cap
            buf = BatchBuffers(shape, caps, self._device(), keep=keep)
            self._buffers[key] = buf
        if host:
            buf.enable_host()
        return buf

    fetch_guess = 6       # records per image copied with the header; grows to what the batches need
    # batches in flight: serialise the splat kernels of the lanes (pp_gate).  Off: measured on B200 the
    # gate keeps a splat running 100 % of the time but the step does not get shorter (0.425 vs 0.420 ms):
    # the other stages' kernels occupy whole SMs while they run, so their time adds to the splat's
    # either way (batches in flight)
    use_splat_gate = False

    def gate(self):
        if getattr(self, '_gate', None) is None:
            self._gate = _lib.lib().pp_gate_create()
            if not self._gate:
                raise _lib.PifPafError(_lib.PP_ERR_CUDA, 'pp_gate_create failed')
        return self._gate

    def collect(self, buf):
        """Results of a ``run_device(..., fetch=True)`` batch whose stream / event has been waited for."""
        res, guess = buf.finish_fetch()
        if guess > self.fetch_guess:
            DeviceDecoder.fetch_guess = guess
        return res.check()

    def run_device(self, cif, caf, params: DecoderParams, *, keep=('ann',), fetch=False):
        """cif (B,F,5,H,W), caf (B,L,9,H,W): CUDA float32 tensors. Enqueues only.  ``fetch``: also
        enqueue the copy of the packed results to the pinned mirror (finish with ``collect``)."""
        assert cif.is_cuda and caf.is_cuda and cif.dtype == torch.float32 and caf.dtype == torch.float32
        cif, caf = cif.contiguous(), caf.contiguous()
        B, F, five, H, W = cif.shape
        assert five == 5 and caf.shape == (B, len(self.skeleton), 9, H, W), (cif.shape, caf.shape)
        buf = self.buffers(B, F, H, W, keep=keep)
        p = params.to_c()
        rc = _lib.lib().pp_decode_batch(_ptr(cif), _ptr(caf), C.byref(self.c_skeleton),
                                        C.byref(buf.d_out), _ptr(buf.ws), buf.ws_bytes,
                                        C.byref(buf.shape), C.byref(buf.caps), C.byref(p),
                                        _stream_ptr())
        _lib.check(rc)
        if fetch:
            buf.enqueue_fetch(self.fetch_guess)
        return buf

    # ---- several batches in flight -------------------------------------------------------
    # The latency-bound stages (assembly, seed sort, NMS: one CTA per image) leave most SMs and all
    # of the HBM bandwidth idle; a second and third batch on their own streams and workspaces fill
    # them with their HBM-bound stages (measured at batch 64 x 641 px: 85 k -> 136 k -> 141 k img/s
    # for 1 / 3 / 4 lanes).
    def lanes(self, n):
        """``n`` independent (engine, stream) lanes sharing this decoder's configuration."""
        cur = getattr(self, '_lanes', [])
        while len(cur) < n:
            child = DeviceDecoder(self.skeleton, confidence_scales=self._confidence_scales,
                                  stride=self.stride, ann_cap=self.ann_cap, device=self._device())
            cur.append(Lane(child, torch.cuda.Stream(device=self._device())))
        self._lanes = cur
        return cur[:n]

    def submit(self, cif, caf, params: DecoderParams, *, n_lanes=8, keep=('ann',), fetch=False):
        """Enqueue one batch on the next lane and return a ``Ticket``; at most ``n_lanes`` batches
        are in flight, so the ticket that used this lane last must have been waited for."""
        lanes = self.lanes(n_lanes)
        self._next_lane = (getattr(self, '_next_lane', -1) + 1) % n_lanes
        lane = lanes[self._next_lane]
        if lane.pending is not None and not lane.pending.done:
            raise RuntimeError('lane still holds an uncollected batch: wait() for the ticket issued '
                               '%d submits ago before submitting more' % n_lanes)
        lane.stream.wait_stream(torch.cuda.current_stream(self._device()))
        if n_lanes > 1 and params.in_flight != n_lanes:
            params = dataclasses.replace(params, in_flight=n_lanes)
        if n_lanes > 1 and self.use_splat_gate and params.splat_gate is None:
            params = dataclasses.replace(params, splat_gate=self.gate())
        with torch.cuda.stream(lane.stream):
            buf = lane.engine.run_device(cif, caf, params, keep=keep, fetch=fetch)
            done = torch.cuda.Event()
            done.record(lane.stream)
        lane.pending = Ticket(buf, done, (cif, caf))
        return lane.pending

    def run_device_multiscale(self, sets, params: DecoderParams, *, hflip_pairs=False, keep=('ann',),
                              fetch=False):
        """Multi-scale FieldConfig: ``sets`` = list of ``(cif, caf, stride, cif_min_scale,
        caf_min_distance, caf_max_distance)`` with CUDA float32 tensors ``(B,F,5,H_k,W_k)`` /
        ``(B,L,9,H_k,W_k)``; set 0 defines the hr shape.  Enqueues only."""
        arr = (_lib.FieldSet * len(sets))()
        alive = []
        seed_cap = caf_cap = src_cap = 0
        B = F = None
        for k, (cif, caf, stride, min_scale, min_d, max_d) in enumerate(sets):
            assert cif.is_cuda and caf.is_cuda and cif.dtype == torch.float32 and caf.dtype == torch.float32
            cif, caf = cif.contiguous(), caf.contiguous()
            alive += [cif, caf]
            B, F, five, H, W = cif.shape
            assert five == 5 and caf.shape == (B, len(self.skeleton), 9, H, W), (cif.shape, caf.shape)
            arr[k] = _lib.FieldSet(cif.data_ptr(), caf.data_ptr(), H, W, int(stride), float(min_scale or 0.0),
                                   float(min_d or 0.0), float(max_d or 0.0))
            seed_cap += F * H * W
            caf_cap += H * W
            src_cap = max(src_cap, (2 if hflip_pairs else 1) * H * W)
        key = ('ms', B, F, tuple((a.H, a.W, a.stride) for a in arr), bool(hflip_pairs), tuple(sorted(keep)))
        buf = self._buffers.get(key)
        if buf is None:
            shape = _lib.Shape(B, F, len(self.skeleton), arr[0].H, arr[0].W, arr[0].stride)
            caps = _lib.Caps(min(seed_cap, (1 << 24) - 1), caf_cap, self.ann_cap, src_cap)
            buf = BatchBuffers(shape, caps, self._device(), keep=keep)
            self._buffers[key] = buf
        p = params.to_c()
        _lib.check(_lib.lib().pp_decode_batch_multiscale(
            arr, len(sets), int(bool(hflip_pairs)), C.byref(self.c_skeleton), C.byref(buf.d_out),
            _ptr(buf.ws), buf.ws_bytes, C.byref(buf.shape), C.byref(buf.caps), C.byref(p), _stream_ptr()))
        buf.alive = alive           # the field tensors must outlive the enqueued kernels
        if fetch:
            buf.enqueue_fetch(self.fetch_guess)
        return buf

    def run_host(self, cif, caf, params: DecoderParams, *, keep=('ann',)):
        """cif, caf: host float32 arrays/tensors (ideally pinned). Synchronous."""
        cif_t = torch.as_tensor(cif)
        caf_t = torch.as_tensor(caf)
        assert cif_t.dtype == torch.float32 and caf_t.dtype == torch.float32
        cif_t, caf_t = cif_t.contiguous(), caf_t.contiguous()
        B, F, five, H, W = cif_t.shape
        assert five == 5 and caf_t.shape == (B, len(self.skeleton), 9, H, W), (cif_t.shape, caf_t.shape)
        buf = self.buffers(B, F, H, W, keep=keep, host=True)
        p = params.to_c()
        with torch.cuda.device(buf.device):
            rc = _lib.lib().pp_decode_batch_host(
                _ptr(cif_t), _ptr(caf_t), _ptr(buf.d_cif), _ptr(buf.d_caf),
                C.byref(self.c_skeleton), C.byref(buf.d_out), C.byref(buf.h_out), _ptr(buf.ws),
                buf.ws_bytes, C.byref(buf.shape), C.byref(buf.caps), C.byref(p), _stream_ptr())
        if rc == _lib.PP_ERR_CAPACITY:
            raise _lib.PifPafError(rc, describe_status(buf.host_result().status))
        _lib.check(rc)
        return buf

    def enqueue_host(self, cif_t, caf_t, params, chunk, *, keep=('ann',), p_c=None, stream_ptr=None):
        """Enqueue the pipelined host decode of contiguous float32 host tensors (chunked H2D on the
        pipeline's copy stream, kernels on the current stream, D2H on its copy-out stream).  Returns
        ``(buf, chunk, n_chunks)``; collect with ``wait_host_chunk``.  ``p_c`` / ``stream_ptr``: an
        already converted parameter struct and the stream to issue on (callers that enqueue many
        sub-batches convert once and skip the stream context)."""
        assert cif_t.dtype == torch.float32 and caf_t.dtype == torch.float32
        B, F, five, H, W = cif_t.shape
        assert five == 5 and caf_t.shape == (B, len(self.skeleton), 9, H, W), (cif_t.shape, caf_t.shape)
        buf = self.buffers(B, F, H, W, keep=keep, host=True)
        if self._pipeline is None:
            self._pipeline = HostPipeline()
        chunk = max(1, min(int(chunk), B))
        if (B + chunk - 1) // chunk > HostPipeline.MAX_CHUNKS:
            chunk = (B + HostPipeline.MAX_CHUNKS - 1) // HostPipeline.MAX_CHUNKS
        p = params.to_c() if p_c is None else p_c
        lib = _lib.lib()
        args = (self._pipeline.handle, C.c_void_p(cif_t.data_ptr()), C.c_void_p(caf_t.data_ptr()),
                _ptr(buf.d_cif), _ptr(buf.d_caf), C.byref(self.c_skeleton), C.byref(buf.d_out),
                C.byref(buf.h_out), _ptr(buf.ws), buf.ws_bytes, C.byref(buf.shape), C.byref(buf.caps),
                C.byref(p), chunk)
        if stream_ptr is not None and torch.cuda.current_device() == buf.device.index:
            _lib.check(lib.pp_decode_batch_host_async(*args, stream_ptr))
        else:
            with torch.cuda.device(buf.device):
                _lib.check(lib.pp_decode_batch_host_async(
                    *args, _stream_ptr() if stream_ptr is None else stream_ptr))
        buf.alive = (cif_t, caf_t, p)
        buf.chunk = chunk
        return buf, chunk, (B + chunk - 1) // chunk

    def wait_host_chunk(self, buf, k):
        lib = _lib.lib()
        rc = lib.pp_pipeline_wait(self._pipeline.handle, k)
        if rc == _lib.PP_ERR_CAPACITY:
            lib.pp_pipeline_wait(self._pipeline.handle, -1)
            B, chunk = buf.shape.B, buf.chunk
            status = np.concatenate([buf.host_result(b0, min(chunk, B - b0)).status
                                     for b0 in range(0, B, chunk)])
            raise _lib.PifPafError(rc, describe_status(status))
        _lib.check(rc)

    def run_host_chunked(self, cif, caf, params, chunk, *, keep=('ann',)):
        """Generator: enqueue the pipelined host decode, then yield ``(buf, b0, b1)`` as each
        chunk's results land in the pinned host mirrors (later chunks still in flight)."""
        cif_t = torch.as_tensor(cif).contiguous()
        caf_t = torch.as_tensor(caf).contiguous()
        B = cif_t.shape[0]
        buf, chunk, n_chunks = self.enqueue_host(cif_t, caf_t, params, chunk, keep=keep)
        for k in range(n_chunks):
            self.wait_host_chunk(buf, k)
            yield buf, k * chunk, min((k + 1) * chunk, B)

    def run_host_lanes(self, cif, caf, params, sub, n_lanes, *, keep=('ann',)):
        """Like ``run_host_chunked`` but every sub-batch of ``sub`` images runs on its own lane
        (stream + workspace + copy streams), ``n_lanes`` in flight: the kernels of one sub-batch
        overlap those of the next.  Worth it when the copies are short (host_transfer='mapped').
        Yields ``(buf, 0, n_images)`` per sub-batch, in order."""
        cif_t = torch.as_tensor(cif).contiguous()
        caf_t = torch.as_tensor(caf).contiguous()
        B = cif_t.shape[0]
        sub = max(1, min(int(sub), B))
        ranges = [(b0, min(b0 + sub, B)) for b0 in range(0, B, sub)]
        lanes = self.lanes(n_lanes)
        cur = torch.cuda.current_stream(self._device())
        pending = []
        if n_lanes > 1 and len(ranges) > 1:
            params = dataclasses.replace(params, in_flight=min(n_lanes, len(ranges)),
                                         splat_gate=self.gate() if self.use_splat_gate else None)

        sliced = params.inverse is not None or params.initial is not None
        p_all = None if sliced else params.to_c()      # one conversion for all sub-batches

        def submit(i):
            b0, b1 = ranges[i]
            lane = lanes[i % n_lanes]
            if i < n_lanes:                     # order the lane after the caller's stream once per call
                lane.stream.wait_stream(cur)
            sub_params = params
            if params.inverse is not None:      # per-image records: this sub-batch's slice
                meta, target = params.inverse
                sub_params = dataclasses.replace(params, inverse=(meta[b0:b1].contiguous(), target))
            if params.initial is not None:
                data, scales, n = params.initial
                sub_params = dataclasses.replace(
                    sub_params, initial=(data[b0:b1].contiguous(), scales[b0:b1].contiguous(),
                                         n[b0:b1].contiguous()))
            if sliced:                          # the slices are made on the lane's stream
                with torch.cuda.stream(lane.stream):
                    buf, _, _ = lane.engine.enqueue_host(cif_t[b0:b1], caf_t[b0:b1], sub_params, b1 - b0,
                                                         keep=keep)
            else:
                buf, _, _ = lane.engine.enqueue_host(cif_t[b0:b1], caf_t[b0:b1], sub_params, b1 - b0,
                                                     keep=keep, p_c=p_all, stream_ptr=lane.stream_ptr)
            pending.append((lane, buf, b1 - b0))

        nxt = min(n_lanes, len(ranges))
        for i in range(nxt):
            submit(i)
        while pending:
            lane, buf, nb = pending.pop(0)
            lane.engine.wait_host_chunk(buf, 0)
            yield buf, 0, nb               # the caller copies what it needs out of the pinned mirrors
            if nxt < len(ranges):          # ... so the lane (and its buffers) can take the next sub-batch
                submit(nxt)
                nxt += 1


    def run_host_stream(self, batches, params, sub, n_lanes, *, keep=('ann',), stats=None):
        """``run_host_lanes`` over a stream of host batches: the sub-batches of successive batches follow
        each other through the lanes without a gap, so the copies of batch k + 1 run while the kernels of
        the last sub-batches of batch k are still in flight.  ``batches``: iterable of ``(cif, caf)`` host
        tensors / arrays.  Yields ``(batch index, buf, n images, last sub-batch of its batch)`` in order;
        the caller copies what it needs out of ``buf.host`` before asking for the next item."""
        if params.inverse is not None or params.initial is not None:
            raise NotImplementedError('per-batch records ride in the batch tuples: (cif, caf, inverse); '
                                      'initial annotations go through decode_batch')
        lanes = self.lanes(n_lanes)
        cur = torch.cuda.current_stream(self._device())
        if n_lanes > 1:
            params = dataclasses.replace(params, in_flight=n_lanes,
                                         splat_gate=self.gate() if self.use_splat_gate else None)
        p_all = params.to_c()
        pending = []
        n_sub = [0]

        def items():
            # a batch is (cif, caf) or (cif, caf, inverse) with inverse = (packed metas (B, .) on the
            # device, hflip targets) for annotations_inverse on the device (DecoderParams.inverse)
            for bid, batch in enumerate(batches):
                cif, caf = batch[0], batch[1]
                inverse = batch[2] if len(batch) > 2 else None
                cif_t = torch.as_tensor(cif).contiguous()
                caf_t = torch.as_tensor(caf).contiguous()
                B = cif_t.shape[0]
                step = max(1, min(int(sub), B))
                for b0 in range(0, B, step):
                    yield bid, cif_t, caf_t, b0, min(b0 + step, B), b0 + step >= B, inverse

        def submit(item):
            bid, cif_t, caf_t, b0, b1, last, inverse = item
            lane = lanes[n_sub[0] % n_lanes]
            if n_sub[0] < n_lanes:
                lane.stream.wait_stream(cur)
            if inverse is None:
                buf, _, _ = lane.engine.enqueue_host(cif_t[b0:b1], caf_t[b0:b1], params, b1 - b0, keep=keep,
                                                     p_c=p_all, stream_ptr=lane.stream_ptr)
            else:       # this sub-batch's slice of the per-image records
                meta, target = inverse
                sub_params = dataclasses.replace(params, inverse=(meta[b0:b1].contiguous(), target))
                with torch.cuda.stream(lane.stream):
                    buf, _, _ = lane.engine.enqueue_host(cif_t[b0:b1], caf_t[b0:b1], sub_params, b1 - b0,
                                                         keep=keep)
            pending.append((lane, buf, bid, b1 - b0, last))
            n_sub[0] += 1

        clock = time.perf_counter
        stats = stats if stats is not None else {}
        it = items()
        t0 = clock()
        for item in it:
            submit(item)
            if len(pending) == n_lanes:
                break
        stats['submit_s'] = stats.get('submit_s', 0.0) + clock() - t0
        while pending:
            lane, buf, bid, nb, last = pending.pop(0)
            t0 = clock()
            lane.engine.wait_host_chunk(buf, 0)
            stats['wait_s'] = stats.get('wait_s', 0.0) + clock() - t0
            yield bid, buf, nb, last
            t0 = clock()
            nxt = next(it, None)           # the lane just collected is the one the next sub-batch takes
            if nxt is not None:
                submit(nxt)
            stats['submit_s'] += clock() - t0


class Lane:
    def __init__(self, engine, stream):
        self.engine, self.stream, self.pending = engine, stream, None
        self.stream_ptr = C.c_void_p(stream.cuda_stream)


class Ticket:
    """One submitted batch: ``wait()`` blocks until its kernels finished and returns the
    ``BatchBuffers`` holding the results (valid until the lane is reused)."""

    def __init__(self, buf, event, inputs):
        self.buf, self.event, self.inputs, self.done = buf, event, inputs, False

    def wait(self):
        self.event.synchronize()
        self.done = True
        self.inputs = None
        return self.buf


class HostPipeline:
    """pp_pipeline handle: chunked H2D / kernels / D2H overlap for host inputs."""
    MAX_CHUNKS = 64

    def __init__(self):
        self.handle = _lib.lib().pp_pipeline_create(self.MAX_CHUNKS)
        if not self.handle:
            raise _lib.PifPafError(_lib.PP_ERR_CUDA, 'pp_pipeline_create failed')

    def close(self):
        if self.handle:
            _lib.lib().pp_pipeline_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:  # pylint: disable=broad-except
            pass



def describe_status(status):
    bad = {}
    for b, word in enumerate(np.asarray(status).tolist()):
        if word & _lib.PP_STATUS_OVERFLOW_MASK:
            bad[b] = [name for bit, name in _lib.STATUS_BITS.items() if word & bit]
    return 'capacity exceeded in images %s' % bad
