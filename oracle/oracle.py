"""ctypes front end of the CPU oracle (``oracle/cifcaf_oracle.c``).

TEST INFRASTRUCTURE: imported by ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` only.  The product
package ``openpifpaf_b200`` never imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, '_build', 'libcifcaf_oracle.so')

f32p = np.ctypeslib.ndpointer(np.float32, flags='C_CONTIGUOUS')
i32p = np.ctypeslib.ndpointer(np.int32, flags='C_CONTIGUOUS')
u8p = np.ctypeslib.ndpointer(np.uint8, flags='C_CONTIGUOUS')


class Config(C.Structure):
    """Mirror of ``orc_config``; defaults = SURVEY.md 8d decoder settings."""
    _fields_ = [
        ('nms_instance_th', C.c_double),
        ('cif_th', C.c_float), ('seed_th', C.c_float), ('seed_score_scale', C.c_float),
        ('caf_th', C.c_float), ('cif_floor', C.c_float), ('keypoint_th', C.c_float),
        ('nms_keypoint_th', C.c_float), ('nms_suppression', C.c_float),
        ('cif_min_scale', C.c_float), ('caf_min_distance', C.c_float),
        ('caf_max_distance', C.c_float),
        ('connection_method', C.c_int), ('greedy', C.c_int), ('force_complete', C.c_int),
        ('do_nms', C.c_int), ('exp_mode', C.c_int), ('stride', C.c_int), ('cif_neighbors', C.c_float),
    ]

    def __init__(self, **kw):
        super().__init__()
        base = dict(nms_instance_th=0.0, cif_th=0.1, seed_th=0.2, seed_score_scale=1.0,
                    caf_th=0.1, cif_floor=0.1, keypoint_th=0.001, nms_keypoint_th=0.001,
                    nms_suppression=0.0, cif_min_scale=0.0, caf_min_distance=0.0,
                    caf_max_distance=0.0, connection_method=0, greedy=0, force_complete=0,
                    do_nms=1, exp_mode=0, stride=8, cif_neighbors=16.0)
        base.update(kw)
        for k, v in base.items():
            if not any(k == name for name, _ in self._fields_):
                raise TypeError('unknown oracle option %r' % k)
            setattr(self, k, v)


class DetConfig(C.Structure):
    """Mirror of ``orc_det_config``; defaults = the reference's class attributes with the
    thresholds the test suite configures (CifHr.v_threshold 0.1, CifSeeds.threshold 0.2)."""
    _fields_ = [
        ('cif_th', C.c_float), ('seed_th', C.c_float), ('seed_score_scale', C.c_float),
        ('cif_min_scale', C.c_float), ('suppression', C.c_float), ('suppression_soft', C.c_float),
        ('instance_th', C.c_float), ('iou_th', C.c_float), ('iou_th_soft', C.c_float),
        ('do_nms', C.c_int), ('stride', C.c_int),
    ]

    def __init__(self, **kw):
        super().__init__()
        base = dict(cif_th=0.1, seed_th=0.2, seed_score_scale=1.0, cif_min_scale=0.0,
                    suppression=0.1, suppression_soft=0.3, instance_th=0.1, iou_th=0.7,
                    iou_th_soft=0.5, do_nms=1, stride=8)
        base.update(kw)
        for k, v in base.items():
            if not any(k == name for name, _ in self._fields_):
                raise TypeError('unknown oracle option %r' % k)
            setattr(self, k, v)


def build(force=False):
    src = os.path.join(HERE, 'cifcaf_oracle.c')
    hdr = os.path.join(HERE, '..', 'include', 'pifpaf_detmath.h')
    stale = (not os.path.exists(LIB_PATH)
             or (os.path.exists(src) and os.path.getmtime(src) > os.path.getmtime(LIB_PATH))
             or (os.path.exists(hdr) and os.path.getmtime(hdr) > os.path.getmtime(LIB_PATH)))
    if stale or force:
        subprocess.run(['make', '-C', HERE] + (['-B'] if force else []), check=True,
                       stdout=subprocess.DEVNULL)
    return LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(LIB_PATH)
        assert _lib.orc_config_size() == C.sizeof(Config), 'orc_config layout mismatch'
        assert _lib.orc_det_config_size() == C.sizeof(DetConfig), 'orc_det_config layout mismatch'
        _lib.orc_seeds.restype = C.c_long
        _lib.orc_caf_center_s.restype = C.c_long
        _lib.orc_annotation_score.restype = C.c_double
        _lib.orc_detexp.restype = C.c_float
        _lib.orc_detexp.argtypes = [C.c_float]
    return _lib


def hr_shape(cif, stride=8):
    return (cif.shape[0], (cif.shape[2] - 1) * stride + 1, (cif.shape[3] - 1) * stride + 1)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def cifhr(cif, cfg=None):
    cfg = cfg or Config()
    cif = _f32(cif)
    F, _, H, W = cif.shape
    hr = np.empty(hr_shape(cif, cfg.stride), np.float32)
    rc = lib().orc_cifhr(cif.ctypes.data_as(C.c_void_p), F, H, W, C.byref(cfg),
                         hr.ctypes.data_as(C.c_void_p))
    assert rc == 0, rc
    return hr


def seeds(cif, hr, cfg=None):
    cfg = cfg or Config()
    cif, hr = _f32(cif), _f32(hr)
    F, _, H, W = cif.shape
    cap = F * H * W
    out = np.empty((cap, 5), np.float32)
    n = lib().orc_seeds(cif.ctypes.data_as(C.c_void_p), F, H, W, hr.ctypes.data_as(C.c_void_p),
                        C.byref(cfg), out.ctypes.data_as(C.c_void_p), C.c_long(cap))
    assert n >= 0, n
    return out[:n].copy()


def caf_scored(caf, hr, skeleton, cfg=None, score_th=None):
    cfg = cfg or Config()
    caf, hr = _f32(caf), _f32(hr)
    L, _, H, W = caf.shape
    cap = H * W
    skel = np.ascontiguousarray(np.asarray(skeleton, dtype=np.int32).reshape(-1))
    fwd = np.zeros((L, 9, cap), np.float32)
    bwd = np.zeros((L, 9, cap), np.float32)
    nf = np.zeros(L, np.int32)
    nb = np.zeros(L, np.int32)
    th = cfg.caf_th if score_th is None else score_th
    rc = lib().orc_caf_scored(caf.ctypes.data_as(C.c_void_p), L, H, W,
                              hr.ctypes.data_as(C.c_void_p), hr.shape[0],
                              skel.ctypes.data_as(C.c_void_p), C.byref(cfg), C.c_float(th),
                              fwd.ctypes.data_as(C.c_void_p), nf.ctypes.data_as(C.c_void_p),
                              bwd.ctypes.data_as(C.c_void_p), nb.ctypes.data_as(C.c_void_p),
                              C.c_long(cap))
    assert rc == 0, rc
    return ([fwd[l, :, :nf[l]].copy() for l in range(L)],
            [bwd[l, :, :nb[l]].copy() for l in range(L)])


class Decoded:
    """Everything the oracle observed for one image."""
    __slots__ = ('hr', 'seeds', 'forward', 'backward', 'data', 'joint_scales', 'decoding_order',
                 'frontier_order', 'pre_data', 'pre_joint_scales', 'tie')


class FieldSet(C.Structure):
    """Mirror of ``orc_set``: one entry of a multi-scale FieldConfig."""
    _fields_ = [('cif', C.c_void_p), ('caf', C.c_void_p), ('H', C.c_int), ('W', C.c_int),
                ('stride', C.c_int), ('cif_min_scale', C.c_float), ('caf_min_distance', C.c_float),
                ('caf_max_distance', C.c_float)]


def decode(cif, caf, skeleton, cfg=None, *, confidence_scales=None, ann_cap=None, initial=None):
    """Full CifCaf decode of one image. Returns a ``Decoded``."""
    cfg = cfg or Config()
    return decode_ms([(cif, caf, cfg.stride, cfg.cif_min_scale, cfg.caf_min_distance,
                       cfg.caf_max_distance)], skeleton, cfg, confidence_scales=confidence_scales,
                     ann_cap=ann_cap, initial=initial)


def decode_ms(sets, skeleton, cfg=None, *, hflip_pairs=False, confidence_scales=None, ann_cap=None,
              initial=None):
    """CifCaf decode over several field sets (multi-scale FieldConfig).  ``sets``: list of
    ``(cif, caf, stride, cif_min_scale, caf_min_distance, caf_max_distance or 0)``; set 0 defines
    the hr shape.  Returns a ``Decoded``."""
    cfg = cfg or Config()
    keep = []
    arr = (FieldSet * len(sets))()
    F = L = None
    seed_cap = caf_cap = 0
    for k, (cif, caf, stride, min_scale, min_d, max_d) in enumerate(sets):
        cif, caf = _f32(cif), _f32(caf)
        keep += [cif, caf]
        F, _, H, W = cif.shape
        L = caf.shape[0]
        arr[k] = FieldSet(cif.ctypes.data, caf.ctypes.data, H, W, int(stride), float(min_scale or 0.0),
                          float(min_d or 0.0), float(max_d or 0.0))
        seed_cap += F * H * W
        caf_cap += H * W
    skel = np.ascontiguousarray(np.asarray(skeleton, dtype=np.int32).reshape(-1))
    assert skel.size == 2 * L
    hr = np.empty((F, (arr[0].H - 1) * arr[0].stride + 1, (arr[0].W - 1) * arr[0].stride + 1), np.float32)
    ann_cap = ann_cap or 4096
    seeds_ = np.empty((seed_cap, 5), np.float32)
    n_seeds = C.c_long(0)
    fwd = np.zeros((L, 9, caf_cap), np.float32)
    bwd = np.zeros((L, 9, caf_cap), np.float32)
    nf = np.zeros(L, np.int32)
    nb = np.zeros(L, np.int32)
    data = np.zeros((ann_cap, F, 3), np.float32)
    js = np.zeros((ann_cap, F), np.float32)
    ndec = np.zeros(ann_cap, np.int32)
    dec = np.zeros((ann_cap, F, 2), np.int32)
    nfront = np.zeros(ann_cap, np.int32)
    front = np.zeros((ann_cap, 4 * L, 2), np.int32)
    pre_n = C.c_int(0)
    pre_data = np.zeros((ann_cap, F, 3), np.float32)
    pre_js = np.zeros((ann_cap, F), np.float32)
    tie = C.c_int(0)
    cs = None
    if confidence_scales is not None:
        cs = _f32(confidence_scales)
    init_n, init_data, init_js = 0, None, None
    if initial is not None:          # (data (n, F, 3), joint_scales (n, F))
        init_data, init_js = _f32(initial[0]), _f32(initial[1])
        init_n = len(init_data)
    vp = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
    n = lib().orc_decode_ms(arr, len(sets), int(bool(hflip_pairs)), F, L, vp(skel),
                            vp(cs) if cs is not None else None, C.byref(cfg),
                            vp(hr), vp(seeds_), C.c_long(seed_cap), C.byref(n_seeds),
                            vp(fwd), vp(nf), vp(bwd), vp(nb), C.c_long(caf_cap),
                            ann_cap, vp(data), vp(js), vp(ndec), vp(dec), vp(nfront), vp(front),
                            C.byref(pre_n), vp(pre_data), vp(pre_js), C.byref(tie),
                            init_n, vp(init_data) if init_n else None, vp(init_js) if init_n else None)
    if n < 0:
        raise RuntimeError('oracle decode failed with code %d' % n)
    out = Decoded()
    out.hr = hr
    out.seeds = seeds_[:n_seeds.value].copy()
    out.forward = [fwd[l, :, :nf[l]].copy() for l in range(L)]
    out.backward = [bwd[l, :, :nb[l]].copy() for l in range(L)]
    out.data = data[:n].copy()
    out.joint_scales = js[:n].copy()
    out.decoding_order = [dec[a, :ndec[a]].copy() for a in range(n)]
    out.frontier_order = [front[a, :min(nfront[a], 4 * L)].copy() for a in range(n)]
    out.pre_data = pre_data[:pre_n.value].copy()
    out.pre_joint_scales = pre_js[:pre_n.value].copy()
    out.tie = bool(tie.value)
    return out


def annotation_score(data):
    data = _f32(data)
    return lib().orc_annotation_score(data.ctypes.data_as(C.c_void_p), data.shape[0])


def detexp(x):
    return lib().orc_detexp(C.c_float(float(x)))


class Detected:
    """Everything the oracle observed for one image of the CifDet path."""
    __slots__ = ('hr', 'seeds', 'pre_field', 'pre_score', 'pre_box', 'field', 'score', 'box')


def cifdet_decode(field, cfg=None, *, ann_cap=4096):
    """CifDet decode of one image: ``field (C, 7, H, W)`` -> ``Detected``."""
    cfg = cfg or DetConfig()
    field = _f32(field)
    Cn, seven, H, W = field.shape
    assert seven == 7
    hr = np.empty((Cn, (H - 1) * cfg.stride + 1, (W - 1) * cfg.stride + 1), np.float32)
    seed_cap = Cn * H * W
    seeds_ = np.empty((seed_cap, 6), np.float32)
    n_seeds, n_pre = C.c_long(0), C.c_int(0)
    pre_field = np.zeros(ann_cap, np.int32)
    pre_score = np.zeros(ann_cap, np.float32)
    pre_box = np.zeros((ann_cap, 4), np.float32)
    det_field = np.zeros(ann_cap, np.int32)
    det_score = np.zeros(ann_cap, np.float32)
    det_box = np.zeros((ann_cap, 4), np.float32)
    vp = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
    n = lib().orc_cifdet_decode(vp(field), Cn, H, W, C.byref(cfg), vp(hr), vp(seeds_),
                                C.c_long(seed_cap), C.byref(n_seeds), ann_cap, vp(pre_field),
                                vp(pre_score), vp(pre_box), C.byref(n_pre), vp(det_field),
                                vp(det_score), vp(det_box))
    if n < 0:
        raise RuntimeError('oracle cifdet decode failed with code %d' % n)
    out = Detected()
    out.hr = hr
    out.seeds = seeds_[:n_seeds.value].copy()
    out.pre_field, out.pre_score = pre_field[:n_pre.value].copy(), pre_score[:n_pre.value].copy()
    out.pre_box = pre_box[:n_pre.value].copy()
    out.field, out.score, out.box = det_field[:n].copy(), det_score[:n].copy(), det_box[:n].copy()
    return out
