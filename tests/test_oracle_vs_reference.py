"""The C restatement against the RUNNING reference (oracle/_ref, staged from
/root/reference by oracle/build_ref.py).  Skipped where the staged copy is
absent.  Complements tests/test_golden.py with inputs that are not committed:
heavier jitter (partial poses), per-primitive known answers for the five live
functional.pyx functions, and the reference's edge cases."""
import numpy as np
import pytest

from conftest import TOL_SCORE, TOL_XY, fields

pytestmark = pytest.mark.reference


@pytest.fixture(scope='module')
def ref():
    from oracle import build_ref
    if not build_ref.staged():
        pytest.skip('oracle/_ref not staged (run python oracle/build_ref.py in the build container)')
    return build_ref.import_reference()


def configure(dec):
    dec.CifHr.v_threshold = 0.1
    dec.CifSeeds.threshold = 0.2
    dec.CafScored.default_score_th = 0.1
    dec.CifCaf.connection_method = 'blend'
    dec.CifCaf.greedy = False
    dec.CifCaf.force_complete = False
    dec.CifCaf.keypoint_threshold = 0.001
    dec.nms.Keypoints.keypoint_threshold = 0.001
    dec.nms.Keypoints.instance_threshold = 0.0


def heavier_jitter(cif, caf, seed):
    """Regression noise sigma 0.3 cells and +-30 % scale noise on top of the generator."""
    rng = np.random.RandomState(seed)
    cif, caf = cif.copy(), caf.copy()
    cif[:, 1:3] += rng.normal(0, 0.25, cif[:, 1:3].shape).astype(np.float32) * (cif[:, 0:1] > 0.1)
    cif[:, 4] *= rng.uniform(0.7, 1.3, cif[:, 4].shape).astype(np.float32)
    for rows in ((1, 2), (5, 6)):
        caf[:, rows] += rng.normal(0, 0.25, caf[:, rows].shape).astype(np.float32) * (caf[:, 0:1] > 0.1)
    return cif, caf


@pytest.mark.parametrize('hw,n,seed', [((321, 321), 6, 21), ((481, 321), 10, 22), ((161, 241), 3, 23)])
def test_full_decode_heavy_jitter(ref, oracle, coco_skeleton, hw, n, seed):
    from openpifpaf_b200.datasets.constants import COCO_KEYPOINTS
    dec = ref.decoder
    configure(dec)
    cif, caf, _ = fields(hw, n, pose_seed=seed, jitter_seed=seed + 1)
    cif, caf = heavier_jitter(cif, caf, seed)
    d = dec.CifCaf(dec.FieldConfig(), keypoints=COCO_KEYPOINTS,
                   skeleton=[list(l) for l in coco_skeleton])
    try:
        anns = d([cif.copy(), caf.copy()])
    except TypeError:
        pytest.skip('the reference raises on an exact frontier tie for this input (appendix B)')
    o = oracle.decode(cif, caf, coco_skeleton, oracle.Config(exp_mode=1))
    hr = dec.CifHr(dec.FieldConfig()).fill([cif, caf]).accumulated
    assert np.array_equal(hr, o.hr)
    assert len(anns) == len(o.data)
    for a, data, js, order in zip(anns, o.data, o.joint_scales, o.decoding_order):
        assert np.array_equal(a.data[:, 2] != 0, data[:, 2] != 0)
        assert np.abs(a.data[:, :2] - data[:, :2]).max() <= TOL_XY
        assert np.abs(a.data[:, 2] - data[:, 2]).max() <= TOL_SCORE
        assert np.abs(a.joint_scales - js).max() <= TOL_XY
        assert [(s, e) for s, e, _, _ in a.decoding_order] == [tuple(x) for x in order.tolist()]
        assert abs(a.score() - oracle.annotation_score(data)) <= TOL_SCORE


def test_primitive_scalar_square_add_gauss_with_max(ref, oracle):
    import ctypes as C
    rng = np.random.RandomState(3)
    n = 400
    x = rng.uniform(-10, 110, n).astype(np.float32)
    y = rng.uniform(-10, 90, n).astype(np.float32)
    s = rng.uniform(0.3, 9, n).astype(np.float32)
    v = rng.uniform(0.01, 0.2, n).astype(np.float32)
    # appendix B edge cases: NaN coordinate, infinite sigma, far out of bounds
    x[5], s[9], x[11] = np.nan, np.inf, 1e9
    a = np.zeros((80, 100), np.float32)
    b = a.copy()
    ref.functional.scalar_square_add_gauss_with_max(a, x, y, s, v, truncate=1.0)
    vp = lambda t: t.ctypes.data_as(C.c_void_p)  # noqa: E731
    oracle.lib().orc_scalar_square_add_gauss_with_max(vp(b), 80, 100, vp(x), vp(y), vp(s), vp(v),
                                                      C.c_long(n), C.c_float(1.0), C.c_float(1.0))
    assert np.array_equal(a, b)


def test_primitive_scalar_values_and_caf_center(ref, oracle):
    import ctypes as C
    rng = np.random.RandomState(4)
    field = rng.uniform(0, 1, (50, 70)).astype(np.float32)
    x = rng.uniform(-5, 75, 300).astype(np.float32)
    y = rng.uniform(-5, 55, 300).astype(np.float32)
    want = ref.functional.scalar_values(field, x, y, default=-1.0)
    got = np.empty(300, np.float32)
    vp = lambda t: t.ctypes.data_as(C.c_void_p)  # noqa: E731
    oracle.lib().orc_scalar_values(vp(field), 50, 70, vp(x), vp(y), C.c_long(300), C.c_float(-1.0),
                                   vp(got))
    assert np.array_equal(want, got)
    caf = rng.uniform(0, 60, (9, 500)).astype(np.float32)
    want = ref.functional.caf_center_s(caf, 30.0, 25.0, 8.0)
    out = np.zeros((9, 500), np.float32)
    k = oracle.lib().orc_caf_center_s(vp(caf), C.c_long(500), C.c_float(30.0), C.c_float(25.0),
                                      C.c_float(8.0), vp(out))
    assert k == want.shape[1] and np.array_equal(out[:, :k], want)


def test_primitive_occupancy(ref, oracle):
    import ctypes as C
    rng = np.random.RandomState(5)
    occ_ref = ref.decoder.Occupancy((1, 120, 160), 2, min_scale=4)
    mine = np.zeros((60, 80), np.uint8)
    vp = lambda t: t.ctypes.data_as(C.c_void_p)  # noqa: E731
    for _ in range(300):
        x, y, s = (np.float32(rng.uniform(-20, 180)), np.float32(rng.uniform(-20, 140)),
                   np.float32(rng.uniform(0, 30)))
        occ_ref.set(0, x, y, s)
        oracle.lib().orc_occupancy_set(vp(mine), C.c_long(60), C.c_long(80), C.c_float(x),
                                       C.c_float(y), C.c_float(s))
    assert np.array_equal(occ_ref.occupancy[0], mine)
    for _ in range(300):
        x, y = np.float32(rng.uniform(-20, 180)), np.float32(rng.uniform(-20, 140))
        got = oracle.lib().orc_occupancy_get(vp(mine), C.c_long(60), C.c_long(80), C.c_float(x),
                                             C.c_float(y), C.c_float(2.0))
        assert bool(got) == bool(occ_ref.get(0, x, y))
    # 256 sets wrap the uint8 counter back to "free" (appendix B)
    m2 = np.zeros((60, 80), np.uint8)
    for _ in range(256):
        oracle.lib().orc_occupancy_set(vp(m2), C.c_long(60), C.c_long(80), C.c_float(40.0),
                                       C.c_float(40.0), C.c_float(4.0))
    assert m2.max() == 0


@pytest.mark.parametrize('seed', [0, 1, 2])
def test_random_init_like_fields(ref, oracle, coco_skeleton, seed):
    """Config-2 statistics (heavy-tailed exp() scales: whole-map and out-of-bounds sources,
    hundreds of overlapping seeds): the restatement still tracks the reference."""
    import warnings
    from openpifpaf_b200.datasets.constants import COCO_KEYPOINTS
    from openpifpaf_b200.encoder import random_init_like_fields
    dec = ref.decoder
    configure(dec)
    cif, caf = random_init_like_fields((13, 17), seed)
    cfg = dec.FieldConfig()
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        hr = dec.CifHr(cfg).fill([cif.copy(), caf.copy()]).accumulated
        seeds = dec.CifSeeds(hr, cfg).fill([cif, caf]).get()
        d = dec.CifCaf(cfg, keypoints=COCO_KEYPOINTS, skeleton=[list(l) for l in coco_skeleton])
        try:
            anns = d([cif.copy(), caf.copy()])
        except TypeError:
            pytest.skip('the reference raises on an exact frontier tie for this input')
    o = oracle.decode(cif, caf, coco_skeleton, oracle.Config(exp_mode=1))
    assert np.array_equal(hr, o.hr)
    assert np.array_equal(np.array([list(s) for s in seeds], np.float32).reshape(-1, 5), o.seeds)
    assert len(anns) == len(o.data) >= 20
    for a, data, js in zip(anns, o.data, o.joint_scales):
        assert np.array_equal(a.data[:, 2] != 0, data[:, 2] != 0)
        assert np.abs(a.data[:, :2] - data[:, :2]).max() <= TOL_XY
        assert np.abs(a.data[:, 2] - data[:, 2]).max() <= TOL_SCORE


def test_nonfinite_sources_stage_outputs(ref, oracle, coco_skeleton):
    """NaN / inf / 1e30 coordinates and scales (appendix B): CifHr, seeds and CafScored of
    the restatement equal the reference's.  (Assembly on such inputs is not pinned: the
    reference itself raises ValueError / OverflowError on part of them.)"""
    import warnings
    from openpifpaf_b200.encoder import random_init_like_fields
    dec = ref.decoder
    configure(dec)
    for seed in range(4):
        cif, caf = random_init_like_fields((11, 14), seed, n_nonfinite=5)
        cfg = dec.FieldConfig()
        with warnings.catch_warnings():
            warnings.simplefilter('ignore')
            hr = dec.CifHr(cfg).fill([cif.copy(), caf.copy()]).accumulated
            seeds = dec.CifSeeds(hr, cfg).fill([cif, caf]).get()
            sc = dec.CafScored(hr, cfg, [list(l) for l in coco_skeleton]).fill([cif, caf])
        o_hr = oracle.cifhr(cif)
        assert np.array_equal(hr, o_hr, equal_nan=True)
        o_seeds = oracle.seeds(cif, o_hr)
        assert np.array_equal(np.array([list(s) for s in seeds], np.float32).reshape(-1, 5), o_seeds,
                              equal_nan=True)
        f, b = oracle.caf_scored(caf, o_hr, coco_skeleton)
        for l in range(len(coco_skeleton)):
            assert np.array_equal(sc.forward[l], f[l], equal_nan=True)
            assert np.array_equal(sc.backward[l], b[l], equal_nan=True)


def test_initial_annotations(ref, oracle, coco_skeleton):
    """cifcaf.py:67-71,95-98: initial annotations are grown and marked occupied before the seeds."""
    from openpifpaf_b200.datasets.constants import COCO_KEYPOINTS
    dec = ref.decoder
    configure(dec)
    cif, caf, poses = fields((321, 481), 6, pose_seed=3, jitter_seed=4)
    poses = np.asarray(poses)
    init_data = np.zeros((2, 17, 3), np.float32)
    init_js = np.zeros((2, 17), np.float32)
    init_data[0, 0] = (poses[0][0][0], poses[0][0][1], 0.9)
    init_js[0, 0] = 4.0
    for j in (5, 6, 11):
        init_data[1, j] = (poses[2][j][0], poses[2][j][1], 0.8)
        init_js[1, j] = 6.0
    initial = [ref.Annotation(COCO_KEYPOINTS, coco_skeleton).set(init_data[i].copy(), init_js[i].copy())
               for i in range(2)]
    d = dec.CifCaf(dec.FieldConfig(), keypoints=COCO_KEYPOINTS, skeleton=[list(l) for l in coco_skeleton])
    anns = d([cif.copy(), caf.copy()], initial_annotations=initial)
    o = oracle.decode(cif, caf, coco_skeleton, oracle.Config(exp_mode=1), initial=(init_data, init_js))
    assert len(anns) == len(o.data)
    for a, data in zip(anns, o.data):
        assert np.array_equal(a.data[:, 2] != 0, data[:, 2] != 0)
        assert np.abs(a.data[:, :2] - data[:, :2]).max() <= TOL_XY
        assert np.abs(a.data[:, 2] - data[:, 2]).max() <= TOL_SCORE


@pytest.mark.parametrize('batch,h,w,seed', [(1, 4, 6, 31), (3, 11, 9, 32)])
def test_head_epilogue(ref, batch, h, w, seed):
    """oracle/head_oracle.py == CompositeFieldFused.forward (eval) + CifCafCollector.forward
    (network/heads.py:406-448, 65-88) on seeded raw convolution outputs, and the decoder accepts
    the collected layout."""
    from oracle.gen_golden_heads import raw_head_outputs, reference_fields
    from oracle.head_oracle import head_epilogue
    import openpifpaf.network.heads  # noqa: F401
    x_cif, x_caf = raw_head_outputs(batch, h, w, seed)
    cif, caf = reference_fields(ref, x_cif, x_caf)
    for x, want, nf, nv in ((x_cif, cif, 17, 1), (x_caf, caf, 19, 2)):
        got = head_epilogue(x, nf, nv, quad=1)
        assert got.shape == want.shape == (batch, nf, 1 + 4 * nv, 2 * h - 1, 2 * w - 1)
        np.testing.assert_allclose(got, want, rtol=2e-6, atol=0)


def _metas(ref):
    """Meta dicts as the reference's transforms build them (transforms/*.py)."""
    from openpifpaf_b200.datasets.constants import COCO_KEYPOINTS, HFLIP
    import openpifpaf.transforms.hflip as ref_hflip
    swap = ref_hflip._HorizontalSwap(COCO_KEYPOINTS, HFLIP)    # pylint: disable=protected-access
    plain = {'offset': np.array((0.0, 0.0)), 'scale': np.array((1.0, 1.0)), 'hflip': False,
             'width_height': np.array((640, 480)), 'rotation': {'angle': 0.0, 'width': None, 'height': None}}
    scaled = dict(plain, offset=np.array((-12.0, 31.5)), scale=np.array((0.8013, 0.7977)))
    flipped = dict(scaled, hflip=True, horizontal_swap=swap)
    rotated = dict(flipped, rotation={'angle': 17.5, 'width': 513, 'height': 385})
    return {'plain': plain, 'scaled': scaled, 'flipped': flipped, 'rotated': rotated}


@pytest.mark.parametrize('kind', ['plain', 'scaled', 'flipped', 'rotated'])
def test_annotations_inverse(ref, oracle, coco_skeleton, kind):
    """oracle/inverse_oracle.py == Preprocess.annotations_inverse (transforms/preprocess.py:35-82)
    on decoded annotations, bit for bit (data, joint_scales, decoding_order coordinates)."""
    from oracle.inverse_oracle import annotations_inverse
    from openpifpaf_b200.datasets.constants import COCO_KEYPOINTS
    from openpifpaf_b200.transforms.preprocess import hflip_target
    import openpifpaf.transforms  # noqa: F401
    meta = _metas(ref)[kind]
    dec = ref.decoder
    configure(dec)
    cif, caf, _ = fields((321, 481), 5, pose_seed=3, jitter_seed=4)
    anns = dec.CifCaf(dec.FieldConfig(), keypoints=COCO_KEYPOINTS, skeleton=coco_skeleton)([cif, caf])
    assert len(anns) >= 3
    data = np.stack([a.data for a in anns])
    js = np.stack([a.joint_scales for a in anns])
    dxy = np.array([c[:2] for a in anns for _, __, c1, c2 in a.decoding_order for c in (c1, c2)], np.float32)
    target = None
    if meta['hflip']:
        target = hflip_target([meta], 17, 'cpu').numpy()
    got_data, got_js, got_dxy = annotations_inverse(data, js, dxy, meta, target)
    want = ref.transforms.Preprocess.annotations_inverse(anns, meta)
    assert np.array_equal(got_data, np.stack([a.data for a in want]))
    assert np.array_equal(got_js, np.stack([a.joint_scales for a in want]))
    want_dxy = np.array([c[:2] for a in want for _, __, c1, c2 in a.decoding_order for c in (c1, c2)],
                        np.float32)
    assert np.array_equal(got_dxy, want_dxy)


def test_json_data_batch(ref, coco_skeleton):
    """annotation.json_data_batch == the reference's Annotation.json_data per annotation."""
    from openpifpaf_b200.annotation import Annotation, json_data_batch
    from openpifpaf_b200.datasets.constants import COCO_KEYPOINTS
    dec = ref.decoder
    configure(dec)
    cif, caf, _ = fields((321, 481), 6, pose_seed=3, jitter_seed=4)
    anns = dec.CifCaf(dec.FieldConfig(), keypoints=COCO_KEYPOINTS, skeleton=coco_skeleton)([cif, caf])
    ours = [Annotation(COCO_KEYPOINTS, coco_skeleton).set(a.data.copy(), a.joint_scales.copy()) for a in anns]
    assert json_data_batch(ours) == [a.json_data() for a in anns]


@pytest.mark.parametrize('hw,nb,seed,ncat,min_scale', [((241, 321), 6, 31, 4, 0.0), ((481, 481), 25, 32, 8, 0.0),
                                                       ((321, 321), 10, 33, 91, 48.0)])
def test_cifdet_decode(ref, oracle, hw, nb, seed, ncat, min_scale):
    """orc_cifdet_decode == CifDetHr / CifDetSeeds / CifDet / nms.Detection of the running reference
    (cif_hr.py:84-100, cif_seeds.py:67-90, cifdet.py:26-52, nms.py:60-102), bit for bit."""
    from oracle.gen_golden_det import run_reference
    from openpifpaf_b200.encoder import synthetic_det_fields
    field, _ = synthetic_det_fields(hw, nb, box_seed=seed, jitter_seed=seed + 100, n_categories=ncat)
    hr, seeds, anns = run_reference(ref, field, min_scale)
    o = oracle.cifdet_decode(field, oracle.DetConfig(cif_min_scale=min_scale))
    assert np.array_equal(o.hr, hr)
    assert np.array_equal(o.seeds, seeds)
    assert len(anns) == len(o.field) and (nb == 0 or len(anns) >= 2)
    assert np.array_equal(o.field, np.array([a.field_i for a in anns], np.int32))
    assert np.array_equal(o.score, np.array([a.score for a in anns], np.float32))
    assert np.array_equal(o.box, np.array([a.bbox for a in anns], np.float32).reshape(-1, 4))


@pytest.mark.parametrize('kind', ['plain', 'scaled', 'flipped', 'rotated'])
def test_detections_inverse(ref, kind):
    """oracle detections_inverse == Preprocess.annotations_inverse on AnnotationDet objects
    (anndet_inverse, transforms/preprocess.py:84-94), bit for bit."""
    from oracle.inverse_oracle import detections_inverse
    import openpifpaf.transforms  # noqa: F401
    meta = _metas(ref)[kind]
    rng = np.random.RandomState(5)
    boxes = np.concatenate([rng.uniform(0, 400, (12, 2)), rng.uniform(5, 200, (12, 2))], axis=1).astype(np.float32)
    anns = [ref.annotation.AnnotationDet(['a', 'b']).set(i % 2, np.float32(0.5), b.copy()) for i, b in enumerate(boxes)]
    want = ref.transforms.Preprocess.annotations_inverse(anns, meta)
    got = detections_inverse(boxes, meta)
    assert np.array_equal(got, np.stack([np.asarray(a.bbox, np.float32) for a in want]))
    assert all(np.asarray(a.bbox).dtype == np.float32 for a in want)
