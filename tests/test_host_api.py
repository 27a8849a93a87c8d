"""Host-side mirror of the reference interface: flag surface, class-attribute
configuration, FieldConfig and Annotation behave like the reference's
(decoder/factory.py, decoder/field_config.py, annotation.py)."""
import argparse
import types

import numpy as np
import pytest


def _ref():
    from oracle import build_ref
    if not build_ref.staged():
        pytest.skip('oracle/_ref not staged')
    return build_ref.import_reference()


def parse(decoder, argv, **cli_kw):
    parser = argparse.ArgumentParser()
    decoder.cli(parser, **cli_kw)
    return parser.parse_args(argv)


@pytest.mark.parametrize('argv,cli_kw', [
    ([], {}),
    (['--seed-threshold', '0.5', '--no-force-complete-pose'], {}),
    (['--force-complete-pose', '--greedy', '--connection-method', 'max'],
     {'force_complete_pose': False}),
    (['--instance-threshold', '0.2', '--cif-th', '0.3', '--caf-th', '0.05', '--dense-connections'],
     {'force_complete_pose': False, 'seed_threshold': 0.3}),
])
def test_cli_and_configure_match_reference(configured_decoder, argv, cli_kw):
    ref = _ref()
    mine = configured_decoder
    a_ref = parse(ref.decoder, argv, **cli_kw)
    a_mine = parse(mine, argv, **cli_kw)
    assert vars(a_ref) == vars(a_mine)
    saved = (ref.decoder.CifSeeds.threshold, ref.decoder.CifCaf.force_complete,
             ref.decoder.CifCaf.keypoint_threshold, ref.decoder.CifCaf.greedy,
             ref.decoder.CifCaf.connection_method, ref.decoder.CifHr.v_threshold,
             ref.decoder.CafScored.default_score_th, ref.decoder.nms.Keypoints.instance_threshold,
             ref.decoder.nms.Keypoints.keypoint_threshold)
    try:
        ref.decoder.configure(a_ref)
        mine.configure(a_mine)
        for cls, attr in (('CifHr', 'v_threshold'), ('CifSeeds', 'threshold'),
                          ('CafScored', 'default_score_th'), ('CifCaf', 'force_complete'),
                          ('CifCaf', 'keypoint_threshold'), ('CifCaf', 'greedy'),
                          ('CifCaf', 'connection_method')):
            assert getattr(getattr(mine, cls), attr) == getattr(getattr(ref.decoder, cls), attr), (cls, attr)
        assert mine.nms.Keypoints.instance_threshold == ref.decoder.nms.Keypoints.instance_threshold
        assert mine.nms.Keypoints.keypoint_threshold == ref.decoder.nms.Keypoints.keypoint_threshold
    finally:
        (ref.decoder.CifSeeds.threshold, ref.decoder.CifCaf.force_complete,
         ref.decoder.CifCaf.keypoint_threshold, ref.decoder.CifCaf.greedy,
         ref.decoder.CifCaf.connection_method, ref.decoder.CifHr.v_threshold,
         ref.decoder.CafScored.default_score_th, ref.decoder.nms.Keypoints.instance_threshold,
         ref.decoder.nms.Keypoints.keypoint_threshold) = saved


def test_field_config_defaults_match_reference():
    ref = _ref()
    from openpifpaf_b200 import decoder
    a, b = ref.decoder.FieldConfig(), decoder.FieldConfig()
    for name in ('cif_indices', 'caf_indices', 'cif_strides', 'caf_strides', 'cif_min_scales',
                 'caf_min_distances', 'caf_max_distances', 'seed_mask', 'confidence_scales'):
        assert getattr(a, name) == getattr(b, name), name


def test_annotation_matches_reference():
    ref = _ref()
    from openpifpaf_b200.annotation import Annotation
    from openpifpaf_b200.datasets.constants import COCO_KEYPOINTS, COCO_PERSON_SKELETON
    rng = np.random.RandomState(0)
    for _ in range(10):
        data = np.zeros((17, 3), np.float32)
        data[:, :2] = rng.uniform(0, 600, (17, 2))
        data[:, 2] = rng.uniform(0, 1, 17) * (rng.uniform(0, 1, 17) > 0.3)
        js = rng.uniform(1, 20, 17).astype(np.float32)
        a = ref.Annotation(COCO_KEYPOINTS, COCO_PERSON_SKELETON).set(data.copy(), js.copy())
        b = Annotation(COCO_KEYPOINTS, COCO_PERSON_SKELETON).set(data.copy(), js.copy())
        assert a.score() == b.score()
        assert a.scale() == b.scale()
        assert np.array_equal(np.asarray(a.bbox()), np.asarray(b.bbox()))
        assert a.json_data() == b.json_data()
        a.rescale(0.5)
        b.rescale(0.5)
        assert np.array_equal(a.data, b.data) and np.array_equal(a.joint_scales, b.joint_scales)


def test_factory_decode_builds_cifcaf(configured_decoder):
    decoder = configured_decoder
    from openpifpaf_b200.datasets.constants import (COCO_KEYPOINTS, COCO_PERSON_SKELETON,
                                                    DENSER_COCO_PERSON_CONNECTIONS)

    def head(name, **meta):
        return types.SimpleNamespace(meta=types.SimpleNamespace(name=name, **meta),
                                     stride=lambda basenet_stride: basenet_stride // 2)
    before = list(COCO_PERSON_SKELETON)
    heads = [head('cif', keypoints=COCO_KEYPOINTS),
             head('caf', keypoints=COCO_KEYPOINTS, skeleton=COCO_PERSON_SKELETON),
             head('caf25', keypoints=COCO_KEYPOINTS, skeleton=DENSER_COCO_PERSON_CONNECTIONS)]
    dec = decoder.factory_decode(heads[:2], basenet_stride=16)
    assert isinstance(dec, decoder.CifCaf)
    assert dec.field_config.cif_strides == [8] and dec.field_config.caf_strides == [8]
    assert list(dec.skeleton) == before
    dense = decoder.factory_decode(heads, basenet_stride=16, dense_connections=True,
                                   dense_coupling=0.01)
    assert len(dense.skeleton) == 44 and dense.confidence_scales is None
    assert dense.field_config.confidence_scales == [1.0] * 19 + [0.01] * 25
    assert list(COCO_PERSON_SKELETON) == before, 'global skeleton must not be mutated'
    with pytest.raises(AssertionError):
        decoder.factory_decode(heads[:2], basenet_stride=16, caf_seeds=True)
    with pytest.raises(Exception):
        decoder.factory_decode([head('foo')], basenet_stride=16)


def test_unconfigured_seed_threshold_raises(configured_decoder):
    decoder = configured_decoder
    from openpifpaf_b200.datasets.constants import COCO_KEYPOINTS, COCO_PERSON_SKELETON
    decoder.CifSeeds.threshold = None
    dec = decoder.CifCaf(decoder.FieldConfig(), keypoints=COCO_KEYPOINTS,
                         skeleton=COCO_PERSON_SKELETON)
    with pytest.raises(TypeError):
        dec.params()


def test_json_data_batch_equals_per_annotation():
    """json_data_batch == [a.json_data() ...] exactly (keypoints, bbox, score), including empty
    annotations, fixed scores and ids."""
    from openpifpaf_b200.annotation import Annotation, json_data_batch
    from openpifpaf_b200.datasets.constants import COCO_KEYPOINTS, COCO_PERSON_SKELETON
    rng = np.random.RandomState(3)
    anns = []
    for i in range(40):
        a = Annotation(COCO_KEYPOINTS, COCO_PERSON_SKELETON)
        data = rng.uniform(0, 640, (17, 3)).astype(np.float32)
        data[:, 2] = rng.uniform(0, 1, 17).astype(np.float32) * (rng.uniform(size=17) > 0.3)
        if i == 5:
            data[:, 2] = 0.0
        a.set(data, rng.uniform(0.5, 9, 17).astype(np.float32))
        if i == 7:
            a.fixed_score = 0.25
        if i == 9:
            a.id_ = 4
        anns.append(a)
    assert json_data_batch(anns) == [a.json_data() for a in anns]
    assert json_data_batch([]) == []


def test_pack_metas_and_hflip_target_layout():
    """Host packing of the reference's meta dicts into pp_image_meta records (no GPU needed)."""
    import ctypes as C
    import math
    from conftest import HorizontalSwap
    from openpifpaf_b200 import _lib
    from openpifpaf_b200.datasets.constants import COCO_KEYPOINTS, HFLIP
    from openpifpaf_b200.transforms.preprocess import hflip_target, pack_metas
    target = np.array([COCO_KEYPOINTS.index(HFLIP.get(k, k)) for k in COCO_KEYPOINTS], np.int32)
    metas = [
        {'offset': np.array((0.0, 0.0)), 'scale': np.array((1.0, 1.0)), 'hflip': False,
         'width_height': np.array((640, 480)), 'rotation': {'angle': 0.0, 'width': None, 'height': None}},
        {'offset': np.array((-12.0, 31.5)), 'scale': np.array((0.8, 0.7)), 'hflip': True,
         'width_height': np.array((640, 480)), 'horizontal_swap': HorizontalSwap(target),
         'rotation': {'angle': 30.0, 'width': 513, 'height': 385}},
    ]
    raw = pack_metas(metas, 'cpu').numpy()
    assert raw.shape == (2, C.sizeof(_lib.ImageMeta))
    recs = [_lib.ImageMeta.from_buffer_copy(raw[i].tobytes()) for i in range(2)]
    assert (recs[0].rotate, recs[0].hflip, recs[0].scale_x, recs[0].offset_y) == (0, 0, 1.0, 0.0)
    r = recs[1]
    assert (r.rotate, r.hflip, r.flip_w1) == (1, 1, 639.0)
    assert (r.offset_x, r.offset_y, r.scale_x, r.scale_y) == (-12.0, 31.5, 0.8, 0.7)
    assert r.rot_cos == np.float32(math.cos(-30.0 / 180.0 * math.pi))       # cos / sin of MINUS the angle
    assert r.rot_sin == np.float32(math.sin(-30.0 / 180.0 * math.pi))
    assert (r.rot_cx, r.rot_cy) == (256.0, 192.0)
    assert hflip_target(metas[:1], 17, 'cpu') is None
    assert np.array_equal(hflip_target(metas, 17, 'cpu').numpy(), target)


def test_fused_head_epilogue_shapes_and_errors():
    """FusedHeadEpilogue validates shapes on the host and refuses CPU tensors (no CPU path)."""
    import torch
    from openpifpaf_b200 import _lib
    from openpifpaf_b200.network import FusedHeadEpilogue, index_field_torch
    ep = FusedHeadEpilogue(17, 1)
    x = torch.zeros(2, 17 * 5 * 4, 6, 9)
    assert ep.out_shape(x) == (2, 17, 5, 11, 17)
    assert FusedHeadEpilogue(19, 2, quad=0).out_shape(torch.zeros(1, 19 * 9, 6, 9)) == (1, 19, 9, 6, 9)
    with pytest.raises(ValueError):
        ep.out_shape(torch.zeros(2, 17 * 5, 6, 9))
    with pytest.raises(_lib.PifPafError):
        ep(x)
    with pytest.raises(NotImplementedError):
        FusedHeadEpilogue(17, 1, quad=2)
    idx = index_field_torch((3, 4))
    assert idx.shape == (2, 3, 4) and idx[0, 1, 3] == 3 and idx[1, 2, 0] == 2     # channel 0 = x, 1 = y


def test_factory_builds_cifdet_and_multiscale_config():
    """factory_decode: a DetectionMeta head gives a CifDet; multi_scale keeps the reference's
    index arithmetic (and therefore its IndexError for a plain three-head network)."""
    from openpifpaf_b200 import decoder

    class Head:
        def __init__(self, meta):
            self.meta = meta

        @staticmethod
        def stride(basenet_stride):
            return basenet_stride // 2

    class DetMeta:
        name, categories = 'cifdet', ['person', 'car']

    class CifMeta:
        name, keypoints = 'cif', ['a', 'b', 'c']

    class CafMeta:
        name, keypoints, skeleton = 'caf', ['a', 'b', 'c'], [(1, 2), (2, 3)]

    det = decoder.factory_decode([Head(DetMeta)], basenet_stride=16)
    assert isinstance(det, decoder.CifDet) and det.categories == ['person', 'car']
    heads = [Head(CifMeta), Head(CafMeta), Head(CafMeta)]
    with pytest.raises(IndexError):
        decoder.factory_decode(heads, basenet_stride=16, multi_scale=True)
    many = [Head(CifMeta), Head(CafMeta), Head(CafMeta)] * 10
    dec = decoder.factory_decode(many, basenet_stride=16, multi_scale=True)
    cfg = dec.field_config
    assert cfg.cif_indices == [3 * v for v in range(10)] and cfg.caf_indices == [3 * v + 1 for v in range(10)]
    assert cfg.cif_min_scales == [0.0, 12.0, 16.0, 24.0, 40.0] * 2
    assert cfg.caf_max_distances == [160.0, 240.0, 320.0, 480.0, None] * 2
    cfg.verify()


def test_wrap_batch_equals_from_buffers():
    """Annotation.wrap_batch (one gather per result array, lazy score_weights / decoding order) builds
    the same objects as the per-annotation constructor."""
    import copy
    import pickle
    from openpifpaf_b200.annotation import Annotation
    from openpifpaf_b200.datasets.constants import COCO_KEYPOINTS, COCO_PERSON_SKELETON
    rng = np.random.RandomState(0)
    n, F = 6, len(COCO_KEYPOINTS)
    data = rng.rand(n, F, 3).astype(np.float32)
    data[:, ::3, 2] = 0.0
    scales = rng.rand(n, F).astype(np.float32)
    pre = rng.rand(n, F, 3).astype(np.float32)
    order = np.full((n, F, 2), -1, np.int32)
    order[:, 0] = (5, 7)
    order[:, 1] = (7, 9)
    graph = {j: [] for j in range(F)}
    graph[5], graph[7] = [7], [9]
    anns = Annotation.wrap_batch(COCO_KEYPOINTS, COCO_PERSON_SKELETON, data.copy(), scales.copy(), pre, order,
                                 graph)
    assert len(anns) == n
    for k, ann in enumerate(anns):
        want = Annotation.from_buffers(COCO_KEYPOINTS, COCO_PERSON_SKELETON, data[k], scales[k],
                                       pre_data=pre[k], order=order[k], graph=graph)
        assert np.array_equal(ann.data, want.data) and np.array_equal(ann.joint_scales, want.joint_scales)
        assert ann.score() == want.score()
        assert np.array_equal(ann.score_weights, want.score_weights)
        assert [(s, e) for s, e, _, _ in ann.decoding_order] == [(5, 7), (7, 9)]
        for (_, _, a1, a2), (_, _, b1, b2) in zip(ann.decoding_order, want.decoding_order):
            assert np.array_equal(a1, b1) and np.array_equal(a2, b2)
        assert ann.frontier_order == want.frontier_order
        assert ann.json_data() == want.json_data()
    # rows are private to their annotation; copies and pickles keep working before the lazy parts exist
    fresh = Annotation.wrap_batch(COCO_KEYPOINTS, COCO_PERSON_SKELETON, data.copy(), scales.copy(), pre, order,
                                  graph)
    fresh[0].data[0, 0] = -1.0
    assert fresh[1].data[0, 0] != -1.0
    clone = pickle.loads(pickle.dumps(fresh[2]))
    assert clone.score() == fresh[2].score() and len(clone.decoding_order) == 2
    assert copy.deepcopy(fresh[3]).score() == fresh[3].score()
    with pytest.raises(AttributeError):
        fresh[4].no_such_attribute  # pylint: disable=pointless-statement
