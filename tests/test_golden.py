"""The CPU oracle (oracle/cifcaf_oracle.c) against the golden vectors the
reference itself produced (tests/golden/*.npz, oracle/gen_golden.py).

This is the pin of the oracle: hr, seeds and CafScored bit-identical to the
reference; annotation integer outputs identical; coordinates and scores inside
the north_star tolerances (the only arithmetic the oracle cannot restate bit
for bit is numpy's SIMD float32 exp, cifcaf.py:139).
"""
import numpy as np
import pytest

from conftest import GOLDEN_NAMES, HEAD_GOLDEN_NAMES, TOL_HEAD_REL, golden, head_golden, sha256


def pad_order(orders, F):
    out = np.full((len(orders), F, 2), -1, np.int32)
    for i, o in enumerate(orders):
        out[i, :len(o)] = o
    return out


@pytest.mark.parametrize('name', GOLDEN_NAMES)
@pytest.mark.parametrize('exp_mode', [0, 1])
def test_oracle_matches_reference_golden(oracle, name, exp_mode):
    g = golden(name)
    z = g.z
    o = oracle.decode(g.cif, g.caf, g.skeleton, oracle.Config(exp_mode=exp_mode, **g.oracle_kwargs()))
    assert not o.tie
    # CifHr: bit-exact (functional.pyx:105-141 restated operation for operation)
    assert o.hr.shape == tuple(z['hr_shape'])
    assert sha256(o.hr) == str(z['hr_sha256'])
    # seeds: bit-exact values and order (cif_seeds.py:23-54)
    assert np.array_equal(o.seeds, z['seeds'])
    # CafScored: bit-exact, order preserved (caf_scored.py:32-87)
    assert [f.shape[1] for f in o.forward] == z['caf_nf'].tolist()
    assert [b.shape[1] for b in o.backward] == z['caf_nb'].tolist()
    assert [sha256(f) for f in o.forward] == z['caf_f_sha256'].tolist()
    assert [sha256(b) for b in o.backward] == z['caf_b_sha256'].tolist()
    if 'caf_f' in z.files:
        assert np.array_equal(np.concatenate(o.forward, axis=1), z['caf_f'])
        assert np.array_equal(np.concatenate(o.backward, axis=1), z['caf_b'])
    # annotations
    F = g.cif.shape[0]
    g.check_annotations(o.data, o.joint_scales, pad_order(o.decoding_order, F), key='ann')
    if not g.overrides.get('force_complete'):
        # the oracle exports its pre list before force-complete; the reference's
        # nms=None run includes it
        g.check_annotations(o.pre_data, o.pre_joint_scales, None, key='pre')
    scores = np.array([oracle.annotation_score(d) for d in o.data])
    assert np.abs(scores - z['ann_score']).max(initial=0.0) <= 1e-4


def test_golden_covers_the_configs():
    assert {'c1_641_5p', 'crowd_1281_50p', 'dense_321_4p', 'empty_161', 'mid_force_complete',
            'mid_greedy', 'mid_max'} <= set(GOLDEN_NAMES)


@pytest.mark.parametrize('name', HEAD_GOLDEN_NAMES)
def test_head_oracle_matches_reference_golden(name):
    """oracle/head_oracle.py against the reference's CompositeFieldFused (eval) + CifCafCollector."""
    from oracle.head_oracle import head_epilogue
    x_cif, x_caf, cif, caf = head_golden(name)
    for x, want, nf, nv in ((x_cif, cif, 17, 1), (x_caf, caf, 19, 2)):
        got = head_epilogue(x, nf, nv, quad=1)
        assert got.shape == want.shape
        exact = [c for c in range(want.shape[2]) if c % 4 in (1, 2, 3)]      # x, y, logb rows: adds / copies
        assert np.array_equal(got[:, :, exact], want[:, :, exact])
        np.testing.assert_allclose(got, want, rtol=TOL_HEAD_REL, atol=0)


@pytest.mark.parametrize('kind', ['plain', 'scaled', 'flipped', 'rotated'])
def test_inverse_oracle_matches_reference_golden(kind):
    """oracle/inverse_oracle.py against the reference's Preprocess.annotations_inverse outputs."""
    from conftest import GOLDEN_DIR, HorizontalSwap, inverse_inputs
    from oracle.gen_golden_inverse import make_meta
    from oracle.inverse_oracle import annotations_inverse
    import os
    g = np.load(os.path.join(GOLDEN_DIR, 'inverse_mid.npz'))
    data, js, _, dxy, target = inverse_inputs()
    meta = make_meta(kind, HorizontalSwap(target))
    got = annotations_inverse(data, js, dxy, meta, target if meta['hflip'] else None)
    assert np.array_equal(got[0], g[kind + '_data'])
    assert np.array_equal(got[1], g[kind + '_scales'])
    assert np.array_equal(got[2], g[kind + '_dxy'])
    from oracle.gen_golden_inverse import detection_boxes
    from oracle.inverse_oracle import detections_inverse
    assert np.array_equal(detections_inverse(detection_boxes(), meta), g[kind + '_detbox'])


from conftest import DET_GOLDEN_NAMES, det_golden  # noqa: E402


@pytest.mark.parametrize('name', DET_GOLDEN_NAMES)
def test_det_oracle_matches_reference_golden(oracle, name):
    """orc_cifdet_decode against the reference's CifDet outputs: everything bit-identical."""
    field, g = det_golden(name)
    o = oracle.cifdet_decode(field, oracle.DetConfig(cif_min_scale=float(g['cif_min_scale'])))
    assert sha256(o.hr) == str(g['hr_sha256'])
    assert np.array_equal(o.seeds, g['seeds'])
    assert np.array_equal(o.field, g['det_field'])
    assert np.array_equal(o.score, g['det_score'])
    assert np.array_equal(o.box, g['det_box'])


def test_det_golden_covers_the_cases():
    assert len(DET_GOLDEN_NAMES) >= 5


from conftest import MS_GOLDEN_NAMES, ms_golden  # noqa: E402


@pytest.mark.parametrize('name', MS_GOLDEN_NAMES)
def test_multiscale_oracle_matches_reference_golden(oracle, coco_skeleton, name):
    """orc_decode_ms under a multi-scale FieldConfig: hr (np.maximum merge, hflip pairs), the seed
    list and the concatenated CafScored lists bit-identical; annotations inside the tolerances."""
    from conftest import TOL_SCORE, TOL_XY
    sets, g = ms_golden(name)
    o = oracle.decode_ms([(c, a, st, ms, mind, maxd or 0.0) for c, a, st, ms, mind, maxd in sets],
                         coco_skeleton, oracle.Config(), hflip_pairs=len(sets) == 10)
    assert sha256(o.hr) == str(g['hr_sha256'])
    assert np.array_equal(o.seeds, g['seeds'])
    assert [f.shape[1] for f in o.forward] == g['caf_nf'].tolist()
    assert [sha256(f) for f in o.forward] == g['caf_f_sha256'].tolist()
    assert [sha256(b) for b in o.backward] == g['caf_b_sha256'].tolist()
    assert len(o.data) == len(g['ann_data']) >= 3
    assert np.array_equal(o.data[:, :, 2] > 0, g['ann_data'][:, :, 2] > 0)
    assert np.abs(o.data[:, :, :2] - g['ann_data'][:, :, :2]).max() <= TOL_XY
    assert np.abs(o.data[:, :, 2] - g['ann_data'][:, :, 2]).max() <= TOL_SCORE
