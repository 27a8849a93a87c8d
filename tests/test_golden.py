"""The CPU oracle (oracle/cifcaf_oracle.c) against the golden vectors the
reference itself produced (tests/golden/*.npz, oracle/gen_golden.py).

This is the pin of the oracle: hr, seeds and CafScored bit-identical to the
reference; annotation integer outputs identical; coordinates and scores inside
the north_star tolerances (the only arithmetic the oracle cannot restate bit
for bit is numpy's SIMD float32 exp, cifcaf.py:139).
"""
import numpy as np
import pytest

from conftest import GOLDEN_NAMES, golden, sha256


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
