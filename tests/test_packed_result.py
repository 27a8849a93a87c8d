"""Host logic of the packed result block (include/pifpaf_b200.h, "Packed results"): the parser the Python layer
uses (engine.PackedResult) against a block laid out by hand from the header's description, the status bits,
and the Annotation lists built from it.  CPU only: the block the GPU writes is compared with the fixed-capacity
arrays in tests/test_gpu_parity.py::test_host_entry_equals_device_entry."""
import logging

import numpy as np
import pytest


def make_block(n_per_image, F, status=None, pad_records=3):
    nb = len(n_per_image)
    rw = (4 + 9 * F + 1) & ~1
    hdr = (4 + 3 * nb + 1) & ~1
    total = int(sum(n_per_image))
    words = np.zeros(hdr + (total + pad_records) * rw, np.int32)
    words[0:4] = (nb, F, total, rw)
    words[4:4 + nb] = n_per_image
    words[4 + nb:4 + 2 * nb] = status if status is not None else 0
    words[4 + 2 * nb:4 + 3 * nb] = [n + 1 for n in n_per_image]
    rng = np.random.RandomState(7)
    want = []
    for r in range(total):
        rec = words[hdr + r * rw: hdr + (r + 1) * rw]
        score = float(rng.rand())
        data = rng.rand(F, 3).astype(np.float32)
        scales = rng.rand(F).astype(np.float32)
        pre = rng.rand(F, 3).astype(np.float32)
        order = rng.randint(-1, F, size=(F, 2)).astype(np.int32)
        rec[0:2] = np.array([score], np.float64).view(np.int32)
        rec[2] = r % 7
        rec[4:4 + 3 * F] = data.reshape(-1).view(np.int32)
        rec[4 + 3 * F:4 + 4 * F] = scales.view(np.int32)
        rec[4 + 4 * F:4 + 7 * F] = pre.reshape(-1).view(np.int32)
        rec[4 + 7 * F:4 + 9 * F] = order.reshape(-1)
        want.append((score, r % 7, data, scales, pre, order))
    words[hdr + total * rw:] = -123456          # whatever lies behind the live records is never parsed
    return words, want


@pytest.mark.parametrize('F,n_per_image', [(17, [2, 0, 3]), (5, [1]), (18, [0, 0]), (24, [4, 1, 0, 2, 7])])
def test_packed_result_layout(F, n_per_image):
    from openpifpaf_b200.decoder.engine import PackedResult
    words, want = make_block(n_per_image, F)
    res = PackedResult(words, len(n_per_image), F)
    assert res.total == sum(n_per_image) and res.n_ann.tolist() == n_per_image
    assert res.n_pre.tolist() == [n + 1 for n in n_per_image]
    assert res.data.shape == (res.total, F, 3) and res.order.shape == (res.total, F, 2)
    for r, (score, source, data, scales, pre, order) in enumerate(want):
        assert res.score[r] == score and res.source[r] == source
        assert np.array_equal(res.data[r], data) and np.array_equal(res.scales[r], scales)
        assert np.array_equal(res.pre[r], pre) and np.array_equal(res.order[r], order)
    # the parsed arrays are a private copy: the (reused, pinned) block may be overwritten afterwards
    words[:] = 0
    if want:
        assert np.array_equal(res.data[0], want[0][2])


def test_packed_result_rejects_a_block_of_another_call():
    from openpifpaf_b200.decoder.engine import PackedResult
    words, _ = make_block([1, 1], 17)
    with pytest.raises(AssertionError):
        PackedResult(words, 3, 17)
    with pytest.raises(AssertionError):
        PackedResult(words, 2, 18)


def test_packed_result_status_bits(caplog):
    from openpifpaf_b200 import _lib
    from openpifpaf_b200.decoder.engine import PackedResult
    words, _ = make_block([1, 2, 1], 17, status=[0, _lib.PP_STATUS_SCORE_ABOVE_ONE, 0])
    with caplog.at_level(logging.WARNING):
        PackedResult(words, 3, 17).check()                   # a warning, not an error
    assert any('above 1' in r.getMessage() and '[1]' in r.getMessage() for r in caplog.records)
    for bit in (1, 2, 4, 8):
        assert bit & _lib.PP_STATUS_OVERFLOW_MASK
        words, _ = make_block([1, 2], 17, status=[0, bit | _lib.PP_STATUS_SCORE_ABOVE_ONE])
        with pytest.raises(_lib.PifPafError) as e:
            PackedResult(words, 2, 17).check()
        assert e.value.code == _lib.PP_ERR_CAPACITY


def test_annotation_lists_from_a_packed_block():
    from openpifpaf_b200.annotation import Annotation
    from openpifpaf_b200.datasets.constants import COCO_KEYPOINTS, COCO_PERSON_SKELETON
    from openpifpaf_b200.decoder.engine import PackedResult
    n_per_image = [2, 0, 3, 1]
    words, want = make_block(n_per_image, 17)
    res = PackedResult(words, 4, 17)
    anns = Annotation.wrap_batch(list(COCO_KEYPOINTS), list(COCO_PERSON_SKELETON), res.data, res.scales, res.pre,
                                 res.order, None)
    assert len(anns) == 6
    for ann, (_, _, data, scales, pre, order) in zip(anns, want):
        assert np.array_equal(ann.data, data) and np.array_equal(ann.joint_scales, scales)
        steps = [(int(s), int(e)) for s, e in order if s >= 0]
        assert [(s, e) for s, e, _, _ in ann.decoding_order] == steps
        for (s, e, xs, xe) in ann.decoding_order:
            assert np.array_equal(xs, pre[s]) and np.array_equal(xe, pre[e])
        assert ann.score() == pytest.approx(float(np.sum(ann.score_weights * np.sort(data[:, 2])[::-1])), rel=1e-6)


def test_decoder_params_tuning_knobs_reach_the_struct():
    """DecoderParams.tuning -> pp_params (the schedule knobs are struct fields, nothing is read from the
    environment); an unknown knob is refused."""
    from openpifpaf_b200.decoder.engine import DecoderParams
    knobs = {'asm_heavy_seeds': -1, 'asm_stage_kb': 48, 'cifhr_group': 16, 'cifhr_ctas_per_sm': 2,
             'cifhr_rshift': 3, 'cifhr_warps': 5, 'cifhr_depth': 12}
    p = DecoderParams(tuning=knobs).to_c()
    for name, value in knobs.items():
        assert getattr(p, name) == value
    q = DecoderParams().to_c()
    assert all(getattr(q, name) == 0 for name in knobs)          # 0 = the library's choice
    with pytest.raises(KeyError):
        DecoderParams(tuning={'cifhr_wraps': 2}).to_c()
