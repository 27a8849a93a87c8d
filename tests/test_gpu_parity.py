"""GPU parity: CUDA path (through the C ABI) vs the CPU oracle on identical inputs.

Bar: bit-exact everywhere.  The oracle runs in its deterministic-exp mode
(exp_mode=1, include/pifpaf_detmath.h), which the kernels reproduce operation
for operation; hr / seeds / CafScored involve no exp at all and are also
bit-identical to the reference itself (tests/test_oracle_vs_reference.py).
"""
import numpy as np
import pytest
import torch

from conftest import fields

pytestmark = pytest.mark.gpu

CASES = [
    ((641, 641), 5, 0),
    ((321, 481), 8, 3),
    ((161, 161), 2, 5),
    ((97, 129), 1, 7),
]


def run_gpu(decoder, cifs, cafs, skeleton, keep=('ann', 'hr', 'seeds', 'caf'), **param_kw):
    from openpifpaf_b200.decoder.engine import DecoderParams, DeviceDecoder
    eng = DeviceDecoder(skeleton)
    cif = torch.from_numpy(np.stack(cifs)).cuda()
    caf = torch.from_numpy(np.stack(cafs)).cuda()
    buf = eng.run_device(cif, caf, DecoderParams(**param_kw), keep=keep)
    torch.cuda.synchronize()
    return {k: v.cpu().numpy() for k, v in buf.dev.items()}


def check_image(orc, out, b, cif, caf, skeleton, **cfg_kw):
    o = orc.decode(cif, caf, skeleton, orc.Config(exp_mode=1, **cfg_kw))
    assert out['status'][b] == 0
    if 'hr' in out:
        assert np.array_equal(out['hr'][b], o.hr), 'hr differs'
    if 'seeds' in out:
        n = out['n_seeds'][b]
        assert n == len(o.seeds)
        assert np.array_equal(out['seeds'][b, :n], o.seeds), 'seed list differs'
    if 'caf_fwd' in out:
        for l in range(caf.shape[0]):
            nf, nb = out['n_caf_fwd'][b, l], out['n_caf_bwd'][b, l]
            assert np.array_equal(out['caf_fwd'][b, l, :, :nf], o.forward[l]), 'caf forward %d' % l
            assert np.array_equal(out['caf_bwd'][b, l, :, :nb], o.backward[l]), 'caf backward %d' % l
    npre = out['n_pre'][b]
    assert npre == len(o.pre_data)
    assert np.array_equal(out['pre_data'][b, :npre], o.pre_data), 'grown annotations differ'
    assert np.array_equal(out['pre_scales'][b, :npre], o.pre_joint_scales)
    n = out['n_ann'][b]
    assert n == len(o.data), (n, len(o.data))
    assert np.array_equal(out['ann_data'][b, :n], o.data), 'final annotations differ'
    assert np.array_equal(out['ann_scales'][b, :n], o.joint_scales)
    for r in range(n):
        src = out['ann_source'][b, r]
        order = out['pre_order'][b, src]
        order = order[order[:, 0] >= 0]
        assert np.array_equal(order, o.decoding_order[r]), 'decoding order differs'
        assert out['ann_score'][b, r] == orc.annotation_score(o.data[r])
    return o


@pytest.mark.parametrize('hw,n_poses,seed', CASES)
def test_single_image_bit_exact(oracle, coco_skeleton, configured_decoder, hw, n_poses, seed):
    cif, caf, _ = fields(hw, n_poses, pose_seed=seed, jitter_seed=seed + 1)
    out = run_gpu(configured_decoder, [cif], [caf], coco_skeleton)
    o = check_image(oracle, out, 0, cif, caf, coco_skeleton)
    assert len(o.data) >= 1


def test_batch_of_different_images(oracle, coco_skeleton, configured_decoder):
    items = [fields((321, 321), n, pose_seed=10 + n, jitter_seed=20 + n) for n in (1, 2, 3, 4, 6, 0)]
    out = run_gpu(configured_decoder, [i[0] for i in items], [i[1] for i in items], coco_skeleton)
    for b, (cif, caf, _) in enumerate(items):
        check_image(oracle, out, b, cif, caf, coco_skeleton)


def test_crowded_1281(oracle, coco_skeleton, configured_decoder):
    cif, caf, _ = fields((1281, 1281), 50)
    out = run_gpu(configured_decoder, [cif], [caf], coco_skeleton)
    o = check_image(oracle, out, 0, cif, caf, coco_skeleton)
    assert len(o.data) >= 45


def test_no_nms_and_max_method(oracle, coco_skeleton, configured_decoder):
    cif, caf, _ = fields((321, 481), 8, pose_seed=3, jitter_seed=4)
    out = run_gpu(configured_decoder, [cif], [caf], coco_skeleton, nms=False)
    check_image(oracle, out, 0, cif, caf, coco_skeleton, do_nms=0)
    out = run_gpu(configured_decoder, [cif], [caf], coco_skeleton, connection_method='max')
    check_image(oracle, out, 0, cif, caf, coco_skeleton, connection_method=1)


def test_public_api_matches_oracle(oracle, coco_skeleton, configured_decoder):
    from openpifpaf_b200.datasets.constants import COCO_KEYPOINTS
    decoder = configured_decoder
    cif, caf, _ = fields((641, 641), 5)
    dec = decoder.CifCaf(decoder.FieldConfig(), keypoints=COCO_KEYPOINTS, skeleton=coco_skeleton)
    o = oracle.decode(cif, caf, coco_skeleton, oracle.Config(exp_mode=1))
    for anns in (dec([cif, caf]), dec([torch.from_numpy(cif).cuda(), torch.from_numpy(caf).cuda()])):
        assert len(anns) == len(o.data)
        for ann, d, js, order in zip(anns, o.data, o.joint_scales, o.decoding_order):
            assert np.array_equal(ann.data, d)
            assert np.array_equal(ann.joint_scales, js)
            assert [(s, e) for s, e, _, _ in ann.decoding_order] == [tuple(x) for x in order.tolist()]
            assert ann.score() == oracle.annotation_score(d)
