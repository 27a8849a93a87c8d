"""Tie class T4, searched instead of argued: the one arithmetic step the CUDA path and
its oracle mode (exp_mode=1, include/pifpaf_detmath.h) do not share bit for bit with the reference is
float32 exp (numpy's SIMD kernel there, cifcaf.py:139).  A last-ulp difference could flip one of the
compares it feeds (`> 0.01`, `< 0.5 s1`, `> s / 2`, `> xy_scale`, `> keypoint_threshold`) and with it an
INTEGER output.  This test decodes 1000 random scenes with libm exp (exp_mode=0, the reference-like
mode the goldens pin) and with the deterministic exp, and counts scenes whose integer outputs --
annotation count, which joints are set, decoding order -- differ.  Expected and asserted: none; the
float outputs stay far inside the north_star tolerances (xy 1e-3 px, scores 1e-4)."""
import numpy as np

N_SCENES = 1000


def test_exp_mode_never_flips_an_integer_output(oracle, coco_skeleton):
    from openpifpaf_b200.encoder import synthetic_fields
    sizes = [(161, 161), (161, 241), (241, 241), (97, 129)]
    flipped, max_dxy, max_dv, n_ann, n_edges = [], 0.0, 0.0, 0, 0
    for seed in range(N_SCENES):
        hw = sizes[seed % len(sizes)]
        cif, caf, _ = synthetic_fields(hw, 1 + seed % 5, pose_seed=10_000 + seed, jitter_seed=20_000 + seed)
        a = oracle.decode(cif, caf, coco_skeleton, oracle.Config(exp_mode=0))
        b = oracle.decode(cif, caf, coco_skeleton, oracle.Config(exp_mode=1))
        same = (len(a.data) == len(b.data)
                and all(np.array_equal(x[:, 2] != 0, y[:, 2] != 0) for x, y in zip(a.data, b.data))
                and all(np.array_equal(x, y) for x, y in zip(a.decoding_order, b.decoding_order))
                and len(a.pre_data) == len(b.pre_data))
        if not same:
            flipped.append(seed)
            continue
        n_ann += len(a.data)
        n_edges += sum(len(o) for o in a.decoding_order)
        for x, y in zip(a.data, b.data):
            max_dxy = max(max_dxy, float(np.abs(x[:, :2] - y[:, :2]).max()))
            max_dv = max(max_dv, float(np.abs(x[:, 2] - y[:, 2]).max()))
    print('T4 search: %d scenes, %d annotations, %d decoding steps, %d scenes with a flipped integer output; '
          'max |dxy| %.2e px, max |dscore| %.2e' % (N_SCENES, n_ann, n_edges, len(flipped), max_dxy, max_dv))
    assert n_ann > 2 * N_SCENES and n_edges > 20 * N_SCENES
    assert not flipped, 'exp last-ulp differences flipped integer outputs in scenes %s' % flipped[:10]
    assert max_dxy <= 1e-3 and max_dv <= 1e-4
