import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box)')
    config.addinivalue_line('markers', 'reference: needs the staged reference in oracle/_ref')


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:  # pylint: disable=broad-except
        has_gpu = False
    skip_gpu = pytest.mark.skip(reason='no CUDA device in this container')
    for item in items:
        if 'gpu' in item.keywords and not has_gpu:
            item.add_marker(skip_gpu)


_FIELD_CACHE = {}


def fields(image_hw=(641, 641), n_poses=5, pose_seed=0, jitter_seed=1, dense=False):
    """Deterministic synthetic decoder inputs, cached per process."""
    from openpifpaf_b200.encoder import synthetic_fields
    key = (tuple(image_hw), n_poses, pose_seed, jitter_seed, dense)
    if key not in _FIELD_CACHE:
        cif, caf, poses = synthetic_fields(image_hw, n_poses, pose_seed=pose_seed,
                                           jitter_seed=jitter_seed, dense=dense)
        _FIELD_CACHE[key] = (cif, caf, poses)
    return _FIELD_CACHE[key]


@pytest.fixture(scope='session')
def oracle():
    from oracle import oracle as orc
    orc.lib()
    return orc


@pytest.fixture(scope='session')
def coco_skeleton():
    from openpifpaf_b200.datasets.constants import COCO_PERSON_SKELETON
    return COCO_PERSON_SKELETON


@pytest.fixture()
def configured_decoder():
    """Decoder class attributes set to the SURVEY 8d profile, restored afterwards."""
    from openpifpaf_b200 import decoder
    saved = (decoder.CifHr.v_threshold, decoder.CifSeeds.threshold, decoder.CifSeeds.score_scale,
             decoder.CafScored.default_score_th, decoder.CifCaf.keypoint_threshold,
             decoder.CifCaf.force_complete, decoder.CifCaf.greedy,
             decoder.CifCaf.connection_method, decoder.nms.Keypoints.keypoint_threshold,
             decoder.nms.Keypoints.instance_threshold, decoder.nms.Keypoints.suppression)
    decoder.CifHr.v_threshold = 0.1
    decoder.CifSeeds.threshold = 0.2
    decoder.CifSeeds.score_scale = 1.0
    decoder.CafScored.default_score_th = 0.1
    decoder.CifCaf.keypoint_threshold = 0.001
    decoder.CifCaf.force_complete = False
    decoder.CifCaf.greedy = False
    decoder.CifCaf.connection_method = 'blend'
    decoder.nms.Keypoints.keypoint_threshold = 0.001
    decoder.nms.Keypoints.instance_threshold = 0.0
    decoder.nms.Keypoints.suppression = 0.0
    yield decoder
    (decoder.CifHr.v_threshold, decoder.CifSeeds.threshold, decoder.CifSeeds.score_scale,
     decoder.CafScored.default_score_th, decoder.CifCaf.keypoint_threshold,
     decoder.CifCaf.force_complete, decoder.CifCaf.greedy,
     decoder.CifCaf.connection_method, decoder.nms.Keypoints.keypoint_threshold,
     decoder.nms.Keypoints.instance_threshold, decoder.nms.Keypoints.suppression) = saved


# ---------------------------------------------------------------------------
# golden fixtures (tests/golden/*.npz, written by oracle/gen_golden.py from the reference)

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')
_ALL_GOLDEN = sorted(n[:-4] for n in os.listdir(GOLDEN_DIR) if n.endswith('.npz')) \
    if os.path.isdir(GOLDEN_DIR) else []
GOLDEN_NAMES = [n for n in _ALL_GOLDEN if not n.startswith(('head_', 'inverse_', 'det_', 'ms_'))]   # pose decoder cases
DET_GOLDEN_NAMES = [n for n in _ALL_GOLDEN if n.startswith('det_')]            # CifDet cases
MS_GOLDEN_NAMES = [n for n in _ALL_GOLDEN if n.startswith('ms_')]              # multi-scale FieldConfig cases
HEAD_GOLDEN_NAMES = [n for n in _ALL_GOLDEN if n.startswith('head_')]       # head epilogue cases
TOL_HEAD_REL = 2e-6  # sigmoid / exp of the head epilogue against torch's CPU kernels


def head_golden(name):
    """(x_cif, x_caf, cif, caf) of a head-epilogue fixture; raw inputs regenerated and hash-checked."""
    import hashlib
    from oracle.gen_golden_heads import raw_head_outputs
    g = np.load(os.path.join(GOLDEN_DIR, name + '.npz'))
    x_cif, x_caf = raw_head_outputs(int(g['batch']), int(g['h']), int(g['w']), int(g['seed']))
    assert hashlib.sha256(x_cif.tobytes()).hexdigest() == str(g['x_cif_sha256'])
    assert hashlib.sha256(x_caf.tobytes()).hexdigest() == str(g['x_caf_sha256'])
    return x_cif, x_caf, g['cif'], g['caf']

# north_star tolerances against the reference
TOL_XY = 1e-3       # px
TOL_SCORE = 1e-4
TOL_HR_REL = 1e-5


def sha256(a):
    import hashlib
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


class Golden:
    """One fixture: regenerated inputs (hash-checked) + the reference's observations."""

    def __init__(self, name):
        import ast
        from openpifpaf_b200.datasets.constants import (COCO_PERSON_SKELETON,
                                                        DENSER_COCO_PERSON_CONNECTIONS)
        self.name = name
        self.z = np.load(os.path.join(GOLDEN_DIR, name + '.npz'))
        z = self.z
        self.dense = bool(z['dense'])
        self.overrides = dict(ast.literal_eval(str(z['overrides'])))
        self.cif, self.caf, _ = fields(tuple(int(v) for v in z['image_hw']), int(z['n_poses']),
                                       pose_seed=int(z['pose_seed']),
                                       jitter_seed=int(z['jitter_seed']), dense=self.dense)
        assert sha256(self.cif) == str(z['cif_sha256']), 'synthetic encoder drifted (cif) for ' + name
        assert sha256(self.caf) == str(z['caf_sha256']), 'synthetic encoder drifted (caf) for ' + name
        self.skeleton = [tuple(l) for l in COCO_PERSON_SKELETON]
        if self.dense:
            self.skeleton += [tuple(l) for l in DENSER_COCO_PERSON_CONNECTIONS]

    def oracle_kwargs(self):
        """Overrides in the oracle Config's vocabulary."""
        o = self.overrides
        kw = {}
        if o.get('connection_method') == 'max':
            kw['connection_method'] = 1
        if o.get('greedy'):
            kw['greedy'] = 1
        if o.get('force_complete'):
            kw['force_complete'] = 1
        if 'keypoint_threshold' in o:
            kw['keypoint_th'] = o['keypoint_threshold']
            kw['nms_keypoint_th'] = o['keypoint_threshold']
        return kw

    def engine_kwargs(self):
        """Overrides in DecoderParams' vocabulary."""
        o = self.overrides
        kw = {k: o[k] for k in ('connection_method', 'greedy', 'force_complete') if k in o}
        if 'keypoint_threshold' in o:
            kw['keypoint_threshold'] = o['keypoint_threshold']
            kw['nms_keypoint_threshold'] = o['keypoint_threshold']
        return kw

    def check_annotations(self, data, scales, order, key='ann', exact=False):
        """Compare annotations with the reference's: integer outputs identical
        (count, which joints are set, decoding order), xy / scores in tolerance."""
        z = self.z
        ref_data, ref_scales, ref_order = z[key + '_data'], z[key + '_scales'], z[key + '_order']
        assert len(data) == len(ref_data), '%s: %d annotations, reference %d' % (
            self.name, len(data), len(ref_data))
        if len(data) == 0:
            return
        assert np.array_equal(data[..., 2] != 0, ref_data[..., 2] != 0), 'joint masks differ'
        if order is not None:
            assert np.array_equal(order, ref_order), 'decoding order differs'
        if exact:
            assert np.array_equal(data, ref_data) and np.array_equal(scales, ref_scales)
            return
        assert np.abs(data[..., :2] - ref_data[..., :2]).max() <= TOL_XY
        assert np.abs(data[..., 2] - ref_data[..., 2]).max() <= TOL_SCORE
        assert np.abs(scales - ref_scales).max() <= TOL_XY


_GOLDEN_CACHE = {}


def golden(name):
    if name not in _GOLDEN_CACHE:
        _GOLDEN_CACHE[name] = Golden(name)
    return _GOLDEN_CACHE[name]


def inverse_inputs():
    """Annotations of the golden case mid_321x481_8p as packed arrays + the flip permutation:
    (data, joint_scales, decoding (s, e) lists, decoding xy rows, hflip target)."""
    from openpifpaf_b200.datasets.constants import COCO_KEYPOINTS, HFLIP
    g = np.load(os.path.join(GOLDEN_DIR, 'mid_321x481_8p.npz'))
    data, js = g['ann_data'], g['ann_scales']
    orders = [[(int(s), int(e)) for s, e in o if s >= 0] for o in g['ann_order']]
    dxy = np.array([d[j][:2] for d, o in zip(data, orders) for s, e in o for j in (s, e)], np.float32)
    target = np.array([COCO_KEYPOINTS.index(HFLIP.get(k, k)) for k in COCO_KEYPOINTS], np.int32)
    return data, js, orders, dxy, target


class HorizontalSwap:
    """Stand-in for the reference's hflip._HorizontalSwap (rows move to their mirrored joint)."""

    def __init__(self, target):
        self.target = np.asarray(target)

    def __call__(self, keypoints):
        out = np.zeros(keypoints.shape)
        out[self.target] = keypoints
        return out


def det_golden(name):
    """(field, golden npz) of a CifDet fixture; the field is regenerated and hash-checked."""
    import hashlib
    from openpifpaf_b200.encoder import synthetic_det_fields
    g = np.load(os.path.join(GOLDEN_DIR, name + '.npz'))
    field, _ = synthetic_det_fields(tuple(int(v) for v in g['image_hw']), int(g['n_boxes']),
                                    box_seed=int(g['box_seed']), jitter_seed=int(g['jitter_seed']),
                                    n_categories=int(g['n_categories']))
    assert hashlib.sha256(field.tobytes()).hexdigest() == str(g['field_sha256'])
    return field, g


def ms_golden(name):
    """(sets, golden npz) of a multi-scale fixture; inputs regenerated and hash-checked."""
    import hashlib
    from oracle.gen_golden_ms import CASES, make_sets
    g = np.load(os.path.join(GOLDEN_DIR, name + '.npz'))
    sets = make_sets(*CASES[name])
    blob = np.concatenate([np.concatenate([s[0].ravel(), s[1].ravel()]) for s in sets])
    assert hashlib.sha256(np.ascontiguousarray(blob).tobytes()).hexdigest() == str(g['inputs_sha256'])
    return sets, g
