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
