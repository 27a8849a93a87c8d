"""COCO person tables used by the decoder, the encoder and the tests.

These are DATA (keypoint names, 1-based skeleton edges, per-keypoint sigmas and
the canonical upright pose) and therefore take the same values as the
reference's ``openpifpaf/datasets/constants.py:4-62,89-147``; the layout below is
this repo's own (tuples, derived tables computed instead of listed).
"""
import numpy as np

COCO_KEYPOINTS = (
    'nose',
    'left_eye', 'right_eye',
    'left_ear', 'right_ear',
    'left_shoulder', 'right_shoulder',
    'left_elbow', 'right_elbow',
    'left_wrist', 'right_wrist',
    'left_hip', 'right_hip',
    'left_knee', 'right_knee',
    'left_ankle', 'right_ankle',
)
COCO_KEYPOINTS = list(COCO_KEYPOINTS)

# 19 CAF links, 1-based joint ids, in head-channel order.
COCO_PERSON_SKELETON = [
    (16, 14), (14, 12), (17, 15), (15, 13), (12, 13), (6, 12), (7, 13),
    (6, 7), (6, 8), (7, 9), (8, 10), (9, 11), (2, 3), (1, 2), (1, 3),
    (2, 4), (3, 5), (4, 6), (5, 7),
]

# Dense 44-edge graph; the 25 edges that are not part of the 19-link skeleton
# are the extra CAF links of the ``caf25`` head (order matters: it is the
# channel order of that head).
DENSER_COCO_PERSON_SKELETON = [
    (1, 2), (1, 3), (2, 3), (1, 4), (1, 5), (4, 5),
    (1, 6), (1, 7), (2, 6), (3, 7),
    (2, 4), (3, 5), (4, 6), (5, 7), (6, 7),
    (6, 12), (7, 13), (6, 13), (7, 12), (12, 13),
    (6, 8), (7, 9), (8, 10), (9, 11), (6, 10), (7, 11),
    (8, 9), (10, 11),
    (10, 12), (11, 13),
    (10, 14), (11, 15),
    (14, 12), (15, 13), (12, 15), (13, 14),
    (12, 16), (13, 17),
    (16, 14), (17, 15), (14, 17), (15, 16),
    (14, 15), (16, 17),
]
DENSER_COCO_PERSON_CONNECTIONS = [
    edge for edge in DENSER_COCO_PERSON_SKELETON if edge not in COCO_PERSON_SKELETON
]
assert len(DENSER_COCO_PERSON_CONNECTIONS) == 25

# OKS sigmas per keypoint (nose, eyes x2, ears x2, shoulders x2, ...).
_SIGMA_BY_PART = (('nose', .026), ('eye', .025), ('ear', .035), ('shoulder', .079),
                  ('elbow', .072), ('wrist', .062), ('hip', .107), ('knee', .087),
                  ('ankle', .089))
COCO_PERSON_SIGMAS = [
    next(s for part, s in _SIGMA_BY_PART if name.endswith(part)) for name in COCO_KEYPOINTS
]

# Canonical standing pose: (x, y up, visibility).
COCO_UPRIGHT_POSE = np.array([
    (0.0, 9.3, 2.0),
    (-0.35, 9.7, 2.0), (0.35, 9.7, 2.0),
    (-0.7, 9.5, 2.0), (0.7, 9.5, 2.0),
    (-1.4, 8.0, 2.0), (1.4, 8.0, 2.0),
    (-1.75, 6.0, 2.0), (1.75, 6.2, 2.0),
    (-1.75, 4.0, 2.0), (1.75, 4.2, 2.0),
    (-1.26, 4.0, 2.0), (1.26, 4.0, 2.0),
    (-1.4, 2.0, 2.0), (1.4, 2.1, 2.0),
    (-1.4, 0.0, 2.0), (1.4, 0.1, 2.0),
])

# left <-> right keypoint names under a horizontal flip (reference datasets/constants.py:86-103)
HFLIP = {}
for _name in COCO_KEYPOINTS:
    if _name.startswith('left_'):
        HFLIP[_name] = 'right_' + _name[len('left_'):]
    elif _name.startswith('right_'):
        HFLIP[_name] = 'left_' + _name[len('right_'):]
