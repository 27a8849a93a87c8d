"""Output types of the decoder.

API-compatible with the reference's ``openpifpaf/annotation.py:9-146`` (same
attribute and method names, same numeric results) because every caller of the
decoder consumes these objects: ``transforms/preprocess.py:35-82``
(``annotations_inverse``), ``show/painters.py:344-383``, ``eval_coco.py:108-158``.

Differences in construction only: instances are normally built from the CUDA
decoder's result buffers by ``Annotation.from_buffers`` and ``decoding_order`` /
``frontier_order`` are materialised lazily from the compact device records.
"""
import numpy as np

NOTSET = '__notset__'


_WEIGHTS = {}


def _score_weights(n, suppress_last=False):
    key = (n, suppress_last)
    if key not in _WEIGHTS:
        w = np.ones((n,))
        if suppress_last:
            w[-1] = 0.0
        w[:3] = 3.0
        _WEIGHTS[key] = w / np.sum(w)
    return _WEIGHTS[key].copy()


class Annotation:
    """One person: ``data[j] = (x, y, v)`` per keypoint and a scale per joint."""

    def __init__(self, keypoints, skeleton, *, category_id=1, suppress_score_index=None):
        n = len(keypoints)
        self.keypoints = keypoints
        self.skeleton = skeleton
        self.category_id = category_id
        self.suppress_score_index = suppress_score_index
        self.data = np.zeros((n, 3), dtype=np.float32)
        self.joint_scales = np.zeros((n,), dtype=np.float32)
        self.fixed_score = NOTSET
        self._decoding_order = []
        self._frontier_order = []
        self._lazy = None
        self.score_weights = _score_weights(n, bool(suppress_score_index))

    # -- construction -----------------------------------------------------
    @classmethod
    def from_buffers(cls, keypoints, skeleton, data, joint_scales, *, pre_data=None,
                     order=None, graph=None):
        """Wrap rows of the decoder's result buffers (copied).

        ``pre_data``/``order`` are the annotation as grown (before NMS) and its
        (start, end) decoding order; ``graph`` the by_source neighbour lists.
        """
        ann = cls(keypoints, skeleton)
        ann.data = np.array(data, dtype=np.float32)
        ann.joint_scales = np.array(joint_scales, dtype=np.float32)
        if order is not None:
            ann._lazy = (np.array(pre_data, dtype=np.float32), np.array(order), graph)
        return ann

    @classmethod
    def wrap(cls, keypoints, skeleton, data, joint_scales, lazy):
        """Fast path of ``from_buffers``: adopts the given arrays (no copies, no zero fill).
        ``data`` / ``joint_scales`` must be arrays this annotation may own and mutate."""
        ann = cls.__new__(cls)
        ann.keypoints = keypoints
        ann.skeleton = skeleton
        ann.category_id = 1
        ann.suppress_score_index = None
        ann.data = data
        ann.joint_scales = joint_scales
        ann.fixed_score = NOTSET
        ann._decoding_order = []
        ann._frontier_order = []
        ann._lazy = lazy
        ann.score_weights = _score_weights(len(keypoints))
        return ann

    @classmethod
    def wrap_batch(cls, keypoints, skeleton, data, joint_scales, pre, order, graph):
        """``wrap`` for the rows of ``data (n, F, 3)``, ``joint_scales (n, F)``: n annotations that own
        row views of these (private) arrays.  ``pre (n, F, 3)`` / ``order (n, F, 2)`` stay whole; row k is
        looked up when ``decoding_order`` / ``frontier_order`` are first read, and ``score_weights`` is
        created on first access."""
        new = cls.__new__
        out = []
        k = 0
        for d, js in zip(data, joint_scales):
            ann = new(cls)
            ann.__dict__ = {
                'keypoints': keypoints, 'skeleton': skeleton, 'category_id': 1,
                'suppress_score_index': None, 'data': d, 'joint_scales': js, 'fixed_score': NOTSET,
                '_decoding_order': None, '_frontier_order': None, '_lazy': (pre, order, graph, k)}
            out.append(ann)
            k += 1
        return out

    def __getattr__(self, name):
        # only reached when the attribute is missing: score_weights of wrap_batch instances
        if name == 'score_weights':
            w = _score_weights(len(self.keypoints), bool(self.suppress_score_index))
            self.__dict__['score_weights'] = w
            return w
        raise AttributeError(name)

    def _materialise(self):
        if self._lazy is None:
            return
        pre, order, graph = self._lazy[:3]
        if len(self._lazy) == 4:                 # wrap_batch: row k of the batch arrays
            k = self._lazy[3]
            pre, order = pre[k], order[k]
        self._lazy = None
        steps = [(int(s), int(e)) for s, e in order if s >= 0]
        # force_complete marker (-2, n1): the first n1 steps belong to the first _grow
        first = [int(e) for s, e in order if s == -2]
        self._decoding_order = [(s, e, np.copy(pre[s]), np.copy(pre[e])) for s, e in steps]
        self._frontier_order = []
        if graph is not None:
            self._frontier_order = _replay_frontier(pre, steps, graph,
                                                    first[0] if first else None)

    @property
    def decoding_order(self):
        self._materialise()
        return self._decoding_order

    @decoding_order.setter
    def decoding_order(self, value):
        self._materialise()
        self._decoding_order = value

    @property
    def frontier_order(self):
        self._materialise()
        return self._frontier_order

    @frontier_order.setter
    def frontier_order(self, value):
        self._materialise()
        self._frontier_order = value

    @property
    def skeleton_m1(self):
        return (np.asarray(self.skeleton) - 1).tolist()

    # -- reference API ----------------------------------------------------
    def add(self, joint_i, xyv):
        self.data[joint_i] = xyv
        return self

    def set(self, data, joint_scales=None, *, fixed_score=NOTSET):
        self.data = data
        if joint_scales is None:
            self.joint_scales[:] = 0.0
        else:
            self.joint_scales = joint_scales
        self.fixed_score = fixed_score
        return self

    def rescale(self, scale_factor):
        self.data[:, 0:2] *= scale_factor
        if self.joint_scales is not None:
            self.joint_scales *= scale_factor
        for _, __, source, target in self.decoding_order:
            source[:2] *= scale_factor
            target[:2] *= scale_factor
        return self

    def score(self):
        """Weighted sum of the descending-sorted confidences, in float64."""
        if self.fixed_score != NOTSET:
            return self.fixed_score
        v = self.data[:, 2]
        if self.suppress_score_index is not None:
            v = np.copy(v)
            v[self.suppress_score_index] = 0.0
        return np.sum(self.score_weights * np.sort(v)[::-1])

    def scale(self, v_th=0.5):
        visible = self.data[:, 2] > v_th
        if not np.any(visible):
            return 0.0
        xy = self.data[visible, :2]
        extent = xy.max(axis=0) - xy.min(axis=0)
        return max(extent[0], extent[1])

    def bbox(self):
        return self.bbox_from_keypoints(self.data, self.joint_scales)

    @staticmethod
    def bbox_from_keypoints(kps, joint_scales):
        visible = kps[:, 2] > 0
        if not np.any(visible):
            return [0, 0, 0, 0]
        lo_x = np.min(kps[:, 0][visible] - joint_scales[visible])
        lo_y = np.min(kps[:, 1][visible] - joint_scales[visible])
        hi_x = np.max(kps[:, 0][visible] + joint_scales[visible])
        hi_y = np.max(kps[:, 1][visible] + joint_scales[visible])
        return [lo_x, lo_y, hi_x - lo_x, hi_y - lo_y]

    def json_data(self):
        """COCO-style dict. Visible keypoints keep v >= 0.01 after rounding."""
        kps = np.copy(self.data)
        seen = kps[:, 2] > 0.0
        kps[seen, 2] = np.maximum(0.01, kps[seen, 2])
        kps = np.around(kps.astype(np.float64), 2)
        out = {
            'keypoints': kps.reshape(-1).tolist(),
            'bbox': [round(float(c), 2) for c in self.bbox()],
            'score': max(0.001, round(self.score(), 3)),
            'category_id': self.category_id,
        }
        id_ = getattr(self, 'id_', None)
        if id_:
            out['id_'] = id_
        return out


def _replay_frontier(pre, steps, graph, n_first=None):
    """Recompute ``frontier_order`` (cifcaf.py:251-263) from the decoding order.

    The order in which (start, end) pairs were queued is a function of the seed
    joint, the by_source neighbour order and the sequence of joints that were
    filled; all three are known on the host.
    """
    n = len(pre)
    filled = np.zeros(n, dtype=bool)
    targets = {e for _, e in steps}
    for j in range(n):
        if pre[j, 2] != 0.0 and j not in targets:
            filled[j] = True      # seed joint(s)
    queued, out = set(), []

    def queue(start):
        for end in graph[start]:
            if filled[end] or (start, end) in queued:
                continue
            queued.add((start, end))
            out.append((start, end))

    for j in range(n):
        if filled[j]:
            queue(j)
    for i, (_, end) in enumerate(steps):
        if n_first is not None and i == n_first:
            n_first = None
            queued.clear()         # complete_annotations runs a second _grow (cifcaf.py:341)
            for j in range(n):
                if filled[j]:
                    queue(j)
        filled[end] = True
        queue(end)
    if n_first is not None and n_first == len(steps):
        queued.clear()
        for j in range(n):
            if filled[j]:
                queue(j)
    return out


def json_data_batch(annotations):
    """``[ann.json_data() for ann in annotations]`` (annotation.py:80-102) with the array work
    done once for the whole list (e.g. every person of a batch) instead of per annotation.
    Results are identical, element for element."""
    anns = list(annotations)
    if not anns:
        return []
    if len({len(a.data) for a in anns}) != 1:
        return [a.json_data() for a in anns]
    data = np.stack([a.data for a in anns]).astype(np.float32, copy=False)
    js = np.stack([a.joint_scales for a in anns]).astype(np.float32, copy=False)
    v = data[:, :, 2]
    seen = v > 0.0
    kps = data.copy()
    kps[:, :, 2] = np.where(seen, np.maximum(np.float32(0.01), v), v)
    kps = np.around(kps.astype(np.float64), 2).reshape(len(anns), -1).tolist()
    # bbox over visible joints, float32 arithmetic like bbox_from_keypoints
    inf = np.float32(np.inf)
    lo_x = np.min(np.where(seen, data[:, :, 0] - js, inf), axis=1)
    lo_y = np.min(np.where(seen, data[:, :, 1] - js, inf), axis=1)
    hi_x = np.max(np.where(seen, data[:, :, 0] + js, -inf), axis=1)
    hi_y = np.max(np.where(seen, data[:, :, 1] + js, -inf), axis=1)
    any_seen = seen.any(axis=1)
    boxes = np.stack([lo_x, lo_y, hi_x - lo_x, hi_y - lo_y], axis=1)
    boxes[~any_seen] = 0.0
    boxes = boxes.tolist()
    # score: weights * descending sort, summed per row in numpy's pairwise order (rows are contiguous)
    sv = v.astype(np.float32, copy=True)
    for i, a in enumerate(anns):
        if a.suppress_score_index is not None:
            sv[i, a.suppress_score_index] = 0.0
    weights = np.stack([a.score_weights for a in anns])
    scores = np.sum(np.ascontiguousarray(weights * np.sort(sv, axis=1)[:, ::-1]), axis=1)
    out = []
    for a, k, box, score, ok in zip(anns, kps, boxes, scores, any_seen.tolist()):
        if a.fixed_score != NOTSET:
            score = a.fixed_score
        item = {
            'keypoints': k,
            'bbox': [round(c, 2) for c in box] if ok else [0, 0, 0, 0],
            'score': max(0.001, round(score, 3)),
            'category_id': a.category_id,
        }
        id_ = getattr(a, 'id_', None)
        if id_:
            item['id_'] = id_
        out.append(item)
    return out


class AnnotationDet:
    """Detection output type (annotation.py:122-146); unused by the CifCaf path."""

    def __init__(self, categories):
        self.categories = categories
        self.field_i = None
        self.score = None
        self.bbox = None

    def set(self, field_i, score, bbox):
        self.field_i = field_i
        self.score = score
        self.bbox = np.asarray(bbox)
        return self

    def json_data(self):
        return {
            'detection_id': self.field_i,
            'category': self.categories[self.field_i],
            'score': max(0.001, round(float(self.score), 3)),
            'bbox': [round(float(c), 2) for c in self.bbox],
        }
