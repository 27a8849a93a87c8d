"""Synthetic CIF/CAF field generator ("the repo's own encoder").

Host-side numpy code that turns keypoint sets into the composite fields the
decoder consumes.  It follows the *semantics* of the reference's training
target encoders -- ``encoder/cif.py:103-145`` (4x4 patch around every keypoint,
nearest keypoint wins a cell), ``encoder/caf.py:162-222`` (3x3 patches walked
along every limb, smallest perpendicular distance wins a cell),
``encoder/annrescaler.py:77-109`` (instance scale = sqrt(bbox area) x pose
factor, times the per-keypoint sigma) -- and then converts targets to the
decoder's input layout the way ``network/heads.py:65-88`` does at inference
(absolute coordinates = regression + cell index):

    cif (F, 5, H, W) = [conf, x, y, logb, scale]
    caf (L, 9, H, W) = [conf, x1, y1, logb1, scale1, x2, y2, logb2, scale2]

Un-jittered targets make the reference decoder raise on exact score ties
(SURVEY.md appendix B), so ``jitter()`` is part of the workload definition.
This module is workload tooling for tests and ``bench.py``; it is not on the
decode path.
"""
import numpy as np

from .datasets.constants import (COCO_PERSON_SIGMAS, COCO_PERSON_SKELETON,
                                 COCO_UPRIGHT_POSE, DENSER_COCO_PERSON_CONNECTIONS)

CIF_SIDE = 4      # encoder/cif.py:21
CAF_SIDE = 3      # encoder/caf.py:27 (min_size, aspect_ratio 0)
PAD = 10          # encoder/cif.py:22, encoder/caf.py:30


def _bbox_area(xy):
    return float((xy[:, 0].max() - xy[:, 0].min()) * (xy[:, 1].max() - xy[:, 1].min()))


class _PoseScale:
    """Instance scale of a keypoint set (encoder/annrescaler.py:9-27,77-109)."""

    def __init__(self, pose=COCO_UPRIGHT_POSE):
        self.pose = np.asarray(pose, dtype=np.float64)
        c = np.cos(np.pi / 4.0)
        rot = np.array(((c, -c), (c, c)))
        self.pose45 = self.pose.copy()
        self.pose45[:, :2] = self.pose[:, :2] @ rot.T
        self.total = _bbox_area(self.pose)
        self.total45 = _bbox_area(self.pose45)

    def __call__(self, kps):
        vis = kps[:, 2] > 0
        if vis.sum() < 3:
            return np.nan
        area = _bbox_area(kps[vis])
        ref = _bbox_area(self.pose[vis])
        ref45 = _bbox_area(self.pose45[vis])
        factor = np.sqrt(min(self.total / ref if ref > 0.1 else np.inf,
                             self.total45 / ref45 if ref45 > 0.1 else np.inf))
        if np.isinf(factor):
            return np.nan
        scale = np.sqrt(area) * min(5.0, factor)
        return scale if scale >= 0.1 else np.nan


def _sink(side):
    """Vectors from every cell centre of a side x side patch to its centre."""
    ramp = np.linspace((side - 1.0) / 2.0, -(side - 1.0) / 2.0, num=side, dtype=np.float32)
    return np.stack((np.tile(ramp[None, :], (side, 1)), np.tile(ramp[:, None], (1, side))))


def _quadrant_radius(xyv, others):
    """Distance to the nearest same-type keypoint of another person, per quadrant
    (encoder/cif.py:78-101)."""
    out = np.full(4, np.inf, dtype=np.float32)
    if len(others) == 0:
        return out
    diff = np.asarray(others)[:, :2] - xyv[None, :2]
    quad = (diff[:, 0] < 0.0).astype(int) + 2 * (diff[:, 1] < 0.0).astype(int)
    dist = np.sqrt((diff ** 2).sum(axis=1))
    for q in range(4):
        sel = quad == q
        if sel.any():
            out[q] = dist[sel].min()
    return out


class FieldEncoder:
    """Keypoint sets (image pixels) -> (cif, caf) decoder inputs at a given stride."""

    def __init__(self, image_hw, *, stride=8, skeleton=None, sigmas=None,
                 sparse_skeleton=None, dense_split=None):
        self.stride = stride
        self.h = (image_hw[0] - 1) // stride + 1
        self.w = (image_hw[1] - 1) // stride + 1
        self.skeleton = [tuple(e) for e in (skeleton or COCO_PERSON_SKELETON)]
        self.sigmas = list(sigmas or COCO_PERSON_SIGMAS)
        # links with index >= dense_split use the caf25 rules (sparse_skeleton
        # check + only_in_field_of_view), datasets/headmeta.py:28-35
        self.dense_split = len(self.skeleton) if dense_split is None else dense_split
        self.sparse_skeleton = sparse_skeleton
        self.pose_scale = _PoseScale()

    @classmethod
    def dense(cls, image_hw, *, stride=8):
        skeleton = list(COCO_PERSON_SKELETON) + list(DENSER_COCO_PERSON_CONNECTIONS)
        return cls(image_hw, stride=stride, skeleton=skeleton,
                   sparse_skeleton=list(COCO_PERSON_SKELETON),
                   dense_split=len(COCO_PERSON_SKELETON))

    # ------------------------------------------------------------------ CIF
    def _encode_cif(self, sets):
        n_f = sets.shape[1] if len(sets) else len(self.sigmas)
        hp, wp = self.h + 2 * PAD, self.w + 2 * PAD
        conf = np.zeros((n_f, hp, wp), np.float32)
        reg = np.zeros((n_f, 2, hp, wp), np.float32)
        scl = np.zeros((n_f, hp, wp), np.float32)
        best = np.full((n_f, hp, wp), np.inf, np.float32)
        sink = _sink(CIF_SIDE)
        half = (CIF_SIDE - 1.0) / 2.0
        for pi, kps in enumerate(sets):
            inst = self.pose_scale(kps)
            for f, xyv in enumerate(kps):
                if xyv[2] <= 0:
                    continue
                others = [o[f] for oi, o in enumerate(sets) if oi != pi and o[f, 2] > 0]
                radius = _quadrant_radius(xyv, others)
                js = min(inst * self.sigmas[f], radius.min() * 0.25)
                ij = np.round(xyv[:2] - half).astype(int) + PAD
                x0, y0 = int(ij[0]), int(ij[1])
                x1, y1 = x0 + CIF_SIDE, y0 + CIF_SIDE
                if x0 < 0 or y0 < 0 or x1 > wp or y1 > hp:
                    continue
                off = (xyv[:2] - (ij + half - PAD)).reshape(2, 1, 1)
                vec = sink + off
                dist = np.sqrt((vec ** 2).sum(axis=0)).astype(np.float32)
                win = dist < best[f, y0:y1, x0:x1]
                best[f, y0:y1, x0:x1][win] = dist[win]
                conf[f, y0:y1, x0:x1][win] = 1.0
                reg[f, :, y0:y1, x0:x1][:, win] = vec[:, win]
                scl[f, y0:y1, x0:x1][win] = js if np.isfinite(js) else 0.0
        sl = slice(PAD, -PAD)
        return conf[:, sl, sl], reg[:, :, sl, sl], scl[:, sl, sl]

    # ------------------------------------------------------------------ CAF
    def _shortest_sparse(self, j, kps):
        best = np.inf
        for a, b in self.sparse_skeleton:
            a, b = a - 1, b - 1
            if j not in (a, b) or kps[a, 2] <= 0 or kps[b, 2] <= 0:
                continue
            best = min(best, float(np.linalg.norm(kps[a, :2] - kps[b, :2])))
        return best

    def _encode_caf(self, sets):
        n_l = len(self.skeleton)
        hp, wp = self.h + 2 * PAD, self.w + 2 * PAD
        conf = np.zeros((n_l, hp, wp), np.float32)
        reg1 = np.zeros((n_l, 2, hp, wp), np.float32)
        reg2 = np.zeros((n_l, 2, hp, wp), np.float32)
        sc1 = np.zeros((n_l, hp, wp), np.float32)
        sc2 = np.zeros((n_l, hp, wp), np.float32)
        best = np.full((n_l, hp, wp), np.inf, np.float32)
        sink = _sink(CAF_SIDE)
        half = (CAF_SIDE - 1.0) / 2.0
        for pi, kps in enumerate(sets):
            inst = self.pose_scale(kps)
            for li, (a, b) in enumerate(self.skeleton):
                a, b = a - 1, b - 1
                j1, j2 = kps[a], kps[b]
                if j1[2] <= 0 or j2[2] <= 0:
                    continue
                if li >= self.dense_split:
                    if self.sparse_skeleton is not None:
                        d = np.linalg.norm(j1[:2] - j2[:2]) / 2.0
                        if self._shortest_sparse(a, kps) < d and self._shortest_sparse(b, kps) < d:
                            continue
                    if (min(j1[0], j2[0], j1[1], j2[1]) < 0
                            or max(j1[0], j2[0]) > self.w - 1 or max(j1[1], j2[1]) > self.h - 1):
                        continue
                o1 = [o[a] for oi, o in enumerate(sets) if oi != pi and o[a, 2] > 0]
                o2 = [o[b] for oi, o in enumerate(sets) if oi != pi and o[b, 2] > 0]
                s1 = min(inst * self.sigmas[a], _quadrant_radius(j1, o1).min() * 0.25)
                s2 = min(inst * self.sigmas[b], _quadrant_radius(j2, o2).min() * 0.25)
                limb = j2[:2] - j1[:2]
                length = float(np.linalg.norm(limb))
                steps = max(2, int(np.ceil(length)))
                margin = min(0.4, (half + 1) / (length + np.spacing(1)))
                for t in np.linspace(margin, 1.0 - margin, num=steps):
                    fij = np.round(j1[:2] + t * limb - half) + PAD
                    x0, y0 = int(fij[0]), int(fij[1])
                    x1, y1 = x0 + CAF_SIDE, y0 + CAF_SIDE
                    if x0 < 0 or y0 < 0 or x1 > wp or y1 > hp:
                        continue
                    centre = fij - PAD + half
                    v1 = sink + (j1[:2] - centre).reshape(2, 1, 1)
                    v2 = sink + (j2[:2] - centre).reshape(2, 1, 1)
                    perp = (np.fabs(limb[1] * v1[0] - limb[0] * v1[1]) / (length + 0.01))
                    perp = perp.astype(np.float32)
                    win = perp < best[li, y0:y1, x0:x1]
                    best[li, y0:y1, x0:x1][win] = perp[win]
                    conf[li, y0:y1, x0:x1][win] = 1.0
                    reg1[li, :, y0:y1, x0:x1][:, win] = v1[:, win]
                    reg2[li, :, y0:y1, x0:x1][:, win] = v2[:, win]
                    sc1[li, y0:y1, x0:x1][win] = s1 if np.isfinite(s1) else 0.0
                    sc2[li, y0:y1, x0:x1][win] = s2 if np.isfinite(s2) else 0.0
        sl = slice(PAD, -PAD)
        return (conf[:, sl, sl], reg1[:, :, sl, sl], reg2[:, :, sl, sl],
                sc1[:, sl, sl], sc2[:, sl, sl])

    # --------------------------------------------------------------- public
    def __call__(self, keypoint_sets):
        """keypoint_sets: (P, F, 3) image-pixel (x, y, visibility). Returns (cif, caf)."""
        sets = np.array(keypoint_sets, dtype=np.float64, copy=True).reshape(-1, len(self.sigmas), 3)
        sets[:, :, :2] /= self.stride
        yy, xx = np.indices((self.h, self.w), dtype=np.float32)

        c, r, s = self._encode_cif(sets)
        cif = np.zeros((c.shape[0], 5, self.h, self.w), np.float32)
        cif[:, 0] = c
        cif[:, 1] = r[:, 0] + xx
        cif[:, 2] = r[:, 1] + yy
        cif[:, 4] = s

        c, r1, r2, s1, s2 = self._encode_caf(sets)
        caf = np.zeros((c.shape[0], 9, self.h, self.w), np.float32)
        caf[:, 0] = c
        caf[:, 1] = r1[:, 0] + xx
        caf[:, 2] = r1[:, 1] + yy
        caf[:, 4] = s1
        caf[:, 5] = r2[:, 0] + xx
        caf[:, 6] = r2[:, 1] + yy
        caf[:, 8] = s2
        return cif, caf


def synthetic_poses(n_poses, image_hw, rng):
    """Upright COCO poses at random positions / sizes."""
    h_img, w_img = image_hw
    base = COCO_UPRIGHT_POSE.copy()
    base[:, 1] = base[:, 1].max() - base[:, 1]          # y grows downwards in images
    base[:, 0] -= 0.5 * (base[:, 0].max() + base[:, 0].min())
    extent = base[:, 1].max() - base[:, 1].min()
    poses = np.zeros((n_poses, base.shape[0], 3), np.float32)
    for i in range(n_poses):
        height = rng.uniform(0.15, 0.45) * h_img
        cx = rng.uniform(0.15, 0.85) * w_img
        cy = rng.uniform(0.25, 0.75) * h_img
        k = height / extent
        poses[i, :, 0] = cx + k * base[:, 0]
        poses[i, :, 1] = cy + k * (base[:, 1] - 0.5 * extent)
        poses[i, :, 2] = 2.0
    return poses


def jitter(cif, caf, rng):
    """Network-like noise; breaks the exact ties the reference cannot decode."""
    cif, caf = cif.copy(), caf.copy()
    for field, regs, scales in ((cif, (1, 2), (4,)), (caf, (1, 2, 5, 6), (4, 8))):
        active = field[:, 0] > 0.5
        field[:, 0] = np.where(active,
                               field[:, 0] * rng.uniform(0.55, 0.98, active.shape),
                               rng.uniform(0.0, 0.02, active.shape)).astype(np.float32)
        for ch in regs:
            field[:, ch] += rng.normal(0.0, 0.05, active.shape).astype(np.float32)
        for ch in scales:
            field[:, ch] *= rng.uniform(0.9, 1.1, active.shape).astype(np.float32)
    return cif, caf


def synthetic_fields(image_hw=(641, 641), n_poses=5, *, pose_seed=0, jitter_seed=1,
                     stride=8, dense=False):
    """One image worth of decoder inputs. Deterministic in the two seeds."""
    enc = FieldEncoder.dense(image_hw, stride=stride) if dense else FieldEncoder(image_hw, stride=stride)
    poses = synthetic_poses(n_poses, image_hw, np.random.RandomState(pose_seed))
    cif, caf = enc(poses)
    cif, caf = jitter(cif, caf, np.random.RandomState(jitter_seed))
    return cif, caf, poses


def random_init_like_fields(field_hw, seed=0, n_links=19, n_keypoints=17, scale_sigma=3.0,
                            n_nonfinite=0):
    """Fields with the statistics of a randomly initialised head:
    sigmoid confidences with ~25 % of the cells above 0.1, regressions a few cells off
    their cell, and ``exp()`` scales with a heavy tail (sigmas from 1e-4 to 1e4 cells, so
    some sources cover the whole hr map and many fall outside it).  ``n_nonfinite`` cells
    additionally get NaN / inf coordinates or scales (SURVEY appendix B edge cases).

    Workload tooling for parity tests and stress runs; not on the decode path.
    """
    H, W = field_hw
    rng = np.random.RandomState(seed)
    yy, xx = np.mgrid[0:H, 0:W].astype(np.float32)

    def block(n, pairs):
        f = np.zeros((n, 1 + 4 * pairs, H, W), np.float32)
        f[:, 0] = 1.0 / (1.0 + np.exp(-rng.normal(-1.0, 1.5, (n, H, W))))
        for p in range(pairs):
            a = 1 + 4 * p
            f[:, a] = xx + rng.normal(0, 3, (n, H, W))
            f[:, a + 1] = yy + rng.normal(0, 3, (n, H, W))
            f[:, a + 2] = rng.normal(0, 1, (n, H, W))
            f[:, a + 3] = np.exp(rng.normal(0, scale_sigma, (n, H, W)))
        return f

    cif = block(n_keypoints, 1)
    caf = block(n_links, 2)
    specials = (np.nan, np.inf, -np.inf, 1e30, -1e30)
    for k in range(n_nonfinite):
        for f, chans in ((cif, (1, 2, 4)), (caf, (1, 2, 4, 5, 6, 8))):
            i = rng.randint(f.shape[0])
            y, x = rng.randint(H), rng.randint(W)
            f[i, chans[rng.randint(len(chans))], y, x] = specials[k % len(specials)]
            f[i, 0, y, x] = max(f[i, 0, y, x], 0.5)        # make sure the cell is active
    return cif.astype(np.float32), caf.astype(np.float32)


def synthetic_det_fields(image_hw=(641, 641), n_boxes=8, *, box_seed=0, jitter_seed=1,
                         n_categories=5, stride=8, duplicates=True):
    """Synthetic CifDet head output ``(C, 7, H, W)``, channels ``[v, x, y, _, w, h, _]`` in
    field units (the layout ``CifDet`` reads, generator/cifdet.py + cif_hr.py:84-100).

    Every box lights up the cells around its centre (confidence falling off with distance,
    regressions pointing at the centre, jittered sizes); with ``duplicates`` some boxes get a
    shifted twin of the same category so that the IoU soft-NMS has work to do.  Returns
    ``(field, boxes)`` with ``boxes`` rows ``(category, cx, cy, w, h)`` in pixels.
    """
    H, W = (image_hw[0] - 1) // stride + 1, (image_hw[1] - 1) // stride + 1
    rng = np.random.RandomState(box_seed)
    boxes = []
    for _ in range(n_boxes):
        w, h = rng.uniform(40, 0.45 * image_hw[1]), rng.uniform(40, 0.45 * image_hw[0])
        cx, cy = rng.uniform(0.5 * w, image_hw[1] - 0.5 * w), rng.uniform(0.5 * h, image_hw[0] - 0.5 * h)
        cat = rng.randint(n_categories)
        boxes.append((cat, cx, cy, w, h))
        if duplicates and rng.uniform() < 0.5:
            shift = rng.choice([0.08, 0.2, 0.5])
            boxes.append((cat, cx + shift * w, cy, w * rng.uniform(0.9, 1.1), h))
    jr = np.random.RandomState(jitter_seed)
    field = np.zeros((n_categories, 7, H, W), np.float32)
    field[:, 0] = jr.uniform(0.0, 0.05, (n_categories, H, W))
    field[:, 4:6] = 1.0
    ys, xs = np.mgrid[0:H, 0:W].astype(np.float32)
    for cat, cx, cy, w, h in boxes:
        fx, fy = cx / stride, cy / stride
        d2 = (xs - fx) ** 2 + (ys - fy) ** 2
        near = d2 < 2.5 ** 2
        peak = jr.uniform(0.55, 0.95)
        conf = (peak * np.exp(-0.5 * d2 / 1.2 ** 2) * jr.uniform(0.9, 1.0, d2.shape)).astype(np.float32)
        upd = near & (conf > field[cat, 0])
        field[cat, 0][upd] = conf[upd]
        field[cat, 1][upd] = (fx + jr.normal(0, 0.05, d2.shape))[upd]
        field[cat, 2][upd] = (fy + jr.normal(0, 0.05, d2.shape))[upd]
        field[cat, 3][upd] = jr.uniform(-1, 1, d2.shape)[upd]
        field[cat, 4][upd] = (w / stride * jr.uniform(0.95, 1.05, d2.shape))[upd]
        field[cat, 5][upd] = (h / stride * jr.uniform(0.95, 1.05, d2.shape))[upd]
        field[cat, 6][upd] = jr.uniform(-1, 1, d2.shape)[upd]
    return field, np.array(boxes, np.float64)


def synthetic_multiscale_fields(image_hw=(641, 641), n_poses=5, *, strides=(8, 16, 24), pose_seed=0,
                                jitter_seed=1):
    """The same scene at several head strides, for multi-scale FieldConfig tests: one (cif, caf)
    pair per stride, made by running the generator on the image shrunk by ``stride / 8`` (the
    generator places poses relative to the image, so equal seeds give the same scene).
    Returns ``[(cif, caf, stride), ...]``; set 0 defines the hr shape."""
    out = []
    for k, stride in enumerate(strides):
        hw = tuple((side - 1) * 8 // stride + 1 for side in image_hw)
        cif, caf, _ = synthetic_fields(hw, n_poses, pose_seed=pose_seed, jitter_seed=jitter_seed + 17 * k)
        out.append((cif, caf, stride))
    return out
