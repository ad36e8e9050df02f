"""Seeded input generators for the Snake matcher parity tests. One `case_<matcher>(seed)` per reference function;
the same cases feed (a) the live pin of the CPU restatement against the reference's own compiled translation units
(test_oracle_snake_live.py, thousands of seeds), (b) tests/golden/make_golden_snake.py, which freezes the compiled
reference's outputs for a subset of the seeds, and (c) the GPU parity tests, which regenerate the inputs from the seed
(checked against a stored digest) and compare the kernels with those frozen outputs.

Inputs are synthetic on purpose: random keypoints on a 20-px feature grid and descriptors drawn from a small pool with
a few flipped bits, so that equal distances (ties), crowded search windows and second-best decisions are the norm, not
the exception. Everything is derived from numpy's default_rng(seed) only."""
import hashlib

import numpy as np

import matchdata as M
import snakeorb
from snakeorb import synthetic as S

KP64 = S.KP64
LEVELS = 8


def _scales():
    s = [1.0]
    f = float(np.float32(1.2))
    for _ in range(LEVELS - 1):
        s.append(f * s[-1])  # ScalePyramid.cpp: scale[i] = scale_factor * scale[i-1], scale_factor a widened float
    return np.array(s)


SCALES = _scales()


def pyramid():
    return snakeorb.scale_pyramid_c(1.2, SCALES)


def pool_descriptors(rng, n, pool, max_flips):
    """n descriptors: noisy copies of `pool` random ones (pool small => many exact ties)."""
    base = rng.integers(0, 2**63, (pool, 4)).astype(np.uint64) * np.uint64(2) + rng.integers(0, 2, (pool, 4)).astype(np.uint64)
    return S.flip_bits(base[rng.integers(0, pool, n)], rng, max_flips), base


def keypoints(rng, n, w, h, border=0.0):
    k = np.zeros(n, KP64)
    k["x"] = rng.uniform(-border, w - 1 + border, n)
    k["y"] = rng.uniform(-border, h - 1 + border, n)
    if rng.random() < 0.5:  # integer-valued level-0 coordinates scaled by the level, as the extractor produces them
        k["octave"] = rng.integers(0, LEVELS, n)
        k["x"] = np.float32(np.round(k["x"] / SCALES[k["octave"]])) * np.float32(SCALES[k["octave"]])
        k["y"] = np.float32(np.round(k["y"] / SCALES[k["octave"]])) * np.float32(SCALES[k["octave"]])
    else:
        k["octave"] = rng.integers(0, LEVELS, n)
    k["angle"] = np.float32(rng.uniform(0, 360, n))
    k["size"] = 31.0 * SCALES[k["octave"]]
    k["response"] = rng.integers(7, 120, n)
    return k


def frame(rng, n, w, h, pool=24, flips=6, pose=None, prematched=0.15, right_frac=0.5, bf=40.0, border=3.0):
    """A FrameData (tests/matchdata.py) from synthetic keypoints, grid-ordered like Preprocess::computeFeatureGrid leaves them."""
    k = keypoints(rng, n, w, h, border)
    d, base = pool_descriptors(rng, n, pool, flips)
    bounds, cells, perm = S.feature_grid(k, w, h)
    k, d = S.apply_permutation(perm, k, np.ascontiguousarray(d))
    f = M.FrameData()
    f.kp, f.desc, f.base = np.ascontiguousarray(k), np.ascontiguousarray(d), base
    z = rng.uniform(0.5, 30, n)
    f.right = np.where(rng.random(n) < right_frac, f.kp["x"] - bf / z, -1.0).astype(np.float32)
    f.depth = np.where(f.right > 0, bf / np.maximum(f.kp["x"] - f.right, 1e-3), -1.0).astype(np.float32)
    f.right = np.where(f.depth > 0, f.right, -1.0).astype(np.float32)
    f.k2p = np.where(rng.random(n) < prematched, rng.integers(100000, 100000 + 4 * n + 1, n), -1).astype(np.int32)
    f.cells = np.ascontiguousarray(cells)
    q, t = pose if pose is not None else (np.array([0.0, 0.0, 0.0, 1.0]), np.zeros(3))
    v = snakeorb.FrameView()
    v.n = n
    v.kp_un, v.desc, v.right_points, v.kp_to_point = f.kp.ctypes.data, f.desc.ctypes.data, f.right.ctypes.data, f.k2p.ctypes.data
    v.depth = f.depth.ctypes.data
    for i in range(4):
        v.pose_q[i] = q[i]
    for i in range(3):
        v.pose_t[i] = t[i]
    g = v.grid
    g.bmin_x, g.bmin_y, g.bmax_x, g.bmax_y = bounds["bmin_x"], bounds["bmin_y"], bounds["bmax_x"], bounds["bmax_y"]
    g.cell_size_inv_x = g.cell_size_inv_y = bounds["inv"]
    g.rows, g.cols = bounds["rows"], bounds["cols"]
    g.cells = f.cells.ctypes.data
    f.view, f.w, f.h = v, w, h
    return f


def small_pose(rng, max_deg=1.0, max_t=0.1):
    q = S.quat_from_axis_angle(rng.normal(size=3), np.radians(rng.uniform(0, max_deg)))
    return q, rng.uniform(-max_t, max_t, 3)


def map_points(rng, f, K, n_points, flips=10, depth=(0.6, 40.0), off_view=0.1):
    """Map points back-projected from the frame's keypoints (world == camera frame before the pose perturbation);
    a fraction faces away / sits behind the camera / outside the scale range so that every cull fires."""
    mp = S.map_points_from_keypoints(f.kp, f.desc, K, rng, n_points, SCALES, max_flips=flips, depth_range=depth)
    n = n_points
    bad = rng.random(n) < off_view
    mp["normal"][bad] = rng.normal(size=(int(bad.sum()), 3))
    mp["normal"] /= np.linalg.norm(mp["normal"], axis=1, keepdims=True)
    behind = rng.random(n) < 0.03
    mp["pos"][behind, 2] *= -1
    far = rng.random(n) < 0.08
    mp["ref_depth"][far] *= np.float32(rng.choice([0.15, 6.0]))
    mp["ref_level"] = np.clip(mp["ref_level"] + rng.integers(-1, 2, n), 0, LEVELS - 1).astype(np.int32)
    mp["octave"] = np.clip(mp["octave"] + rng.integers(-1, 2, n), 0, LEVELS - 1).astype(np.int32)
    return mp


def cam_of(rng, w, h):
    fx = rng.uniform(0.7, 1.3) * w * 0.7
    K = (fx, fx * rng.uniform(0.97, 1.03), w / 2 + rng.uniform(-20, 20), h / 2 + rng.uniform(-20, 20))
    bf = fx * rng.uniform(0.05, 0.55)
    return K, bf, M.camera(*K, bf)


def _size(rng, big):
    if big:
        return int(rng.integers(900, 2100)), (752, 480) if rng.random() < 0.5 else (1241, 376)
    return int(rng.integers(40, 260)), (int(rng.integers(200, 420)), int(rng.integers(160, 300)))


# ------------------------------------------------------------------------------------------------------------------
# one generator per matcher: returns a dict of everything the three implementations need
# ------------------------------------------------------------------------------------------------------------------
def case_stereo(seed, big=False):
    rng = np.random.default_rng([seed, 1])
    n, (w, h) = _size(rng, big)
    left = keypoints(rng, n, w, h)
    dl, _ = pool_descriptors(rng, n, max(4, n // 6), 5)
    keep = rng.random(n) < 0.85
    right = left[keep].copy()
    disp = rng.uniform(-2.0, 60.0, len(right))
    right["x"] = right["x"] - disp
    right["y"] += rng.normal(0, 1.2, len(right))
    right["octave"] = np.clip(right["octave"] + rng.choice([-2, -1, 0, 0, 0, 1, 2], len(right)), 0, LEVELS - 1)
    right["angle"] = np.float32((right["angle"] + rng.normal(0, 9, len(right))) % 360)
    dr = S.flip_bits(dl[keep], rng, 30)
    extra = keypoints(rng, int(rng.integers(0, n // 3 + 1)), w, h)
    de, _ = pool_descriptors(rng, len(extra), 8, 5)
    right = np.concatenate([right, extra])
    dr = np.concatenate([dr, de])
    order = rng.permutation(len(right))
    right, dr = np.ascontiguousarray(right[order]), np.ascontiguousarray(dr[order])
    fx = rng.uniform(300, 800)
    rl, rr = snakeorb.Rectify(), snakeorb.Rectify()
    identity = rng.random() < 0.5
    for r in (rl, rr):
        r.fx, r.fy, r.cx, r.cy = fx, fx * 1.01, w / 2 + rng.uniform(-5, 5), h / 2 + rng.uniform(-5, 5)
        q = np.array([0.0, 0.0, 0.0, 1.0]) if identity else S.quat_from_axis_angle(rng.normal(size=3), np.radians(rng.uniform(0, 0.4)))
        for i in range(4):
            r.quat[i] = q[i]
        dist = np.zeros(8) if identity else np.array([rng.uniform(-0.3, 0.1), rng.uniform(-0.1, 0.1), 0, 0, 0, 0,
                                                        rng.uniform(-1e-3, 1e-3), rng.uniform(-1e-3, 1e-3)])
        for i in range(8):
            r.dist[i] = dist[i]
        r.dfx, r.dfy, r.dcx, r.dcy, r.ds = (r.fx, r.fy, r.cx, r.cy, 0.0) if identity else (fx * 0.98, fx * 0.98, w / 2, h / 2, 0.0)
    bf = float(rng.uniform(20, 130))
    return dict(left=left, dl=np.ascontiguousarray(dl), right=right, dr=dr, rl=rl, rr=rr, bf=bf, relaxed=bool(rng.random() < 0.7),
                w=w, h=h)


def case_rgbd(seed, big=False):
    rng = np.random.default_rng([seed, 2])
    n, (w, h) = _size(rng, big)
    kp = keypoints(rng, n, w, h, border=-2.0)  # strictly inside: the reference asserts the lookup lands in the image
    p = snakeorb.RgbdParams()
    fx = rng.uniform(400, 600)
    p.fx, p.fy, p.cx, p.cy = fx, fx * 1.01, w / 2, h / 2
    p.dfx, p.dfy, p.dcx, p.dcy, p.ds = fx, fx * 1.01, w / 2, h / 2, 0.0
    dist = np.zeros(8) if rng.random() < 0.5 else np.array([rng.uniform(-0.05, 0.05), rng.uniform(-0.02, 0.02), 0, 0, 0, 0, 1e-4, -1e-4])
    for i in range(8):
        p.dist[i] = dist[i]
    p.bf = float(rng.uniform(20, 60))
    depth = rng.uniform(0.4, 8.0, (h + 2, w + 2)).astype(np.float32)
    depth[rng.random(depth.shape) < 0.15] = 0.0
    return dict(kp=kp, p=p, depth=np.ascontiguousarray(depth))


def _projection_inputs(rng, big, th_choices):
    n, (w, h) = _size(rng, big)
    K, bf, cam = cam_of(rng, w, h)
    f = frame(rng, n, w, h, pool=max(4, n // 8), flips=6, pose=small_pose(rng), bf=bf)
    n_points = int(rng.integers(0, 4 * n)) if not big else int(rng.integers(5000, 20000))
    mp = map_points(rng, f, K[:4], n_points) if n_points else dict(pos=np.zeros((0, 3)), normal=np.zeros((0, 3)),
                                                                   desc=np.zeros((0, 4), np.uint64), ref_depth=np.zeros(0, np.float32),
                                                                   ref_level=np.zeros(0, np.int32), octave=np.zeros(0, np.int32),
                                                                   angle=np.zeros(0, np.float32))
    return rng, f, mp, cam, float(rng.choice(th_choices))


def case_fine(seed, big=False):
    rng, f, mp, cam, th = _projection_inputs(np.random.default_rng([seed, 3]), big, [1.0, 4.0, 5.0, 2.5])
    return dict(f=f, pts=M.fine_points(mp, rng), cam=cam, th=th, ratio=float(rng.choice([0.8, 0.6, 0.9])))


def case_coarse(seed, big=False):
    rng, f, mp, cam, th = _projection_inputs(np.random.default_rng([seed, 4]), big, [15.0, 10.0, 30.0, 7.0])
    return dict(f=f, pts=M.coarse_points(mp), cam=cam, th=th, fe=int(rng.choice([75, 50, 100])), motion=int(rng.integers(0, 3)))


def case_kf(seed, big=False):
    rng, f, mp, cam, th = _projection_inputs(np.random.default_rng([seed, 5]), big, [5.0, 10.0, 15.0])
    pts = M.kf_points(mp, rng)
    # some keyframe points are the very points the frame already holds (sAlreadyFound)
    held = f.k2p[f.k2p >= 0]
    if len(held) and len(pts):
        sel = rng.integers(0, len(pts), min(len(pts) // 10 + 1, len(held)))
        pts["id"][sel] = held[rng.integers(0, len(held), len(sel))]
    # ids must be unique among the keyframe's points (one MapPoint object per feature) except the nulls
    _, first = np.unique(pts["id"], return_index=True)
    dup = np.ones(len(pts), bool)
    dup[first] = False
    pts["id"][dup & (pts["id"] >= 0)] = -1
    return dict(f=f, pts=pts, cam=cam, th=th, fe=int(rng.choice([50, 75, 60])))


def case_bow(seed, big=False):
    rng = np.random.default_rng([seed, 6])
    n1 = int(rng.integers(30, 300)) if not big else int(rng.integers(900, 1300))
    n2 = int(rng.integers(30, 300)) if not big else int(rng.integers(900, 1300))
    pool = max(4, n1 // 5)
    base = rng.integers(0, 2**63, (pool, 4)).astype(np.uint64)
    d1 = S.flip_bits(base[rng.integers(0, pool, n1)], rng, 12)
    d2 = S.flip_bits(base[rng.integers(0, pool, n2)], rng, 12)
    nodes = int(rng.integers(3, 60)) if not big else 100
    fv1, fv2 = M.featvec(d1, rng, nodes), M.featvec(d2, rng, nodes)
    # 1 = good map point, 0 = none, 2 = bad map point
    v1 = rng.choice([0, 1, 1, 1, 2], n1).astype(np.uint8)
    v2 = rng.choice([0, 1, 1, 1, 2], n2).astype(np.uint8)
    return dict(d1=np.ascontiguousarray(d1), d2=np.ascontiguousarray(d2), fv1=fv1, fv2=fv2, v1=v1, v2=v2,
                th=int(rng.choice([50, 40, 70])), ratio=float(rng.choice([0.75, 0.6, 0.9])))


def case_bow_batch(seed, big=False):
    """One keyframe against several loop candidates (LoopDetector.cpp:384-391): candidates of different sizes from the
    same descriptor pool, now and then one without features or one that shares no vocabulary node."""
    rng = np.random.default_rng([seed, 14])
    n1 = int(rng.integers(30, 300)) if not big else int(rng.integers(900, 1300))
    pool = max(4, n1 // 5)
    base = rng.integers(0, 2**63, (pool, 4)).astype(np.uint64)
    nodes = int(rng.integers(3, 60)) if not big else 100
    d1 = np.ascontiguousarray(S.flip_bits(base[rng.integers(0, pool, n1)], rng, 12))
    fv1 = M.featvec(d1, rng, nodes)
    v1 = rng.choice([0, 1, 1, 1, 2], n1).astype(np.uint8)
    cands = []
    for k in range(int(rng.integers(1, 7)) if not big else 12):
        kind = rng.random()
        n2 = 0 if kind < 0.1 else (int(rng.integers(1, 300)) if not big else int(rng.integers(900, 1300)))
        d2 = np.ascontiguousarray(S.flip_bits(base[rng.integers(0, pool, n2)], rng, 12)).reshape(-1, 4)
        fv2 = M.featvec(d2, rng, nodes)
        if 0.1 <= kind < 0.2:  # node ids the first keyframe does not have
            fv2._keep[0][:] += 100000
        cands.append((d2, fv2, rng.choice([0, 1, 1, 1, 2], n2).astype(np.uint8)))
    return dict(d1=d1, fv1=fv1, v1=v1, cands=cands, th=int(rng.choice([50, 40, 70])), ratio=float(rng.choice([0.75, 0.6, 0.9])))


def case_fuse(seed, big=False):
    rng, f, mp, cam, th = _projection_inputs(np.random.default_rng([seed, 7]), big, [3.0, 2.5, 4.0, 6.0])
    pts = M.fusion_points(mp, rng)
    mask = (rng.random(len(pts)) < 0.8).astype(np.uint8) if rng.random() < 0.6 else None
    return dict(f=f, pts=pts, cam=cam, th=th, obs=float(rng.choice([2.0, 1.0, 1.5])), fth=int(rng.choice([50, 60, 40])), mask=mask)


def case_fuse_loop(seed, big=False):
    rng, f, mp, cam, th = _projection_inputs(np.random.default_rng([seed, 8]), big, [4.0, 3.0, 10.0])
    pts = M.fusion_points(mp, rng)
    skip = (rng.random(len(pts)) < 0.1).astype(np.uint8)  # 1 = bad point
    # the keyframe's map points are unique objects: one id per occupied feature
    occ = np.nonzero(f.k2p >= 0)[0]
    f.k2p[occ] = 100000 + np.arange(len(occ), dtype=np.int32)
    return dict(f=f, pts=pts, cam=cam, th=th, skip=skip)


def _two_keyframes(rng, big):
    n, (w, h) = _size(rng, big)
    K, bf, cam = cam_of(rng, w, h)
    base_t = rng.uniform(0.05, 0.3)
    q = np.array([0.0, 0.0, 0.0, 1.0])
    f1 = frame(rng, n, w, h, pool=max(4, n // 8), flips=6, pose=(q, np.zeros(3)), prematched=0.2, bf=bf)
    # second view: the same points moved by a disparity, plus new ones
    k2 = f1.kp.copy()
    k2["x"] -= rng.uniform(0, 25, n)
    k2["y"] += rng.normal(0, 0.8, n)
    d2 = S.flip_bits(f1.desc, rng, 14)
    f2 = frame(rng, n, w, h, pool=max(4, n // 8), flips=6, pose=(q, np.array([-base_t, 0.0, 0.0])), prematched=0.2, bf=bf)
    bounds, cells, perm = S.feature_grid(k2, w, h)
    k2, d2 = S.apply_permutation(perm, k2, np.ascontiguousarray(d2))
    f2.kp[:], f2.desc[:], f2.cells[:] = k2, d2, cells
    E = np.array([[0.0, 0.0, 0.0], [0.0, 0.0, -1.0], [0.0, 1.0, 0.0]])
    if rng.random() < 0.5:  # a general essential matrix: small rotation on top
        a = rng.normal(0, 0.02, 3)
        R = np.eye(3) + np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
        E = E @ R
    return rng, f1, f2, M.normalized_points(f1, K[:4]), M.normalized_points(f2, K[:4]), np.ascontiguousarray(E), cam, K


def case_tri_bow(seed, big=False):
    rng, f1, f2, np1, np2, E, cam, K = _two_keyframes(np.random.default_rng([seed, 9]), big)
    nodes = int(rng.integers(3, 50))
    return dict(f1=f1, f2=f2, np1=np1, np2=np2, E=E, cam=cam, fv1=M.featvec(f1.desc, rng, nodes), fv2=M.featvec(f2.desc, rng, nodes),
                epi=float(rng.choice([3.0, 1.5, 6.0])), fd=int(rng.choice([50, 40, 75])))


def case_tri_project(seed, big=False):
    rng, f1, f2, np1, np2, E, cam, K = _two_keyframes(np.random.default_rng([seed, 10]), big)
    g = f1.view.grid
    grid = rng.uniform(0.8, 20.0, ((g.rows + 3) // 4 + 1, (g.cols + 3) // 4 + 1))
    return dict(f1=f1, f2=f2, np1=np1, np2=np2, E=E, cam=cam, grid=np.ascontiguousarray(grid),
                epi=float(rng.choice([3.0, 1.5, 6.0])), fd=int(rng.choice([50, 40, 75])))


def case_prediction(seed, big=False):
    rng = np.random.default_rng([seed, 11])
    n, (w, h) = _size(rng, big)
    f = frame(rng, n, w, h, pool=max(4, n // 8), flips=8)
    nl = int(rng.integers(1, n + 40))
    src = rng.integers(0, n, nl)
    dl = S.flip_bits(f.desc[src], rng, 25)
    pos = np.stack([f.kp["x"][src], f.kp["y"][src]], 1) + rng.normal(0, 15.0, (nl, 2))
    return dict(f=f, dl=np.ascontiguousarray(dl), pos=np.ascontiguousarray(pos))


def digest(*arrays):
    h = hashlib.sha256()
    for a in arrays:
        h.update(np.ascontiguousarray(a).tobytes())
    return h.hexdigest()[:16]
