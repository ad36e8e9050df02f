"""Builds the host-side views (snk_frame_view & co.) the projection / BoW matchers take, from extractor output.
Shared by the CPU and GPU matcher tests; keeps every numpy buffer alive on the returned object."""
import numpy as np

import snakeorb
from snakeorb import synthetic as S


class FrameData:
    pass


def make_frame(kps32, desc, w, h, rng, pose_q, pose_t, prematched_frac=0.15, right_frac=0.5, bf=386.1):
    k64 = snakeorb.kp32_to_kp64(kps32)
    bounds, cells, perm = S.feature_grid(k64, w, h)
    k64, desc = S.apply_permutation(perm, k64, np.ascontiguousarray(desc))
    n = len(k64)
    f = FrameData()
    f.kp = np.ascontiguousarray(k64)
    f.desc = np.ascontiguousarray(desc)
    f.right = np.where(rng.random(n) < right_frac, f.kp["x"] - bf / rng.uniform(3, 60, n), -1000.0).astype(np.float32)
    f.depth = np.where(f.right > 0, bf / np.maximum(f.kp["x"] - f.right, 1e-3), -1.0).astype(np.float32)
    f.right = np.where(f.depth > 0, f.right, -1.0).astype(np.float32)  # right_points and depth are set together
    f.k2p = np.where(rng.random(n) < prematched_frac, rng.integers(100000, 200000, n), -1).astype(np.int32)
    f.cells = np.ascontiguousarray(cells)
    v = snakeorb.FrameView()
    v.n = n
    v.kp_un, v.desc, v.right_points, v.kp_to_point = f.kp.ctypes.data, f.desc.ctypes.data, f.right.ctypes.data, f.k2p.ctypes.data
    v.depth = f.depth.ctypes.data
    for i in range(4):
        v.pose_q[i] = pose_q[i]
    for i in range(3):
        v.pose_t[i] = pose_t[i]
    g = v.grid
    g.bmin_x, g.bmin_y, g.bmax_x, g.bmax_y = bounds["bmin_x"], bounds["bmin_y"], bounds["bmax_x"], bounds["bmax_y"]
    g.cell_size_inv_x = g.cell_size_inv_y = bounds["inv"]
    g.rows, g.cols = bounds["rows"], bounds["cols"]
    g.cells = f.cells.ctypes.data
    f.view = v
    return f


def camera(fx, fy, cx, cy, bf):
    c = snakeorb.Camera()
    c.fx, c.fy, c.cx, c.cy, c.s, c.bf = fx, fy, cx, cy, 0.0, bf
    return c


def fine_points(mp, rng, invalid_frac=0.05):
    p = np.zeros(len(mp["pos"]), snakeorb.POINT_FINE)
    p["desc"], p["pos"], p["normal"] = mp["desc"], mp["pos"], mp["normal"]
    p["ref_depth"], p["ref_level"] = mp["ref_depth"], mp["ref_level"]
    p["valid"] = (rng.random(len(p)) >= invalid_frac).astype(np.int32)
    p["id"] = np.arange(len(p), dtype=np.int32) + 7
    return p


def coarse_points(mp):
    p = np.zeros(len(mp["pos"]), snakeorb.POINT_COARSE)
    p["desc"], p["pos"], p["normal"] = mp["desc"], mp["pos"], mp["normal"]
    p["octave"], p["angle"] = mp["octave"], mp["angle"]
    p["id"] = np.arange(len(p), dtype=np.int32) + 7
    return p


def kf_points(mp, rng, null_frac=0.1):
    p = np.zeros(len(mp["pos"]), snakeorb.POINT_KF)
    p["desc"], p["pos"] = mp["desc"], mp["pos"]
    ids = np.arange(len(p), dtype=np.int32) + 7
    ids[rng.random(len(p)) < null_frac] = -1
    p["id"] = ids
    return p


def featvec(desc, rng, n_nodes=100):
    if len(desc) == 0:  # a keyframe without features: no nodes, offsets = [0]
        ids, off, idx = np.zeros(1, np.int32), np.zeros(1, np.int32), np.zeros(1, np.int32)
        fv = snakeorb.FeatVecView()
        fv.n_nodes = 0
        fv.node_ids, fv.offsets, fv.indices = ids.ctypes.data, off.ctypes.data, idx.ctypes.data
        fv._keep = (ids, off, idx)
        return fv
    ids, off, idx = S.bow_feature_vector(desc, rng, n_nodes)
    fv = snakeorb.FeatVecView()
    fv.n_nodes = len(ids)
    fv.node_ids, fv.offsets, fv.indices = ids.ctypes.data, off.ctypes.data, idx.ctypes.data
    fv._keep = (ids, off, idx)
    return fv


def fusion_points(mp, rng):
    p = np.zeros(len(mp["pos"]), snakeorb.POINT_FUSION)
    p["desc"], p["pos"], p["normal"] = mp["desc"], mp["pos"], mp["normal"]
    p["ref_depth"], p["ref_level"] = mp["ref_depth"], mp["ref_level"]
    p["observations"] = rng.integers(1, 7, len(p))
    p["id"] = np.arange(len(p), dtype=np.int32) * 3 + 5
    return p


def normalized_points(frame, K):
    fx, fy, cx, cy = K
    return np.ascontiguousarray(np.stack([(frame.kp["x"] - cx) / fx, (frame.kp["y"] - cy) / fy], 1))


def stereo_keyframes(feats, rng, K, w, h, baseline):
    """Two keyframes looking at the synthetic stereo pair: kf1 at the identity, kf2 translated by `baseline` along x
    (camera-frame translation -baseline), E = [t]x for t = (1, 0, 0): epipolar lines are image rows."""
    kl, dl = feats["l"]
    kr, dr = feats["r"]
    q = np.array([0.0, 0.0, 0.0, 1.0])
    f1 = make_frame(kl, dl, w, h, rng, q, np.zeros(3), prematched_frac=0.2)
    f2 = make_frame(kr, dr, w, h, rng, q, np.array([-baseline, 0.0, 0.0]), prematched_frac=0.2)
    E = np.array([[0.0, 0.0, 0.0], [0.0, 0.0, -1.0], [0.0, 1.0, 0.0]])
    return f1, f2, normalized_points(f1, K), normalized_points(f2, K), E
