"""TEST INFRASTRUCTURE ONLY — ctypes view of oracle/liborb_oracle.so (the CPU restatement) and, when present,
oracle/_ref/libsaiga_orb_ref.so (the reference's own translation units). Imported by tests/,
__graft_entry__.smoke() and the cpu_baseline / --impl reference legs of bench.py — never by the product."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
KP32 = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4")])
KP64 = np.dtype([("x", "<f8"), ("y", "<f8"), ("size", "<f8"), ("angle", "<f8"), ("response", "<f8"),
                 ("octave", "<i4"), ("_pad", "<i4")])
KNN2 = np.dtype([("dist0", "<i4"), ("idx0", "<i4"), ("dist1", "<i4"), ("idx1", "<i4")])
assert KP32.itemsize == 24 and KP64.itemsize == 48


def build(force=False):
    so = os.path.join(HERE, "liborb_oracle.so")
    if force or not os.path.exists(so) or os.path.isdir("/root/reference"):
        subprocess.check_call(["make", "-s", "-C", HERE, "all"])
    return so


_lib = None
_ref = None


def lib():
    global _lib
    if _lib is None:
        so = os.environ.get("SNK_ORACLE_SO") or os.path.join(HERE, "liborb_oracle.so")  # tools/cpu_anchor.py swaps the build
        if not os.path.exists(so):
            build()
        _lib = C.CDLL(so)
        _lib.orc_extractor_create.restype = C.c_void_p
    return _lib


def ref():
    """The compiled reference pieces, or None when oracle/_ref was not built (no /root/reference)."""
    global _ref
    if _ref is None:
        so = os.path.join(HERE, "_ref", "libsaiga_orb_ref.so")
        if not os.path.exists(so):
            return None
        _ref = C.CDLL(so)
    return _ref


def _p(a):
    return C.c_void_p(a.ctypes.data)


def resize(src, dh, dw):
    src = np.ascontiguousarray(src)
    dst = np.zeros((dh, dw), np.uint8)
    lib().orc_resize(_p(src), src.shape[0], src.shape[1], C.c_size_t(src.shape[1]), _p(dst), dh, dw, C.c_size_t(dw))
    return dst


def rgba_to_gray(src):
    assert src.dtype == np.uint8 and src.ndim == 3 and src.shape[2] == 4 and src.strides[1:] == (4, 1)
    dst = np.zeros(src.shape[:2], np.uint8)
    lib().orc_rgba_to_gray(_p(src), src.shape[0], src.shape[1], C.c_size_t(src.strides[0]), _p(dst),
                           C.c_size_t(src.shape[1]))
    return dst


def blur(src):
    src = np.ascontiguousarray(src)
    dst = np.zeros_like(src)
    lib().orc_blur(_p(src), src.shape[0], src.shape[1], C.c_size_t(src.shape[1]), _p(dst), C.c_size_t(src.shape[1]))
    return dst


def fast(roi_parent, y0, x0, h, w, th):
    """FAST on the ROI [y0:y0+h, x0:x0+w] of a C-contiguous parent; returns int32 (n,3) x,y,score."""
    out = np.zeros((h * w, 3), np.int32)
    ptr = C.c_void_p(roi_parent.ctypes.data + y0 * roi_parent.strides[0] + x0)
    n = lib().orc_fast(ptr, h, w, C.c_size_t(roi_parent.strides[0]), th, _p(out), h * w)
    return out[:n]


def scale_pyramid(levels, scale_factor, nfeatures):
    s = np.zeros(levels, np.float64)
    f = np.zeros(levels, np.int32)
    lib().orc_scale_pyramid(levels, C.c_float(scale_factor), nfeatures, _p(s), _p(f))
    return s, f


def cell_grid(rows, cols):
    o = np.zeros(7, np.int32)
    lib().orc_cell_grid(rows, cols, _p(o))
    return dict(zip(["nCols", "nRows", "wCell", "hCell", "minB", "maxBX", "maxBY"], [int(v) for v in o]))


def angles(img, xy):
    img = np.ascontiguousarray(img)
    xy = np.ascontiguousarray(xy, np.float32)
    out = np.zeros(len(xy), np.float32)
    lib().orc_angles(_p(img), C.c_size_t(img.shape[1]), _p(xy), len(xy), _p(out))
    return out


def descriptors(img, xy, ang):
    img = np.ascontiguousarray(img)
    xy = np.ascontiguousarray(xy, np.float32)
    ang = np.ascontiguousarray(ang, np.float32)
    out = np.zeros((len(xy), 4), np.uint64)
    lib().orc_descriptors(_p(img), C.c_size_t(img.shape[1]), _p(xy), _p(ang), len(xy), _p(out))
    return out


def distribute(kps, mn, mx, target, which="oracle"):
    """Returns (selected records, permuted input)."""
    kps = np.array(kps, dtype=KP32, copy=True)
    out = np.zeros(len(kps) + 8, KP32)
    if which == "oracle":
        fn = lib().orc_distribute
    else:
        fn = ref().ref_distribute
    n = fn(_p(kps), len(kps), C.c_float(mn[0]), C.c_float(mn[1]), C.c_float(mx[0]), C.c_float(mx[1]), target,
           _p(out), len(out))
    return out[:n].copy(), kps


class Extractor:
    """orbref::OrbExtractor == Saiga::ORBExtractor restated (oracle/orb_ref.cpp)."""

    def __init__(self, nfeatures=1000, scale=1.2, nlevels=8, th=20, thmin=7, threads=1):
        self.h = C.c_void_p(lib().orc_extractor_create(nfeatures, C.c_float(scale), nlevels, th, thmin, threads))
        self.nlevels = nlevels
        self.cap = nfeatures + 8 * nlevels + 64

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_extractor_destroy(self.h)
            self.h = None

    def detect(self, img):
        assert img.dtype == np.uint8 and img.ndim == 2 and img.strides[1] == 1
        kps = np.zeros(self.cap, KP32)
        desc = np.zeros((self.cap, 4), np.uint64)
        n = lib().orc_extractor_detect(self.h, _p(img), img.shape[0], img.shape[1], C.c_size_t(img.strides[0]),
                                       _p(kps), _p(desc), self.cap)
        assert n <= self.cap
        return kps[:n].copy(), desc[:n].copy()

    def level(self, l, blurred=False):
        r, c = C.c_int32(), C.c_int32()
        lib().orc_extractor_level_dims(self.h, l, C.byref(r), C.byref(c))
        out = np.zeros((r.value, c.value), np.uint8)
        lib().orc_extractor_level(self.h, l, int(blurred), _p(out))
        return out

    def level_kps(self, l, which):
        out = np.zeros(200000, KP32)
        n = lib().orc_extractor_level_kps(self.h, l, which, _p(out), len(out))
        return out[:n].copy()


def hamming(a, b):
    return lib().orc_hamming(_p(np.ascontiguousarray(a, np.uint64)), _p(np.ascontiguousarray(b, np.uint64)))


def match_knn2(d1, d2, threads=1):
    d1 = np.ascontiguousarray(d1, np.uint64)
    d2 = np.ascontiguousarray(d2, np.uint64)
    out = np.zeros(len(d1), KNN2)
    lib().orc_match_knn2(_p(d1), len(d1), _p(d2), len(d2), _p(out), threads)
    return out


def filter_matches(knn2, th, ratio):
    out = np.zeros((len(knn2), 2), np.int32)
    n = lib().orc_filter_matches(_p(knn2), len(knn2), th, C.c_float(ratio), _p(out))
    return out[:n].copy()


class StereoParams(C.Structure):
    _fields_ = [("bf", C.c_double), ("relaxed", C.c_int32), ("nlevels", C.c_int32), ("scale", C.c_double * 16)]


def stereo_params(bf, relaxed, scales):
    p = StereoParams()
    p.bf = bf
    p.relaxed = int(relaxed)
    p.nlevels = len(scales)
    for i, s in enumerate(scales):
        p.scale[i] = s
    return p


def kp32_to_kp64(k):
    o = np.zeros(len(k), KP64)
    for f in ("x", "y", "size", "angle", "response", "octave"):
        o[f] = k[f]
    return o


def stereo_match(params, left, dl, right, dr):
    left = np.ascontiguousarray(left, KP64)
    right = np.ascontiguousarray(right, KP64)
    dl = np.ascontiguousarray(dl, np.uint64)
    dr = np.ascontiguousarray(dr, np.uint64)
    rp = np.zeros(len(left), np.float32)
    dp = np.zeros(len(left), np.float32)
    ops = C.c_longlong(0)
    n = lib().orc_stereo_match(C.byref(params), _p(left), _p(dl), len(left), _p(right), _p(dr), len(right), _p(rp),
                               _p(dp), C.byref(ops))
    return n, rp, dp, ops.value


# ---- projection / BoW matchers: the caller passes ctypes structs laid out as in include/snakeorb.h ----------------
def search_projection_fine(frame, points, cam, pyr, th, ratio):
    pts = np.array(points, copy=True)
    n = len(pts)
    k2p = np.zeros(max(frame.n, 1), np.int32)
    bi = np.zeros(max(n, 1), np.int32)
    bd = np.zeros(max(n, 1), np.int32)
    vis = np.zeros(max(n, 1), np.uint8)
    ops = C.c_longlong(0)
    m = lib().orc_search_projection_fine(C.byref(frame), _p(pts), n, C.byref(cam), C.byref(pyr), C.c_float(th),
                                         C.c_float(ratio), _p(k2p), _p(bi), _p(bd), _p(vis), C.byref(ops))
    return m, k2p[:frame.n], bi[:n], bd[:n], vis[:n], pts, ops.value


def search_projection_coarse(frame, points, cam, pyr, th, feature_error, motion):
    pts = np.ascontiguousarray(points)
    k2p = np.zeros(max(frame.n, 1), np.int32)
    ops = C.c_longlong(0)
    m = lib().orc_search_projection_coarse(C.byref(frame), _p(pts), len(pts), C.byref(cam), C.byref(pyr),
                                           C.c_float(th), feature_error, motion, _p(k2p), C.byref(ops))
    return m, k2p[:frame.n], ops.value


def search_projection_frame_to_kf(frame, points, cam, th, feature_error):
    pts = np.ascontiguousarray(points)
    k2p = np.zeros(max(frame.n, 1), np.int32)
    ops = C.c_longlong(0)
    m = lib().orc_search_projection_frame_to_kf(C.byref(frame), _p(pts), len(pts), C.byref(cam), C.c_float(th),
                                                feature_error, _p(k2p), C.byref(ops))
    return m, k2p[:frame.n], ops.value


def match_bow(d1, d2, fv1, fv2, valid1, valid2, th, ratio):
    d1 = np.ascontiguousarray(d1, np.uint64).reshape(-1, 4)
    d2 = np.ascontiguousarray(d2, np.uint64).reshape(-1, 4)
    m12 = np.zeros(max(len(d1), 1), np.int32)
    ops = C.c_longlong(0)
    m = lib().orc_match_bow(_p(d1), len(d1), _p(d2), len(d2), C.byref(fv1), C.byref(fv2), _p(valid1), _p(valid2), th,
                            C.c_float(ratio), _p(m12), C.byref(ops))
    return m, m12[:len(d1)], ops.value


def stereo_from_rgbd(params, kp_un, depth_img):
    kp_un = np.ascontiguousarray(kp_un, KP64)
    rp = np.zeros(max(len(kp_un), 1), np.float32)
    dp = np.zeros(max(len(kp_un), 1), np.float32)
    n = lib().orc_stereo_from_rgbd(C.byref(params), _p(kp_un), len(kp_un), _p(depth_img), depth_img.shape[0],
                                   depth_img.shape[1], C.c_size_t(depth_img.strides[0]), _p(rp), _p(dp))
    return n, rp[:len(kp_un)], dp[:len(kp_un)]


def undistort_keypoints(rect, kps):
    kps = np.ascontiguousarray(kps, KP64)
    out = np.zeros(len(kps), KP64)
    nrm = np.zeros((len(kps), 2), np.float64)
    lib().orc_undistort_keypoints(C.byref(rect), _p(kps), len(kps), _p(out), _p(nrm))
    return out, nrm


def feature_grid(bounds, kp_un):
    kp_un = np.ascontiguousarray(kp_un, KP64)
    perm = np.zeros(max(len(kp_un), 1), np.int32)
    cells = np.zeros((bounds.rows * bounds.cols, 2), np.int32)
    lib().orc_feature_grid(C.byref(bounds), _p(kp_un), len(kp_un), _p(perm), _p(cells))
    return perm[:len(kp_un)], cells


class Vocabulary:
    """bowref::Vocabulary == MiniBow2::TemplatedVocabulary restated (oracle/bow_ref.cpp)."""

    def __init__(self, path=None, arrays=None):
        lib().orc_voc_load_raw.restype = C.c_void_p
        lib().orc_voc_from_arrays.restype = C.c_void_p
        lib().orc_bow_score.restype = C.c_float
        if path is not None:
            self.h = C.c_void_p(lib().orc_voc_load_raw(path.encode()))
        else:
            k, L, parent, weight, word_id, desc = arrays
            parent = np.ascontiguousarray(parent, np.int32)
            weight = np.ascontiguousarray(weight, np.float64)
            word_id = np.ascontiguousarray(word_id, np.int32)
            desc = np.ascontiguousarray(desc, np.uint64)
            self.h = C.c_void_p(lib().orc_voc_from_arrays(k, L, len(parent), _p(parent), _p(weight), _p(word_id), _p(desc)))
        if not self.h:
            raise ValueError("bad vocabulary")
        o = np.zeros(4, np.int32)
        lib().orc_voc_info(self.h, _p(o))
        self.k, self.L, self.n_nodes, self.n_words = [int(x) for x in o]

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_voc_destroy(self.h)
            self.h = None

    def save_raw(self, path):
        assert lib().orc_voc_save_raw(self.h, path.encode()) == 0

    def transform(self, desc, levelsup):
        """-> bow_ids, bow_vals, fv_nodes, fv_offsets, fv_indices"""
        desc = np.ascontiguousarray(desc, np.uint64).reshape(-1, 4)
        n = len(desc)
        ids, vals = np.zeros(n + 1, np.int32), np.zeros(n + 1, np.float32)
        fn, fo, fi = np.zeros(n + 1, np.int32), np.zeros(n + 2, np.int32), np.zeros(n + 1, np.int32)
        nb, nf = C.c_int(), C.c_int()
        lib().orc_bow_transform(self.h, _p(desc), n, levelsup, _p(ids), _p(vals), C.byref(nb), _p(fn), _p(fo), _p(fi),
                                C.byref(nf))
        return ids[:nb.value], vals[:nb.value], fn[:nf.value], fo[:nf.value + 1], fi[:fo[nf.value]]


def bow_score(ids1, v1, ids2, v2):
    lib().orc_bow_score.restype = C.c_float
    ids1, ids2 = np.ascontiguousarray(ids1, np.int32), np.ascontiguousarray(ids2, np.int32)
    v1, v2 = np.ascontiguousarray(v1, np.float32), np.ascontiguousarray(v2, np.float32)
    return float(lib().orc_bow_score(_p(ids1), _p(v1), len(ids1), _p(ids2), _p(v2), len(ids2)))


# ---- local mapping / loop closing searches ----
def epipolar_distance_squared(p1, p2, F):
    lib().orc_epipolar_distance_squared.restype = C.c_double
    p1, p2, F = (np.ascontiguousarray(a, np.float64) for a in (p1, p2, F))
    return float(lib().orc_epipolar_distance_squared(_p(p1), _p(p2), _p(F)))


def fuse_search(kf, mask, points, cam, pyr, th, obs_factor, feature_th):
    pts = np.ascontiguousarray(points)
    n = len(pts)
    ck, ci = np.zeros(max(n, 1), np.int32), np.zeros(max(n, 1), np.int32)
    ops = C.c_longlong(0)
    m = lib().orc_fuse_search(C.byref(kf), _p(mask) if mask is not None else None, _p(pts), n, C.byref(cam), C.byref(pyr),
                              C.c_float(th), C.c_float(obs_factor), feature_th, _p(ck), _p(ci), C.byref(ops))
    return ck[:m].copy(), ci[:m].copy(), ops.value


def fuse_search_loop(kf, skip, points, cam, th):
    pts = np.ascontiguousarray(points)
    n = len(pts)
    bi, bd = np.zeros(max(n, 1), np.int32), np.zeros(max(n, 1), np.int32)
    lib().orc_fuse_search_loop(C.byref(kf), _p(skip) if skip is not None else None, _p(pts), n, C.byref(cam),
                               C.c_float(th), _p(bi), _p(bd))
    return bi[:n], bd[:n]


def search_triangulation_bow(kf1, np1, fv1, kf2, np2, fv2, E, cam, epi, fd):
    pairs = np.zeros((max(kf1.n, 1), 2), np.int32)
    ops = C.c_longlong(0)
    E = np.ascontiguousarray(E, np.float64)
    m = lib().orc_search_triangulation_bow(C.byref(kf1), _p(np1), C.byref(fv1), C.byref(kf2), _p(np2), C.byref(fv2), _p(E),
                                           C.byref(cam), C.c_float(epi), fd, _p(pairs), C.byref(ops))
    return pairs[:m].copy(), ops.value


def search_triangulation_project(kf1, np1, kf2, np2, grid, E, cam, epi, fd):
    pairs = np.zeros((max(kf1.n, 1), 2), np.int32)
    ops = C.c_longlong(0)
    E = np.ascontiguousarray(E, np.float64)
    grid = np.ascontiguousarray(grid, np.float64)
    m = lib().orc_search_triangulation_project(C.byref(kf1), _p(np1), C.byref(kf2), _p(np2), _p(grid), grid.shape[0],
                                               grid.shape[1], _p(E), C.byref(cam), C.c_float(epi), fd, _p(pairs),
                                               C.byref(ops))
    return pairs[:m].copy(), ops.value


def match_prediction(desc_left, positions, right, radius, feature_error):
    d = np.ascontiguousarray(desc_left, np.uint64).reshape(-1, 4)
    pos = np.array(positions, np.float64, copy=True)
    matches = np.zeros(max(len(d), 1), np.int32)
    m = lib().orc_match_prediction(_p(d), len(d), _p(pos), C.byref(right), C.c_float(radius), feature_error, _p(matches))
    return m, matches[:len(d)], pos


def features_in_area(frame, px, py, r):
    out = np.zeros(max(frame.n, 1), np.int32)
    c = lib().orc_features_in_area(C.byref(frame), C.c_double(px), C.c_double(py), C.c_float(r), _p(out))
    return out[:c].copy()


# ---- the reference's own Snake matchers (oracle/_ref/libsnake_ref.so, built by the Makefile from the translation units
# where they lie under /root/reference; see ref_snake_shim.cpp). Same arguments and results as the functions above. ----
_snake = None


def snake_ref():
    """The compiled Snake matchers, or None when oracle/_ref/libsnake_ref.so was not built (no /root/reference)."""
    global _snake
    if _snake is None:
        so = os.path.join(HERE, "_ref", "libsnake_ref.so")
        if not os.path.exists(so):
            return None
        _snake = C.CDLL(so)
    return _snake


def snake_adapter():
    """oracle/_ref/libsnake_adapter.so: the same entry points as snake_ref(), but the real Snake objects are driven through
    the drop-in classes of snake-slam_b200/host/snakeorb_snake_adapter.h (libsnakeorb.so, needs a GPU). None if not built."""
    so = os.path.join(HERE, "_ref", "libsnake_adapter.so")
    return C.CDLL(so) if os.path.exists(so) else None


def set_snake_lib(lib):
    """Route the ref_* wrappers below to `lib` (snake_adapter()) or back to the compiled reference (None)."""
    global _snake
    _snake = lib


def ref_stereo_matching(rect_left, rect_right, bf, relaxed, levels, scale_factor, left, dl, right, dr):
    """Preprocess::StereoMatching on raw keypoints (it rectifies them itself)."""
    left = np.ascontiguousarray(left, KP64)
    right = np.ascontiguousarray(right, KP64)
    dl = np.ascontiguousarray(dl, np.uint64)
    dr = np.ascontiguousarray(dr, np.uint64)
    rp = np.zeros(len(left), np.float32)
    dp = np.zeros(len(left), np.float32)
    n = snake_ref().snkref_stereo_matching(C.byref(rect_left), C.byref(rect_right), C.c_double(bf), int(relaxed), levels,
                                           C.c_double(scale_factor), _p(left), _p(dl), len(left), _p(right), _p(dr),
                                           len(right), _p(rp), _p(dp))
    return n, rp, dp


def ref_stereo_from_rgbd(params, kp_un, depth_img):
    kp_un = np.ascontiguousarray(kp_un, KP64)
    rp = np.zeros(max(len(kp_un), 1), np.float32)
    dp = np.zeros(max(len(kp_un), 1), np.float32)
    n = snake_ref().snkref_stereo_from_rgbd(C.byref(params), _p(kp_un), len(kp_un), _p(depth_img), depth_img.shape[0],
                                            depth_img.shape[1], C.c_size_t(depth_img.strides[0]), _p(rp), _p(dp))
    return n, rp[:len(kp_un)], dp[:len(kp_un)]


def ref_undistort_keypoints(rect, kps, rows, cols):
    kps = np.ascontiguousarray(kps, KP64)
    out = np.zeros(len(kps), KP64)
    nrm = np.zeros((len(kps), 2), np.float64)
    snake_ref().snkref_undistort_keypoints(C.byref(rect), _p(kps), len(kps), rows, cols, _p(out), _p(nrm))
    return out, nrm


def ref_search_projection_fine(frame, points, cam, pyr, th, ratio, threads=1):
    pts = np.array(points, copy=True)
    n = len(pts)
    k2p = np.zeros(max(frame.n, 1), np.int32)
    bi = np.zeros(max(n, 1), np.int32)
    bd = np.zeros(max(n, 1), np.int32)
    vis = np.zeros(max(n, 1), np.uint8)
    m = snake_ref().snkref_search_projection_fine(C.byref(frame), _p(pts), n, C.byref(cam), C.byref(pyr), C.c_float(th),
                                                  C.c_float(ratio), _p(k2p), _p(bi), _p(bd), _p(vis), threads)
    return m, k2p[:frame.n], bi[:n], bd[:n], vis[:n], pts


def ref_search_projection_coarse(frame, points, cam, pyr, th, feature_error, motion, threads=1):
    pts = np.ascontiguousarray(points)
    k2p = np.zeros(max(frame.n, 1), np.int32)
    m = snake_ref().snkref_search_projection_coarse(C.byref(frame), _p(pts), len(pts), C.byref(cam), C.byref(pyr),
                                                    C.c_float(th), feature_error, motion, _p(k2p), threads)
    return m, k2p[:frame.n]


def ref_search_projection_frame_to_kf(frame, points, cam, th, feature_error):
    pts = np.ascontiguousarray(points)
    k2p = np.zeros(max(frame.n, 1), np.int32)
    m = snake_ref().snkref_search_projection_frame_to_kf(C.byref(frame), _p(pts), len(pts), C.byref(cam), C.c_float(th),
                                                         feature_error, _p(k2p))
    return m, k2p[:frame.n]


def ref_match_bow(d1, d2, fv1, fv2, valid1, valid2, th, ratio):
    d1 = np.ascontiguousarray(d1, np.uint64).reshape(-1, 4)
    d2 = np.ascontiguousarray(d2, np.uint64).reshape(-1, 4)
    m12 = np.zeros(max(len(d1), 1), np.int32)
    m = snake_ref().snkref_match_bow(_p(d1), len(d1), _p(d2), len(d2), C.byref(fv1), C.byref(fv2), _p(valid1), _p(valid2),
                                     th, C.c_float(ratio), _p(m12))
    return m, m12[:len(d1)]


class BowCandidate(C.Structure):
    """snk_bow_candidate (include/snakeorb.h)"""
    _fields_ = [("desc", C.c_void_p), ("n", C.c_int32), ("fv", C.c_void_p), ("valid", C.c_void_p)]


def ref_match_bow_batch(d1, fv1, valid1, cands, th, ratio):
    """The reference's candidate loop (LoopDetector.cpp:384-391 without the early exit): MatchBoW(kf1, cand) for every
    (d2, fv2, valid2) of cands -> (n_matches[n_cand], match12[n_cand, n1])."""
    d1 = np.ascontiguousarray(d1, np.uint64).reshape(-1, 4)
    arr, keep = (BowCandidate * max(len(cands), 1))(), []
    for k, (d2, fv2, v2) in enumerate(cands):
        d2 = np.ascontiguousarray(d2, np.uint64).reshape(-1, 4)
        v2 = np.ascontiguousarray(v2, np.uint8)
        keep += [d2, v2, fv2]
        arr[k].desc, arr[k].n, arr[k].fv, arr[k].valid = d2.ctypes.data, len(d2), C.addressof(fv2), v2.ctypes.data
    m12 = np.zeros((len(cands), max(len(d1), 1)), np.int32)
    nm = np.zeros(max(len(cands), 1), np.int32)
    snake_ref().snkref_match_bow_batch(_p(d1), len(d1), C.byref(fv1), _p(np.ascontiguousarray(valid1, np.uint8)), arr, len(cands), th,
                                       C.c_float(ratio), _p(m12), _p(nm))
    return nm[:len(cands)], m12[:, :len(d1)]


def ref_fuse_search(kf, mask, points, cam, pyr, th, obs_factor, feature_th):
    pts = np.ascontiguousarray(points)
    n = len(pts)
    ck, ci = np.zeros(max(n, 1), np.int32), np.zeros(max(n, 1), np.int32)
    m = snake_ref().snkref_fuse_search(C.byref(kf), _p(mask) if mask is not None else None, _p(pts), n, C.byref(cam),
                                       C.byref(pyr), C.c_float(th), C.c_float(obs_factor), feature_th, _p(ck), _p(ci))
    return ck[:m].copy(), ci[:m].copy()


def ref_fuse_loop(kf, skip, points, cam, th):
    """LoopORBmatcher::Fuse, search and map edits: (nFused, feature index each point ends up observing or -1,
    per keyframe feature 0 = its original point is still good / 1 = it went bad (replaced) / 2 = it had none)."""
    pts = np.ascontiguousarray(points)
    n = len(pts)
    state = np.zeros(max(n, 1), np.int32)
    bad = np.zeros(max(kf.n, 1), np.uint8)
    m = snake_ref().snkref_fuse_loop(C.byref(kf), _p(skip) if skip is not None else None, _p(pts), n, C.byref(cam),
                                     C.c_float(th), _p(state), _p(bad))
    return m, state[:n], bad[:kf.n]


def ref_search_triangulation_bow(kf1, np1, fv1, kf2, np2, fv2, E, cam, epi, fd):
    pairs = np.zeros((max(kf1.n, 1), 2), np.int32)
    E = np.ascontiguousarray(E, np.float64)
    m = snake_ref().snkref_search_triangulation_bow(C.byref(kf1), _p(np1), C.byref(fv1), C.byref(kf2), _p(np2),
                                                    C.byref(fv2), _p(E), C.byref(cam), C.c_float(epi), fd, _p(pairs))
    return pairs[:m].copy()


def ref_search_triangulation_project(kf1, np1, kf2, np2, grid, E, cam, epi, fd):
    pairs = np.zeros((max(kf1.n, 1), 2), np.int32)
    E = np.ascontiguousarray(E, np.float64)
    grid = np.ascontiguousarray(grid, np.float64)
    m = snake_ref().snkref_search_triangulation_project(C.byref(kf1), _p(np1), C.byref(kf2), _p(np2), _p(grid),
                                                        grid.shape[0], grid.shape[1], _p(E), C.byref(cam),
                                                        C.c_float(epi), fd, _p(pairs))
    return pairs[:m].copy()


def ref_match_prediction(desc_left, positions, right, radius, feature_error, threads=1):
    d = np.ascontiguousarray(desc_left, np.uint64).reshape(-1, 4)
    pos = np.array(positions, np.float64, copy=True)
    matches = np.zeros(max(len(d), 1), np.int32)
    m = snake_ref().snkref_match_prediction(_p(d), len(d), _p(pos), C.byref(right), C.c_float(radius), feature_error,
                                            threads, _p(matches))
    return m, matches[:len(d)], pos


def ref_features_in_area(frame, px, py, r):
    out = np.zeros(max(frame.n, 1), np.int32)
    c = snake_ref().snkref_features_in_area(C.byref(frame), C.c_double(px), C.c_double(py), C.c_float(r), _p(out))
    return out[:c].copy()
