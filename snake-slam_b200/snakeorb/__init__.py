"""ctypes view of libsnakeorb.so (include/snakeorb.h) with the reference's class / method names.

This is test-and-benchmark plumbing: the product is the shared library and its C ABI. There is no fallback —
importing works without a GPU (so symbol tests can run), but every call needs a B200 and raises otherwise."""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(HERE), "libsnakeorb.so")

KP32 = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4")])
KP64 = np.dtype([("x", "<f8"), ("y", "<f8"), ("size", "<f8"), ("angle", "<f8"), ("response", "<f8"),
                 ("octave", "<i4"), ("_pad", "<i4")])
KNN2 = np.dtype([("dist0", "<i4"), ("idx0", "<i4"), ("dist1", "<i4"), ("idx1", "<i4")])
POINT_FINE = np.dtype([("desc", "<u8", 4), ("pos", "<f8", 3), ("normal", "<f8", 3), ("ref_depth", "<f4"),
                       ("ref_level", "<i4"), ("valid", "<i4"), ("id", "<i4")])
POINT_COARSE = np.dtype([("desc", "<u8", 4), ("pos", "<f8", 3), ("normal", "<f8", 3), ("octave", "<i4"),
                         ("angle", "<f4"), ("id", "<i4"), ("_pad", "<i4")])
POINT_FUSION = np.dtype([("desc", "<u8", 4), ("pos", "<f8", 3), ("normal", "<f8", 3), ("ref_depth", "<f4"),
                         ("ref_level", "<i4"), ("observations", "<i4"), ("id", "<i4")])
POINT_KF = np.dtype([("desc", "<u8", 4), ("pos", "<f8", 3), ("id", "<i4"), ("_pad", "<i4")])
assert KP32.itemsize == 24 and KP64.itemsize == 48 and POINT_FINE.itemsize == 96 and POINT_COARSE.itemsize == 96
assert POINT_KF.itemsize == 64


class SnakeOrbError(RuntimeError):
    pass


class OrbParams(C.Structure):
    _fields_ = [("nfeatures", C.c_int32), ("scale_factor", C.c_float), ("nlevels", C.c_int32), ("th_fast", C.c_int32),
                ("th_fast_min", C.c_int32), ("max_batch", C.c_int32), ("device", C.c_int32)]


class StereoParams(C.Structure):
    _fields_ = [("bf", C.c_double), ("relaxed", C.c_int32), ("nlevels", C.c_int32), ("scale", C.c_double * 16)]


class RgbdParams(C.Structure):
    _fields_ = [("fx", C.c_double), ("fy", C.c_double), ("cx", C.c_double), ("cy", C.c_double), ("dfx", C.c_double),
                ("dfy", C.c_double), ("dcx", C.c_double), ("dcy", C.c_double), ("ds", C.c_double),
                ("dist", C.c_double * 8), ("bf", C.c_double)]


class Rectify(C.Structure):
    """snk_rectify == the fields of Saiga::Rectification (Rectify.h:15-33) that undistortKeypoints reads."""
    _fields_ = [("fx", C.c_double), ("fy", C.c_double), ("cx", C.c_double), ("cy", C.c_double),
                ("dist", C.c_double * 8), ("quat", C.c_double * 4), ("dfx", C.c_double), ("dfy", C.c_double),
                ("dcx", C.c_double), ("dcy", C.c_double), ("ds", C.c_double)]

    @classmethod
    def from_vector(cls, v):
        """v: K_src fx fy cx cy s | D_src (8) | quaternion x y z w | K_dst fx fy cx cy s (22 doubles)."""
        r = cls()
        r.fx, r.fy, r.cx, r.cy = v[0:4]
        for i in range(8):
            r.dist[i] = v[5 + i]
        for i in range(4):
            r.quat[i] = v[13 + i]
        r.dfx, r.dfy, r.dcx, r.dcy, r.ds = v[17:22]
        return r


class Camera(C.Structure):
    _fields_ = [("fx", C.c_double), ("fy", C.c_double), ("cx", C.c_double), ("cy", C.c_double), ("s", C.c_double),
                ("bf", C.c_double)]


class GridView(C.Structure):
    _fields_ = [("bmin_x", C.c_double), ("bmin_y", C.c_double), ("bmax_x", C.c_double), ("bmax_y", C.c_double),
                ("cell_size_inv_x", C.c_double), ("cell_size_inv_y", C.c_double), ("rows", C.c_int32),
                ("cols", C.c_int32), ("cells", C.c_void_p)]


class FrameView(C.Structure):
    _fields_ = [("n", C.c_int32), ("kp_un", C.c_void_p), ("desc", C.c_void_p), ("right_points", C.c_void_p),
                ("kp_to_point", C.c_void_p), ("pose_q", C.c_double * 4), ("pose_t", C.c_double * 3),
                ("grid", GridView), ("depth", C.c_void_p)]


class ScalePyramidC(C.Structure):
    _fields_ = [("nlevels", C.c_int32), ("_pad", C.c_int32), ("scale_factor", C.c_double), ("scale", C.c_double * 16)]


class FeatVecView(C.Structure):
    _fields_ = [("n_nodes", C.c_int32), ("node_ids", C.c_void_p), ("offsets", C.c_void_p), ("indices", C.c_void_p)]


class BowCandidate(C.Structure):
    """snk_bow_candidate"""
    _fields_ = [("desc", C.c_void_p), ("n", C.c_int32), ("fv", C.POINTER(FeatVecView)), ("valid", C.c_void_p)]


_lib = None


def lib():
    """Loads libsnakeorb.so; raises if it has not been built (no silent fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise SnakeOrbError(f"{LIB_PATH} is missing: run `make -C snake-slam_b200` (or __graft_entry__.build())")
        _lib = C.CDLL(LIB_PATH)
        _lib.snk_last_error.restype = C.c_char_p
        _lib.snk_version.restype = C.c_char_p
    return _lib


def _check(rc):
    if rc != 0:
        raise SnakeOrbError(f"libsnakeorb error {rc}: {lib().snk_last_error().decode()}")


def _p(a):
    return C.c_void_p(a.ctypes.data) if a is not None else C.c_void_p(0)


def stereo_params(bf, relaxed, scales):
    p = StereoParams()
    p.bf, p.relaxed, p.nlevels = bf, int(relaxed), len(scales)
    for i, s in enumerate(scales):
        p.scale[i] = s
    return p


def scale_pyramid_c(scale_factor_f32, scales):
    p = ScalePyramidC()
    p.nlevels = len(scales)
    p.scale_factor = float(np.float32(scale_factor_f32))
    for i, s in enumerate(scales):
        p.scale[i] = s
    return p


class ORBExtractor:
    """Saiga::ORBExtractor(nfeatures, scaleFactor, num_levels, th_fast, th_fast_min, num_threads) —
    saiga/vision/features/ORBExtractor.h:49-53. `num_threads` is accepted and ignored (the GPU path has no
    per-level OpenMP loop); `max_batch` sizes the batched entry points."""

    def __init__(self, nfeatures=1000, scaleFactor=1.2, num_levels=8, th_fast=20, th_fast_min=7, num_threads=1,
                 max_batch=1, device=0):
        self.h = C.c_void_p()
        prm = OrbParams(nfeatures, scaleFactor, num_levels, th_fast, th_fast_min, max_batch, device)
        _check(lib().snk_orb_create(C.byref(prm), C.byref(self.h)))
        self.nlevels, self.max_batch = num_levels, max_batch
        self.cap = lib().snk_orb_max_keypoints(self.h)

    def close(self):
        if getattr(self, "h", None) and lib is not None and _lib is not None:  # module globals are gone at interpreter exit
            _lib.snk_orb_destroy(self.h)
            self.h = None

    __del__ = close

    def set_stream(self, cuda_stream):
        _check(lib().snk_orb_set_stream(self.h, C.c_void_p(cuda_stream)))

    def getPyramid(self):
        s = np.zeros(self.nlevels)
        f = np.zeros(self.nlevels, np.int32)
        _check(lib().snk_orb_pyramid(self.h, _p(s), _p(f)))
        return s, f

    def Detect(self, img):
        """Detect(ImageView<uchar>, vector<KeyPoint<float>>&, vector<DescriptorORB>&) for a host numpy image."""
        assert img.dtype == np.uint8 and img.ndim == 2 and (img.size == 0 or img.strides[1] == 1)
        kps = np.zeros(self.cap, KP32)
        desc = np.zeros((self.cap, 4), np.uint64)
        n = C.c_int(0)
        _check(lib().snk_orb_detect(self.h, _p(img), img.shape[0], img.shape[1], C.c_size_t(img.strides[0]), 0,
                                    _p(kps), _p(desc), self.cap, C.byref(n)))
        return kps[:n.value].copy(), desc[:n.value].copy()

    def DetectBatch(self, imgs):
        """imgs: (B, rows, cols) uint8 C-contiguous host array. Returns per-image (kps, desc) lists."""
        assert imgs.dtype == np.uint8 and imgs.ndim == 3 and imgs.flags.c_contiguous
        B, rows, cols = imgs.shape
        kps = np.zeros((B, self.cap), KP32)
        desc = np.zeros((B, self.cap, 4), np.uint64)
        n = np.zeros(B, np.int32)
        _check(lib().snk_orb_detect_batch(self.h, _p(imgs), B, rows, cols, C.c_size_t(cols), C.c_size_t(rows * cols), 0,
                                          _p(kps), _p(desc), self.cap, _p(n)))
        return [(kps[i, :n[i]].copy(), desc[i, :n[i]].copy()) for i in range(B)]

    def enqueue_ptr(self, ptr, n_imgs, rows, cols, pitch, stride, on_device):
        _check(lib().snk_orb_enqueue(self.h, C.c_void_p(ptr), n_imgs, rows, cols, C.c_size_t(pitch), C.c_size_t(stride),
                                     int(on_device)))

    def DetectBatchRGBA(self, rgba):
        """rgba: (B, rows, cols, 4) uint8 host array; rows may be padded (strides[1] >= 4*cols). Grey conversion
        (Input::computeGrayscaleImage) fused in front of the extractor. Returns per-image (kps, desc)."""
        assert rgba.dtype == np.uint8 and rgba.ndim == 4 and rgba.shape[3] == 4 and rgba.strides[2:] == (4, 1)
        B, rows, cols, _ = rgba.shape
        _check(lib().snk_orb_enqueue_rgba(self.h, _p(rgba), B, rows, cols, C.c_size_t(rgba.strides[1]),
                                          C.c_size_t(rgba.strides[0]), 0))
        kps, desc, n = self.fetch(B)
        return [(kps[i, :n[i]].copy(), desc[i, :n[i]].copy()) for i in range(B)]

    def fetch(self, n_imgs, kps=None, desc=None, counts=None):
        kps = np.zeros((n_imgs, self.cap), KP32) if kps is None else kps
        desc = np.zeros((n_imgs, self.cap, 4), np.uint64) if desc is None else desc
        counts = np.zeros(n_imgs, np.int32) if counts is None else counts
        _check(lib().snk_orb_fetch(self.h, _p(kps), _p(desc), self.cap, _p(counts)))
        return kps, desc, counts

    def check(self):
        """snk_orb_check: synchronise and raise if a kernel of the asynchronous path reported a fault."""
        _check(lib().snk_orb_check(self.h))

    def last_batch(self):
        return lib().snk_orb_last_batch(self.h)

    def wait_ctx(self, ctx):
        _check(lib().snk_orb_wait_ctx(self.h, ctx.c))

    def device_results(self):
        k, d, c = C.c_void_p(), C.c_void_p(), C.c_void_p()
        s = C.c_int()
        _check(lib().snk_orb_device_results(self.h, C.byref(k), C.byref(d), C.byref(c), C.byref(s)))
        return k.value, d.value, c.value, s.value

    def GetImage(self, level, img=0, blurred=False):
        """ORBExtractor::GetImage(level), copied to the host."""
        r, c = C.c_int(), C.c_int()
        pitch = C.c_size_t()
        ptr = C.c_void_p()
        _check(lib().snk_orb_get_level(self.h, img, level, int(blurred), C.byref(ptr), C.byref(r), C.byref(c), C.byref(pitch)))
        out = np.zeros((r.value, c.value), np.uint8)
        _check(lib().snk_orb_read_level(self.h, img, level, int(blurred), _p(out), C.c_size_t(c.value)))
        return out

    def candidates(self, level, img=0):
        buf = np.zeros(400000, np.uint32)
        n = C.c_int()
        _check(lib().snk_orb_read_candidates(self.h, img, level, _p(buf), len(buf), C.byref(n)))
        p = buf[:n.value]
        return np.stack([p & 0xfff, (p >> 12) & 0xfff, p >> 24], 1).astype(np.int32)

    def set_profiling(self, on):
        _check(lib().snk_orb_set_profiling(self.h, int(on)))

    def stage_ms(self, reset=True):
        ms = np.zeros(8, np.float32)
        ln = np.zeros(8, np.int32)
        _check(lib().snk_orb_stage_ms(self.h, _p(ms), _p(ln), int(reset)))
        return ms, ln


class FrameDev:
    """snk_frame_dev: a frame view kept on the device between searches."""

    def __init__(self, ctx, view):
        self.h, self.ctx, self.n = C.c_void_p(), ctx, view.n
        _check(lib().snk_frame_dev_create(ctx.c, C.byref(view), C.byref(self.h)))

    def set_pose(self, q_xyzw, t):
        q, t = np.ascontiguousarray(q_xyzw, np.float64), np.ascontiguousarray(t, np.float64)
        _check(lib().snk_frame_dev_set_pose(self.h, _p(q), _p(t)))

    def set_assignment(self, k2p):
        k2p = np.ascontiguousarray(k2p, np.int32)
        _check(lib().snk_frame_dev_set_assignment(self.h, _p(k2p)))

    def close(self):
        if getattr(self, "h", None) and _lib is not None and getattr(self.ctx, "c", None):
            _lib.snk_frame_dev_destroy(self.h)
            self.h = None

    __del__ = close


class LocalMapDev:
    """snk_localmap: LocalMap<FineTrackingPoint> kept on the device."""

    def __init__(self, ctx, capacity):
        self.h, self.ctx, self.n = C.c_void_p(), ctx, 0
        _check(lib().snk_localmap_create(ctx.c, capacity, C.byref(self.h)))

    def upload(self, points, first=0, new_size=None):
        pts = np.ascontiguousarray(points, POINT_FINE)
        self.n = first + len(pts) if new_size is None else new_size
        _check(lib().snk_localmap_upload(self.h, _p(pts), first, len(pts), self.n))

    def close(self):
        if getattr(self, "h", None) and _lib is not None and getattr(self.ctx, "c", None):
            _lib.snk_localmap_destroy(self.h)
            self.h = None

    __del__ = close


class Context:
    """snk_ctx: device + stream + scratch for the matchers."""

    def __init__(self, device=0):
        self.c = C.c_void_p()
        _check(lib().snk_ctx_create(device, C.byref(self.c)))

    def close(self):
        if getattr(self, "c", None) and lib is not None and _lib is not None:  # module globals are gone at interpreter exit
            _lib.snk_ctx_destroy(self.c)
            self.c = None

    __del__ = close

    def set_stream(self, cuda_stream):
        _check(lib().snk_ctx_set_stream(self.c, C.c_void_p(cuda_stream)))

    def wait_orb(self, orb):
        """Order this context's stream after the extractor's last enqueue (snk_ctx_wait_orb)."""
        _check(lib().snk_ctx_wait_orb(self.c, orb.h))

    def check(self):
        _check(lib().snk_ctx_check(self.c))

    def distance(self, a, b):
        """Saiga::distance for n pairs."""
        a = np.ascontiguousarray(a, np.uint64).reshape(-1, 4)
        b = np.ascontiguousarray(b, np.uint64).reshape(-1, 4)
        out = np.zeros(len(a), np.int32)
        _check(lib().snk_hamming_pairs(self.c, _p(a), _p(b), len(a), _p(out)))
        return out

    def matchKnn2(self, d1, d2):
        """BruteForceMatcher::matchKnn2 -> knn2 records."""
        d1 = np.ascontiguousarray(d1, np.uint64).reshape(-1, 4)
        d2 = np.ascontiguousarray(d2, np.uint64).reshape(-1, 4)
        out = np.zeros(len(d1), KNN2)
        _check(lib().snk_match_knn2(self.c, _p(d1), len(d1), _p(d2), len(d2), _p(out)))
        return out

    @staticmethod
    def filterMatches(knn2, threshold, ratio):
        out = np.zeros((max(len(knn2), 1), 2), np.int32)
        n = C.c_int()
        _check(lib().snk_filter_matches(_p(knn2), len(knn2), threshold, C.c_float(ratio), _p(out), C.byref(n)))
        return out[:n.value].copy()

    def StereoMatching(self, params, left, dl, right, dr):
        """Preprocess::StereoMatching on rectified keypoints -> (matches, right_points, depth)."""
        left = np.ascontiguousarray(left, KP64)
        right = np.ascontiguousarray(right, KP64)
        dl = np.ascontiguousarray(dl, np.uint64)
        dr = np.ascontiguousarray(dr, np.uint64)
        rp = np.zeros(max(len(left), 1), np.float32)
        dp = np.zeros(max(len(left), 1), np.float32)
        n = C.c_int()
        _check(lib().snk_stereo_match(self.c, C.byref(params), _p(left), _p(dl), len(left), _p(right), _p(dr),
                                      len(right), _p(rp), _p(dp), C.byref(n)))
        return n.value, rp[:len(left)], dp[:len(left)]

    def undistortKeypoints(self, rect, kps):
        """Preprocess::undistortKeypoints -> (undistorted_keypoints, normalized_points[n,2])."""
        kps = np.ascontiguousarray(kps, KP64)
        out = np.zeros(len(kps), KP64)
        nrm = np.zeros((len(kps), 2), np.float64)
        _check(lib().snk_undistort_keypoints(self.c, C.byref(rect), _p(kps), len(kps), _p(out), _p(nrm)))
        return out, nrm

    def computeGrayscaleImage(self, rgba):
        """Input::computeGrayscaleImage: (rows, cols, 4) uint8 -> (rows, cols) uint8."""
        assert rgba.dtype == np.uint8 and rgba.ndim == 3 and rgba.shape[2] == 4 and rgba.strides[1:] == (4, 1)
        out = np.zeros(rgba.shape[:2], np.uint8)
        _check(lib().snk_rgba_to_gray(self.c, _p(rgba), rgba.shape[0], rgba.shape[1], C.c_size_t(rgba.strides[0]),
                                      _p(out), C.c_size_t(out.strides[0] if out.size else 0)))
        return out

    def StereoMatchingRaw(self, rect_left, rect_right, params, left, dl, right, dr):
        """The whole Preprocess::StereoMatching incl. Rectification::Forward of the raw keypoints."""
        left, right = np.ascontiguousarray(left, KP64), np.ascontiguousarray(right, KP64)
        dl, dr = np.ascontiguousarray(dl, np.uint64), np.ascontiguousarray(dr, np.uint64)
        rp = np.zeros(max(len(left), 1), np.float32)
        dp = np.zeros(max(len(left), 1), np.float32)
        n = C.c_int()
        _check(lib().snk_stereo_matching(self.c, C.byref(rect_left), C.byref(rect_right), C.byref(params), _p(left), _p(dl),
                                         len(left), _p(right), _p(dr), len(right), _p(rp), _p(dp), C.byref(n)))
        return n.value, rp[:len(left)], dp[:len(left)]

    def ComputeStereoFromRGBD(self, params, kp_un, depth_img):
        """Preprocess::ComputeStereoFromRGBD -> (matches, right_points, depth)."""
        kp_un = np.ascontiguousarray(kp_un, KP64)
        assert depth_img.dtype == np.float32 and depth_img.ndim == 2 and depth_img.strides[1] == 4
        rp = np.zeros(max(len(kp_un), 1), np.float32)
        dp = np.zeros(max(len(kp_un), 1), np.float32)
        n = C.c_int()
        _check(lib().snk_stereo_from_rgbd(self.c, C.byref(params), _p(kp_un), len(kp_un), _p(depth_img), depth_img.shape[0],
                                          depth_img.shape[1], C.c_size_t(depth_img.strides[0]), _p(rp), _p(dp), C.byref(n)))
        return n.value, rp[:len(kp_un)], dp[:len(kp_un)]

    def knn2_batch_device(self, d1_ptr, n_ptr, stride1, d2_ptr, m_ptr, stride2, n_problems, out_ptr):
        _check(lib().snk_match_knn2_batch_device(self.c, C.c_void_p(d1_ptr), C.c_void_p(n_ptr), stride1, C.c_void_p(d2_ptr),
                                                 C.c_void_p(m_ptr), stride2, n_problems, C.c_void_p(out_ptr)))

    def bow_transform_batch_device(self, voc, desc_ptr, counts_ptr, n_frames, stride, levelsup, bow_ids, bow_vals,
                                   fv_nodes, fv_offsets, fv_indices, out_counts):
        _check(lib().snk_bow_transform_batch_device(self.c, voc.v, C.c_void_p(desc_ptr), C.c_void_p(counts_ptr), n_frames,
                                                    stride, levelsup, C.c_void_p(bow_ids), C.c_void_p(bow_vals),
                                                    C.c_void_p(fv_nodes), C.c_void_p(fv_offsets), C.c_void_p(fv_indices),
                                                    C.c_void_p(out_counts)))

    # ---- local mapping / loop closing searches ----
    def Fuse(self, kf, point_mask, points, cam, pyr, th, obs_factor, feature_th):
        """MappingORBMatcher::Fuse (LocalMap<FusionPoint> overload) -> fuseCandidates as (kp index, point id) arrays."""
        pts = np.ascontiguousarray(points, POINT_FUSION)
        n = len(pts)
        ck, ci = np.zeros(max(n, 1), np.int32), np.zeros(max(n, 1), np.int32)
        m = C.c_int()
        mask = None if point_mask is None else np.ascontiguousarray(point_mask, np.uint8)
        _check(lib().snk_fuse_search(self.c, C.byref(kf), _p(mask) if mask is not None else None, _p(pts), n, C.byref(cam),
                                     C.byref(pyr), C.c_float(th), C.c_float(obs_factor), feature_th, _p(ck), _p(ci),
                                     C.byref(m)))
        return ck[:m.value].copy(), ci[:m.value].copy()

    def FuseLoopSearch(self, kf, skip, points, cam, th):
        """The search of LoopORBmatcher::Fuse -> best_idx, best_dist per point."""
        pts = np.ascontiguousarray(points, POINT_FUSION)
        n = len(pts)
        bi, bd = np.zeros(max(n, 1), np.int32), np.zeros(max(n, 1), np.int32)
        skip = None if skip is None else np.ascontiguousarray(skip, np.uint8)
        _check(lib().snk_fuse_search_loop(self.c, C.byref(kf), _p(skip) if skip is not None else None, _p(pts), n,
                                          C.byref(cam), C.c_float(th), _p(bi), _p(bd)))
        return bi[:n], bd[:n]

    def SearchForTriangulation2(self, kf1, np1, fv1, kf2, np2, fv2, E, cam, epipolar_distance, feature_distance):
        pairs = np.zeros((max(kf1.n, 1), 2), np.int32)
        E = np.ascontiguousarray(E, np.float64)
        m = C.c_int()
        _check(lib().snk_search_triangulation_bow(self.c, C.byref(kf1), _p(np1), C.byref(fv1), C.byref(kf2), _p(np2),
                                                  C.byref(fv2), _p(E), C.byref(cam), C.c_float(epipolar_distance),
                                                  feature_distance, _p(pairs), C.byref(m)))
        return pairs[:m.value].copy()

    def SearchForTriangulationProject(self, kf1, np1, kf2, np2, grid, E, cam, epipolar_distance, feature_distance):
        pairs = np.zeros((max(kf1.n, 1), 2), np.int32)
        E = np.ascontiguousarray(E, np.float64)
        grid = np.ascontiguousarray(grid, np.float64)
        m = C.c_int()
        _check(lib().snk_search_triangulation_project(self.c, C.byref(kf1), _p(np1), C.byref(kf2), _p(np2), _p(grid),
                                                      grid.shape[0], grid.shape[1], _p(E), C.byref(cam),
                                                      C.c_float(epipolar_distance), feature_distance, _p(pairs),
                                                      C.byref(m)))
        return pairs[:m.value].copy()

    def ComputeMatchesPrediction(self, desc_left, positions, right, radius=50.0, feature_error=50):
        d = np.ascontiguousarray(desc_left, np.uint64).reshape(-1, 4)
        pos = np.array(positions, np.float64, copy=True)
        matches = np.zeros(max(len(d), 1), np.int32)
        m = C.c_int()
        _check(lib().snk_match_prediction(self.c, _p(d), len(d), _p(pos), C.byref(right), C.c_float(radius),
                                          feature_error, _p(matches), C.byref(m)))
        return m.value, matches[:len(d)], pos

    def bow_scores(self, q_ids, q_vals, cands):
        """FeatureVector::score of (q_ids, q_vals) against every (ids, vals) pair of `cands` -> float32 array."""
        q_ids, q_vals = np.ascontiguousarray(q_ids, np.int32), np.ascontiguousarray(q_vals, np.float32)
        off = np.zeros(len(cands) + 1, np.int32)
        for k, (i, _) in enumerate(cands):
            off[k + 1] = off[k] + len(i)
        ci = np.ascontiguousarray(np.concatenate([np.asarray(i, np.int32) for i, _ in cands]) if cands else np.zeros(0, np.int32))
        cvv = np.ascontiguousarray(np.concatenate([np.asarray(v, np.float32) for _, v in cands]) if cands else np.zeros(0, np.float32))
        out = np.zeros(max(len(cands), 1), np.float32)
        _check(lib().snk_bow_score_batch(self.c, _p(q_ids), _p(q_vals), len(q_ids), _p(ci), _p(cvv), _p(off), len(cands), _p(out)))
        return out[:len(cands)]

    def stereo_orb_device(self, orb, n_pairs, params, rp_ptr, depth_ptr, nm_ptr):
        _check(lib().snk_stereo_match_orb_device(self.c, orb.h, n_pairs, C.byref(params), C.c_void_p(rp_ptr),
                                                 C.c_void_p(depth_ptr), C.c_void_p(nm_ptr)))

    def SearchByProjection2(self, frame, points, cam, pyr, th, ratio):
        pts = np.array(points, dtype=POINT_FINE, copy=True)
        n = len(pts)
        k2p = np.zeros(max(frame.n, 1), np.int32)
        bi = np.zeros(max(n, 1), np.int32)
        bd = np.zeros(max(n, 1), np.int32)
        vis = np.zeros(max(n, 1), np.uint8)
        m = C.c_int()
        _check(lib().snk_search_projection_fine(self.c, C.byref(frame), _p(pts), n, C.byref(cam), C.byref(pyr),
                                                C.c_float(th), C.c_float(ratio), _p(k2p), _p(bi), _p(bd), _p(vis),
                                                C.byref(m)))
        return m.value, k2p[:frame.n], bi[:n], bd[:n], vis[:n], pts

    def frame_dev(self, view):
        """Device copy of a frame view (snk_frame_dev_create); keep the returned object alive while it is used."""
        return FrameDev(self, view)

    def localmap(self, capacity):
        return LocalMapDev(self, capacity)

    def SearchByProjection2Resident(self, frame_dev, lm, cam, pyr, th, ratio, keep_assignment=True):
        """SearchByProjection2 on a device-resident frame and local map -> (matches, kp_to_point, valid, visible)."""
        k2p = np.zeros(max(frame_dev.n, 1), np.int32)
        valid = np.zeros(max(lm.n, 1), np.int32)
        vis = np.zeros(max(lm.n, 1), np.uint8)
        m = C.c_int()
        _check(lib().snk_search_projection_fine_resident(self.c, frame_dev.h, lm.h, C.byref(cam), C.byref(pyr), C.c_float(th),
                                                         C.c_float(ratio), int(keep_assignment), _p(k2p), _p(valid), _p(vis),
                                                         C.byref(m)))
        return m.value, k2p[:frame_dev.n], valid[:lm.n], vis[:lm.n]

    def SearchByProjectionFrameFrame2(self, frame, points, cam, pyr, th, feature_error, motion):
        pts = np.ascontiguousarray(points, POINT_COARSE)
        k2p = np.zeros(max(frame.n, 1), np.int32)
        m = C.c_int()
        _check(lib().snk_search_projection_coarse(self.c, C.byref(frame), _p(pts), len(pts), C.byref(cam),
                                                  C.byref(pyr), C.c_float(th), feature_error, motion, _p(k2p),
                                                  C.byref(m)))
        return m.value, k2p[:frame.n]

    def SearchByProjectionFrameToKeyframe(self, frame, points, cam, th, feature_error):
        pts = np.ascontiguousarray(points, POINT_KF)
        k2p = np.zeros(max(frame.n, 1), np.int32)
        m = C.c_int()
        _check(lib().snk_search_projection_frame_to_kf(self.c, C.byref(frame), _p(pts), len(pts), C.byref(cam),
                                                       C.c_float(th), feature_error, _p(k2p), C.byref(m)))
        return m.value, k2p[:frame.n]

    def MatchBoW(self, d1, d2, fv1, fv2, valid1, valid2, th, ratio):
        d1 = np.ascontiguousarray(d1, np.uint64).reshape(-1, 4)
        d2 = np.ascontiguousarray(d2, np.uint64).reshape(-1, 4)
        m12 = np.zeros(max(len(d1), 1), np.int32)
        m = C.c_int()
        _check(lib().snk_match_bow(self.c, _p(d1), len(d1), _p(d2), len(d2), C.byref(fv1), C.byref(fv2), _p(valid1),
                                   _p(valid2), th, C.c_float(ratio), _p(m12), C.byref(m)))
        return m.value, m12[:len(d1)]

    def MatchBoWBatch(self, d1, fv1, valid1, cands, th, ratio):
        """MatchBoW of one keyframe against every (d2, fv2, valid2) of `cands` in one call (snk_match_bow_batch) ->
        (n_matches[n_cand], match12[n_cand, n1])."""
        d1 = np.ascontiguousarray(d1, np.uint64).reshape(-1, 4)
        keep, arr = [], (BowCandidate * max(len(cands), 1))()
        for k, (d2, fv2, v2) in enumerate(cands):
            d2 = np.ascontiguousarray(d2, np.uint64).reshape(-1, 4)
            v2 = np.ascontiguousarray(v2, np.uint8)
            keep += [d2, v2, fv2]
            arr[k].desc, arr[k].n, arr[k].fv, arr[k].valid = d2.ctypes.data, len(d2), C.pointer(fv2), v2.ctypes.data
        m12 = np.zeros((len(cands), max(len(d1), 1)), np.int32)
        nm = np.zeros(max(len(cands), 1), np.int32)
        _check(lib().snk_match_bow_batch(self.c, _p(d1), len(d1), C.byref(fv1), _p(np.ascontiguousarray(valid1, np.uint8)), arr,
                                         len(cands), th, C.c_float(ratio), _p(m12), _p(nm)))
        return nm[:len(cands)], m12[:, :len(d1)] if len(d1) else m12[:, :0]

    def computeFeatureGrid(self, bounds, kp_un):
        """FeatureGrid2::create -> (permutation, cells[rows*cols, 2]); `bounds` is a GridView (cells ignored)."""
        kp_un = np.ascontiguousarray(kp_un, KP64)
        perm = np.zeros(max(len(kp_un), 1), np.int32)
        cells = np.zeros((bounds.rows * bounds.cols, 2), np.int32)
        _check(lib().snk_feature_grid(self.c, C.byref(bounds), _p(kp_un), len(kp_un), _p(perm), _p(cells)))
        return perm[:len(kp_un)], cells

    def hamming_peak(self):
        v = C.c_double()
        _check(lib().snk_hamming_peak_selftest(self.c, C.byref(v)))
        return v.value


class Vocabulary:
    """snk_voc == MiniBow2::TemplatedVocabulary on the device; transform() needs a Context."""

    def __init__(self, path=None, arrays=None, device=0):
        self.v = C.c_void_p()
        if path is not None:
            _check(lib().snk_voc_load_raw(device, path.encode(), C.byref(self.v)))
        else:
            k, L, parent, weight, word_id, desc = arrays
            parent = np.ascontiguousarray(parent, np.int32)
            weight = np.ascontiguousarray(weight, np.float64)
            word_id = np.ascontiguousarray(word_id, np.int32)
            desc = np.ascontiguousarray(desc, np.uint64)
            _check(lib().snk_voc_create(device, k, L, len(parent), _p(parent), _p(weight), _p(word_id), _p(desc),
                                        C.byref(self.v)))
        k, L, nn, nw = C.c_int(), C.c_int(), C.c_int(), C.c_int()
        _check(lib().snk_voc_info(self.v, C.byref(k), C.byref(L), C.byref(nn), C.byref(nw)))
        self.k, self.L, self.n_nodes, self.n_words = k.value, L.value, nn.value, nw.value

    def close(self):
        if getattr(self, "v", None) and lib is not None and _lib is not None:  # module globals are gone at interpreter exit
            _lib.snk_voc_destroy(self.v)
            self.v = None

    __del__ = close

    def transform(self, ctx, desc, levelsup):
        """TemplatedVocabulary::transform -> bow_ids, bow_vals, fv_nodes, fv_offsets, fv_indices."""
        desc = np.ascontiguousarray(desc, np.uint64).reshape(-1, 4)
        n = len(desc)
        ids, vals = np.zeros(n + 1, np.int32), np.zeros(n + 1, np.float32)
        fn, fo, fi = np.zeros(n + 1, np.int32), np.zeros(n + 2, np.int32), np.zeros(n + 1, np.int32)
        nb, nf = C.c_int(), C.c_int()
        _check(lib().snk_bow_transform(ctx.c, self.v, _p(desc), n, levelsup, _p(ids), _p(vals), C.byref(nb), _p(fn),
                                       _p(fo), _p(fi), C.byref(nf)))
        return ids[:nb.value], vals[:nb.value], fn[:nf.value], fo[:nf.value + 1], fi[:fo[nf.value]]


def grid_bounds(w, h, cam, dist, resolution=20):
    """FeatureGridBounds2::computeFromIntrinsicsDist -> GridView without cells (host-side helper, no GPU)."""
    d = np.ascontiguousarray(dist, np.float64)
    assert len(d) == 8
    g = GridView()
    _check(lib().snk_grid_bounds(w, h, C.byref(cam), _p(d), resolution, C.byref(g)))
    return g


def kp32_to_kp64(k):
    o = np.zeros(len(k), KP64)
    for f in ("x", "y", "size", "angle", "response", "octave"):
        o[f] = k[f]
    return o


def features_write(path, kps64, desc):
    kps64 = np.ascontiguousarray(kps64, KP64)
    desc = np.ascontiguousarray(desc, np.uint64).reshape(-1, 4)
    _check(lib().snk_features_write(path.encode(), _p(kps64), _p(desc), len(kps64)))


def features_read(path, capacity=100000):
    kps = np.zeros(capacity, KP64)
    desc = np.zeros((capacity, 4), np.uint64)
    n = C.c_int()
    _check(lib().snk_features_read(path.encode(), _p(kps), _p(desc), capacity, C.byref(n)))
    return kps[:n.value].copy(), desc[:n.value].copy()


def keypoints_widen(k32):
    k32 = np.ascontiguousarray(k32, KP32)
    out = np.zeros(max(len(k32), 1), KP64)
    _check(lib().snk_keypoints_widen(_p(k32), len(k32), _p(out)))
    return out[:len(k32)]
