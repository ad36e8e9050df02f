"""Generates the committed golden vectors. Run in the build container only (needs cv2 4.13 and /root/reference):

    python tests/golden/make_golden.py

* golden_cv.npz        — cv2 4.13.0 outputs of the three OpenCV calls on the path (resize / GaussianBlur / FAST)
* golden_ref.npz       — outputs of the reference's own compiled ScalePyramid / AngleUmax / ComputeAngle /
                         ComputeDescriptor / Distribute / BruteForceMatcher (oracle/_ref)
* golden_next.npz      — rows after the extractor: undistortKeypoints' per-point math from the
                         reference's camera-model headers (`python tests/golden/make_golden.py next` makes only this)
* golden_extract.npz   — end-to-end keypoints + descriptors of the tier-1 hybrid (tests/hybrid.py: ORBExtractor
                         driver with real cv2 + reference pieces) on seeded images of every BASELINE shape
"""
import ctypes as C
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [HERE, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "snake-slam_b200")]
import cv2  # noqa: E402
import hybrid  # noqa: E402
import oracle_py as O  # noqa: E402
from snakeorb import synthetic as S  # noqa: E402

cv2.setNumThreads(1)
assert cv2.__version__.startswith("4.13"), cv2.__version__
R = O.ref()
assert R is not None, "build oracle/_ref first (make -C oracle)"


def _p(a):
    return C.c_void_p(a.ctypes.data)


def sha(a):
    return hashlib.sha1(np.ascontiguousarray(a).tobytes()).hexdigest()


def make_cv():
    rng = np.random.default_rng(11)
    out = {}
    srcs = {"noise": rng.integers(0, 256, (83, 101), dtype=np.uint8), "synth": S.image(77, 160, 120)}
    for name, s in srcs.items():
        dh, dw = hybrid.iround_f32(np.float32(s.shape[0]) * np.float32(1 / 1.2)), hybrid.iround_f32(
            np.float32(s.shape[1]) * np.float32(1 / 1.2))
        out[f"{name}_src"] = s
        out[f"{name}_resize"] = cv2.resize(s, (dw, dh), interpolation=cv2.INTER_LINEAR)
        out[f"{name}_blur"] = cv2.GaussianBlur(s, (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
        rois, res = [], []
        for k in range(12):
            h, w = int(rng.integers(7, 46)), int(rng.integers(7, 46))
            y0, x0 = int(rng.integers(0, s.shape[0] - h)), int(rng.integers(0, s.shape[1] - w))
            th = [20, 7, 1, 40][k % 4]
            kps = cv2.FastFeatureDetector_create(threshold=th, nonmaxSuppression=True).detect(s[y0:y0 + h, x0:x0 + w])
            rois.append((y0, x0, h, w, th, len(kps)))
            res += [(int(kp.pt[0]), int(kp.pt[1]), int(kp.response)) for kp in kps]
        out[f"{name}_fast_rois"] = np.array(rois, np.int32)
        out[f"{name}_fast_kps"] = np.array(res, np.int32).reshape(-1, 3)
    np.savez_compressed(os.path.join(HERE, "golden_cv.npz"), **out)


def make_ref():
    rng = np.random.default_rng(12)
    out = {}
    for cfg in ((8, 1000), (8, 1200), (8, 2000), (4, 1000), (7, 2000), (1, 500)):
        s = np.zeros(cfg[0])
        f = np.zeros(cfg[0], np.int32)
        R.ref_scale_pyramid(cfg[0], C.c_float(1.2), cfg[1], _p(s), _p(f))
        out[f"pyr_{cfg[0]}_{cfg[1]}_scale"] = s
        out[f"pyr_{cfg[0]}_{cfg[1]}_features"] = f
    um = np.zeros(16, np.int32)
    R.ref_umax(_p(um))
    out["umax"] = um
    pat = np.zeros(1024, np.int32)
    R.ref_pattern(_p(pat))
    out["pattern_sha1"] = np.array(sha(pat.astype(np.int8)))
    img = S.image(5, 200, 160)
    g = cv2.GaussianBlur(img, (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
    n = 600
    xy = np.stack([rng.integers(19, 200 - 19, n), rng.integers(19, 160 - 19, n)], 1).astype(np.float32)
    ang = np.zeros(n, np.float32)
    R.ref_angles(_p(img), 160, 200, C.c_size_t(200), _p(xy), n, _p(ang))
    d = np.zeros((n, 4), np.uint64)
    R.ref_descriptors(_p(g), 160, 200, C.c_size_t(200), _p(xy), _p(ang), n, _p(d))
    # also descriptors at arbitrary (non-moment) angles incl. 0 — tests/test_cuda_orb.cu:155-222 does the same
    ang2 = rng.uniform(0, 360, n).astype(np.float32)
    ang2[:50] = 0
    d2 = np.zeros((n, 4), np.uint64)
    R.ref_descriptors(_p(g), 160, 200, C.c_size_t(200), _p(xy), _p(ang2), n, _p(d2))
    out.update(od_img_sha1=np.array(sha(img)), od_xy=xy, od_angle=ang, od_desc=d, od_angle2=ang2, od_desc2=d2)
    # distribution: random candidate clouds with many response ties and clustered positions
    for t, (n, w, h, target) in enumerate(((3000, 720, 448, 217), (500, 300, 200, 60), (50, 100, 100, 60),
                                            (9000, 1209, 344, 434), (700, 64, 64, 151))):
        x = rng.integers(3, w - 3, n)
        y = rng.integers(3, h - 3, n)
        if t == 4:
            x = (x // 8) * 8  # heavy clustering: many equal coordinates
        pos = np.unique(np.stack([y, x], 1), axis=0)  # FAST never emits the same pixel twice; row-major order
        k = np.zeros(len(pos), O.KP32)
        k["x"], k["y"] = pos[:, 1], pos[:, 0]
        k["size"], k["angle"], k["response"] = 7, -1, rng.integers(7, 60, len(pos))
        sel, perm = O.distribute(k, (16, 16), (w + 16, h + 16), target, which="ref")
        out[f"dist{t}_in"] = k
        out[f"dist{t}_args"] = np.array([16, 16, w + 16, h + 16, target], np.int32)
        out[f"dist{t}_out"] = sel
        out[f"dist{t}_perm"] = perm
    a = rng.integers(0, 2**63, (300, 4)).astype(np.uint64)
    b = a[rng.integers(0, 300, 400)] ^ (np.uint64(1) << rng.integers(0, 64, (400, 4)).astype(np.uint64))
    k4 = np.zeros((300, 4), np.int32)
    R.ref_knn2(_p(a), 300, _p(np.ascontiguousarray(b)), 400, _p(k4))
    out.update(knn_a=a, knn_b=b, knn_out=k4)
    np.savez_compressed(os.path.join(HERE, "golden_ref.npz"), **out)


from cases import CASES, case_image  # noqa: E402


def make_extract():
    out = {}
    for name, (kind, w, h, nf, nl) in CASES.items():
        img = case_image(kind, w, h)
        k, d, st = hybrid.hybrid_detect(img, nf, 1.2, nl, 20, 7)
        out[f"{name}_sha1"] = np.array(sha(img))
        out[f"{name}_kps"] = k
        out[f"{name}_desc"] = d
        out[f"{name}_nraw"] = np.array([len(s["raw"]) for s in st], np.int32)
        out[f"{name}_level_sha1"] = np.array([sha(s["level"]) for s in st])
        out[f"{name}_gauss_sha1"] = np.array([sha(s["gauss"]) for s in st])
        print(name, len(k), out[f"{name}_nraw"].sum())
    np.savez_compressed(os.path.join(HERE, "golden_extract.npz"), **out)


RECT_CASES = {
    # name: (K_src fx fy cx cy, D_src k1..k6 p1 p2, R quaternion x y z w (normalised below), K_dst fx fy cx cy s, w, h)
    "euroc": ((458.654, 457.296, 367.215, 248.375), (-0.28340811, 0.07395907, 0, 0, 0, 0, 0.00019359, 1.76187114e-05),
              (0.003, -0.007, 0.001, 1.0), (435.2, 435.2, 367.4, 252.2, 0.0), 752, 480),
    "tum": ((517.3, 516.5, 318.6, 255.3), (0.2624, -0.9531, 1.1633, 0, 0, 0, -0.0054, 0.0026), (0, 0, 0, 1.0),
            (517.3, 516.5, 318.6, 255.3, 0.0), 640, 480),
    "rational": ((700.0, 705.0, 620.0, 188.0), (-0.2, 0.05, 0.01, 0.02, -0.01, 0.003, 0.001, -0.002),
                 (0.01, 0.02, -0.03, 1.0), (690.0, 690.0, 610.0, 180.0, 0.5), 1241, 376),
    "identity": ((500.0, 500.0, 320.0, 240.0), (0,) * 8, (0, 0, 0, 1.0), (500.0, 500.0, 320.0, 240.0, 0.0), 640, 480),
}


def rect_vector(case):
    k, d, q, kd, _, _ = RECT_CASES[case]
    q = np.array(q, np.float64)
    q = q / np.sqrt((q * q).sum())
    return np.array(list(k) + [0.0] + list(d) + list(q) + list(kd), np.float64)  # ref_shim.cpp layout, 22 doubles


VOC_FILE = os.path.join(HERE, "voc_k9_L3.bin")


def ref_transform(voc_path, desc, levelsup):
    n = len(desc)
    ids, vals = np.zeros(n + 1, np.int32), np.zeros(n + 1, np.float32)
    fn, fo, fi = np.zeros(n + 1, np.int32), np.zeros(n + 2, np.int32), np.zeros(n + 1, np.int32)
    nb, nf = C.c_int(), C.c_int()
    desc = np.ascontiguousarray(desc, np.uint64)
    R.ref_voc_transform(voc_path.encode(), _p(desc), n, levelsup, _p(ids), _p(vals), C.byref(nb), _p(fn), _p(fo), _p(fi),
                        C.byref(nf))
    return ids[:nb.value], vals[:nb.value], fn[:nf.value], fo[:nf.value + 1], fi[:fo[nf.value]]


def make_voc(out, rng):
    """A small vocabulary trained by the reference's own MiniBow2 create() (k-means++ on descriptors of synthetic
    frames, k=9, L=3, IDF weights) and written by its saveRaw(); plus transform() outputs of the reference for
    descriptor sets of several sizes and levelsup values."""
    train = []
    for i in range(14):
        k, d = O.Extractor(700, 1.2, 8, 20, 7, 2).detect(S.image(900 + i, 640, 480))
        train.append(d)
    counts = np.array([len(d) for d in train], np.int32)
    alld = np.ascontiguousarray(np.concatenate(train))
    nwords = R.ref_voc_create(_p(alld), _p(counts), len(train), 9, 3, 1234, VOC_FILE.encode())
    print("vocabulary:", nwords, "words,", os.path.getsize(VOC_FILE), "bytes")
    sets = {
        "a": O.Extractor(1000, 1.2, 8, 20, 7, 2).detect(S.image(77, 752, 480))[1],
        "b": O.Extractor(2500, 1.2, 8, 20, 7, 2).detect(S.image(78, 1241, 376))[1],
        "c": train[3][:1],
        "d": rng.integers(0, 2**63, (300, 4)).astype(np.uint64),
    }
    for name, d in sets.items():
        out[f"bow_{name}_desc"] = d
        for lu in (0, 1, 2, 3, 5):
            ids, vals, fn, fo, fi = ref_transform(VOC_FILE, d, lu)
            out[f"bow_{name}_{lu}_ids"], out[f"bow_{name}_{lu}_vals"] = ids, vals
            out[f"bow_{name}_{lu}_fv_nodes"], out[f"bow_{name}_{lu}_fv_offsets"], out[f"bow_{name}_{lu}_fv_indices"] = fn, fo, fi
        print(name, len(d), "->", len(out[f"bow_{name}_1_ids"]), "words,", len(out[f"bow_{name}_1_fv_nodes"]), "nodes,",
              len(d) - len(out[f"bow_{name}_1_fv_indices"]), "features on zero-weight words")


def make_next():
    """Golden vectors for the rows after the extractor, from the reference's own headers compiled into
    oracle/_ref: Preprocess::undistortKeypoints' per-point math."""
    rng = np.random.default_rng(21)
    out = {}
    for case, spec in RECT_CASES.items():
        w, h = spec[4], spec[5]
        n = 400
        xy = np.stack([rng.uniform(0, w - 1, n), rng.uniform(0, h - 1, n)], 1)
        xy[:100] = np.round(xy[:100])  # level-0 keypoints are integers
        xy[100] = (0, 0)
        xy[101] = (w - 1, h - 1)
        rect = rect_vector(case)
        nrm, und = np.zeros(2 * n), np.zeros(2 * n)
        R.ref_undistort_points(_p(rect), _p(xy), n, _p(nrm), _p(und))
        out[f"und_{case}_rect"] = rect
        out[f"und_{case}_xy"] = xy
        out[f"und_{case}_normalized"] = nrm.reshape(n, 2)
        out[f"und_{case}_undistorted"] = und.reshape(n, 2)
    # Input::computeGrayscaleImage == cv2.cvtColor(COLOR_RGBA2GRAY): a seeded image with a padded pitch and a colour
    # lattice; the fixed-point formula itself is checked here against cv2 for all 2^24 colours
    v = np.arange(1 << 24, dtype=np.uint32)
    cube = np.zeros((4096, 4096, 4), np.uint8)
    for ch in range(3):
        cube[..., ch] = ((v >> (8 * ch)) & 255).reshape(4096, 4096)
    cube[..., 3] = 77
    want = cv2.cvtColor(cube, cv2.COLOR_RGBA2GRAY)
    c64 = cube.astype(np.int64)
    assert np.array_equal(want, (c64[..., 0] * 9798 + c64[..., 1] * 19235 + c64[..., 2] * 3735 + (1 << 14)) >> 15)
    rgba = rng.integers(0, 256, (61, 97, 4), dtype=np.uint8)
    lat = np.stack(np.meshgrid(*[np.arange(0, 256, 17)] * 3, indexing="ij"), -1).reshape(64, 64, 3).astype(np.uint8)
    lat = np.concatenate([lat, np.full((64, 64, 1), 255, np.uint8)], -1)
    out.update(gray_rgba=rgba, gray_out=cv2.cvtColor(rgba, cv2.COLOR_RGBA2GRAY), gray_lattice=lat,
               gray_lattice_out=cv2.cvtColor(lat, cv2.COLOR_RGBA2GRAY))
    make_voc(out, rng)
    np.savez_compressed(os.path.join(HERE, "golden_next.npz"), **out)


if __name__ == "__main__":
    import sys

    which = sys.argv[1:] or ["cv", "ref", "extract", "next"]
    if "cv" in which:
        make_cv()
    if "ref" in which:
        make_ref()
    if "extract" in which:
        make_extract()
    if "next" in which:
        make_next()
