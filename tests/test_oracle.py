"""CPU tests: the in-repo oracle against the committed golden vectors (cv2 4.13 outputs, outputs of the reference's
own compiled classes, and the tier-1 hybrid end to end), and — where /root/reference was compiled into oracle/_ref —
live against the reference itself."""
import ctypes as C
import hashlib
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
G = os.path.join(HERE, "golden")


def sha(a):
    return hashlib.sha1(np.ascontiguousarray(a).tobytes()).hexdigest()


def _p(a):
    return C.c_void_p(a.ctypes.data)


@pytest.fixture(scope="module")
def gcv():
    return np.load(os.path.join(G, "golden_cv.npz"))


@pytest.fixture(scope="module")
def gref():
    return np.load(os.path.join(G, "golden_ref.npz"))


@pytest.fixture(scope="module")
def gext():
    return np.load(os.path.join(G, "golden_extract.npz"))


@pytest.mark.parametrize("name", ["noise", "synth"])
def test_cvlite_matches_cv2_golden(oracle, gcv, name):
    src = gcv[f"{name}_src"]
    want = gcv[f"{name}_resize"]
    assert np.array_equal(oracle.resize(src, *want.shape), want)
    assert np.array_equal(oracle.blur(src), gcv[f"{name}_blur"])
    kps = gcv[f"{name}_fast_kps"]
    pos = 0
    for (y0, x0, h, w, th, n) in gcv[f"{name}_fast_rois"]:
        got = oracle.fast(src, int(y0), int(x0), int(h), int(w), int(th))
        assert np.array_equal(got, kps[pos:pos + n]), (y0, x0, h, w, th)
        pos += n
    assert pos == len(kps)


def test_scale_pyramid_and_umax_golden(oracle, gref):
    for (nl, nf) in ((8, 1000), (8, 1200), (8, 2000), (4, 1000), (7, 2000), (1, 500)):
        s, f = oracle.scale_pyramid(nl, 1.2, nf)
        assert np.array_equal(s, gref[f"pyr_{nl}_{nf}_scale"])  # doubles, bit-exact
        assert np.array_equal(f, gref[f"pyr_{nl}_{nf}_features"])
    # quotas for C1
    assert list(oracle.scale_pyramid(8, 1.2, 1000)[1]) == [217, 181, 151, 126, 105, 87, 73, 60]
    um = np.zeros(16, np.int32)
    oracle.lib().orc_umax(_p(um))
    assert np.array_equal(um, gref["umax"])
    assert list(um) == [15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3]


def test_pattern_table_is_the_reference_table(gref):
    import re

    txt = open(os.path.join(os.path.dirname(HERE), "include", "snakeorb_pattern.h")).read()
    body = txt[txt.index("kOrbPattern[1024] = {") + len("kOrbPattern[1024] = {"):]
    nums = np.array([int(v) for v in re.findall(r"-?\d+", body[: body.index("};")])], np.int8)
    assert len(nums) == 1024 and sha(nums) == str(gref["pattern_sha1"])


def test_angle_and_descriptor_golden(oracle, gref):
    from snakeorb import synthetic as S

    img = S.image(5, 200, 160)
    assert sha(img) == str(gref["od_img_sha1"]), "numpy PCG64 stream changed: regenerate goldens"
    xy = gref["od_xy"]
    ang = oracle.angles(img, xy)
    assert np.array_equal(ang, gref["od_angle"])  # bit-exact floats (same libm)
    g = oracle.blur(img)
    assert np.array_equal(oracle.descriptors(g, xy, ang), gref["od_desc"])
    assert np.array_equal(oracle.descriptors(g, xy, gref["od_angle2"]), gref["od_desc2"])


@pytest.mark.parametrize("t", range(5))
def test_distribute_golden(oracle, gref, t):
    a = gref[f"dist{t}_args"]
    sel, perm = oracle.distribute(gref[f"dist{t}_in"], (a[0], a[1]), (a[2], a[3]), int(a[4]))
    assert sel.tobytes() == gref[f"dist{t}_out"].tobytes()
    assert perm.tobytes() == gref[f"dist{t}_perm"].tobytes()


def test_knn2_golden(oracle, gref):
    k = oracle.match_knn2(gref["knn_a"], gref["knn_b"])
    got = np.stack([k["dist0"], k["idx0"], k["dist1"], k["idx1"]], 1)
    assert np.array_equal(got, gref["knn_out"])


def _cases():
    import importlib.util

    spec = importlib.util.spec_from_file_location("make_golden_cases", os.path.join(G, "cases.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


@pytest.mark.parametrize("name", ["euroc_c1", "euroc_c2", "kitti_c3", "tum_c4", "noise", "checker", "constant", "small4"])
def test_extractor_end_to_end_golden(oracle, gext, name):
    cases = _cases()
    kind, w, h, nf, nl = cases.CASES[name]
    img = cases.case_image(kind, w, h)
    assert sha(img) == str(gext[f"{name}_sha1"]), "numpy PCG64 stream changed: regenerate goldens"
    ex = oracle.Extractor(nf, 1.2, nl, 20, 7, 1)
    k, d = ex.detect(img)
    assert k.tobytes() == gext[f"{name}_kps"].tobytes()
    assert d.tobytes() == gext[f"{name}_desc"].tobytes()
    assert [len(ex.level_kps(l, 0)) for l in range(nl)] == list(gext[f"{name}_nraw"])
    assert [sha(ex.level(l)) for l in range(nl)] == list(gext[f"{name}_level_sha1"])
    # the reference only blurs levels that kept at least one keypoint (ORBExtractor.cpp:208)
    for l in range(nl):
        if len(ex.level_kps(l, 1)):
            assert sha(ex.level(l, True)) == str(gext[f"{name}_gauss_sha1"][l])


def test_extractor_threads_invariant(oracle):
    from snakeorb import synthetic as S

    img = S.image(321, 640, 480)
    a = oracle.Extractor(1000, 1.2, 8, 20, 7, 1).detect(img)
    b = oracle.Extractor(1000, 1.2, 8, 20, 7, 4).detect(img)
    assert a[0].tobytes() == b[0].tobytes() and a[1].tobytes() == b[1].tobytes()


# ---- live against the compiled reference (skipped on machines without oracle/_ref) -------------------------------
def test_live_angle_descriptor_vs_reference(oracle, refso):
    from snakeorb import synthetic as S

    rng = np.random.default_rng(99)
    img = S.adversarial("noise", 300, 200, 3)
    g = oracle.blur(img)
    n = 4000
    xy = np.stack([rng.integers(19, 300 - 19, n), rng.integers(19, 200 - 19, n)], 1).astype(np.float32)
    ang = np.zeros(n, np.float32)
    refso.ref_angles(_p(img), 200, 300, C.c_size_t(300), _p(xy), n, _p(ang))
    assert np.array_equal(oracle.angles(img, xy), ang)
    d = np.zeros((n, 4), np.uint64)
    refso.ref_descriptors(_p(g), 200, 300, C.c_size_t(300), _p(xy), _p(ang), n, _p(d))
    assert np.array_equal(oracle.descriptors(g, xy, ang), d)


def test_live_distribute_vs_reference(oracle, refso):
    rng = np.random.default_rng(5)
    for t in range(30):
        n = int(rng.integers(2, 6000))
        w, h = int(rng.integers(60, 1300)), int(rng.integers(60, 500))
        pos = np.unique(np.stack([rng.integers(3, h - 3, n), rng.integers(3, w - 3, n)], 1), axis=0)
        k = np.zeros(len(pos), oracle.KP32)
        k["x"], k["y"], k["size"], k["angle"] = pos[:, 1], pos[:, 0], 7, -1
        k["response"] = rng.integers(7, 12 if t % 2 else 200, len(pos))
        target = int(rng.integers(1, 500))
        a = oracle.distribute(k, (16, 16), (w + 16, h + 16), target, "oracle")
        b = oracle.distribute(k, (16, 16), (w + 16, h + 16), target, "ref")
        assert a[0].tobytes() == b[0].tobytes() and a[1].tobytes() == b[1].tobytes()
        assert len(a[0]) <= max(target + 3, 0) or len(pos) <= target


def test_live_hybrid_vs_oracle(oracle, refso):
    cv2 = pytest.importorskip("cv2")
    if not cv2.__version__.startswith("4.13"):
        pytest.skip("goldens are pinned to cv2 4.13")
    import hybrid
    from snakeorb import synthetic as S

    img = S.image(4242, 400, 300)
    k, d, _ = hybrid.hybrid_detect(img, 700, 1.2, 6, 20, 7)
    k2, d2 = oracle.Extractor(700, 1.2, 6, 20, 7, 1).detect(img)
    assert k.tobytes() == k2.tobytes() and d.tobytes() == d2.tobytes()


def test_rgba_to_gray_vs_cv2_golden(oracle):
    """Input::computeGrayscaleImage: the restated fixed-point conversion against cv2 4.13's cvtColor outputs
    (tests/golden/make_golden.py next also checks the formula exhaustively over all 2^24 colours)."""
    import os

    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden_next.npz"))
    assert np.array_equal(oracle.rgba_to_gray(np.ascontiguousarray(g["gray_rgba"])), g["gray_out"])
    assert np.array_equal(oracle.rgba_to_gray(np.ascontiguousarray(g["gray_lattice"])), g["gray_lattice_out"])
    wide = np.zeros((61, 120, 4), np.uint8)
    wide[:, :97] = g["gray_rgba"]
    assert np.array_equal(oracle.rgba_to_gray(wide[:, :97]), g["gray_out"])  # padded pitch
