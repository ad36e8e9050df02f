"""The reference's own data fixtures (saiga data/vision/{0,50}.features: real ORB features of two frames, written by
its BinaryFile) as inputs: the feature-cache reader against real files, and matcher parity on real descriptors."""
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
F0 = os.path.join(HERE, "golden", "saiga_0.features")
F50 = os.path.join(HERE, "golden", "saiga_50.features")


def test_feature_cache_reader_on_reference_files(tmp_path):
    """snk_features_read parses files the reference wrote; snk_features_write reproduces them byte for byte."""
    import snakeorb as so

    for path, n in ((F0, 1007), (F50, 1006)):
        k, d = so.features_read(path)
        assert len(k) == n and d.shape == (n, 4)
        assert np.all((k["octave"] >= 0) & (k["octave"] < 8)) and np.all(k["x"] > 0) and np.all(k["y"] > 0)
        assert np.all((k["angle"] >= 0) & (k["angle"] < 360)) and np.all(k["response"] >= 7)
        # KeyPoint size = 31 * scale(level) truncated (ORBExtractor.cpp:229), scale = 1.2^level in float
        scale = np.cumprod(np.concatenate([[np.float32(1)], np.full(7, np.float32(1.2))])).astype(np.float32)
        assert np.array_equal(k["size"], np.floor(31 * scale[k["octave"]].astype(np.float64)))
        out = str(tmp_path / "copy.features")
        so.features_write(out, k, d)
        assert open(out, "rb").read() == open(path, "rb").read()


def test_knn2_on_real_descriptors_oracle_vs_compiled_reference(oracle, refso):
    """BruteForceMatcher::matchKnn2 + filterMatches(60, 0.8) (TrackingCoarse.cpp:351-352) between the two real frames:
    restatement == the reference's compiled class."""
    import ctypes as C

    import snakeorb as so

    d0, d50 = so.features_read(F0)[1], so.features_read(F50)[1]
    knn = oracle.match_knn2(d0, d50)
    ref = np.zeros((len(d0), 4), np.int32)
    refso.ref_knn2(C.c_void_p(d0.ctypes.data), len(d0), C.c_void_p(d50.ctypes.data), len(d50), C.c_void_p(ref.ctypes.data))
    assert np.array_equal(np.stack([knn["dist0"], knn["idx0"], knn["dist1"], knn["idx1"]], 1), ref)
    m = oracle.filter_matches(knn, 60, 0.8)
    ij = np.zeros((len(d0), 2), np.int32)
    c = refso.ref_knn2_filter(C.c_void_p(d0.ctypes.data), len(d0), C.c_void_p(d50.ctypes.data), len(d50), 60,
                                 C.c_float(0.8), C.c_void_p(ij.ctypes.data))
    assert c == len(m) and np.array_equal(ij[:c], m) and c > 50


@pytest.mark.gpu
def test_gpu_matchers_on_real_descriptors(oracle):
    import snakeorb as so

    ctx = so.Context()
    d0, d50 = so.features_read(F0)[1], so.features_read(F50)[1]
    knn = ctx.matchKnn2(d0, d50)
    want = oracle.match_knn2(d0, d50)
    assert knn.tobytes() == want.tobytes()
    assert np.array_equal(ctx.filterMatches(knn, 60, 0.8), oracle.filter_matches(want, 60, 0.8))
    # vocabulary transform of real descriptors (the vocabulary the reference's create() trained, tests/golden)
    voc = so.Vocabulary(os.path.join(HERE, "golden", "voc_k9_L3.bin"))
    vo = oracle.Vocabulary(os.path.join(HERE, "golden", "voc_k9_L3.bin"))
    for d in (d0, d50):
        got, w = voc.transform(ctx, d, 1), vo.transform(d, 1)
        for a, b in zip(got, w):
            assert a.tobytes() == b.tobytes()


def test_reference_feature_grid_tests_restated(oracle):
    """saiga tests/test_vision_feature_grid.cpp restated. FeatureGridBounds.Create (:48-77): EuRoC intrinsics,
    resolution 5000, bounds within 0.5 of the reference's golden numbers and the grid just covers them.
    FeatureGrid.Create (:79-212): every keypoint sits in the cell it is listed under, and the grid radius search
    returns exactly the brute-force set, for 500 random positions with r = 40."""
    import snakeorb as so
    from snakeorb import synthetic as S

    cam = so.Camera()
    cam.fx, cam.fy, cam.cx, cam.cy, cam.s, cam.bf = 458.654, 457.296, 367.215, 248.375, 0.0, 0.0
    dist = np.array([-0.283408, 0.0739591, 0, 0, 0, 0, 0.00019359, 1.76187e-05])
    g = so.grid_bounds(752, 480, cam, dist, 5000)
    ref_bmin, ref_bmax = np.array([-135.795, -92.8749]), np.array([895.507, 565.497])
    assert np.all(np.abs(np.array([g.bmin_x, g.bmin_y]) - ref_bmin) < 0.5)
    assert np.all(np.abs(np.array([g.bmax_x, g.bmax_y]) - ref_bmax) < 0.5)
    ref_size = ref_bmax - ref_bmin
    grid_size = 20.0 * np.array([g.cols, g.rows])
    assert np.all(grid_size >= ref_size) and np.all(grid_size < ref_size + 20.0)

    rng = np.random.default_rng(953454)
    kp = np.zeros(500, oracle.KP64)
    kp["x"] = rng.uniform(ref_bmin[0] - 50, ref_bmax[0] + 50, 500)
    kp["y"] = rng.uniform(ref_bmin[1] - 50, ref_bmax[1] + 50, 500)
    perm, cells = oracle.feature_grid(g, kp)
    permuted = np.zeros_like(kp)
    permuted[perm] = kp
    inside = (permuted["x"] >= g.bmin_x) & (permuted["y"] >= g.bmin_y) & (permuted["x"] < g.bmax_x) & (permuted["y"] < g.bmax_y)
    for r in range(g.rows):
        for c in range(g.cols):
            b, e = cells[r * g.cols + c]
            for i in range(b, e):
                assert int(np.floor((permuted["x"][i] - g.bmin_x) / 20)) == c and int(np.floor((permuted["y"][i] - g.bmin_y) / 20)) == r
    cells = np.ascontiguousarray(cells)
    view = so.FrameView()
    view.n, view.kp_un, view.grid = len(kp), permuted.ctypes.data, g
    view.grid.cells = cells.ctypes.data
    for _ in range(500):
        x, y = rng.uniform(ref_bmin[0] - 50, ref_bmax[0] + 50), rng.uniform(ref_bmin[1] - 50, ref_bmax[1] + 50)
        d2 = (permuted["x"] - x) ** 2 + (permuted["y"] - y) ** 2
        want = np.nonzero(inside & (d2 < 40.0 * 40.0))[0]
        got = np.sort(oracle.features_in_area(view, x, y, 40.0))
        assert np.array_equal(got, want), (x, y)
