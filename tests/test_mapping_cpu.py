"""Oracle checks for the local-mapping / loop-closing searches: the pieces that compile from the
reference are pinned bit for bit, the search loops are checked against brute-force numpy restatements."""
import ctypes as C

import numpy as np

import matchdata as M

K = (500.0, 510.0, 320.0, 240.0)


def test_epipolar_distance_vs_compiled_reference(oracle, refso):
    """Saiga::EpipolarDistanceSquared compiled from saiga/vision/reconstruction/Epipolar.cpp."""
    refso.ref_epipolar_distance_squared.restype = C.c_double
    rng = np.random.default_rng(0)
    for _ in range(20000):
        p1, p2, F = rng.uniform(-1, 1, 2), rng.uniform(-1, 1, 2), rng.normal(size=9)
        want = refso.ref_epipolar_distance_squared(C.c_void_p(p1.ctypes.data), C.c_void_p(p2.ctypes.data),
                                                   C.c_void_p(F.ctypes.data))
        assert oracle.epipolar_distance_squared(p1, p2, F) == want


def _frames(oracle, seed):
    from snakeorb import synthetic as S

    rng = np.random.default_rng(seed)
    ex = oracle.Extractor(600, 1.2, 8, 20, 7, 2)
    left = S.image(500 + seed, 640, 480)
    right = S.stereo_right(left, 600 + seed, 3.0, 25.0)
    feats = dict(l=ex.detect(left), r=ex.detect(right))
    return rng, M.stereo_keyframes(feats, rng, K, 640, 480, 0.1)


def test_triangulation_bow_vs_bruteforce(oracle):
    rng, (f1, f2, np1, np2, E) = _frames(oracle, 1)
    fv1, fv2 = M.featvec(f1.desc, rng, 40), M.featvec(f2.desc, rng, 40)
    cam = M.camera(*K, 50.0)
    pairs, ops = oracle.search_triangulation_bow(f1.view, np1, fv1, f2.view, np2, fv2, E, cam, 3.0, 50)
    ids1, off1, idx1 = fv1._keep
    ids2, off2, idx2 = fv2._keep
    th2 = (np.float32(3.0) * 2 / K[0]) ** 2
    want = []
    for a, nid in enumerate(ids1):
        b = np.nonzero(ids2 == nid)[0]
        if len(b) == 0:
            continue
        for i1 in idx1[off1[a]:off1[a + 1]]:
            if f1.k2p[i1] >= 0:
                continue
            best, bi = 50, -1
            for i2 in idx2[off2[b[0]]:off2[b[0] + 1]]:
                if f2.k2p[i2] >= 0:
                    continue
                d = oracle.hamming(f1.desc[i1], f2.desc[i2])
                if d > 50 or d > best:
                    continue
                if (np1[i1, 1] - np2[i2, 1]) ** 2 < th2:  # pure x translation: squared row distance
                    best, bi = d, i2
            if bi >= 0:
                want.append((i1, bi))
    assert len(want) > 30 and np.array_equal(pairs, np.array(want, np.int32))


def test_prediction_and_loop_fuse_vs_bruteforce(oracle):
    rng, (f1, f2, np1, np2, E) = _frames(oracle, 2)
    pos = np.stack([f1.kp["x"], f1.kp["y"]], 1) + rng.normal(0, 2.0, (len(f1.kp), 2))
    m, matches, newpos = oracle.match_prediction(f1.desc, pos, f2.view, 30.0, 60)
    xy2 = np.stack([f2.kp["x"], f2.kp["y"]], 1)
    for i in range(0, len(pos), 7):
        d2 = ((xy2 - pos[i]) ** 2).sum(1)
        cand = np.nonzero(d2 < 900.0)[0]
        dist = np.array([oracle.hamming(f1.desc[i], f2.desc[j]) for j in cand])
        if len(cand) and dist.min() < 60:
            assert dist[np.nonzero(cand == matches[i])[0][0]] == dist.min()
            assert np.array_equal(newpos[i], xy2[matches[i]])
        else:
            assert matches[i] == -1 and np.array_equal(newpos[i], pos[i])
    assert m == (matches >= 0).sum() and m > 100


def test_candidate_enumeration_vs_reference_grid_classes(oracle, refso):
    """Features::GetFeaturesInArea, the candidate order of every projection / fuse / prediction search: the oracle's
    enumeration against the reference's own FeatureGridBounds2 / FeatureGrid2 objects (cell ranges with the float
    radius, clamping at the borders, cx-outer / cy-inner / index order, float r*r distance gate)."""
    rng, (f1, f2, np1, np2, E) = _frames(oracle, 3)
    g = f1.view.grid
    out = np.zeros(f1.view.n, np.int32)
    total = 0
    for _ in range(400):
        px, py = rng.uniform(g.bmin_x - 30, g.bmax_x + 30), rng.uniform(g.bmin_y - 30, g.bmax_y + 30)
        r = float(np.float32(rng.choice([2.5, 7.0, 20.0, 50.0, 133.3])))
        c = refso.ref_features_in_area(C.c_double(g.bmin_x), C.c_double(g.bmin_y), C.c_double(g.bmax_x), C.c_double(g.bmax_y),
                                       g.rows, g.cols, C.c_void_p(f1.kp.ctypes.data), f1.view.n, C.c_double(px),
                                       C.c_double(py), C.c_float(r), C.c_void_p(out.ctypes.data))
        assert c >= 0, "keypoints not in grid order"
        assert np.array_equal(oracle.features_in_area(f1.view, px, py, r), out[:c])
        total += c
    assert total > 2000


def test_grid_bounds_product_oracle_reference(oracle, refso):
    """FeatureGridBounds2::computeFromIntrinsicsDist: the library's host helper (no GPU involved), the oracle and the
    reference's own class compiled into oracle/_ref agree to the last bit."""
    import snakeorb as so

    cases = [(752, 480, [458.654, 457.296, 367.215, 248.375, 0.0], [-0.28340811, 0.07395907, 0, 0, 0, 0, 0.00019359, 1.76187114e-05]),
             (640, 480, [517.3, 516.5, 318.6, 255.3, 0.0], [0.2624, -0.9531, 1.1633, 0, 0, 0, -0.0054, 0.0026]),
             (1241, 376, [718.856, 718.856, 607.19, 185.22, 0.0], [0.0] * 8),
             (1280, 720, [700.0, 705.0, 620.0, 350.0, 0.5], [-0.2, 0.05, 0.01, 0.02, -0.01, 0.003, 0.001, -0.002])]
    for w, h, K, D in cases:
        for res in (20, 7):
            Ka, Da = np.array(K), np.array(D, np.float64)
            o6, rc = np.zeros(6), np.zeros(2, np.int32)
            refso.ref_grid_bounds(w, h, C.c_void_p(Ka.ctypes.data), C.c_void_p(Da.ctypes.data), res, C.c_void_p(o6.ctypes.data),
                                  C.c_void_p(rc.ctypes.data))
            go = so.GridView()
            oracle.lib().orc_grid_bounds(w, h, C.c_void_p(Ka.ctypes.data), C.c_void_p(Da.ctypes.data), res, C.byref(go))
            cam = M.camera(K[0], K[1], K[2], K[3], 0.0)
            cam.s = K[4]
            gp = so.grid_bounds(w, h, cam, Da, res)
            for g in (go, gp):
                got = np.array([g.bmin_x, g.bmin_y, g.bmax_x, g.bmax_y, g.cell_size_inv_x, g.cell_size_inv_y])
                assert got.tobytes() == o6.tobytes() and (g.rows, g.cols) == tuple(rc), (w, h, res)
