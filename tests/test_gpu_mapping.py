"""GPU parity of the local-mapping / loop-closing searches: every candidate pair, index and
distance equals the oracle's (oracle/matchers_ref.cpp)."""
import numpy as np
import pytest

import matchdata as M

pytestmark = pytest.mark.gpu
K = (718.856, 718.856, 607.19, 185.22)
BF = 386.1
W, H = 1241, 376


@pytest.fixture(scope="module")
def so():
    import snakeorb

    return snakeorb


@pytest.fixture(scope="module")
def ctx(so):
    return so.Context()


@pytest.fixture(scope="module")
def feats(oracle):
    from snakeorb import synthetic as S

    ex = oracle.Extractor(2000, 1.2, 8, 20, 7, 4)
    left = S.image(3000, W, H)
    right = S.stereo_right(left, 3001, 4.0, 40.0)
    return dict(l=ex.detect(left), r=ex.detect(right), scales=oracle.scale_pyramid(8, 1.2, 2000)[0])


def _scene(so, feats, seed, n_points):
    from snakeorb import synthetic as S

    rng = np.random.default_rng(seed)
    q = S.quat_from_axis_angle(rng.normal(size=3), np.radians(rng.uniform(0, 0.5)))
    t = rng.uniform(-0.05, 0.05, 3)
    frame = M.make_frame(*feats["l"], W, H, rng, q, t)
    mp = S.map_points_from_keypoints(frame.kp, frame.desc, K, rng, n_points, feats["scales"])
    return rng, frame, M.fusion_points(mp, rng), M.camera(*K, BF), so.scale_pyramid_c(1.2, feats["scales"])


@pytest.mark.parametrize("th,obs_factor,with_mask,n_points", [(3.0, 1.5, False, 6000), (2.0, 1.0, True, 2500), (3.0, 2.0, True, 0)])
def test_mapping_fuse(so, ctx, oracle, feats, th, obs_factor, with_mask, n_points):
    rng, frame, pts, cam, pyr = _scene(so, feats, 31, n_points)
    mask = (rng.random(n_points) < 0.8).astype(np.uint8) if with_mask else None
    ck, ci = ctx.Fuse(frame.view, mask, pts, cam, pyr, th, obs_factor, 50)
    ck2, ci2, ops = oracle.fuse_search(frame.view, mask, pts, cam, pyr, th, obs_factor, 50)
    assert np.array_equal(ck, ck2) and np.array_equal(ci, ci2)
    if n_points:
        assert len(ck) > n_points // 20 and ops > len(ck)
        if mask is not None:
            assert set(ci) <= set(pts["id"][mask > 0])


def test_loop_fuse_search(so, ctx, oracle, feats):
    rng, frame, pts, cam, pyr = _scene(so, feats, 32, 4000)
    skip = (rng.random(len(pts)) < 0.25).astype(np.uint8)
    for th in (4.0, 10.0):
        bi, bd = ctx.FuseLoopSearch(frame.view, skip, pts, cam, th)
        bi2, bd2 = oracle.fuse_search_loop(frame.view, skip, pts, cam, th)
        assert np.array_equal(bi, bi2) and np.array_equal(bd, bd2)
        assert np.all(bi[skip > 0] == -1) and (bd <= 50).sum() > 500
    bi, bd = ctx.FuseLoopSearch(frame.view, None, pts, cam, 4.0)
    bi2, bd2 = oracle.fuse_search_loop(frame.view, None, pts, cam, 4.0)
    assert np.array_equal(bi, bi2) and np.array_equal(bd, bd2)


@pytest.mark.parametrize("epi,fd", [(2.0, 50), (4.0, 30)])
def test_search_for_triangulation_bow(so, ctx, oracle, feats, epi, fd):
    rng = np.random.default_rng(41)
    f1, f2, np1, np2, E = M.stereo_keyframes(feats, rng, K, W, H, BF / K[0])
    fv1, fv2 = M.featvec(f1.desc, rng, 60), M.featvec(f2.desc, rng, 60)
    cam = M.camera(*K, BF)
    pairs = ctx.SearchForTriangulation2(f1.view, np1, fv1, f2.view, np2, fv2, E, cam, epi, fd)
    pairs2, ops = oracle.search_triangulation_bow(f1.view, np1, fv1, f2.view, np2, fv2, E, cam, epi, fd)
    assert np.array_equal(pairs, pairs2) and len(pairs) > 100 and ops > 10 * len(pairs)
    assert np.all(f1.k2p[pairs[:, 0]] < 0) and np.all(f2.k2p[pairs[:, 1]] < 0)
    # rows agree within the epipolar gate (pure x translation)
    assert np.all(np.abs(f1.kp["y"][pairs[:, 0]] - f2.kp["y"][pairs[:, 1]]) < 2 * epi + 1e-6)


@pytest.mark.parametrize("epi,fd", [(2.0, 50), (6.0, 40)])
def test_search_for_triangulation_project(so, ctx, oracle, feats, epi, fd):
    rng = np.random.default_rng(42)
    f1, f2, np1, np2, E = M.stereo_keyframes(feats, rng, K, W, H, BF / K[0])
    g = f1.view.grid
    grid = BF / rng.uniform(10.0, 30.0, ((g.rows + 3) // 4, (g.cols + 3) // 4))  # scene depth per coarse cell
    cam = M.camera(*K, BF)
    pairs = ctx.SearchForTriangulationProject(f1.view, np1, f2.view, np2, grid, E, cam, epi, fd)
    pairs2, ops = oracle.search_triangulation_project(f1.view, np1, f2.view, np2, grid, E, cam, epi, fd)
    assert np.array_equal(pairs, pairs2) and len(pairs) > 100 and ops > len(pairs)
    assert np.all(np.diff(pairs[:, 0]) > 0)


def test_compute_matches_prediction(so, ctx, oracle, feats):
    rng = np.random.default_rng(43)
    f1, f2, _, _, _ = M.stereo_keyframes(feats, rng, K, W, H, BF / K[0])
    pos = np.stack([f1.kp["x"], f1.kp["y"]], 1) + rng.normal(0, 3.0, (len(f1.kp), 2))
    for radius, fe in ((50.0, 50), (15.0, 80)):
        m, matches, newpos = ctx.ComputeMatchesPrediction(f1.desc, pos, f2.view, radius, fe)
        m2, matches2, newpos2 = oracle.match_prediction(f1.desc, pos, f2.view, radius, fe)
        assert m == m2 and np.array_equal(matches, matches2) and newpos.tobytes() == newpos2.tobytes() and m > 100
        hit = matches >= 0
        assert np.array_equal(newpos[hit, 0], f2.kp["x"][matches[hit]]) and np.array_equal(newpos[~hit], pos[~hit])
    m, matches, _ = ctx.ComputeMatchesPrediction(np.zeros((0, 4), np.uint64), np.zeros((0, 2)), f2.view)
    assert m == 0
