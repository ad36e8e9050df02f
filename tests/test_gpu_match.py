"""GPU parity tests of the Hamming matchers: every output index and distance must equal the oracle's."""
import os

import numpy as np
import pytest

import matchdata as M

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def so():
    import snakeorb

    return snakeorb


@pytest.fixture(scope="module")
def ctx(so):
    return so.Context()


@pytest.fixture(scope="module")
def feats(oracle):
    """Oracle-extracted features of a KITTI-shaped stereo pair and a second frame (inputs only)."""
    from snakeorb import synthetic as S

    ex = oracle.Extractor(2000, 1.2, 8, 20, 7, 4)
    left = S.image(3000, 1241, 376)
    right = S.stereo_right(left, 3001, 4.0, 40.0)
    other = S.image(3002, 1241, 376)
    return dict(l=ex.detect(left), r=ex.detect(right), o=ex.detect(other), scales=oracle.scale_pyramid(8, 1.2, 2000)[0])


def test_distance_and_knn2(so, ctx, oracle, feats):
    rng = np.random.default_rng(0)
    a, b = feats["l"][1], feats["r"][1]
    n = min(len(a), len(b))
    d = ctx.distance(a[:n], b[:n])
    assert all(d[i] == oracle.hamming(a[i], b[i]) for i in range(0, n, 37))
    for (x, y) in ((a, b), (a[:1], b), (a, b[:1]), (a[:130], a[:130]), (a[:5], b[:0])):
        got = ctx.matchKnn2(x, y)
        want = oracle.match_knn2(x, y)
        assert got.tobytes() == want.tobytes()
        for th, ratio in ((60, 0.8), (256, 1.0), (30, 0.5)):
            assert np.array_equal(so.Context.filterMatches(got, th, ratio), oracle.filter_matches(want, th, ratio))
    # heavy ties: many duplicates of few descriptors — first-minimum semantics must hold
    base = a[rng.integers(0, 8, 700)]
    got = ctx.matchKnn2(a[:300], base)
    assert got.tobytes() == oracle.match_knn2(a[:300], base).tobytes()
    g = np.load(os.path.join(HERE, "golden", "golden_ref.npz"))
    k = ctx.matchKnn2(g["knn_a"], g["knn_b"])
    assert np.array_equal(np.stack([k["dist0"], k["idx0"], k["dist1"], k["idx1"]], 1), g["knn_out"])


@pytest.mark.parametrize("n,m", [(1000, 1000), (1, 1), (1, 2), (31, 33), (33, 31), (257, 129), (1000, 7), (5, 4096), (2000, 1999),
                                 (40000, 64)])
def test_knn2_split_train_path(so, ctx, oracle, n, m):
    """One problem too small for a thread per query runs with the train set split over warps (k_knn2_split: lane per
    train descriptor, REDUX min for best / second best, k_knn2_merge); n = 40000 takes the thread-per-query kernel.
    Tie-heavy inputs: the train set holds many exact copies, so first-minimum and second-best order decide."""
    rng = np.random.default_rng(n * 7919 + m)
    pool = rng.integers(0, 2**63, (max(2, min(m, 40) // 3), 4)).astype(np.uint64)
    d2 = pool[rng.integers(0, len(pool), m)].copy()
    d2[rng.random(m) < 0.3] ^= np.uint64(1) << np.uint64(int(rng.integers(0, 63)))
    d1 = pool[rng.integers(0, len(pool), n)] ^ (np.uint64(1) << rng.integers(0, 63, (n, 4)).astype(np.uint64))
    got = ctx.matchKnn2(d1, d2)
    assert got.tobytes() == oracle.match_knn2(d1, d2, threads=4).tobytes()


@pytest.mark.parametrize("relaxed,bf", [(True, 386.1), (False, 386.1), (True, 47.9)])
def test_stereo_row_search(so, ctx, oracle, feats, relaxed, bf):
    kl, dl = feats["l"]
    kr, dr = feats["r"]
    p = so.stereo_params(bf, relaxed, feats["scales"])
    n, rp, dp = ctx.StereoMatching(p, so.kp32_to_kp64(kl), dl, so.kp32_to_kp64(kr), dr)
    n2, rp2, dp2, ops = oracle.stereo_match(oracle.stereo_params(bf, relaxed, feats["scales"]), oracle.kp32_to_kp64(kl), dl,
                                            oracle.kp32_to_kp64(kr), dr)
    assert n == n2 and n > 0 and ops > 0
    assert rp.tobytes() == rp2.tobytes() and dp.tobytes() == dp2.tobytes()
    # rows outside the counting sort's range (rectified coordinates can leave the image): the general sort path,
    # and many keypoints on one row (long in-row insertion)
    for dy, squash in ((5000.0, False), (-300.0, False), (0.0, True)):
        a, b = so.kp32_to_kp64(kl), so.kp32_to_kp64(kr)
        if squash:
            a["y"], b["y"] = np.round(a["y"] / 40.0) * 40.0, np.round(b["y"] / 40.0) * 40.0
        a["y"] += dy
        b["y"] += dy
        n3, rp3, dp3 = ctx.StereoMatching(p, a, dl, b, dr)
        n4, rp4, dp4, _ = oracle.stereo_match(oracle.stereo_params(bf, relaxed, feats["scales"]), a, dl, b, dr)
        assert n3 == n4 and n3 > 0 and rp3.tobytes() == rp4.tobytes() and dp3.tobytes() == dp4.tobytes(), (dy, squash)
    # empty sides
    n, rp, dp = ctx.StereoMatching(p, so.kp32_to_kp64(kl), dl, so.kp32_to_kp64(kr[:0]), dr[:0])
    assert n == 0 and np.all(rp == -1000) and np.all(dp == -1000)
    n, rp, dp = ctx.StereoMatching(p, so.kp32_to_kp64(kl[:0]), dl[:0], so.kp32_to_kp64(kr), dr)
    assert n == 0 and len(rp) == 0


def test_stereo_device_path_equals_host_path(so, ctx, oracle):
    """Extractor results consumed on the device (images 2p, 2p+1 = stereo pair p), as bench.py does."""
    import ctypes as C

    from snakeorb import synthetic as S

    pairs = []
    for i in range(3):
        l = S.image(700 + i, 752, 480)
        pairs += [l, S.stereo_right(l, 800 + i, 4.0, 20.0)]
    imgs = np.stack(pairs)
    # expected values come from a second extractor: the handle under test must not have seen these images before, so
    # that results left over from an earlier run cannot stand in for an unordered read
    res = so.ORBExtractor(1200, 1.2, 8, 20, 7, max_batch=6).DetectBatch(imgs)
    ex = so.ORBExtractor(1200, 1.2, 8, 20, 7, max_batch=6)
    scales, _ = ex.getPyramid()
    p = so.stereo_params(47.9, True, scales)
    cudart = C.CDLL("libcudart.so")
    warm = np.stack([S.image(990 + i, 752, 480) for i in range(6)])
    ex.enqueue_ptr(warm.ctypes.data, 6, 480, 752, 752, 752 * 480, False)  # sizes the buffers with other images
    stride = ex.device_results()[3]
    bufs = [C.c_void_p() for _ in range(3)]
    for b, nbytes in zip(bufs, (3 * stride * 4, 3 * stride * 4, 3 * 4)):
        assert cudart.cudaMalloc(C.byref(b), C.c_size_t(nbytes)) == 0
    # default handles: the extractor and the context each own a stream; the call itself orders them
    ex.enqueue_ptr(imgs.ctypes.data, 6, 480, 752, 752, 752 * 480, False)
    ctx.stereo_orb_device(ex, 3, p, bufs[0].value, bufs[1].value, bufs[2].value)
    with pytest.raises(so.SnakeOrbError):
        ctx.stereo_orb_device(ex, 4, p, bufs[0].value, bufs[1].value, bufs[2].value)  # 8 images, 6 were enqueued
    # the next enqueue is ordered behind the stereo kernel: it may overwrite the results only after they were read
    ex.enqueue_ptr(warm.ctypes.data, 6, 480, 752, 752, 752 * 480, False)
    ctx.check()
    ex.check()
    rp = np.zeros((3, stride), np.float32)
    dp = np.zeros((3, stride), np.float32)
    nm = np.zeros(3, np.int32)
    assert cudart.cudaDeviceSynchronize() == 0
    for host, dev in ((rp, bufs[0]), (dp, bufs[1]), (nm, bufs[2])):
        assert cudart.cudaMemcpy(C.c_void_p(host.ctypes.data), dev, C.c_size_t(host.nbytes), 2) == 0
    for i in range(3):
        (kl, dl), (kr, dr) = res[2 * i], res[2 * i + 1]
        n2, rp2, dp2, _ = oracle.stereo_match(oracle.stereo_params(47.9, True, scales), oracle.kp32_to_kp64(kl), dl,
                                              oracle.kp32_to_kp64(kr), dr)
        assert nm[i] == n2 and n2 > 0
        assert rp[i, :len(kl)].tobytes() == rp2.tobytes() and dp[i, :len(kl)].tobytes() == dp2.tobytes()
    for b in bufs:
        cudart.cudaFree(b)


def _c3(so, feats, seed, n_points):
    from snakeorb import synthetic as S

    rng = np.random.default_rng(seed)
    K = (718.856, 718.856, 607.19, 185.22)
    q = S.quat_from_axis_angle(rng.normal(size=3), np.radians(rng.uniform(0, 0.5)))
    t = rng.uniform(-0.05, 0.05, 3)
    kl, dl = feats["l"]
    frame = M.make_frame(kl, dl, 1241, 376, rng, q, t)
    mp = S.map_points_from_keypoints(frame.kp, frame.desc, K, rng, n_points, feats["scales"])
    cam = M.camera(*K, 386.1)
    pyr = so.scale_pyramid_c(1.2, feats["scales"])
    return rng, frame, mp, cam, pyr


@pytest.mark.parametrize("th,n_points", [(4.0, 20000), (1.0, 3000), (5.0, 0)])
def test_search_by_projection_fine(so, ctx, oracle, feats, th, n_points):
    rng, frame, mp, cam, pyr = _c3(so, feats, 11, n_points)
    pts = M.fine_points(mp, rng)
    m, k2p, bi, bd, vis, pout = ctx.SearchByProjection2(frame.view, pts, cam, pyr, th, 0.8)
    m2, k2p2, bi2, bd2, vis2, pout2, ops = oracle.search_projection_fine(frame.view, pts, cam, pyr, th, 0.8)
    assert m == m2
    assert np.array_equal(bi, bi2) and np.array_equal(bd, bd2) and np.array_equal(vis, vis2)
    assert np.array_equal(k2p, k2p2) and np.array_equal(pout["valid"], pout2["valid"])
    if n_points:
        assert m > 100 and ops > 1000


def test_search_by_projection_fine_resident(so, ctx, oracle, feats):
    """Device-resident frame + local map (snk_frame_dev / snk_localmap): the search must return what the host-struct entry
    point and the oracle return, a second search on the same frame must see the first one's assignments (the reference's
    Frame::mvpMapPoints persists between searches), and a changed pose / a partly replaced local map must take effect."""
    rng, frame, mp, cam, pyr = _c3(so, feats, 21, 20000)
    pts = M.fine_points(mp, rng)
    want = oracle.search_projection_fine(frame.view, pts, cam, pyr, 4.0, 0.8)
    fd, lm = ctx.frame_dev(frame.view), ctx.localmap(25000)
    lm.upload(pts)
    m, k2p, valid, vis = ctx.SearchByProjection2Resident(fd, lm, cam, pyr, 4.0, 0.8)
    assert m == want[0] and m > 500
    assert np.array_equal(k2p, want[1]) and np.array_equal(vis, want[4]) and np.array_equal(valid, want[5]["valid"])
    # second search, other points, tighter radius: the first search's assignments are taken (device side) and skipped
    pts2 = M.fine_points(S_map(so, frame, feats, rng, 6000), rng)
    pts2["id"] += 100000
    frame2_k2p = want[1].copy()
    frame.k2p[:] = frame2_k2p  # the host view for the oracle: assignments after search 1
    want2 = oracle.search_projection_fine(frame.view, pts2, cam, pyr, 1.0, 0.7)
    lm.upload(pts2)
    m2, k2p2, valid2, vis2 = ctx.SearchByProjection2Resident(fd, lm, cam, pyr, 1.0, 0.7)
    assert m2 == want2[0] and np.array_equal(k2p2, want2[1]) and np.array_equal(valid2, want2[5]["valid"])
    # a new pose and a partial update of the local map
    q = so.synthetic.quat_from_axis_angle(np.array([0.0, 1.0, 0.2]), 0.003)
    t = np.array([0.02, -0.01, 0.03])
    fd.set_pose(q, t)
    fd.set_assignment(np.full(frame.view.n, -1, np.int32))
    frame.k2p[:] = -1
    for i in range(4):
        frame.view.pose_q[i] = q[i]
    for i in range(3):
        frame.view.pose_t[i] = t[i]
    pts3 = want2[5].copy()  # the local map as search 2 left it (valid flags cleared where the reference clears them)
    pts3[1000:2000] = pts[1000:2000]
    lm.upload(pts[1000:2000], first=1000, new_size=len(pts2))
    want3 = oracle.search_projection_fine(frame.view, pts3, cam, pyr, 4.0, 0.8)
    m3, k2p3, valid3, _ = ctx.SearchByProjection2Resident(fd, lm, cam, pyr, 4.0, 0.8, keep_assignment=False)
    assert m3 == want3[0] and np.array_equal(k2p3, want3[1])


def S_map(so, frame, feats, rng, n):
    from snakeorb import synthetic as S

    return S.map_points_from_keypoints(frame.kp, frame.desc, (718.856, 718.856, 607.19, 185.22), rng, n, feats["scales"])


@pytest.mark.parametrize("motion", [0, 1, 2])
def test_search_by_projection_coarse(so, ctx, oracle, feats, motion):
    rng, frame, mp, cam, pyr = _c3(so, feats, 12 + motion, 3000)
    pts = M.coarse_points(mp)
    m, k2p = ctx.SearchByProjectionFrameFrame2(frame.view, pts, cam, pyr, 15.0, 75, motion)
    m2, k2p2, ops = oracle.search_projection_coarse(frame.view, pts, cam, pyr, 15.0, 75, motion)
    assert m == m2 and np.array_equal(k2p, k2p2) and m > 50


def test_search_by_projection_frame_to_keyframe(so, ctx, oracle, feats):
    rng, frame, mp, cam, pyr = _c3(so, feats, 21, 1500)
    pts = M.kf_points(mp, rng)
    # some ids already found in the frame (sAlreadyFound)
    frame.k2p[np.nonzero(frame.k2p < 0)[0][:40]] = pts["id"][pts["id"] >= 0][:40]
    m, k2p = ctx.SearchByProjectionFrameToKeyframe(frame.view, pts, cam, 10.0, 60)
    m2, k2p2, ops = oracle.search_projection_frame_to_kf(frame.view, pts, cam, 10.0, 60)
    assert m == m2 and np.array_equal(k2p, k2p2) and m > 50


@pytest.mark.parametrize("seed", [1, 2])
def test_match_bow(so, ctx, oracle, feats, seed):
    from snakeorb import synthetic as S

    rng = np.random.default_rng(seed)
    d1 = feats["l"][1][:1000]
    d2 = np.concatenate([S.flip_bits(d1[rng.permutation(len(d1))[:600]], rng, 12), feats["o"][1][:400]])
    fv1, fv2 = M.featvec(d1, rng), M.featvec(d2, rng)
    v1 = (rng.random(len(d1)) < 0.7).astype(np.uint8)
    v2 = (rng.random(len(d2)) < 0.7).astype(np.uint8)
    m, m12 = ctx.MatchBoW(d1, d2, fv1, fv2, v1, v2, 50, 0.75)
    m2, m122, ops = oracle.match_bow(d1, d2, fv1, fv2, v1, v2, 50, 0.75)
    assert m == m2 and np.array_equal(m12, m122) and m > 20 and ops > 1000


def test_match_bow_batch_equals_call_by_call(so, ctx, oracle, feats):
    """snk_match_bow_batch on 20 loop candidates of ~1000 features: every row equals the single-candidate call and the
    oracle; errors of one candidate are reported with its number."""
    from snakeorb import synthetic as S

    rng = np.random.default_rng(11)
    d1 = feats["l"][1][:1000]
    fv1 = M.featvec(d1, rng)
    v1 = (rng.random(len(d1)) < 0.7).astype(np.uint8)
    cands = []
    for k in range(20):
        n_true = int(rng.integers(0, 700))
        d2 = np.concatenate([S.flip_bits(d1[rng.permutation(len(d1))[:n_true]], rng, 12), feats["o"][1][:1000 - n_true]])
        cands.append((np.ascontiguousarray(d2), M.featvec(d2, rng), (rng.random(len(d2)) < 0.7).astype(np.uint8)))
    nm, m12 = ctx.MatchBoWBatch(d1, fv1, v1, cands, 50, 0.75)
    assert m12.shape == (20, 1000) and nm.max() > 100
    for k, (d2, fv2, v2) in enumerate(cands):
        m, row = ctx.MatchBoW(d1, d2, fv1, fv2, v1, v2, 50, 0.75)
        mo, rowo, _ = oracle.match_bow(d1, d2, fv1, fv2, v1, v2, 50, 0.75)
        assert nm[k] == m == mo and np.array_equal(m12[k], row) and np.array_equal(row, rowo), k
    nm0, m0 = ctx.MatchBoWBatch(d1, fv1, v1, [], 50, 0.75)
    assert len(nm0) == 0 and m0.shape[0] == 0
    bad = cands[3][1]._keep[2].copy()
    cands[3][1]._keep[2][0] = 5000  # feature index beyond the candidate's size
    with pytest.raises(so.SnakeOrbError, match="candidate 3"):
        ctx.MatchBoWBatch(d1, fv1, v1, cands, 50, 0.75)
    cands[3][1]._keep[2][:] = bad


def test_stereo_from_rgbd_c4(so, ctx, oracle):
    """Config C4: TUM-shaped RGB-D frame, depth planes 0.5-8 m with 5 % holes, depth camera with distortion."""
    from snakeorb import synthetic as S

    rng = np.random.default_rng(4)
    img = S.image(4000, 640, 480)
    k, d = oracle.Extractor(1000, 1.2, 8, 20, 7, 2).detect(img)
    depth = np.empty((480, 640), np.float32)
    for r0 in range(0, 480, 60):
        for c0 in range(0, 640, 80):
            depth[r0:r0 + 60, c0:c0 + 80] = rng.uniform(0.5, 8.0)
    depth[rng.random(depth.shape) < 0.05] = 0.0
    wide = np.zeros((480, 650), np.float32)  # pitch larger than the row
    wide[:, :640] = depth
    p = so.RgbdParams()
    p.fx, p.fy, p.cx, p.cy = 535.4, 539.2, 320.1, 247.6
    p.dfx, p.dfy, p.dcx, p.dcy, p.ds = 535.4, 539.2, 320.1, 247.6, 0.0
    for i, v in enumerate((0.2312, -0.7849, 0.9172, 0.0, 0.0, 0.0, -0.0033, -0.0001)):
        p.dist[i] = v
    p.bf = 40.0
    kp = so.kp32_to_kp64(k)
    for view in (depth, wide[:, :640]):
        n, rp, dp = ctx.ComputeStereoFromRGBD(p, kp, view)
        n2, rp2, dp2 = oracle.stereo_from_rgbd(p, oracle.kp32_to_kp64(k), view)
        assert n == n2 and n > 800 and n < len(k)
        assert rp.tobytes() == rp2.tobytes() and dp.tobytes() == dp2.tobytes()


def test_feature_grid(so, ctx, oracle, feats):
    """Preprocess::computeFeatureGrid on the device == FeatureGrid2::create (oracle restatement, itself checked
    against the reference's compiled class in tests/test_matchers_cpu.py)."""
    rng = np.random.default_rng(6)
    kl, dl = feats["l"]
    kp = so.kp32_to_kp64(kl)
    kp["x"] += rng.normal(0, 0.3, len(kp))  # undistorted coordinates are not integers
    kp["x"][:7] = [-5.0, 1300.0, 3.0, -0.05, 1240.09, 1240.11, 620.0]  # outside / on the border of the bounds
    kp["y"][:7] = [10.0, 10.0, -2.0, 0.0, 375.09, 100.0, 375.2]
    frame = M.make_frame(kl[:10], dl[:10], 1241, 376, rng, (0, 0, 0, 1), (0, 0, 0))
    for sub in (kp, kp[:1], kp[:0], kp[:257]):
        perm, cells = ctx.computeFeatureGrid(frame.view.grid, sub)
        perm2, cells2 = oracle.feature_grid(frame.view.grid, sub)
        assert np.array_equal(perm, perm2) and np.array_equal(cells, cells2)
        assert sorted(perm) == list(range(len(sub)))


def test_hamming_peak_probe(so, ctx):
    v = ctx.hamming_peak()
    assert 1e10 < v < 5e12


def test_undistort_keypoints(so, ctx, oracle):
    """Preprocess::undistortKeypoints on the device: bit-identical doubles to the oracle and to the committed
    vectors of the reference's own headers, for the EuRoC / TUM / rational-model / identity rectifications."""
    import os

    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden_next.npz"))
    rng = np.random.default_rng(9)
    for case in ("euroc", "tum", "rational", "identity"):
        rect = so.Rectify.from_vector(g[f"und_{case}_rect"])
        xy = g[f"und_{case}_xy"]
        kp = np.zeros(len(xy), so.KP64)
        kp["x"], kp["y"] = xy[:, 0], xy[:, 1]
        kp["size"], kp["angle"], kp["response"], kp["octave"] = 31, 12.5, 40, 2
        un, nrm = ctx.undistortKeypoints(rect, kp)
        assert nrm.tobytes() == g[f"und_{case}_normalized"].tobytes(), case
        assert np.stack([un["x"], un["y"]], 1).tobytes() == g[f"und_{case}_undistorted"].tobytes(), case
        assert np.array_equal(un["octave"], kp["octave"]) and np.array_equal(un["size"], kp["size"])
        # a larger random set, points outside the image included (the GN step rejection path)
        n = 20000
        kp = np.zeros(n, so.KP64)
        kp["x"], kp["y"] = rng.uniform(-50, 1300, n), rng.uniform(-50, 520, n)
        un, nrm = ctx.undistortKeypoints(rect, kp)
        un2, nrm2 = oracle.undistort_keypoints(rect, kp)
        assert un.tobytes() == un2.tobytes() and nrm.tobytes() == nrm2.tobytes(), case
    un, nrm = ctx.undistortKeypoints(rect, np.zeros(0, so.KP64))
    assert len(un) == 0


def test_stereo_matching_with_rectification(so, ctx, oracle, feats):
    """The whole Preprocess::StereoMatching: Rectification::Forward of the raw left / right keypoints (distorted
    cameras, a small rectifying rotation each) and the row search, in one call — equal to the oracle's two steps."""
    kl, dl = feats["l"]
    kr, dr = feats["r"]
    K = [718.856, 718.856, 607.19, 185.22, 0.0]
    dist = [-0.12, 0.03, 0, 0, 0, 0, 0.0004, -0.0002]

    def rect(q):
        q = np.array(q, np.float64)
        q /= np.sqrt((q * q).sum())
        return so.Rectify.from_vector(np.array(K + dist + list(q) + K, np.float64))

    rl, rr = rect([0.001, -0.002, 0.0005, 1.0]), rect([-0.0015, 0.001, -0.0005, 1.0])
    p = so.stereo_params(386.1, True, feats["scales"])
    a, b = so.kp32_to_kp64(kl), so.kp32_to_kp64(kr)
    n, rp, dp = ctx.StereoMatchingRaw(rl, rr, p, a, dl, b, dr)
    ua, _ = oracle.undistort_keypoints(rl, a)
    ub, _ = oracle.undistort_keypoints(rr, b)
    n2, rp2, dp2, _ = oracle.stereo_match(oracle.stereo_params(386.1, True, feats["scales"]), ua, dl, ub, dr)
    assert n == n2 and n > 100 and rp.tobytes() == rp2.tobytes() and dp.tobytes() == dp2.tobytes()
    # identity rectification: the plain entry point
    ident = so.Rectify.from_vector(np.array(K + [0.0] * 8 + [0, 0, 0, 1.0] + K, np.float64))
    n3, rp3, dp3 = ctx.StereoMatchingRaw(ident, ident, p, a, dl, b, dr)
    ui, _ = oracle.undistort_keypoints(ident, a)
    uj, _ = oracle.undistort_keypoints(ident, b)
    n4, rp4, dp4, _ = oracle.stereo_match(oracle.stereo_params(386.1, True, feats["scales"]), ui, dl, uj, dr)
    assert n3 == n4 and rp3.tobytes() == rp4.tobytes() and dp3.tobytes() == dp4.tobytes()
