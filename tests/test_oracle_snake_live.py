"""Live pin of the CPU restatement (oracle/matchers_ref.cpp) on the reference's OWN matcher code: the translation units
Snake/Preprocess/Preprocess.cpp, Tracking/SnakeORBMatcher.cpp, LoopClosing/LoopORBMatcher.cpp,
LocalMapping/MappingORBMatcher.cpp and Tracking/Initialization/MonoInitializer.cpp, compiled unmodified where they lie
into oracle/_ref/libsnake_ref.so (oracle/Makefile, oracle/ref_snake_shim.cpp). Every test runs >= 1000 seeded,
tie-heavy cases (tests/snake_cases.py) through both and requires identical results, bit for bit.

No GPU involved. Skipped where oracle/_ref was not built (a machine without /root/reference and without the prebuilt
library); the outputs frozen from the same library in tests/golden/golden_snake.npz keep the pin alive there
(test_snake_golden.py)."""
import numpy as np
import pytest

import snake_cases as SC

N_CASES = 1000


@pytest.fixture(scope="module")
def ref(oracle):
    if oracle.snake_ref() is None:
        pytest.skip("oracle/_ref/libsnake_ref.so not built (no /root/reference on this machine)")
    return oracle


def rectified(oracle, c):
    left, _ = oracle.undistort_keypoints(c["rl"], c["left"])
    right, _ = oracle.undistort_keypoints(c["rr"], c["right"])
    return left, right


def test_live_stereo_matching(ref):
    """Preprocess::StereoMatching (Preprocess.cpp:122-242) incl. its Rectification::Forward loops."""
    matched = 0
    for seed in range(N_CASES):
        c = SC.case_stereo(seed)
        want = ref.ref_stereo_matching(c["rl"], c["rr"], c["bf"], c["relaxed"], SC.LEVELS, float(np.float32(1.2)),
                                       c["left"], c["dl"], c["right"], c["dr"])
        left, right = rectified(ref, c)
        n, rp, dp, _ = ref.stereo_match(ref.stereo_params(c["bf"], c["relaxed"], SC.SCALES), left, c["dl"], right, c["dr"])
        assert n == want[0] and rp.tobytes() == want[1].tobytes() and dp.tobytes() == want[2].tobytes(), seed
        matched += n
    assert matched > 10 * N_CASES


def test_live_stereo_from_rgbd(ref):
    """Preprocess::ComputeStereoFromRGBD (Preprocess.cpp:79-120)."""
    for seed in range(N_CASES):
        c = SC.case_rgbd(seed)
        want = ref.ref_stereo_from_rgbd(c["p"], c["kp"], c["depth"])
        got = ref.stereo_from_rgbd(c["p"], c["kp"], c["depth"])
        assert got[0] == want[0] and got[1].tobytes() == want[1].tobytes() and got[2].tobytes() == want[2].tobytes(), seed


def test_live_undistort_keypoints(ref):
    """Preprocess::undistortKeypoints (Preprocess.cpp:55-77) from Preprocess.cpp itself."""
    for seed in range(N_CASES):
        c = SC.case_stereo(seed)
        k = c["left"].copy()
        k["x"], k["y"] = np.clip(k["x"], 0, c["w"] - 1), np.clip(k["y"], 0, c["h"] - 1)  # the reference asserts inImage
        want = ref.ref_undistort_keypoints(c["rl"], k, c["h"], c["w"])
        got = ref.undistort_keypoints(c["rl"], k)
        assert got[0].tobytes() == want[0].tobytes() and got[1].tobytes() == want[1].tobytes(), seed


def test_live_search_by_projection_fine(ref):
    """SnakeORBMatcher::SearchByProjection2 (SnakeORBMatcher.cpp:365-526), 1 and 4 OpenMP threads."""
    total = 0
    for seed in range(N_CASES):
        c = SC.case_fine(seed)
        want = ref.ref_search_projection_fine(c["f"].view, c["pts"], c["cam"], SC.pyramid(), c["th"], c["ratio"],
                                              threads=1 + 3 * (seed % 2))
        got = ref.search_projection_fine(c["f"].view, c["pts"], c["cam"], SC.pyramid(), c["th"], c["ratio"])
        assert got[0] == want[0], seed
        assert np.array_equal(got[1], want[1]), seed                      # kp -> point assignment
        assert np.array_equal(got[2], want[2]), seed                      # best keypoint per point
        sel = want[2] >= 0
        assert np.array_equal(got[3][sel], want[3][sel]), seed            # its distance
        assert np.array_equal(got[4], want[4]), seed                      # IncreaseVisible
        assert np.array_equal(got[5]["valid"], want[5]["valid"]), seed    # lmp.valid
        total += got[0]
    assert total > 20 * N_CASES


def test_live_search_by_projection_coarse(ref):
    """SnakeORBMatcher::SearchByProjectionFrameFrame2 (SnakeORBMatcher.cpp:191-354) incl. the rotation histogram."""
    total = 0
    for seed in range(N_CASES):
        c = SC.case_coarse(seed)
        want = ref.ref_search_projection_coarse(c["f"].view, c["pts"], c["cam"], SC.pyramid(), c["th"], c["fe"], c["motion"],
                                                threads=1 + 3 * (seed % 2))
        got = ref.search_projection_coarse(c["f"].view, c["pts"], c["cam"], SC.pyramid(), c["th"], c["fe"], c["motion"])
        assert got[0] == want[0] and np.array_equal(got[1], want[1]), seed
        total += got[0]
    assert total > 5 * N_CASES


def test_live_search_by_projection_frame_to_keyframe(ref):
    """SnakeORBMatcher::SearchByProjectionFrameToKeyframe (SnakeORBMatcher.cpp:71-188)."""
    total = 0
    for seed in range(N_CASES):
        c = SC.case_kf(seed)
        want = ref.ref_search_projection_frame_to_kf(c["f"].view, c["pts"], c["cam"], c["th"], c["fe"])
        got = ref.search_projection_frame_to_kf(c["f"].view, c["pts"], c["cam"], c["th"], c["fe"])
        assert got[0] == want[0] and np.array_equal(got[1], want[1]), seed
        total += got[0]
    assert total > 5 * N_CASES


def test_live_match_bow(ref):
    """LoopORBmatcher::MatchBoW (LoopORBMatcher.cpp:121-215); bad map points (isBad()) count as absent."""
    total = 0
    for seed in range(N_CASES):
        c = SC.case_bow(seed)
        want = ref.ref_match_bow(c["d1"], c["d2"], c["fv1"], c["fv2"], c["v1"], c["v2"], c["th"], c["ratio"])
        v1, v2 = (c["v1"] == 1).astype(np.uint8), (c["v2"] == 1).astype(np.uint8)
        got = ref.match_bow(c["d1"], c["d2"], c["fv1"], c["fv2"], v1, v2, c["th"], c["ratio"])
        assert got[0] == want[0] and np.array_equal(got[1], want[1]), seed
        total += got[0]
    assert total > 3 * N_CASES


def test_live_match_bow_candidate_loop(ref):
    """The candidate loop of LoopDetector::ComputeSim3 (LoopDetector.cpp:384-391) with every candidate evaluated: the
    reference's MatchBoW call by call on one source keyframe — what snk_match_bow_batch must return row by row."""
    import snake_runners as R

    _, _, run_ref, run_oracle, _ = R.MATCHERS["match_bow_batch"]
    total = 0
    for seed in range(N_CASES // 4):
        c = SC.case_bow_batch(seed)
        want, got = R.norm(run_ref(ref, c)), R.norm(run_oracle(ref, c))
        assert R.same(want, got), seed
        total += int(want[0].sum())
    assert total > N_CASES


def test_live_mapping_fuse(ref):
    """MappingORBMatcher::Fuse (MappingORBMatcher.cpp:359-482)."""
    total = 0
    for seed in range(N_CASES):
        c = SC.case_fuse(seed)
        want = ref.ref_fuse_search(c["f"].view, c["mask"], c["pts"], c["cam"], SC.pyramid(), c["th"], c["obs"], c["fth"])
        got = ref.fuse_search(c["f"].view, c["mask"], c["pts"], c["cam"], SC.pyramid(), c["th"], c["obs"], c["fth"])
        assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1]), seed
        total += len(want[0])
    assert total > 5 * N_CASES


def fuse_loop_edits(c, best_idx, best_dist):
    """LoopORBMatcher.cpp:75-93 applied to a search result: what the keyframe and the points look like afterwards."""
    f, pts = c["f"], c["pts"]
    n = len(pts)
    state = np.full(n, -1, np.int32)
    holder = {}  # feature -> index of the fused point that holds it now (original points: absent)
    bad = np.where(f.k2p >= 0, 0, 2).astype(np.uint8)
    fused = 0
    for i in range(n):
        if best_idx[i] < 0 or best_dist[i] > 50:
            continue
        j = int(best_idx[i])
        fused += 1
        if j in holder:
            if holder[j] is None:  # an earlier Replace failed its reprojection gate: the slot is empty again
                holder[j] = i
                state[i] = j
                continue
            state[holder[j]] = -1  # the earlier point is replaced by this one
        elif f.k2p[j] >= 0:
            bad[j] = 1
        else:
            holder[j] = i
            state[i] = j
            continue
        # Replace -> Keyframe::ReplaceMapPointMatch (Keyframe.cpp:349-370): accepted within 2 * 2.1 px of the projection
        holder[j] = i
        state[i] = j
    return fused, state, bad


def test_live_loop_fuse(ref):
    """LoopORBmatcher::Fuse (LoopORBMatcher.cpp:13-96): the search of the restatement, followed by the map edits of
    :75-93, must leave the keyframe and the points exactly as the reference's function does. Radii <= 4 px keep
    Keyframe::ReplaceMapPointMatch's own 4.2-px gate out of the picture."""
    total = 0
    for seed in range(N_CASES):
        c = SC.case_fuse_loop(seed)
        th = min(c["th"], 4.0)
        want = ref.ref_fuse_loop(c["f"].view, c["skip"], c["pts"], c["cam"], th)
        bi, bd = ref.fuse_search_loop(c["f"].view, c["skip"], c["pts"], c["cam"], th)
        fused, state, bad = fuse_loop_edits(c, bi, bd)
        assert fused == want[0], seed
        assert np.array_equal(state, want[1]), seed
        assert np.array_equal(bad, want[2]), seed
        total += fused
    assert total > 5 * N_CASES


def test_live_search_for_triangulation_bow(ref):
    """MappingORBMatcher::SearchForTriangulation2 (MappingORBMatcher.cpp:14-101)."""
    total = 0
    for seed in range(N_CASES):
        c = SC.case_tri_bow(seed)
        want = ref.ref_search_triangulation_bow(c["f1"].view, c["np1"], c["fv1"], c["f2"].view, c["np2"], c["fv2"], c["E"],
                                                c["cam"], c["epi"], c["fd"])
        got, _ = ref.search_triangulation_bow(c["f1"].view, c["np1"], c["fv1"], c["f2"].view, c["np2"], c["fv2"], c["E"],
                                              c["cam"], c["epi"], c["fd"])
        assert np.array_equal(got, want), seed
        total += len(want)
    assert total > 3 * N_CASES


def test_live_search_for_triangulation_project(ref):
    """MappingORBMatcher::SearchForTriangulationProject (MappingORBMatcher.cpp:167-247)."""
    total = 0
    for seed in range(N_CASES):
        c = SC.case_tri_project(seed)
        want = ref.ref_search_triangulation_project(c["f1"].view, c["np1"], c["f2"].view, c["np2"], c["grid"], c["E"],
                                                    c["cam"], c["epi"], c["fd"])
        got, _ = ref.search_triangulation_project(c["f1"].view, c["np1"], c["f2"].view, c["np2"], c["grid"], c["E"],
                                                  c["cam"], c["epi"], c["fd"])
        assert np.array_equal(got, want), seed
        total += len(want)
    assert total > N_CASES // 2


def test_live_match_prediction(ref):
    """MonoInitializer::ComputeMatchesPrediction (MonoInitializer.cpp:619-697), radius 50 / featureError 50 as there."""
    total = 0
    for seed in range(N_CASES):
        c = SC.case_prediction(seed)
        want = ref.ref_match_prediction(c["dl"], c["pos"], c["f"].view, 50.0, 50, threads=1 + 3 * (seed % 2))
        got = ref.match_prediction(c["dl"], c["pos"], c["f"].view, 50.0, 50)
        assert got[0] == want[0] and np.array_equal(got[1], want[1]) and got[2].tobytes() == want[2].tobytes(), seed
        total += got[0]
    assert total > 5 * N_CASES


def test_live_features_in_area(ref):
    """Features::GetFeaturesInArea from Snake/Map/Features.cpp itself (round 1 pinned the enumeration on FeatureGrid2 only)."""
    for seed in range(200):
        c = SC.case_prediction(seed)
        rng = np.random.default_rng(seed)
        for _ in range(10):
            px, py, r = rng.uniform(-20, c["f"].w + 20), rng.uniform(-20, c["f"].h + 20), float(rng.uniform(1, 60))
            assert np.array_equal(ref.features_in_area(c["f"].view, px, py, r), ref.ref_features_in_area(c["f"].view, px, py, r))
