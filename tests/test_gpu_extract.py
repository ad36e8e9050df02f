"""GPU parity tests of the extraction path: libsnakeorb.so through its C ABI against the CPU oracle on the same seeded
inputs, stage by stage (pyramid, blur, FAST candidates, selected keypoints, angles, descriptors) and end to end
against the committed golden vectors. Integer/byte/index outputs must be bit-exact; angle and response are compared
bit-exactly as well (the north star allows 1e-4 relative; the device libm ports make them exact)."""
import hashlib
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
G = os.path.join(HERE, "golden")


def sha(a):
    return hashlib.sha1(np.ascontiguousarray(a).tobytes()).hexdigest()


def _cases():
    import importlib.util

    spec = importlib.util.spec_from_file_location("golden_cases", os.path.join(G, "cases.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


@pytest.fixture(scope="module")
def so():
    import snakeorb

    return snakeorb


def assert_same_features(k, d, k2, d2):
    assert len(k) == len(k2), (len(k), len(k2))
    for f in ("x", "y", "size", "octave"):
        assert np.array_equal(k[f], k2[f]), f
    # tolerance stated by the north star: 1e-4 relative; we additionally require bit equality below
    assert np.allclose(k["angle"], k2["angle"], rtol=1e-4, atol=0)
    assert np.allclose(k["response"], k2["response"], rtol=1e-4, atol=0)
    assert k.tobytes() == k2.tobytes()
    assert d.tobytes() == d2.tobytes()


@pytest.mark.parametrize("name", ["euroc_c1", "euroc_c2", "kitti_c3", "tum_c4", "noise", "checker", "constant", "small4"])
def test_extract_matches_golden_and_oracle_stagewise(so, oracle, name):
    cases = _cases()
    gext = np.load(os.path.join(G, "golden_extract.npz"))
    kind, w, h, nf, nl = cases.CASES[name]
    img = cases.case_image(kind, w, h)
    assert sha(img) == str(gext[f"{name}_sha1"])
    ex = so.ORBExtractor(nf, 1.2, nl, 20, 7)
    k, d = ex.Detect(img)
    ref = oracle.Extractor(nf, 1.2, nl, 20, 7, 1)
    k2, d2 = ref.detect(img)
    # stage by stage first, so a failure names the stage
    for l in range(nl):
        assert np.array_equal(ex.GetImage(l), ref.level(l)), f"pyramid level {l}"
    for l in range(nl):
        raw = ref.level_kps(l, 0)
        want = np.stack([raw["x"], raw["y"], raw["response"]], 1).astype(np.int32) if len(raw) else np.zeros((0, 3), np.int32)
        got = ex.candidates(l)
        assert got.shape == want.shape and np.array_equal(got, want), f"FAST candidates level {l}"
    for l in range(nl):
        assert np.array_equal(ex.GetImage(l, blurred=True), oracle.blur(ref.level(l))), f"blur level {l}"
    assert_same_features(k, d, k2, d2)
    assert k.tobytes() == gext[f"{name}_kps"].tobytes()
    assert d.tobytes() == gext[f"{name}_desc"].tobytes()
    s, f = ex.getPyramid()
    s2, f2 = oracle.scale_pyramid(nl, 1.2, nf)
    assert np.array_equal(s, s2) and np.array_equal(f, f2)


def test_batch_equals_single_and_ragged_pitch(so, oracle):
    from snakeorb import synthetic as S

    imgs = np.stack([S.image(500 + i, 752, 480) for i in range(5)] + [S.adversarial("constant", 752, 480)])
    ex = so.ORBExtractor(1200, 1.2, 8, 20, 7, max_batch=6)
    res = ex.DetectBatch(imgs)
    ref = oracle.Extractor(1200, 1.2, 8, 20, 7, 1)
    for i in range(len(imgs)):
        k2, d2 = ref.detect(imgs[i])
        assert_same_features(res[i][0], res[i][1], k2, d2)
    assert len(res[-1][0]) == 0
    # a view with a pitch larger than the row (ImageView pitchBytes), single-image path on the same handle
    wide = np.zeros((480, 800), np.uint8)
    wide[:, :752] = imgs[2]
    k, d = ex.Detect(wide[:, :752])
    assert_same_features(k, d, *ref.detect(imgs[2]))


def test_reuse_handle_across_sizes_and_empty(so, oracle):
    from snakeorb import synthetic as S

    ex = so.ORBExtractor(500, 1.2, 4, 20, 7)
    for (w, h, seed) in ((320, 240, 1), (640, 480, 2), (320, 240, 3)):
        img = S.image(seed, w, h)
        k, d = ex.Detect(img)
        assert_same_features(k, d, *oracle.Extractor(500, 1.2, 4, 20, 7, 1).detect(img))
    # empty image: silently no output (ORBExtractor.cpp:182)
    k, d = ex.Detect(np.zeros((0, 0), np.uint8))
    assert len(k) == 0 and len(d) == 0


def test_too_small_image_is_rejected(so):
    ex = so.ORBExtractor(300, 1.2, 8, 20, 7)
    with pytest.raises(so.SnakeOrbError):
        ex.Detect(np.zeros((152, 200), np.uint8))  # level 7 would have zero FAST cell rows (UB in the reference)


@pytest.mark.parametrize("th,thmin", [(20, 7), (40, 20), (12, 3), (7, 7)])
def test_thresholds_and_dense_noise(so, oracle, th, thmin):
    from snakeorb import synthetic as S

    img = S.adversarial("noise", 400, 300, 100 + th)
    ex = so.ORBExtractor(800, 1.2, 6, th, thmin)
    k, d = ex.Detect(img)
    assert_same_features(k, d, *oracle.Extractor(800, 1.2, 6, th, thmin, 1).detect(img))


def test_reference_own_orb_tests_restated(so, oracle):
    """saiga tests/test_cuda_orb.cu restated against the new kernels with the CPU path as truth: circle image, angle
    within 2 degrees of the analytic one on the rim (ORB.Angle :68-114); descriptor distance 0 to the CPU
    descriptor (ORB.DescriptorRandom :155-222, tightened from <=10 bits to 0)."""
    rng = np.random.default_rng(3)
    img = rng.integers(0, 256, (200, 200), dtype=np.uint8)
    yy, xx = np.mgrid[0:200, 0:200]
    img[(yy - 100) ** 2 + (xx - 100) ** 2 < 50 ** 2] = 255
    ex = so.ORBExtractor(1000, 1.2, 2, 20, 7)
    k, d = ex.Detect(img)
    k2, d2 = oracle.Extractor(1000, 1.2, 2, 20, 7, 1).detect(img)
    assert_same_features(k, d, k2, d2)
    lvl0 = k[k["octave"] == 0]
    rim = lvl0[np.abs(np.hypot(lvl0["x"] - 100, lvl0["y"] - 100) - 50) < 1.5]
    assert len(rim) > 0
    ana = np.degrees(np.arctan2(100 - rim["y"], 100 - rim["x"])) % 360  # bright disc: centroid points to the centre
    diff = np.abs((rim["angle"] - ana + 180) % 360 - 180)
    assert np.all(diff < 8.0)  # the rim keypoints sit on a noisy boundary; the reference's own bound is on synthetic rim points


def test_no_cpu_fallback_symbols(so):
    import subprocess

    out = subprocess.run(["nm", "-D", "--defined-only", so.LIB_PATH], capture_output=True, text=True).stdout
    assert "orc_" not in out and "cvlite" not in out


def test_async_fetch_equals_sync_fetch(so, oracle):
    """snk_orb_enqueue + snk_orb_fetch_async (no host wait) deliver the same records as the blocking fetch."""
    import ctypes as C

    from snakeorb import synthetic as S

    imgs = np.stack([S.image(40 + i, 640, 480) for i in range(3)])
    ex = so.ORBExtractor(1000, 1.2, 8, 20, 7, max_batch=3)
    want = ex.DetectBatch(imgs)
    kps = np.zeros((3, ex.cap), so.KP32)
    desc = np.zeros((3, ex.cap, 4), np.uint64)
    cnt = np.zeros(3, np.int32)
    ex.enqueue_ptr(imgs.ctypes.data, 3, 480, 640, 640, 640 * 480, False)
    so._check(so.lib().snk_orb_fetch_async(ex.h, C.c_void_p(kps.ctypes.data), C.c_void_p(desc.ctypes.data), ex.cap,
                                           C.c_void_p(cnt.ctypes.data)))
    assert C.CDLL("libcudart.so").cudaDeviceSynchronize() == 0
    for i in range(3):
        assert cnt[i] == len(want[i][0])
        assert kps[i, :cnt[i]].tobytes() == want[i][0].tobytes() and desc[i, :cnt[i]].tobytes() == want[i][1].tobytes()


def test_rgba_input(so, oracle):
    """Input::computeGrayscaleImage on the device: standalone conversion equals cv2's committed outputs and the
    oracle on odd sizes / padded pitches; the fused RGBA entry of the extractor returns exactly what the grey entry
    returns for the converted image."""
    import os

    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden_next.npz"))
    ctx = so.Context()
    assert np.array_equal(ctx.computeGrayscaleImage(np.ascontiguousarray(g["gray_rgba"])), g["gray_out"])
    assert np.array_equal(ctx.computeGrayscaleImage(np.ascontiguousarray(g["gray_lattice"])), g["gray_lattice_out"])
    rng = np.random.default_rng(3)
    for rows, cols, padw in ((5, 1, 3), (33, 510, 515), (480, 752, 752), (100, 127, 127)):
        wide = rng.integers(0, 256, (rows, padw, 4), dtype=np.uint8)
        view = wide[:, :cols]
        assert np.array_equal(ctx.computeGrayscaleImage(view), oracle.rgba_to_gray(view)), (rows, cols)
    # fused path: colourised synthetic frames -> same keypoints/descriptors as the grey path
    from snakeorb import synthetic as S

    grey = np.stack([S.image(70 + i, 640, 480) for i in range(3)])
    rgba = np.zeros((3, 480, 648, 4), np.uint8)  # padded rows
    rgba[..., 0] = 255  # padding content must not matter
    noise = rng.integers(-40, 40, (3, 480, 640, 3))
    rgba[:, :, :640, :3] = np.clip(grey[..., None].astype(np.int64) + noise, 0, 255).astype(np.uint8)
    rgba[:, :, :640, 3] = rng.integers(0, 256, (3, 480, 640), dtype=np.uint8)
    view = rgba[:, :, :640]
    conv = np.stack([oracle.rgba_to_gray(view[i]) for i in range(3)])
    ex = so.ORBExtractor(1000, 1.2, 8, 20, 7, max_batch=4)
    a = ex.DetectBatchRGBA(view)
    b = ex.DetectBatch(conv)
    for (ka, da), (kb, db) in zip(a, b):
        assert len(ka) > 500 and ka.tobytes() == kb.tobytes() and da.tobytes() == db.tobytes()
    # and against the oracle's extraction of the converted image
    ko, do = oracle.Extractor(1000, 1.2, 8, 20, 7, 1).detect(conv[1])
    assert np.array_equal(a[1][1], do) and np.array_equal(a[1][0]["x"], ko["x"])


@pytest.mark.parametrize("w,h,nf,scale,nl,kind", [
    (1920, 1080, 3000, 1.2, 8, "image"),    # full HD: wide cell rows, many 4-cell groups plus ragged ends
    (640, 480, 1500, 1.3, 5, "image"),      # largest supported scale step
    (1241, 376, 2000, 1.1, 10, "image"),    # small scale step, 10 levels
    (1241, 376, 2000, 1.2, 8, "noise"),     # every pixel a FAST candidate: list capacities of the cell groups
    (333, 257, 400, 1.2, 4, "image"),       # odd sizes, few cells per row (groups shorter than 4)
    (97, 230, 100, 1.2, 2, "noise"),        # a single cell column
])
def test_varied_geometries(so, oracle, w, h, nf, scale, nl, kind):
    """Cell grids, group shapes and pyramid tables away from the BASELINE shapes, end to end against the oracle."""
    from snakeorb import synthetic as S

    img = S.image(4242 + w, w, h) if kind == "image" else S.adversarial("noise", w, h, 7 + h)
    ex = so.ORBExtractor(nf, scale, nl, 20, 7)
    k, d = ex.Detect(img)
    k2, d2 = oracle.Extractor(nf, scale, nl, 20, 7, 4).detect(img)
    assert len(k2) > 20
    assert_same_features(k, d, k2, d2)


def test_two_threads_two_handles(so, oracle):
    """INTEGRATION.md promises a re-entrant library with one handle per thread (Snake's Preprocess, Tracking and
    LoopClosing threads each own theirs): two host threads drive their own extractor and matcher context at the same
    time, on different image sizes, and get exactly the single-threaded results."""
    import threading

    from snakeorb import synthetic as S

    jobs = [(np.stack([S.image(300 + i, 752, 480) for i in range(4)]), 1200), (np.stack([S.image(400 + i, 640, 480) for i in range(4)]), 1000)]
    want = []
    for imgs, nf in jobs:
        ex = so.ORBExtractor(nf, 1.2, 8, 20, 7, max_batch=4)
        want.append(ex.DetectBatch(imgs))
    got = [None, None]
    errors = []

    def work(slot):
        try:
            imgs, nf = jobs[slot]
            ex = so.ORBExtractor(nf, 1.2, 8, 20, 7, max_batch=4)
            ctx = so.Context()
            for _ in range(6):
                res = ex.DetectBatch(imgs)
                knn = ctx.matchKnn2(res[0][1], res[1][1])
            got[slot] = (res, knn)
        except Exception as exc:  # noqa: BLE001
            errors.append(exc)

    threads = [threading.Thread(target=work, args=(i,)) for i in range(2)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
    for slot in range(2):
        res, knn = got[slot]
        for (k, d), (k2, d2) in zip(res, want[slot]):
            assert k.tobytes() == k2.tobytes() and d.tobytes() == d2.tobytes()
        assert knn.tobytes() == oracle.match_knn2(want[slot][0][1], want[slot][1][1]).tobytes()


def test_handles_of_different_sizes_interleaved(so, oracle):
    """Several extractor handles alive at once, with different image sizes and feature counts, used alternately: the
    kernels' shared-memory opt-in is per kernel, not per handle, so a later handle must not shrink it under an
    earlier one (regression: the second handle's set-up made the first handle's launches fail)."""
    from snakeorb import synthetic as S

    big = S.image(61, 1241, 376)
    small = S.image(62, 333, 257)
    a = so.ORBExtractor(2000, 1.2, 8, 20, 7)
    ka, da = a.Detect(big)
    b = so.ORBExtractor(300, 1.2, 4, 20, 7)
    kb, db = b.Detect(small)
    for _ in range(3):
        k, d = a.Detect(big)
        assert k.tobytes() == ka.tobytes() and d.tobytes() == da.tobytes()
        k, d = b.Detect(small)
        assert k.tobytes() == kb.tobytes() and d.tobytes() == db.tobytes()
    assert_same_features(ka, da, *oracle.Extractor(2000, 1.2, 8, 20, 7, 2).detect(big))
    assert_same_features(kb, db, *oracle.Extractor(300, 1.2, 4, 20, 7, 1).detect(small))


def test_no_device_memory_leak_over_handle_lifetimes(so):
    """Handles, contexts and vocabularies give back everything they allocated: free device memory after 25 create /
    use / destroy cycles (two image sizes, so the buffers are re-allocated inside a handle as well) is what it was."""
    import ctypes as C

    from snakeorb import synthetic as S

    rt = C.CDLL("libcudart.so")

    def free_bytes():
        f, t = C.c_size_t(), C.c_size_t()
        assert rt.cudaMemGetInfo(C.byref(f), C.byref(t)) == 0
        return f.value

    a, b = S.image(1, 752, 480), S.image(2, 640, 400)
    d = np.random.default_rng(0).integers(0, 2**63, (500, 4)).astype(np.uint64)

    def cycle():
        ex = so.ORBExtractor(1000, 1.2, 8, 20, 7, max_batch=2)
        ex.Detect(a)
        ex.Detect(b)
        ctx = so.Context()
        ctx.matchKnn2(d, d)
        ex.close() if hasattr(ex, "close") else None
        ctx.close()
        del ex, ctx

    for _ in range(3):
        cycle()  # first cycles: module load, lazily created library state
    import gc

    gc.collect()
    before = free_bytes()
    for _ in range(25):
        cycle()
    gc.collect()
    after = free_bytes()
    assert abs(before - after) <= 8 << 20, (before, after)


def test_random_geometries_fuzz(so, oracle):
    """Seeded sweep over image sizes, feature counts, level counts, scale factors and thresholds (cell grids with
    ragged last columns / rows, groups of 1-4 cells, tiny top levels): every case bit-exact against the oracle."""
    from snakeorb import synthetic as S

    rng = np.random.default_rng(2024)
    done = 0
    for case in range(40):
        w, h = int(rng.integers(96, 1000)), int(rng.integers(96, 700))
        nl = int(rng.integers(1, 9))
        scale = float(np.float32(rng.choice([1.1, 1.15, 1.2, 1.25, 1.3])))
        nf = int(rng.integers(50, 2500))
        th = int(rng.integers(8, 30))
        thmin = int(rng.integers(3, th + 1))
        # every level needs at least one FAST cell row and column (the reference divides by zero otherwise)
        if min(w, h) / scale ** (nl - 1) < 70:
            continue
        kind = "noise" if case % 5 == 4 else "image"
        img = S.image(7000 + case, w, h) if kind == "image" else S.adversarial("noise", w, h, case)
        ex = so.ORBExtractor(nf, scale, nl, th, thmin)
        k, d = ex.Detect(img)
        k2, d2 = oracle.Extractor(nf, scale, nl, th, thmin, 2).detect(img)
        assert_same_features(k, d, k2, d2)
        done += 1
    assert done >= 25


@pytest.mark.parametrize("mode", ["2", "1", "0"])
def test_blur_every_level_byte_equal_on_random_sizes(so, oracle, monkeypatch, mode):
    """cv::GaussianBlur of every pyramid level (SNK_BLUR_MMA, read when an extractor allocates its pyramid: unset or 2 = the
    tcgen05 kernel k_blur_umma, 1 = the mma.sync tensor-core kernel k_blur_mma, 0 = the dp4a kernel k_blur): byte-equal to the oracle on random image sizes — widths and heights of every residue mod 4 / 8 /
    16 / 64, so that partial last strips, partial last words and the folded taps at all four borders are hit — on noise
    (every byte matters) and in batches (the image index only offsets the pointers)."""
    from snakeorb import synthetic as S

    monkeypatch.setenv("SNK_BLUR_MMA", mode)
    rng = np.random.default_rng(99)
    for case in range(24):
        w, h = int(rng.integers(80, 900)), int(rng.integers(80, 600))
        nl = int(rng.integers(1, 6))
        if min(w, h) / 1.2 ** (nl - 1) < 70:
            nl = 1
        imgs = np.stack([S.adversarial("noise", w, h, 100 + case), S.image(8000 + case, w, h)])
        ex = so.ORBExtractor(300, 1.2, nl, 20, 7, max_batch=2)
        ex.DetectBatch(imgs)
        for b in range(2):
            ref = oracle.Extractor(300, 1.2, nl, 20, 7, 1)
            ref.detect(imgs[b])
            for l in range(nl):
                assert np.array_equal(ex.GetImage(l, img=b, blurred=True), ref.level(l, blurred=True)), (case, w, h, b, l)


def test_more_than_eight_levels_take_the_fallback_kernels(so, oracle):
    """The tcgen05 blur and the TMA descriptor kernel carry their tensor maps in the kernel parameters, eight levels'
    worth; a deeper pyramid runs k_blur_mma and k_describe (and still the TMA resize) — same bytes."""
    from snakeorb import synthetic as S

    img = S.image(4242, 1600, 1200)
    ex = so.ORBExtractor(2000, 1.2, 10, 20, 7)
    k, d = ex.Detect(img)
    ref = oracle.Extractor(2000, 1.2, 10, 20, 7, 2)
    k2, d2 = ref.detect(img)
    assert len(k) > 1500 and int(k["octave"].max()) == 9
    assert_same_features(k, d, k2, d2)
    for l in (0, 5, 9):
        assert np.array_equal(ex.GetImage(l, blurred=True), ref.level(l, blurred=True)), l


@pytest.mark.parametrize("env", [{"SNK_RESIZE_TMA": "0"}, {"SNK_DESCRIBE_TMA": "0"}, {"SNK_BLUR_MMA": "1"}, {"SNK_FAST_CHUNK": "1"},
                                 {"SNK_RESIZE_TMA": "0", "SNK_DESCRIBE_TMA": "0", "SNK_BLUR_MMA": "0", "SNK_FAST_PER_CELL": "1"}])
def test_kernel_variants_selected_by_environment_give_the_same_bytes(so, oracle, monkeypatch, env):
    """Every A/B kernel the environment can select (read when an extractor allocates its pyramid) against the oracle, on a
    batch of two different images — the plain-load resize, the global-memory descriptor kernel, the mma.sync and dp4a blurs,
    the chunk-wise FAST compass pass, the per-cell FAST kernel."""
    from snakeorb import synthetic as S

    for name, value in env.items():
        monkeypatch.setenv(name, value)
    imgs = np.stack([S.image(5151, 752, 480), S.adversarial("noise", 752, 480, 7)])
    ex = so.ORBExtractor(1000, 1.2, 8, 20, 7, max_batch=2)
    res = ex.DetectBatch(imgs)
    for b in range(2):
        k2, d2 = oracle.Extractor(1000, 1.2, 8, 20, 7, 2).detect(imgs[b])
        assert_same_features(res[b][0], res[b][1], k2, d2)


def test_uhd_image_through_the_tma_and_tcgen05_kernels(so, oracle):
    """3840 x 2160: 30 column strips and 45 row tiles at level 0 of the tcgen05 blur, TMA windows of the resize and
    descriptor kernels far from the origin, 16-bit coordinates everywhere — same bytes as the oracle."""
    from snakeorb import synthetic as S

    img = S.image(777, 3840, 2160)
    ex = so.ORBExtractor(5000, 1.2, 8, 20, 7)
    k, d = ex.Detect(img)
    ref = oracle.Extractor(5000, 1.2, 8, 20, 7, 4)
    k2, d2 = ref.detect(img)
    assert len(k) > 4000
    assert_same_features(k, d, k2, d2)
    for l in (0, 3, 7):
        assert np.array_equal(ex.GetImage(l), ref.level(l)), l
        assert np.array_equal(ex.GetImage(l, blurred=True), ref.level(l, blurred=True)), l
