"""Error behaviour of the C ABI (include/snakeorb.h): bad arguments come back as status codes with a message in
snk_last_error(), never as a crash, and a handle stays usable after a rejected call. The reference asserts
(SAIGA_ASSERT) in the same places; the adapter header turns a non-zero status into the same abort."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

OK, INVALID, CUDA, CAPACITY, UNSUPPORTED = 0, -1, -2, -3, -4


@pytest.fixture(scope="module")
def so():
    import snakeorb

    return snakeorb


def last(so):
    return so.lib().snk_last_error().decode()


def create(so, **kw):
    p = so.OrbParams(nfeatures=500, scale_factor=1.2, nlevels=4, th_fast=20, th_fast_min=7, max_batch=2, device=0)
    for k, v in kw.items():
        setattr(p, k, v)
    h = C.c_void_p()
    rc = so.lib().snk_orb_create(C.byref(p), C.byref(h))
    return rc, h


def test_create_rejects_bad_parameters(so):
    L = so.lib()
    for kw in (dict(nfeatures=0), dict(nlevels=0), dict(nlevels=99), dict(scale_factor=1.0), dict(scale_factor=1.31),
               dict(th_fast=0), dict(th_fast=255), dict(th_fast_min=-1), dict(max_batch=0)):
        rc, h = create(so, **kw)
        assert rc == INVALID and not h.value, kw
        assert last(so), kw
    rc, h = create(so, device=4711)
    assert rc in (INVALID, CUDA) and not h.value
    assert L.snk_orb_create(None, C.byref(C.c_void_p())) == INVALID
    p = so.OrbParams(nfeatures=500, scale_factor=1.2, nlevels=4, th_fast=20, th_fast_min=7, max_batch=1, device=0)
    assert L.snk_orb_create(C.byref(p), None) == INVALID
    L.snk_orb_destroy(None)  # like free(NULL)


def test_detect_rejects_bad_arguments_and_the_handle_survives(so, oracle):
    from snakeorb import synthetic as S

    L = so.lib()
    rc, h = create(so)
    assert rc == OK
    cap = L.snk_orb_max_keypoints(h)
    img = S.image(77, 320, 240)
    kps = np.zeros(cap, so.KP32)
    desc = np.zeros((cap, 4), np.uint64)
    n = C.c_int(-5)
    P = lambda a: C.c_void_p(a.ctypes.data)
    good = lambda: L.snk_orb_detect(h, P(img), 240, 320, C.c_size_t(320), 0, P(kps), P(desc), cap, C.byref(n))
    assert L.snk_orb_detect(None, P(img), 240, 320, C.c_size_t(320), 0, P(kps), P(desc), cap, C.byref(n)) == INVALID
    assert L.snk_orb_detect(h, P(img), 240, 320, C.c_size_t(319), 0, P(kps), P(desc), cap, C.byref(n)) == INVALID  # pitch < cols
    assert L.snk_orb_detect(h, P(img), 240, 320, C.c_size_t(320), 0, None, P(desc), cap, C.byref(n)) == INVALID
    assert L.snk_orb_detect(h, P(img), 240, 320, C.c_size_t(320), 0, P(kps), P(desc), cap, None) == INVALID
    # fewer slots than keypoints found: SNK_ERR_CAPACITY (a capacity below the bound is fine while the image has fewer)
    assert L.snk_orb_detect(h, P(img), 240, 320, C.c_size_t(320), 0, P(kps), P(desc), 10, C.byref(n)) == CAPACITY
    assert L.snk_orb_detect(h, P(img), 4001, 320, C.c_size_t(320), 0, P(kps), P(desc), cap, C.byref(n)) == INVALID  # > 4000
    assert L.snk_orb_detect(h, P(img), 40, 40, C.c_size_t(40), 0, P(kps), P(desc), cap, C.byref(n)) == INVALID      # no cell rows
    assert "snk" in last(so) or last(so)
    # more images than max_batch
    nb = np.zeros(3, np.int32)
    assert L.snk_orb_detect_batch(h, P(img), 3, 80, 320, C.c_size_t(320), C.c_size_t(80 * 320), 0, P(kps), P(desc), cap // 3,
                                  P(nb)) == INVALID
    # after all that the handle still extracts, and correctly
    assert good() == OK
    k2, d2 = oracle.Extractor(500, 1.2, 4, 20, 7, 1).detect(img)
    assert n.value == len(k2) > 50 and np.array_equal(desc[:n.value], d2)
    # an empty image is not an error (ORBExtractor.cpp:182)
    assert L.snk_orb_detect(h, None, 0, 0, C.c_size_t(0), 0, P(kps), P(desc), cap, C.byref(n)) == OK and n.value == 0
    L.snk_orb_destroy(h)


def test_matcher_entry_points_reject_bad_arguments(so, tmp_path):
    L = so.lib()
    ctx = so.Context()
    c = ctx.c
    P = lambda a: C.c_void_p(a.ctypes.data)
    d = np.arange(40, dtype=np.uint64).reshape(10, 4)
    out = np.zeros((10, 4), np.int32)
    # kNN-2: null context, negative sizes, null descriptor arrays with rows, and the empty problems that ARE fine
    assert L.snk_match_knn2(None, P(d), 10, P(d), 10, P(out)) == INVALID
    assert L.snk_match_knn2(c, P(d), -1, P(d), 10, P(out)) == INVALID
    assert L.snk_match_knn2(c, None, 10, P(d), 10, P(out)) == INVALID
    assert L.snk_match_knn2(c, P(d), 10, P(d), 10, None) == INVALID
    assert L.snk_match_knn2(c, None, 0, P(d), 10, P(out)) == OK
    assert L.snk_match_knn2(c, P(d), 10, None, 0, P(out)) == OK
    assert (out[:, 1] == -1).all() and (out[:, 0] == 1000).all()  # {1000, -1} when there is nothing to match (Features.h:330)
    nm = C.c_int()
    ij = np.zeros(32, np.int32)
    assert L.snk_filter_matches(None, 10, 50, C.c_float(0.8), P(ij), C.byref(nm)) == INVALID
    assert L.snk_filter_matches(P(out), -1, 50, C.c_float(0.8), P(ij), C.byref(nm)) == INVALID
    # stereo: level count and octaves are checked before anything is launched
    sp = so.stereo_params(47.9, True, [1.2 ** i for i in range(8)])
    kp = np.zeros(4, so.KP64)
    kp["octave"] = [0, 1, 2, 9]
    rp, dp = np.zeros(4, np.float32), np.zeros(4, np.float32)
    args = (P(kp), P(d), 4, P(kp), P(d), 4, P(rp), P(dp), C.byref(nm))
    assert L.snk_stereo_match(c, C.byref(sp), *args) == INVALID and "octave" in last(so)
    kp["octave"] = [0, 1, 2, 3]
    assert L.snk_stereo_match(c, C.byref(sp), *args) == OK
    sp.nlevels = 17
    assert L.snk_stereo_match(c, C.byref(sp), *args) == INVALID
    assert L.snk_stereo_match(c, None, *args) == INVALID
    # feature cache files: a missing file and a capacity that is too small (the count is still reported)
    n = C.c_int(-1)
    assert L.snk_features_read(str(tmp_path / "nope.features").encode(), None, None, 0, C.byref(n)) == INVALID
    k = np.zeros(5, so.KP64)
    so.features_write(str(tmp_path / "five.features"), k, d[:5])
    assert L.snk_features_read(str(tmp_path / "five.features").encode(), None, None, 0, C.byref(n)) == CAPACITY and n.value == 5
    # the context still works
    got = ctx.matchKnn2(d, d)
    assert (got["idx0"] == np.arange(10)).all() and (got["dist0"] == 0).all()


def test_device_fault_flag_is_reported_once_on_the_async_and_the_sync_path(so):
    """The kernels' fault flag (capacity guards, bounded TMA wait) cannot be tripped with valid input, so it is injected
    (snk_orb_debug_set_error): snk_orb_check after an enqueue + fetch_async reports it as SNK_ERR_CAPACITY and clears it;
    the synchronous fetch does the same; a clean run afterwards is clean."""
    import ctypes as C

    from snakeorb import synthetic as S

    img = S.image(77, 320, 240)
    ex = so.ORBExtractor(500, 1.2, 4, 20, 7, max_batch=1)
    want_k, want_d = ex.Detect(img)
    lib = so.lib()
    cap = ex.cap
    kps, desc, cnt = np.zeros((1, cap), so.KP32), np.zeros((1, cap, 4), np.uint64), np.zeros(1, np.int32)
    # asynchronous path
    ex.enqueue_ptr(img.ctypes.data, 1, 240, 320, 320, 320 * 240, False)
    assert lib.snk_orb_debug_set_error(ex.h, 4) == 0
    assert lib.snk_orb_fetch_async(ex.h, C.c_void_p(kps.ctypes.data), C.c_void_p(desc.ctypes.data), cap, C.c_void_p(cnt.ctypes.data)) == 0
    assert lib.snk_orb_check(ex.h) == -3 and b"capacity error 4" in lib.snk_last_error()  # SNK_ERR_CAPACITY
    assert lib.snk_orb_check(ex.h) == 0  # reported once
    # synchronous path
    ex.enqueue_ptr(img.ctypes.data, 1, 240, 320, 320, 320 * 240, False)
    assert lib.snk_orb_debug_set_error(ex.h, 2) == 0
    with pytest.raises(so.SnakeOrbError):
        ex.fetch(1)
    k2, d2 = ex.Detect(img)
    assert k2.tobytes() == want_k.tobytes() and d2.tobytes() == want_d.tobytes()
