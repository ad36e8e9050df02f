"""MiniBow2: the CPU restatement against the reference's own outputs.

tests/golden/voc_k9_L3.bin was trained and written by the reference's TemplatedVocabulary::create / saveRaw;
golden_next.npz holds what its loadRaw + transform returned (tests/golden/make_golden.py next)."""
import ctypes as C
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
VOC = os.path.join(HERE, "golden", "voc_k9_L3.bin")
SETS = ("a", "b", "c", "d")
LEVELSUP = (0, 1, 2, 3, 5)


@pytest.fixture(scope="module")
def golden():
    return np.load(os.path.join(HERE, "golden", "golden_next.npz"))


def check_against_golden(g, name, lu, got):
    ids, vals, fn, fo, fi = got
    assert np.array_equal(ids, g[f"bow_{name}_{lu}_ids"]), (name, lu)
    assert vals.tobytes() == g[f"bow_{name}_{lu}_vals"].tobytes(), (name, lu)  # float bits
    assert np.array_equal(fn, g[f"bow_{name}_{lu}_fv_nodes"]), (name, lu)
    assert np.array_equal(fo, g[f"bow_{name}_{lu}_fv_offsets"]), (name, lu)
    assert np.array_equal(fi, g[f"bow_{name}_{lu}_fv_indices"]), (name, lu)  # order inside a node included


def test_voc_file_and_transform_vs_reference_golden(oracle, golden):
    v = oracle.Vocabulary(VOC)
    assert (v.k, v.L, v.n_words) == (9, 3, 729) and v.n_nodes == 1 + 9 + 81 + 729
    for name in SETS:
        for lu in LEVELSUP:
            check_against_golden(golden, name, lu, v.transform(golden[f"bow_{name}_desc"], lu))
    ids, vals, fn, fo, fi = v.transform(np.zeros((0, 4), np.uint64), 1)
    assert len(ids) == 0 and len(fn) == 0 and list(fo) == [0]


def test_voc_save_raw_round_trip(oracle, tmp_path):
    v = oracle.Vocabulary(VOC)
    out = str(tmp_path / "copy.bin")
    v.save_raw(out)
    assert open(out, "rb").read() == open(VOC, "rb").read()  # byte-identical to what the reference's saveRaw wrote


def test_bow_invariants(oracle, golden):
    v = oracle.Vocabulary(VOC)
    d = golden["bow_b_desc"]
    ids, vals, fn, fo, fi = v.transform(d, 1)
    assert np.all(np.diff(ids) > 0) and np.all(np.diff(fn) > 0)
    assert abs(vals.sum() - 1.0) < 1e-5 and np.all(vals > 0)
    assert len(np.unique(fi)) == len(fi) and fo[-1] == len(fi) and len(fi) <= len(d)
    # levelsup >= L: every feature hangs under node 0
    _, _, fn5, fo5, fi5 = v.transform(d, 5)
    assert list(fn5) == [0] and len(fi5) == len(fi)
    # score of a vector with itself is 1 (L1, MiniBow2.h:134-170)
    assert abs(oracle.bow_score(ids, vals, ids, vals) - 1.0) < 1e-5


def test_transform_vs_compiled_reference_random(oracle, refso, golden):
    """Fresh random descriptor sets through the reference's transform and through the restatement."""
    v = oracle.Vocabulary(VOC)
    rng = np.random.default_rng(5)
    base = golden["bow_b_desc"]
    for n in (17, 400, 3000):
        d = base[rng.integers(0, len(base), n)] ^ (np.uint64(1) << rng.integers(0, 64, (n, 4)).astype(np.uint64))
        d = np.ascontiguousarray(d)
        ids, vals = np.zeros(n + 1, np.int32), np.zeros(n + 1, np.float32)
        fn, fo, fi = np.zeros(n + 1, np.int32), np.zeros(n + 2, np.int32), np.zeros(n + 1, np.int32)
        nb, nf = C.c_int(), C.c_int()
        refso.ref_voc_transform(VOC.encode(), C.c_void_p(d.ctypes.data), n, 2, C.c_void_p(ids.ctypes.data),
                                C.c_void_p(vals.ctypes.data), C.byref(nb), C.c_void_p(fn.ctypes.data),
                                C.c_void_p(fo.ctypes.data), C.c_void_p(fi.ctypes.data), C.byref(nf))
        got = v.transform(d, 2)
        assert np.array_equal(got[0], ids[:nb.value]) and got[1].tobytes() == vals[:nb.value].tobytes()
        assert np.array_equal(got[2], fn[:nf.value]) and np.array_equal(got[3], fo[:nf.value + 1])
        assert np.array_equal(got[4], fi[:fo[nf.value]])


def test_score_vs_compiled_reference(oracle, refso, golden):
    """FeatureVector::score: the oracle's merge against the reference's compiled function on real BowVectors."""
    refso.ref_bow_score.restype = C.c_float
    v = oracle.Vocabulary(VOC)
    vecs = [v.transform(golden[f"bow_{name}_desc"], 1)[:2] for name in SETS]
    for i1, v1 in vecs:
        for i2, v2 in vecs:
            a1, a2 = np.ascontiguousarray(i1, np.int32), np.ascontiguousarray(i2, np.int32)
            b1, b2 = np.ascontiguousarray(v1, np.float32), np.ascontiguousarray(v2, np.float32)
            want = refso.ref_bow_score(C.c_void_p(a1.ctypes.data), C.c_void_p(b1.ctypes.data), len(a1), C.c_void_p(a2.ctypes.data),
                                       C.c_void_p(b2.ctypes.data), len(a2))
            assert np.float32(oracle.bow_score(a1, b1, a2, b2)).tobytes() == np.float32(want).tobytes()
