"""MiniBow2::transform on the device against the reference's committed outputs and the oracle.

Everything is compared exactly: word ids, float bits of the normalised weights, node ids, CSR offsets and the order
of the feature indices inside every node (the permutation libstdc++'s unstable std::sort leaves)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
VOC = os.path.join(HERE, "golden", "voc_k9_L3.bin")


@pytest.fixture(scope="module")
def so():
    import snakeorb

    return snakeorb


@pytest.fixture(scope="module")
def ctx(so):
    return so.Context()


def same(a, b, what):
    assert np.array_equal(a[0], b[0]), ("bow ids", what)
    assert a[1].tobytes() == b[1].tobytes(), ("bow values", what)
    assert np.array_equal(a[2], b[2]), ("fv nodes", what)
    assert np.array_equal(a[3], b[3]), ("fv offsets", what)
    assert np.array_equal(a[4], b[4]), ("fv indices", what)


def test_transform_vs_reference_golden(so, ctx):
    g = np.load(os.path.join(HERE, "golden", "golden_next.npz"))
    v = so.Vocabulary(VOC)
    assert (v.k, v.L, v.n_nodes, v.n_words) == (9, 3, 820, 729)
    for name in ("a", "b", "c", "d"):
        for lu in (0, 1, 2, 3, 5):
            got = v.transform(ctx, g[f"bow_{name}_desc"], lu)
            want = tuple(g[f"bow_{name}_{lu}_{k}"] for k in ("ids", "vals", "fv_nodes", "fv_offsets", "fv_indices"))
            same(got, want, (name, lu))
    got = v.transform(ctx, np.zeros((0, 4), np.uint64), 1)
    assert len(got[0]) == 0 and len(got[2]) == 0 and list(got[3]) == [0]


@pytest.mark.parametrize("k,L,seed", [(10, 4, 1), (3, 6, 2), (20, 2, 3), (2, 9, 4)])
def test_transform_random_vocabularies(so, ctx, oracle, k, L, seed):
    """Ragged trees (short branches, fewer than k children, k > 16), zero-weight words, every size class of
    std::sort (<= 16 insertion only, one partition, deep recursion) and heavy node-id ties (large levelsup)."""
    from snakeorb import synthetic as S

    arr = S.vocabulary_arrays(seed, k, L)
    vo = oracle.Vocabulary(arrays=arr)
    vd = so.Vocabulary(arrays=arr)
    assert (vd.n_nodes, vd.n_words) == (vo.n_nodes, vo.n_words)
    rng = np.random.default_rng(seed)
    nodes = arr[5]
    for n in (1, 2, 16, 17, 18, 33, 100, 257, 1000, 2047, 2048, 4099, 8192):
        d = nodes[rng.integers(1, len(nodes), n)].copy()
        flips = rng.integers(0, 6, n)
        for j in range(5):
            m = flips > j
            d[m, rng.integers(0, 4)] ^= np.uint64(1) << rng.integers(0, 64, m.sum()).astype(np.uint64)
        for lu in (0, 1, L - 1, L, L + 2):
            same(vd.transform(ctx, d, lu), vo.transform(d, lu), (k, L, n, lu))


def test_sort_permutation_adversarial(so, ctx, oracle):
    """Inputs built to stress the exact-std::sort emulation: all features on one node, two alternating nodes,
    ascending / descending / organ-pipe node sequences (the last drives introsort towards its depth limit)."""
    from snakeorb import synthetic as S

    arr = S.vocabulary_arrays(11, 8, 3, ragged=False, zero_weight_fraction=0.0)
    vo = oracle.Vocabulary(arrays=arr)
    vd = so.Vocabulary(arrays=arr)
    nodes = arr[5]
    leaves = np.array([i for i in range(1, len(nodes)) if i not in set(arr[2][1:])])
    rng = np.random.default_rng(0)
    for n in (40, 500, 3000, 8192):
        seqs = {
            "one": np.zeros(n, np.int64),
            "two": np.arange(n) % 2,
            "asc": (np.arange(n) * len(leaves)) // n,
            "desc": ((n - 1 - np.arange(n)) * len(leaves)) // n,
            "pipe": np.where(np.arange(n) % 2 == 0, np.arange(n) // 2, n - 1 - np.arange(n) // 2) % len(leaves),
            "rand3": rng.integers(0, 3, n),
        }
        for name, seq in seqs.items():
            d = nodes[leaves[seq]]
            for lu in (0, 1, 2):
                same(vd.transform(ctx, d, lu), vo.transform(d, lu), (name, n, lu))


def test_bow_errors(so, ctx, tmp_path):
    v = so.Vocabulary(VOC)
    with pytest.raises(so.SnakeOrbError):
        v.transform(ctx, np.zeros((8193, 4), np.uint64), 1)
    with pytest.raises(so.SnakeOrbError):
        so.Vocabulary(str(tmp_path / "missing.bin"))
    bad = tmp_path / "bad.bin"
    bad.write_bytes(open(VOC, "rb").read()[:1000])
    with pytest.raises(so.SnakeOrbError):
        so.Vocabulary(str(bad))


def test_device_chain_extract_bow_knn2(so, ctx, oracle):
    """Device-resident chain: extractor results -> batched BoW transform and batched kNN-2 (frame i against
    frame i+1) without leaving the GPU; every frame equals the host-call results / the oracle."""
    from devmem import DevBuf
    from snakeorb import synthetic as S

    B = 5
    imgs = np.stack([S.image(40 + i // 2 if i < 4 else 99, 752, 480) for i in range(B)])  # frames 0/1 and 2/3 equal
    ex = so.ORBExtractor(1200, 1.2, 8, 20, 7, max_batch=B)
    # default handles (each owns a stream): the raw-pointer consumers are ordered behind the extraction explicitly
    ex.enqueue_ptr(imgs.ctypes.data, B, 480, 752, 752, 752 * 480, False)
    ctx.wait_orb(ex)
    kps_ptr, desc_ptr, cnt_ptr, stride = ex.device_results()
    v = so.Vocabulary(VOC)
    vo = oracle.Vocabulary(VOC)
    bi, bv, fn, fi = (DevBuf(4 * B * stride) for _ in range(4))
    fo = DevBuf(4 * B * (stride + 1))
    oc = DevBuf(8 * B)
    ctx.bow_transform_batch_device(v, desc_ptr, cnt_ptr, B, stride, 1, bi.ptr, bv.ptr, fn.ptr, fo.ptr, fi.ptr, oc.ptr)
    # kNN-2: problem p = frame p against frame p+1 (views into the same device arrays)
    out = DevBuf(16 * (B - 1) * stride)
    ctx.knn2_batch_device(desc_ptr, cnt_ptr, stride, desc_ptr + 32 * stride, cnt_ptr + 4, stride, B - 1, out.ptr)
    kps, desc, counts = ex.fetch(B)
    occ = oc.get(np.int32, (B, 2))
    bih, bvh = bi.get(np.int32, (B, stride)), bv.get(np.float32, (B, stride))
    fnh, fih, foh = fn.get(np.int32, (B, stride)), fi.get(np.int32, (B, stride)), fo.get(np.int32, (B, stride + 1))
    knn = out.get(so.KNN2, (B - 1, stride))
    for f in range(B):
        d = desc[f, :counts[f]]
        want = vo.transform(d, 1)
        nb, nn = occ[f]
        got = (bih[f, :nb], bvh[f, :nb], fnh[f, :nn], foh[f, :nn + 1], fih[f, :foh[f, nn]])
        same(got, want, f)
        same(v.transform(ctx, d, 1), want, ("host call", f))
    for p in range(B - 1):
        want = oracle.match_knn2(desc[p, :counts[p]], desc[p + 1, :counts[p + 1]])
        assert knn[p, :counts[p]].tobytes() == want.tobytes(), p
    assert np.all(knn[0, :counts[0]]["dist0"] == 0)  # identical frames


def test_bow_score_batch(so, ctx, oracle, refso=None):
    """FeatureVector::score of one frame against many: float bits equal to the oracle's serial merge (itself equal to
    the reference's compiled score on the same vectors, tests/test_bow_cpu.py)."""
    from snakeorb import synthetic as S

    arr = S.vocabulary_arrays(21, 10, 4)
    vo = oracle.Vocabulary(arrays=arr)
    rng = np.random.default_rng(2)
    nodes = arr[5]
    vecs = []
    for k in range(60):
        n = int(rng.integers(1, 1500))
        d = S.flip_bits(nodes[rng.integers(1, len(nodes), n)], rng, 6)
        ids, vals, *_ = vo.transform(d, 2)
        vecs.append((ids, vals))
    vecs.append((np.zeros(0, np.int32), np.zeros(0, np.float32)))  # a frame without words
    q = vecs[7]
    got = ctx.bow_scores(q[0], q[1], vecs)
    want = np.array([oracle.bow_score(q[0], q[1], i, v) for i, v in vecs], np.float32)
    assert got.tobytes() == want.tobytes()
    assert abs(got[7] - 1.0) < 1e-5 and got[-1] == 0.0 and np.all(got[:-1] >= 0)
    assert len(ctx.bow_scores(q[0], q[1], [])) == 0


def test_c4_chain_transform_then_match_bow(so, ctx, oracle):
    """Config C4 as SURVEY 8(d) words it: two TUM-shaped keyframes, their FeatureVectors from the vocabulary transform
    (levelsup as Frame::computeBoW passes it — the intra-node order is the std::sort-dependent one), validity masks
    70 % true, MatchBoW(th = 50, ratio = 0.75); plus the relocalisation form matchKnn2 + filterMatches(60, 0.8).
    The whole device chain equals the whole oracle chain."""
    from snakeorb import synthetic as S

    rng = np.random.default_rng(44)
    ex = so.ORBExtractor(1000, 1.2, 8, 20, 7)
    k1, d1 = ex.Detect(S.image(4000, 640, 480))
    # second keyframe: the same scene shifted by a few pixels plus a different frame's features
    img2 = np.roll(S.image(4000, 640, 480), (3, 5), (0, 1))
    k2, d2 = ex.Detect(img2)
    v, vo = so.Vocabulary(VOC), oracle.Vocabulary(VOC)

    def fv_view(nodes, offs, idx):
        fv = so.FeatVecView()
        nodes, offs, idx = (np.ascontiguousarray(a, np.int32) for a in (nodes, offs, idx))
        fv.n_nodes, fv.node_ids, fv.offsets, fv.indices = len(nodes), nodes.ctypes.data, offs.ctypes.data, idx.ctypes.data
        fv._keep = (nodes, offs, idx)
        return fv

    v1 = (rng.random(len(d1)) < 0.7).astype(np.uint8)
    v2 = (rng.random(len(d2)) < 0.7).astype(np.uint8)
    for levelsup in (1, 2):  # the test vocabulary has L = 3 (ORBvoc: L = 6, levelsup = 4): same node depths above the words
        g1, g2 = v.transform(ctx, d1, levelsup), v.transform(ctx, d2, levelsup)
        o1, o2 = vo.transform(d1, levelsup), vo.transform(d2, levelsup)
        same(g1, o1, "kf1"), same(g2, o2, "kf2")
        m, m12 = ctx.MatchBoW(d1, d2, fv_view(*g1[2:]), fv_view(*g2[2:]), v1, v2, 50, 0.75)
        mo, m12o, ops = oracle.match_bow(d1, d2, fv_view(*o1[2:]), fv_view(*o2[2:]), v1, v2, 50, 0.75)
        assert m == mo and np.array_equal(m12, m12o) and m > 50
        # BoW similarity of the two keyframes, as the loop detector would rank them
        assert ctx.bow_scores(g1[0], g1[1], [(g2[0], g2[1])])[0] == np.float32(oracle.bow_score(o1[0], o1[1], o2[0], o2[1]))
    knn = ctx.matchKnn2(d1, d2)
    assert knn.tobytes() == oracle.match_knn2(d1, d2).tobytes()
    assert np.array_equal(ctx.filterMatches(knn, 60, 0.8), oracle.filter_matches(oracle.match_knn2(d1, d2), 60, 0.8))
