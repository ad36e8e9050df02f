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
