"""Full BASELINE sizes on the GPU through properties that do not need the (slow) oracle on every frame:
frame independence, determinism, permutation equivariance, and a checksum of checksums over the C5 sequence that
must not depend on how the frames are sharded and batched."""
import hashlib

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
W, H = 752, 480


@pytest.fixture(scope="module")
def so():
    import snakeorb

    return snakeorb


def frame_digest(k, d):
    return hashlib.sha1(k.tobytes() + d.tobytes()).digest()


def test_c2_full_batch_independence_determinism_permutation(so, oracle):
    """C2 at the bench's batch: 512 images (256 stereo pairs) in one launch sequence. The batch holds 8 distinct
    pairs tiled 32 times: every replica must be bit-identical (frames are independent), the distinct ones equal the
    oracle, a second run repeats the bytes, and permuting the batch permutes the results."""
    from snakeorb import synthetic as S

    distinct = []
    for i in range(8):
        left = S.image(9000 + i, W, H)
        distinct += [left, S.stereo_right(left, 9100 + i, 4.0, 20.0)]
    distinct = np.stack(distinct)
    batch = np.ascontiguousarray(np.tile(distinct, (32, 1, 1)))
    assert batch.shape == (512, H, W)
    ex = so.ORBExtractor(1200, 1.2, 8, 20, 7, max_batch=512)
    res = ex.DetectBatch(batch)
    dig = [frame_digest(k, d) for k, d in res]
    for i in range(512):
        assert dig[i] == dig[i % 16], i
    ref = oracle.Extractor(1200, 1.2, 8, 20, 7, 4)
    for i in range(16):
        k2, d2 = ref.detect(distinct[i])
        assert len(k2) > 1000 and np.array_equal(res[i][1], d2)
        for f in ("x", "y", "octave", "response", "size"):
            assert np.array_equal(res[i][0][f], k2[f]), (i, f)
        assert np.allclose(res[i][0]["angle"], k2["angle"], rtol=1e-4, atol=1e-4)
    assert [frame_digest(k, d) for k, d in ex.DetectBatch(batch)] == dig  # determinism
    perm = np.random.default_rng(0).permutation(512)
    res_p = ex.DetectBatch(np.ascontiguousarray(batch[perm]))
    assert [frame_digest(k, d) for k, d in res_p] == [dig[j] for j in perm]


@pytest.mark.parametrize("world", [1, 2, 4, 8])
def test_c5_sequence_checksum_is_sharding_invariant(so, world):
    """C5: a 3680-pair sequence cut into contiguous blocks per rank (snakeorb.sharding) and into batches of 256 pairs
    per rank. The sequence is 16 distinct pairs in a seeded random order; the per-frame digests — keypoints and
    descriptors of both images AND the stereo row search's right_points / depth / match count of the pair — gathered in
    frame order must equal the digests of the distinct pairs looked up by the same order, whatever the world size: no
    frame lost, duplicated, reordered or influenced by its batch. The expected stereo results come from the host entry
    point (snk_stereo_match on fetched keypoints), the sharded run uses the device-resident one fed by the extractor."""
    from devmem import DevBuf
    from snakeorb import sharding
    from snakeorb import synthetic as S

    F, PAIRS = 3680, 16
    rng = np.random.default_rng(5)
    order = rng.integers(0, PAIRS, F)
    pairs = []
    for i in range(PAIRS):
        left = S.image(9500 + i, W, H)
        pairs.append(np.stack([left, S.stereo_right(left, 9600 + i, 4.0, 20.0)]))
    pairs = np.stack(pairs)  # (16, 2, H, W)
    ex = so.ORBExtractor(1200, 1.2, 8, 20, 7, max_batch=512)
    ctx = so.Context()
    scales, _ = ex.getPyramid()
    sp = so.stereo_params(47.9, True, scales)
    base = ex.DetectBatch(np.ascontiguousarray(pairs.reshape(-1, H, W)))

    def pair_digest(kl, dl, kr, dr, nm, rp, dp):
        return hashlib.sha1(kl.tobytes() + dl.tobytes() + kr.tobytes() + dr.tobytes() + np.int32(nm).tobytes() + rp.tobytes() +
                            dp.tobytes()).digest()

    base_dig = []
    for p in range(PAIRS):
        (kl, dl), (kr, dr) = base[2 * p], base[2 * p + 1]
        nm, rp, dp = ctx.StereoMatching(sp, so.kp32_to_kp64(kl), dl, so.kp32_to_kp64(kr), dr)
        assert nm > 100
        base_dig.append(pair_digest(kl, dl, kr, dr, nm, rp, dp))
    want = hashlib.sha1(b"".join(base_dig[p] for p in order)).hexdigest()
    stride = ex.device_results()[3]
    d_rp, d_dp, d_nm = DevBuf(4 * 256 * stride), DevBuf(4 * 256 * stride), DevBuf(4 * 256)
    per_rank = []
    for rank in range(world):
        lo, hi = sharding.shard_range(F, rank, world)
        mine = []
        for b0, b1 in sharding.batches(lo, hi, 256):
            imgs = np.ascontiguousarray(pairs[order[b0:b1]].reshape(-1, H, W))
            n = b1 - b0
            ex.enqueue_ptr(imgs.ctypes.data, 2 * n, H, W, W, W * H, False)
            ctx.stereo_orb_device(ex, n, sp, d_rp.ptr, d_dp.ptr, d_nm.ptr)
            kps, desc, cnt = ex.fetch(2 * n)
            rp, dp, nm = d_rp.get(np.float32, (256, stride)), d_dp.get(np.float32, (256, stride)), d_nm.get(np.int32, (256,))
            for j in range(n):
                cl, cr = cnt[2 * j], cnt[2 * j + 1]
                mine.append(pair_digest(kps[2 * j, :cl], desc[2 * j, :cl], kps[2 * j + 1, :cr], desc[2 * j + 1, :cr], nm[j], rp[j, :cl],
                                        dp[j, :cl]))
        per_rank.append(mine)
    got = sharding.gather_in_frame_order(per_rank)
    assert len(got) == F
    assert hashlib.sha1(b"".join(got)).hexdigest() == want
