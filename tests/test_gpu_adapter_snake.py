"""The Snake-struct drop-in classes (snake-slam_b200/host/snakeorb_snake_adapter.h: SnakeORBMatcherB200,
LoopORBmatcherB200, PreprocessB200) driven with REAL Snake::Frame / Keyframe / MapPoint / LocalMap objects on the GPU.

oracle/_ref/libsnake_adapter.so is oracle/ref_snake_shim.cpp compiled against the reference's own headers with
SNK_USE_ADAPTER: it builds the same Snake objects as for the reference's functions, but calls the adapter classes,
which go through the C ABI of libsnakeorb.so. Their effects on the Snake objects (Frame::mvpMapPoints,
LocalMap::points[i].valid, MapPoint::mnVisible, Frame::right_points / depth, vpMatches12) must equal what the
reference's own compiled matchers left behind for the same inputs (tests/golden/golden_snake.npz).
The library is built where /root/reference exists (oracle/Makefile) and travels to the GPU box with the snapshot."""
import numpy as np
import pytest

import snake_cases as SC
import snake_runners as R
from test_snake_golden import GOLD, frozen

pytestmark = pytest.mark.gpu
COVERED = ["stereo", "rgbd", "fine", "coarse", "frame_to_kf", "match_bow", "match_bow_batch"]


@pytest.fixture(scope="module")
def adapter(oracle):
    lib = oracle.snake_adapter()
    if lib is None:
        pytest.skip("oracle/_ref/libsnake_adapter.so not built (needs /root/reference at build time)")
    oracle.set_snake_lib(lib)
    yield oracle
    oracle.set_snake_lib(None)


@pytest.mark.parametrize("name", COVERED)
def test_adapter_classes_reproduce_the_reference_on_real_snake_objects(adapter, name):
    gold = np.load(GOLD)
    gen, inputs, run_ref, _, _ = R.MATCHERS[name]
    for seed, big in R.GOLDEN_CASES:
        c = gen(seed, big)
        assert R.same(R.norm(run_ref(adapter, c)), frozen(gold, name, seed, big, c, inputs)), (name, seed, big)


def test_adapter_undistort_keypoints(adapter):
    for seed in range(20):
        c = SC.case_stereo(seed)
        k = c["left"].copy()
        k["x"], k["y"] = np.clip(k["x"], 0, c["w"] - 1), np.clip(k["y"], 0, c["h"] - 1)
        got = adapter.ref_undistort_keypoints(c["rl"], k, c["h"], c["w"])
        want = adapter.undistort_keypoints(c["rl"], k)
        assert got[0].tobytes() == want[0].tobytes() and got[1].tobytes() == want[1].tobytes(), seed
