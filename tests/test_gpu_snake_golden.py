"""GPU parity of every Snake matcher against the outputs of the reference's OWN compiled code: the CUDA library is
called through the C ABI on the seeded cases of tests/snake_runners.py and must reproduce, bit for bit, what
/root/reference's Preprocess.cpp / SnakeORBMatcher.cpp / LoopORBMatcher.cpp / MappingORBMatcher.cpp /
MonoInitializer.cpp returned for the same inputs (frozen in tests/golden/golden_snake.npz), and — on 300 further seeds
per matcher — what the CPU restatement returns (itself pinned live on 1000 seeds, test_oracle_snake_live.py)."""
import os

import numpy as np
import pytest

import snake_runners as R
from test_snake_golden import GOLD, frozen

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLD)


@pytest.fixture(scope="module")
def so():
    import snakeorb

    return snakeorb


@pytest.fixture(scope="module")
def ctx(so):
    return so.Context()


@pytest.mark.parametrize("name", list(R.MATCHERS))
def test_gpu_equals_frozen_reference_outputs(so, ctx, gold, name):
    gen, inputs, _, _, run_gpu = R.MATCHERS[name]
    for seed, big in R.GOLDEN_CASES:
        c = gen(seed, big)
        assert R.same(R.norm(run_gpu(so, ctx, c)), frozen(gold, name, seed, big, c, inputs)), (name, seed, big)


@pytest.mark.parametrize("name", list(R.MATCHERS))
def test_gpu_equals_oracle_on_more_seeds(so, ctx, oracle, name):
    gen, _, _, run_oracle, run_gpu = R.MATCHERS[name]
    for seed in range(2000, 2300):
        c = gen(seed)
        assert R.same(R.norm(run_gpu(so, ctx, c)), R.norm(run_oracle(oracle, c))), (name, seed)
