"""The CPU restatement of the Snake matchers against the outputs frozen from the reference's own compiled code
(tests/golden/golden_snake.npz, written by tests/golden/make_golden_snake.py). Runs without /root/reference and
without oracle/_ref: this is the pin that travels."""
import os

import numpy as np
import pytest

import snake_cases as SC
import snake_runners as R

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden_snake.npz")


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLD)


def frozen(gold, name, seed, big, c, inputs):
    key = f"{name}/{seed}/{int(big)}"
    assert bytes(gold[key + "/digest"]).decode() == SC.digest(*inputs(c)), "seeded inputs drifted from the frozen ones"
    out, i = [], 0
    while f"{key}/{i}" in gold:
        out.append(gold[f"{key}/{i}"])
        i += 1
    return out


@pytest.mark.parametrize("name", list(R.MATCHERS))
def test_oracle_equals_frozen_reference_outputs(oracle, gold, name):
    gen, inputs, _, run_oracle, _ = R.MATCHERS[name]
    for seed, big in R.GOLDEN_CASES:
        c = gen(seed, big)
        assert R.same(R.norm(run_oracle(oracle, c)), frozen(gold, name, seed, big, c, inputs)), (name, seed, big)
