#!/usr/bin/env python3
"""Freezes the outputs of the reference's OWN compiled Snake matchers (oracle/_ref/libsnake_ref.so: Preprocess.cpp,
SnakeORBMatcher.cpp, LoopORBMatcher.cpp, MappingORBMatcher.cpp, MonoInitializer.cpp built unmodified from
/root/reference by oracle/Makefile) for the seeded cases of tests/snake_runners.py into tests/golden/golden_snake.npz.
Needs /root/reference (run in the build container):  python tests/golden/make_golden_snake.py
Inputs are not stored: they are regenerated from the seed (tests/snake_cases.py) and checked against the stored digest."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "snake-slam_b200")):
    sys.path.insert(0, p)
import oracle_py as O  # noqa: E402
import snake_cases as SC  # noqa: E402
import snake_runners as R  # noqa: E402

O.build()
assert O.snake_ref() is not None, "oracle/_ref/libsnake_ref.so missing: run make -C oracle (needs /root/reference)"
out = {}
for name, (gen, inputs, run_ref, run_oracle, _) in R.MATCHERS.items():
    for seed, big in R.GOLDEN_CASES:
        c = gen(seed, big)
        res = R.norm(run_ref(O, c))
        assert R.same(res, R.norm(run_oracle(O, c))), (name, seed)  # the restatement agrees while we are at it
        key = f"{name}/{seed}/{int(big)}"
        out[key + "/digest"] = np.frombuffer(SC.digest(*inputs(c)).encode(), np.uint8)
        for i, a in enumerate(res):
            out[f"{key}/{i}"] = a
    print(name, "frozen")
np.savez_compressed(os.path.join(HERE, "golden_snake.npz"), **out)
print(os.path.getsize(os.path.join(HERE, "golden_snake.npz")), "bytes")
