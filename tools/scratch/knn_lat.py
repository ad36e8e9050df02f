"""Host-to-host latency of one kNN-2 problem through the C ABI (snk_match_knn2)."""
import sys
import time

sys.path[:0] = ["snake-slam_b200"]
import numpy as np

import snakeorb as so

ctx = so.Context()
rng = np.random.default_rng(1)
for n, m in ((1000, 1000), (2000, 2000), (300, 1000), (5000, 5000)):
    a = rng.integers(0, 2**63, (n, 4)).astype(np.uint64)
    b = rng.integers(0, 2**63, (m, 4)).astype(np.uint64)
    for _ in range(20):
        ctx.matchKnn2(a, b)
    t0 = time.perf_counter()
    for _ in range(200):
        ctx.matchKnn2(a, b)
    print(n, m, "host-to-host ms", (time.perf_counter() - t0) / 200 * 1e3)
