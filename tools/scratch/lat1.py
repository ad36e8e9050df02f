"""Single-image / stereo-pair latency and its per-stage split (stage events serialise the side stream)."""
import sys, time, os
sys.path[:0] = [os.path.join(os.path.dirname(__file__), "..", "..", "snake-slam_b200")]
import numpy as np
import snakeorb
from snakeorb import synthetic as S

for nb in (1, 2):
    imgs = np.stack([S.image(9000 + i, 752, 480) for i in range(nb)])
    ex = snakeorb.ORBExtractor(1000, 1.2, 8, 20, 7, max_batch=nb)
    for _ in range(20): ex.DetectBatch(imgs)
    ts = []
    for _ in range(200):
        t = time.perf_counter(); ex.DetectBatch(imgs); ts.append(time.perf_counter() - t)
    print("batch", nb, "median ms", round(np.median(ts) * 1e3, 4), "min", round(min(ts) * 1e3, 4))
    ex.set_profiling(True); ex.stage_ms(reset=True)
    for _ in range(100): ex.DetectBatch(imgs)
    ms, ln = ex.stage_ms(reset=True)
    print("   stage us per call:", [round(float(x) * 10, 1) for x in ms], "launches", list(ln // 100))
