import sys, time
sys.path[:0] = ["snake-slam_b200"]
import numpy as np
import snakeorb
from snakeorb import synthetic as S
ex = snakeorb.ORBExtractor(1000, 1.2, 8, 20, 7)
img = S.image(1000, 752, 480)
for _ in range(20): ex.Detect(img)
ex.set_profiling(True); ex.stage_ms(reset=True)
for _ in range(100): ex.Detect(img)
ms, ln = ex.stage_ms(reset=True)
names = ["copy-in", "pyramid", "fast", "distribute", "blur", "describe", "-", "total"]
print({n: round(float(m) / 100 * 1e3, 1) for n, m in zip(names, ms)}, "us per frame (device events)")
ex.set_profiling(False)
t0 = time.perf_counter()
for _ in range(200): ex.Detect(img)
print("wall per Detect: %.1f us" % ((time.perf_counter() - t0) / 200 * 1e6))
