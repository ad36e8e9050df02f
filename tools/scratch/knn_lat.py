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

# C3: SearchByProjection2 of 20 000 points, host structs per call vs device-resident frame + local map
sys.path[:0] = ["tests", "oracle"]
import matchdata as MD
from snakeorb import synthetic as S

rngn = np.random.default_rng(5)
exk = so.ORBExtractor(2000, 1.2, 8, 20, 7)
kk, dk = exk.Detect(S.image(3000, 1241, 376))
sc_k, _ = exk.getPyramid()
Kk = (718.856, 718.856, 607.19, 185.22)
fr = MD.make_frame(kk, dk, 1241, 376, rngn, S.quat_from_axis_angle(rngn.normal(size=3), 0.004), rngn.uniform(-0.05, 0.05, 3))
mp = S.map_points_from_keypoints(fr.kp, fr.desc, Kk, rngn, 20000, sc_k)
pts = MD.fine_points(mp, rngn)
cam, pyr = MD.camera(*Kk, 386.1), so.scale_pyramid_c(1.2, sc_k)
for _ in range(5):
    ctx.SearchByProjection2(fr.view, pts, cam, pyr, 4.0, 0.8)
t0 = time.perf_counter()
for _ in range(50):
    ctx.SearchByProjection2(fr.view, pts, cam, pyr, 4.0, 0.8)
print("C3 host structs per call ms", (time.perf_counter() - t0) / 50 * 1e3)
fd, lm = ctx.frame_dev(fr.view), ctx.localmap(20000)
lm.upload(pts)
for _ in range(5):
    ctx.SearchByProjection2Resident(fd, lm, cam, pyr, 4.0, 0.8, keep_assignment=False)
t0 = time.perf_counter()
for _ in range(50):
    lm.upload(pts)  # valid flags back to the start (the search clears some)
    r = ctx.SearchByProjection2Resident(fd, lm, cam, pyr, 4.0, 0.8, keep_assignment=False)
t1 = time.perf_counter()
for _ in range(50):
    r = ctx.SearchByProjection2Resident(fd, lm, cam, pyr, 4.0, 0.8, keep_assignment=False)
print("C3 resident ms incl. re-upload of the points", (t1 - t0) / 50 * 1e3, "search only", (time.perf_counter() - t1) / 50 * 1e3, "matches", r[0])
