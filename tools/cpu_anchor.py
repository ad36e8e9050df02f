#!/usr/bin/env python
"""Two more CPU figures for the bench line, beside `cpu_baseline`, each as one JSON line:

  --variant native   the CPU oracle rebuilt with -O3 -march=native on THIS host (oracle/_native/, `make -C oracle native`),
                     one extractor per thread — what a release build of the port gives on the box's own cores
  --variant hybrid   tests/hybrid.py: the ORBExtractor driver executed with the real cv2 4.13 kernels (cv::resize, cv::FAST,
                     cv::GaussianBlur — OpenCV's SIMD code, as the reference binary calls them at ORBExtractor.cpp:133,213,282)
                     plus the reference's own compiled Distribute / ComputeAngle / ComputeDescriptor (oracle/_ref), one
                     process per core. Closest to the real reference binary; the ~700 Python-level FAST calls per frame
                     cost it something, so it is a lower bound of that binary's speed.

Same workload as bench.py (C2: EuRoC stereo pair, 1200 kp, 8 levels, stereo row search). TEST / BENCH infrastructure only."""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "snake-slam_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
W, H, NFEAT, NLEVELS, SCALE, TH, THMIN, BF = 752, 480, 1200, 8, 1.2, 20, 7, 47.9


def pairs(n):
    import numpy as np

    from snakeorb import synthetic as S

    out = []
    for i in range(n):
        l = S.image(10000 + i, W, H)
        out.append((l, S.stereo_right(l, 15000 + i, 4.0, BF / 2 - 2.0)))
    return out


def _hybrid_worker(args):
    k, budget = args
    import hybrid
    import oracle_py as O

    sp = O.stereo_params(BF, True, O.scale_pyramid(NLEVELS, SCALE, NFEAT)[0])
    data = pairs(2)
    done, t0 = 0, time.perf_counter()
    while time.perf_counter() - t0 < budget:
        l, r = data[(k + done) % 2]
        kl, dl, _ = hybrid.hybrid_detect(l, NFEAT, SCALE, NLEVELS, TH, THMIN)
        kr, dr, _ = hybrid.hybrid_detect(r, NFEAT, SCALE, NLEVELS, TH, THMIN)
        O.stereo_match(sp, O.kp32_to_kp64(kl), dl, O.kp32_to_kp64(kr), dr)
        done += 1
    return done, time.perf_counter() - t0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--variant", required=True, choices=["native", "hybrid"])
    ap.add_argument("--threads", type=int, default=len(os.sched_getaffinity(0)))
    ap.add_argument("--budget", type=float, default=8.0)
    a = ap.parse_args()
    if a.variant == "native":
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "native"])
        os.environ["SNK_ORACLE_SO"] = os.path.join(ROOT, "oracle", "_native", "liborb_oracle_native.so")
        import oracle_py as O

        sp = O.stereo_params(BF, True, O.scale_pyramid(NLEVELS, SCALE, NFEAT)[0])
        data = pairs(4)
        exs = [O.Extractor(NFEAT, SCALE, NLEVELS, TH, THMIN, 1) for _ in range(a.threads)]
        counts = [0] * a.threads

        def work(t, stop):
            while time.perf_counter() < stop:
                l, r = data[(t + counts[t]) % len(data)]
                kl, dl = exs[t].detect(l)
                kr, dr = exs[t].detect(r)
                O.stereo_match(sp, O.kp32_to_kp64(kl), dl, O.kp32_to_kp64(kr), dr)
                counts[t] += 1

        for t in range(a.threads):  # warm-up: allocations
            work(t, time.perf_counter() + 0.05)
        counts = [0] * a.threads
        t0 = time.perf_counter()
        ths = [threading.Thread(target=work, args=(t, t0 + a.budget)) for t in range(a.threads)]
        [t.start() for t in ths]
        [t.join() for t in ths]
        dt = time.perf_counter() - t0
        print(json.dumps({"value": sum(counts) / dt, "unit": "frames/s", "cores": a.threads, "kind": "port",
                          "build": "-O3 -march=native -ffp-contract=off (this host)",
                          "sample": f"{sum(counts)} stereo pairs in {dt:.1f} s, one extractor per thread"}))
    else:
        import cv2  # noqa: F401
        import multiprocessing as mp

        import oracle_py as O

        if O.ref() is None:
            raise SystemExit("oracle/_ref not built")
        with mp.get_context("fork").Pool(a.threads) as pool:
            res = pool.map(_hybrid_worker, [(k, a.budget) for k in range(a.threads)])
        n, dt = sum(r[0] for r in res), max(r[1] for r in res)
        print(json.dumps({"value": n / dt, "unit": "frames/s", "cores": a.threads, "kind": "reference kernels (cv2 " + cv2.__version__ +
                          " SIMD resize / FAST / GaussianBlur + the reference's compiled Distribute, ComputeAngle, ComputeDescriptor), "
                          "Python driver", "sample": f"{n} stereo pairs in {dt:.1f} s, one process per core"}))


if __name__ == "__main__":
    main()
