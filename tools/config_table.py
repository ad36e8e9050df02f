#!/usr/bin/env python
"""One-GPU measurements for the result table of BASELINE.md section 4 (configs C1, C3, C4 and the matcher kernels;
C2 / C5 are bench.py's own workload). Device-resident, batched, CUDA events on the launching stream; the CPU column is
the oracle port on the host cores (bounded sample). Prints one JSON line.

    python tools/config_table.py [--batch 512] [--steps 10]
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "snake-slam_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)

CONFIGS = {  # name: (width, height, nfeatures)
    "C1 EuRoC mono 752x480, 1000 kp": (752, 480, 1000),
    "C3 KITTI 1241x376, 2000 kp": (1241, 376, 2000),
    "C4 TUM 640x480, 1000 kp": (640, 480, 1000),
}
NLEVELS, SCALE, TH, THMIN = 8, 1.2, 20, 7


def level_sizes(w, h):
    """ScalePyramid's level sizes (float32 arithmetic, as bench.py)."""
    out, sc = [], 1.0
    for l in range(NLEVELS):
        if l:
            sc = float(np.float32(SCALE)) * sc
        s = np.float32(1.0 / sc)
        r = lambda v: int(np.float32(np.float32(v) * s) + np.float32(0.5))
        out.append((r(h), r(w)) if l else (h, w))
    return out


def algorithmic_bytes(w, h, n_kp, n_cand):
    """SURVEY.md 8(d), same accounting as bench.py."""
    S = [a * b for a, b in level_sizes(w, h)]
    sig = sum(S)
    return (sig - S[-1]) + (sig - S[0]) + sig + 8 * n_cand + 2 * sig + 2386 * n_kp


def cpu_images_per_s(imgs, nf, threads, budget_s):
    import oracle_py

    oracle_py.build()
    done = [0] * threads
    stop = time.perf_counter() + budget_s

    def work(t):
        ex = oracle_py.Extractor(nf, SCALE, NLEVELS, TH, THMIN, 1)
        i = t
        while time.perf_counter() < stop:
            ex.detect(imgs[i % len(imgs)])
            i += threads
            done[t] += 1

    t0 = time.perf_counter()
    ths = [threading.Thread(target=work, args=(t,)) for t in range(threads)]
    [t.start() for t in ths]
    [t.join() for t in ths]
    return sum(done) / (time.perf_counter() - t0)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=512)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--cpu-budget", type=float, default=4.0)
    ap.add_argument("--threads", type=int, default=len(os.sched_getaffinity(0)))
    args = ap.parse_args()
    import torch

    import snakeorb
    from snakeorb import synthetic as S

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm = peaks.get("hbm_gbs", 6650.0)
    st = torch.cuda.Stream()
    torch.cuda.set_stream(st)
    rows = {}
    for name, (w, h, nf) in CONFIGS.items():
        B = args.batch if w * h < 400000 else args.batch // 2 * 2
        distinct = np.stack([S.image(20000 + i, w, h) for i in range(8)])
        batch = np.empty((B, h, w), np.uint8)
        for i in range(B):
            batch[i] = np.roll(distinct[i % 8], 3 * (i // 8), axis=0)
        dev = torch.from_numpy(batch).cuda()
        ex = snakeorb.ORBExtractor(nf, SCALE, NLEVELS, TH, THMIN, max_batch=B)
        ex.set_stream(st.cuda_stream)
        for _ in range(3):
            ex.enqueue_ptr(dev.data_ptr(), B, h, w, w, w * h, True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(args.steps):
            ex.enqueue_ptr(dev.data_ptr(), B, h, w, w, w * h, True)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / args.steps
        res = ex.DetectBatch(batch[:8])
        n_kp = float(np.mean([len(k) for k, _ in res]))
        n_cand = sum(len(ex.candidates(l, 0)) for l in range(NLEVELS))
        ab = algorithmic_bytes(w, h, n_kp, n_cand)
        cpu = cpu_images_per_s(distinct, nf, args.threads, args.cpu_budget)
        # latency: one image, host in, host out, synchronous Detect; median of 200 after 20 warm-ups
        ex1 = snakeorb.ORBExtractor(nf, SCALE, NLEVELS, TH, THMIN, max_batch=1)
        for _ in range(20):
            ex1.Detect(distinct[0])
        lat = []
        for i in range(200):
            t0 = time.perf_counter()
            ex1.Detect(distinct[i % 8])
            lat.append(time.perf_counter() - t0)
        # the reference's own threading: one extractor, fd_threads = 2 (Settings.h:90)
        import oracle_py

        oe = oracle_py.Extractor(nf, SCALE, NLEVELS, TH, THMIN, 2)
        oe.detect(distinct[0])
        t0 = time.perf_counter()
        for i in range(10):
            oe.detect(distinct[i % 8])
        cpu2_ms = (time.perf_counter() - t0) / 10 * 1e3
        del ex1
        rows[name] = {"batch": B, "ms_per_batch": ms, "images_per_s": B / (ms * 1e-3), "keypoints_per_s": B * n_kp / (ms * 1e-3),
                      "keypoints_per_image": n_kp, "raw_candidates_per_image": n_cand, "extraction_bytes_per_image": ab,
                      "hbm_frac": ab * B / (ms * 1e-3) / 1e9 / hbm, "cpu_port_images_per_s": cpu, "cpu_threads": args.threads,
                      "latency_ms_median": float(np.median(lat)) * 1e3, "latency_ms_p95": float(np.percentile(lat, 95)) * 1e3,
                      "cpu_port_ms_per_image_fd_threads_2": cpu2_ms}
        del ex, dev

    # ---- matcher kernels on the C2 shape: Hamming pair evaluations per second against the measured popc issue peak
    import ctypes as C

    import oracle_py

    w, h, nf, P = 752, 480, 1200, args.batch // 2
    pairs = []
    for i in range(8):
        l = S.image(10000 + i, w, h)
        pairs += [l, S.stereo_right(l, 15000 + i, 4.0, 47.9 / 2 - 2.0)]
    pairs = np.stack(pairs)
    batch = np.empty((2 * P, h, w), np.uint8)
    for i in range(0, 2 * P, 2):
        src = (i // 2) % 8
        batch[i], batch[i + 1] = np.roll(pairs[2 * src], i // 16, axis=0), np.roll(pairs[2 * src + 1], i // 16, axis=0)
    dev = torch.from_numpy(batch).cuda()
    ex = snakeorb.ORBExtractor(nf, SCALE, NLEVELS, TH, THMIN, max_batch=2 * P)
    ctx = snakeorb.Context()
    ex.set_stream(st.cuda_stream)
    ctx.set_stream(st.cuda_stream)
    scales, _ = ex.getPyramid()
    sp = snakeorb.stereo_params(47.9, True, scales)
    cap = ex.cap
    rp = torch.empty((P, cap), dtype=torch.float32, device="cuda")
    dp = torch.empty((P, cap), dtype=torch.float32, device="cuda")
    nm = torch.empty((P,), dtype=torch.int32, device="cuda")
    ex.enqueue_ptr(dev.data_ptr(), 2 * P, h, w, w, w * h, True)

    def timed(fn):
        for _ in range(3):
            fn()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record()
        for _ in range(args.steps):
            fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / args.steps

    ms_stereo = timed(lambda: ctx.stereo_orb_device(ex, P, sp, rp.data_ptr(), dp.data_ptr(), nm.data_ptr()))
    # pair evaluations of the row search: counted by the oracle on the 8 distinct pairs
    orc = oracle_py.Extractor(nf, SCALE, NLEVELS, TH, THMIN, 1)
    ops = []
    for i in range(8):
        kl, dl = orc.detect(pairs[2 * i])
        kr, dr = orc.detect(pairs[2 * i + 1])
        ops.append(oracle_py.stereo_match(oracle_py.stereo_params(47.9, True, scales), oracle_py.kp32_to_kp64(kl), dl,
                                          oracle_py.kp32_to_kp64(kr), dr)[3])
    ops_pair = float(np.mean(ops))
    peak = ctx.hamming_peak()
    matchers = {"popc_peak_pair_ops_per_s": peak,
                "stereo_row_search": {"pairs": P, "ms": ms_stereo, "pair_ops_per_pair": ops_pair,
                                      "pair_ops_per_s": ops_pair * P / (ms_stereo * 1e-3),
                                      "popc_frac": ops_pair * P / (ms_stereo * 1e-3) / peak}}
    # kNN-2 of every left image's descriptors against its right image's (BruteForceMatcher::matchKnn2, batched)
    kps_ptr, desc_ptr, cnt_ptr, stride = ex.device_results()
    cnt = np.zeros(2 * P, np.int32)
    torch.cuda.synchronize()
    C.CDLL("libcudart.so").cudaMemcpy(C.c_void_p(cnt.ctypes.data), C.c_void_p(cnt_ptr), C.c_size_t(8 * P), 2)
    n_dev, m_dev = torch.from_numpy(cnt[0::2].copy()).cuda(), torch.from_numpy(cnt[1::2].copy()).cuda()
    out = torch.empty((P, 2 * stride, 4), dtype=torch.int32, device="cuda")
    ms_knn = timed(lambda: ctx.knn2_batch_device(desc_ptr, n_dev.data_ptr(), 2 * stride, desc_ptr + 32 * stride, m_dev.data_ptr(),
                                                 2 * stride, P, out.data_ptr()))
    knn_ops = float((cnt[0::2].astype(np.int64) * cnt[1::2]).sum())
    matchers["knn2_batch"] = {"problems": P, "ms": ms_knn, "pair_ops": knn_ops, "pair_ops_per_s": knn_ops / (ms_knn * 1e-3),
                              "popc_frac": knn_ops / (ms_knn * 1e-3) / peak}
    # relocalisation form of C4: matchKnn2 1000 x 1000 + filterMatches(60, 0.8), one call, host arrays in and out
    kq, dq = orc.detect(pairs[0])
    kt, dt = orc.detect(pairs[2])
    dq, dt = dq[:1000], dt[:1000]
    for _ in range(5):
        ctx.matchKnn2(dq, dt)
    ts = []
    for _ in range(50):
        t0 = time.perf_counter()
        k2 = ctx.matchKnn2(dq, dt)
        snakeorb.Context.filterMatches(k2, 60, 0.8)
        ts.append(time.perf_counter() - t0)
    t0 = time.perf_counter()
    for _ in range(3):
        oracle_py.filter_matches(oracle_py.match_knn2(dq, dt), 60, 0.8)
    matchers["knn2_1000x1000_single_call"] = {"ms_median": float(np.median(ts)) * 1e3, "cpu_port_ms_1_thread": (time.perf_counter() - t0) / 3 * 1e3}
    print(json.dumps({"hbm_peak_gbs": hbm, "extraction": rows, "matchers": matchers}))


if __name__ == "__main__":
    main()
