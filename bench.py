#!/usr/bin/env python
"""Benchmark of the ORB front-end hot path on B200 (contract: one JSON line on stdout from rank 0).

Workload (BASELINE.json configs[1], "C2"): EuRoC-shaped stereo pairs 752x480, ORB extraction on both images
(1200 keypoints, 8 levels, scale 1.2, FAST 20/7) + stereo row-search Hamming matching. One step = one batch of
`--pairs` synthetic stereo pairs through libsnakeorb.so's C ABI:

  value : pairs/s with the batch already resident in HBM (snk_orb_enqueue on device memory + snk_stereo_match_orb_device),
          timed with CUDA events on the launching stream, max over ranks
  e2e   : same metric with HOST (pinned) buffers: H2D of the images, all kernels, D2H of keypoints, descriptors,
          right_points, depth and match counts inside the timed region
  roofline     : dominant extraction kernel, algorithmic bytes / its event-timed duration
  cpu_baseline : the CPU oracle (port of the reference's CPU ORBExtractor + StereoMatching) on the host cores

`--impl reference` times the reference's CPU path (the oracle port; the reference's own ORBExtractor.cpp needs OpenCV
C++ and cannot be compiled here) on all host threads, same config/metric.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "snake-slam_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

W, H, NFEAT, NLEVELS, SCALE, TH, THMIN = 752, 480, 1200, 8, 1.2, 20, 7
BF = 47.9  # EuRoC-like fx*baseline: max disparity bf/2
METRIC = "frames/s (EuRoC stereo pair 752x480: ORB extract x2, 1200 kp, 8 levels + stereo row-search Hamming match)"


def level_sizes():
    out = []
    sc = 1.0
    for l in range(NLEVELS):
        if l:
            sc = float(np.float32(SCALE)) * sc
        s = np.float32(1.0 / sc)
        r = lambda v: int(np.float32(np.float32(v) * s) + np.float32(0.5))
        out.append((r(H), r(W)) if l else (H, W))
    return out


def algorithmic_bytes(n_kp_img, n_cand_img):
    """per-image algorithmic bytes of each extraction stage."""
    S = [h * w for h, w in level_sizes()]
    sig = sum(S)
    return {
        "pyramid": (sig - S[-1]) + (sig - S[0]),
        "fast": sig + 8 * n_cand_img,
        "blur": 2 * sig,
        "describe": 2386 * n_kp_img,
    }


def synth_pairs(n_distinct, seed0=10000):
    from snakeorb import synthetic as S

    imgs = []
    for i in range(n_distinct):
        l = S.image(seed0 + i, W, H)
        imgs += [l, S.stereo_right(l, seed0 + 5000 + i, 4.0, BF / 2 - 2.0)]
    return np.stack(imgs)


def fill_batch(distinct, n_imgs):
    """Tile the distinct pairs to the batch size with a per-copy circular row shift, so every image of the batch has
    different bytes (no accidental cache reuse) while generation stays cheap."""
    out = np.empty((n_imgs, H, W), np.uint8)
    nd = len(distinct)
    for i in range(0, n_imgs, 2):
        src = (i // 2) % (nd // 2)
        shift = (i // 2) // (nd // 2) * 7
        out[i] = np.roll(distinct[2 * src], shift, axis=0)
        out[i + 1] = np.roll(distinct[2 * src + 1], shift, axis=0)
    return out


class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows = []
        self.proc = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        rows = [r for (t, r) in self.rows if t0 <= t <= t1 + 0.2] or [r for (_, r) in self.rows]
        for r in rows:
            f = [x.strip() for x in r.split(",")]
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def cpu_pairs_per_s(distinct, threads, budget_s, stereo_params_fn):
    """The CPU oracle (oracle/orb_ref.cpp + matchers_ref.cpp) on `threads` host threads, one extractor per thread
    (the reference's class is not thread-safe), frames statically partitioned. Returns (pairs/s, pairs timed)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle_py as O

    O.build()
    n_pairs_avail = len(distinct) // 2
    scales, _ = O.scale_pyramid(NLEVELS, SCALE, NFEAT)
    sp = O.stereo_params(BF, True, scales)
    exs = [O.Extractor(NFEAT, SCALE, NLEVELS, TH, THMIN, 1) for _ in range(threads)]

    def work(tid, lo, hi, out):
        kp = 0
        for p in range(lo, hi):
            q = p % n_pairs_avail
            kl, dl = exs[tid].detect(distinct[2 * q])
            kr, dr = exs[tid].detect(distinct[2 * q + 1])
            O.stereo_match(sp, O.kp32_to_kp64(kl), dl, O.kp32_to_kp64(kr), dr)
            kp += len(kl) + len(kr)
        out[tid] = kp

    def run(n):
        out = [0] * threads
        ts = []
        t0 = time.perf_counter()
        for t in range(threads):
            lo, hi = n * t // threads, n * (t + 1) // threads
            th = threading.Thread(target=work, args=(t, lo, hi, out))
            th.start()
            ts.append(th)
        for th in ts:
            th.join()
        return time.perf_counter() - t0, sum(out)

    run(threads)  # warm-up: allocations, page-in
    dt, _ = run(2 * threads)
    n = int(max(2 * threads, min(4096, budget_s / (dt / (2 * threads)))))
    dt, kp = run(n)
    return n / dt, n, kp / dt


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--pairs", type=int, default=256, help="stereo pairs per step per GPU")
    ap.add_argument("--impl", default="snakeorb", choices=["snakeorb", "reference"])
    ap.add_argument("--cpu-budget", type=float, default=12.0, help="seconds of CPU work for cpu_baseline")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-next", action="store_true", help="skip the side measurements of the rows after the extractor")
    ap.add_argument("--chunks", type=int, default=2, help="e2e path: chunks per step (sweep on B200, 55 GB/s link, chunks/lanes: 1/2 64.3k, 2/3 73.0k, 2/6 72.9k, 4/4 66.4k, 4/6 66.8k, 8/8 62.3k pairs/s)")
    ap.add_argument("--device-lanes", type=int, default=1, help="device-resident path: the step's batch cut into this many parts, each on its own handle and stream, all in flight together (measured: 1: 3.33 ms/step, 2: 4.02, 4: 4.03, 8: 4.07 - kernels with different shared-memory carve-outs do not share an SM, parts that start together only get in each other's way)")
    ap.add_argument("--depth", type=int, default=2, help="e2e path: steps in flight (pinned result sets); the host waits for step s-depth+1")
    ap.add_argument("--lanes", type=int, default=3, help="e2e path: streams/handles the stream of chunks goes round robin over")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    config = {"workload": "C2 EuRoC stereo 752x480, 1200 kp/image, 8 levels, scale 1.2, FAST 20/7, stereo row search "
                          f"(bf={BF}, relaxed)", "pairs_per_step_per_gpu": args.pairs, "images_per_step_per_gpu": 2 * args.pairs,
              "parallelism": f"frame-sharded x{world}, no collective",
              "l2": (f"inputs {2 * args.pairs * W * H / 1e6:.0f} MB per step > 126 MB L2 (no flush needed)"
                     if 2 * args.pairs * W * H > 126e6 else
                     f"inputs {2 * args.pairs * W * H / 1e6:.0f} MB per step fit the 126 MB L2 and are NOT flushed: a reduced "
                     "--pairs run is a smoke run, not a bench value")}

    if args.impl == "reference":
        if rank != 0:
            return 0
        distinct = synth_pairs(8)
        K, Wm = args.steps, args.warmup
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import oracle_py as O

        O.build()
        scales, _ = O.scale_pyramid(NLEVELS, SCALE, NFEAT)
        sp = O.stereo_params(BF, True, scales)
        threads = cores
        exs = [O.Extractor(NFEAT, SCALE, NLEVELS, TH, THMIN, 1) for _ in range(threads)]
        per_step = 2 * threads  # bounded sample of the workload per step

        def step():
            def work(t):
                for p in range(2):
                    q = (t * 2 + p) % (len(distinct) // 2)
                    kl, dl = exs[t].detect(distinct[2 * q])
                    kr, dr = exs[t].detect(distinct[2 * q + 1])
                    O.stereo_match(sp, O.kp32_to_kp64(kl), dl, O.kp32_to_kp64(kr), dr)

            ts = [threading.Thread(target=work, args=(t,)) for t in range(threads)]
            [t.start() for t in ts]
            [t.join() for t in ts]

        for _ in range(Wm):
            step()
        t0 = time.perf_counter()
        for _ in range(K):
            step()
        dt = time.perf_counter() - t0
        v = per_step * K / dt
        line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "frames/s", "n_gpus": args.gpus, "steps": K,
                "warmup": Wm, "ms_per_step": dt / K * 1e3, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "u8", "data": "synthetic", "config": dict(config, pairs_per_step=per_step),
                "cpu_baseline": {"value": v, "unit": "frames/s", "cores": threads, "kind": "port",
                                 "sample": f"{per_step} stereo pairs per step x {K} steps, one extractor per thread"},
                "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        print(json.dumps(line))
        return 0

    import torch
    import torch.distributed as dist

    import snakeorb

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (libsnakeorb has no CPU fallback)")
    torch.cuda.set_device(local_rank)
    # keep this rank's host threads and its pinned buffers (first touch) on the NUMA node of its GPU: with several ranks
    # per box the host-to-device copies otherwise cross the socket interconnect
    numa = "unset"
    try:
        import pynvml

        pynvml.nvmlInit()
        pynvml.nvmlDeviceSetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(local_rank))
        numa = "cpu affinity of GPU %d (%d cpus)" % (local_rank, len(os.sched_getaffinity(0)))
    except Exception as exc:  # noqa: BLE001
        numa = "unset (%s)" % type(exc).__name__
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    P, NI = args.pairs, 2 * args.pairs
    distinct = synth_pairs(16, seed0=10000 + 100 * rank)
    batch = fill_batch(distinct, NI)
    host = torch.empty((NI, H, W), dtype=torch.uint8, pin_memory=True)
    host.numpy()[:] = batch
    dev = host.cuda(non_blocking=False)

    ex = snakeorb.ORBExtractor(NFEAT, SCALE, NLEVELS, TH, THMIN, max_batch=NI, device=local_rank)
    ctx = snakeorb.Context(device=local_rank)
    # a dedicated (non-default) torch stream: handle 0 means "the library's own stream" in the C ABI, and
    # torch.cuda.Event only sees the stream it is recorded on
    tstream = torch.cuda.Stream()
    torch.cuda.set_stream(tstream)
    stream = tstream.cuda_stream
    assert stream != 0
    ex.set_stream(stream)
    ctx.set_stream(stream)
    scales, _ = ex.getPyramid()
    sp = snakeorb.stereo_params(BF, True, scales)
    cap = ex.cap
    rp = torch.empty((P, cap), dtype=torch.float32, device="cuda")
    dp = torch.empty((P, cap), dtype=torch.float32, device="cuda")
    nm = torch.empty((P,), dtype=torch.int32, device="cuda")

    # the whole batch on one handle (also the path of the per-stage timing pass), or cut into DL parts on DL handles and
    # streams that run concurrently: the thin kernels of one part (quadtree, stereo, the small pyramid levels) then
    # share the SMs with the wide ones of another
    DL = args.device_lanes if (args.device_lanes > 1 and P % args.device_lanes == 0) else 1
    dev_lanes = []
    for _ in range(DL if DL > 1 else 0):
        st = torch.cuda.Stream()
        exk = snakeorb.ORBExtractor(NFEAT, SCALE, NLEVELS, TH, THMIN, max_batch=NI // DL, device=local_rank)
        cxk = snakeorb.Context(device=local_rank)
        exk.set_stream(st.cuda_stream)
        cxk.set_stream(st.cuda_stream)
        dev_lanes.append((st, exk, cxk))

    def step_device(single=False):
        if DL == 1 or single:
            ex.enqueue_ptr(dev.data_ptr(), NI, H, W, W, W * H, True)
            ctx.stereo_orb_device(ex, P, sp, rp.data_ptr(), dp.data_ptr(), nm.data_ptr())
            return
        ci, cp = NI // DL, P // DL
        for k, (st, exk, cxk) in enumerate(dev_lanes):
            st.wait_stream(tstream)
            exk.enqueue_ptr(dev.data_ptr() + k * ci * W * H, ci, H, W, W, W * H, True)
            cxk.stereo_orb_device(exk, cp, sp, rp.data_ptr() + 4 * k * cp * cap, dp.data_ptr() + 4 * k * cp * cap, nm.data_ptr() + 4 * k * cp)
        for st, _, _ in dev_lanes:
            tstream.wait_stream(st)

    # ---- end-to-end path: host (pinned) buffers in and out, the batch cut into chunks that alternate between two
    # extractor handles on two streams so that the H2D copy of chunk k+1 and the D2H of chunk k-1 overlap chunk k's
    # kernels (every byte still crosses PCIe inside the timed region)
    import ctypes as C

    NCH = args.chunks if (NI % (2 * args.chunks) == 0) else 1
    CI, CP = NI // NCH, P // NCH
    DEPTH = max(2, args.depth)
    # DEPTH sets of pinned result buffers: step s+1 is enqueued while the results of step s are still arriving (a streaming
    # consumer reads set s % 2 while set (s + 1) % 2 is being filled)
    h_kps = torch.empty((DEPTH, NI, cap, 24), dtype=torch.uint8, pin_memory=True)
    h_desc = torch.empty((DEPTH, NI, cap, 4), dtype=torch.int64, pin_memory=True)
    h_cnt = torch.empty((DEPTH, NI), dtype=torch.int32, pin_memory=True)
    h_rp = torch.empty((DEPTH, P, cap), dtype=torch.float32, pin_memory=True)
    h_dp = torch.empty((DEPTH, P, cap), dtype=torch.float32, pin_memory=True)
    h_nm = torch.empty((DEPTH, P), dtype=torch.int32, pin_memory=True)
    e2e_lanes = []
    if not args.no_e2e:
        for _ in range(args.lanes):
            st = torch.cuda.Stream()
            exc = snakeorb.ORBExtractor(NFEAT, SCALE, NLEVELS, TH, THMIN, max_batch=CI, device=local_rank)
            cxc = snakeorb.Context(device=local_rank)
            exc.set_stream(st.cuda_stream)
            cxc.set_stream(st.cuda_stream)
            e2e_lanes.append((st, exc, cxc, torch.empty((CP, cap), dtype=torch.float32, device="cuda"),
                              torch.empty((CP, cap), dtype=torch.float32, device="cuda"),
                              torch.empty((CP,), dtype=torch.int32, device="cuda")))

    def e2e_enqueue(k, par, lane):
        st, exc, cxc, rpc, dpc, nmc = e2e_lanes[lane]
        exc.enqueue_ptr(host.data_ptr() + k * CI * W * H, CI, H, W, W, W * H, False)
        cxc.stereo_orb_device(exc, CP, sp, rpc.data_ptr(), dpc.data_ptr(), nmc.data_ptr())
        with torch.cuda.stream(st):
            h_rp[par, k * CP:(k + 1) * CP].copy_(rpc, non_blocking=True)
            h_dp[par, k * CP:(k + 1) * CP].copy_(dpc, non_blocking=True)
            h_nm[par, k * CP:(k + 1) * CP].copy_(nmc, non_blocking=True)

    def e2e_fetch(k, par, lane):
        _, exc, _, _, _, _ = e2e_lanes[lane]
        snakeorb._check(snakeorb.lib().snk_orb_fetch_async(exc.h, C.c_void_p(h_kps[par].data_ptr() + k * CI * cap * 24),
                                                           C.c_void_p(h_desc[par].data_ptr() + k * CI * cap * 32), cap,
                                                           C.c_void_p(h_cnt[par].data_ptr() + k * CI * 4)))

    e2e_done = [[] for _ in range(DEPTH)]  # per result set: one event per lane, recorded behind the set's last copy

    def step_e2e(s):
        # every chunk: H2D, kernels, D2H on its lane's stream, no host wait in between. The host then waits for the
        # PREVIOUS step's results (set (s - 1) % 2), so the copies of consecutive steps overlap like a stream of frames.
        par = s % DEPTH
        for k in range(NCH):
            # lanes are taken round robin over the stream of chunks: with more lanes than chunks a chunk's upload
            # does not queue behind the same lane's previous step
            lane = (s * NCH + k) % len(e2e_lanes)
            e2e_enqueue(k, par, lane)
            e2e_fetch(k, par, lane)
        e2e_done[par] = []
        for st, *_ in e2e_lanes:
            ev = torch.cuda.Event()
            ev.record(st)
            e2e_done[par].append(ev)
        for ev in e2e_done[(s + 1) % DEPTH]:  # the oldest step still in flight
            ev.synchronize()

    # ---- device-resident timing --------------------------------------------------------------------------------
    for _ in range(max(args.warmup, 3)):
        step_device()
    barrier()
    sampler = ClockSampler(local_rank)
    time.sleep(0.25)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t_wall0 = time.time()
    e0.record()
    for _ in range(args.steps):
        step_device()
    e1.record()
    barrier()
    t_wall1 = time.time()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop(t_wall0, t_wall1)
    # second pass over the same K steps with per-stage events (cudaEvents on the launching stream between the
    # kernels; the library serialises its side stream while they are on, so each stage's time is its own)
    ex.set_profiling(True)
    ex.stage_ms(reset=True)
    p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    p0.record()
    for _ in range(args.steps):
        step_device(single=True)
    p1.record()
    barrier()
    ms_profiled = p0.elapsed_time(p1)
    stage_ms, stage_launches = ex.stage_ms(reset=True)
    ex.set_profiling(False)
    kps_ptr, desc_ptr, cnt_ptr, stride = ex.device_results()
    cudart = C.CDLL("libcudart.so")  # noqa: F841
    cnt_host = np.zeros(NI, np.int32)
    cudart.cudaMemcpy(C.c_void_p(cnt_host.ctypes.data), C.c_void_p(cnt_ptr), C.c_size_t(4 * NI), 2)
    n_kp_step = int(cnt_host.sum())
    n_match_step = int(nm.sum().item())

    # ---- side measurement, not part of `value`: MiniBow2::transform of every image of the step on the
    # extractor's device results, ORBvoc-shaped synthetic vocabulary (k=10, L=6), levelsup 4 as Frame::computeBoW uses
    next_rows = None
    if rank == 0 and not args.no_next:
        from snakeorb import synthetic as S

        arr = S.vocabulary_arrays_bfs(7, 10, 6)
        voc = snakeorb.Vocabulary(arrays=arr, device=local_rank)
        b_ids = torch.empty((NI, stride), dtype=torch.int32, device="cuda")
        b_val = torch.empty((NI, stride), dtype=torch.float32, device="cuda")
        f_nod = torch.empty((NI, stride), dtype=torch.int32, device="cuda")
        f_off = torch.empty((NI, stride + 1), dtype=torch.int32, device="cuda")
        f_idx = torch.empty((NI, stride), dtype=torch.int32, device="cuda")
        o_cnt = torch.empty((NI, 2), dtype=torch.int32, device="cuda")

        def step_bow():
            ctx.bow_transform_batch_device(voc, desc_ptr, cnt_ptr, NI, stride, 4, b_ids.data_ptr(), b_val.data_ptr(),
                                           f_nod.data_ptr(), f_off.data_ptr(), f_idx.data_ptr(), o_cnt.data_ptr())

        for _ in range(3):
            step_bow()
        b0, b1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        b0.record()
        for _ in range(args.steps):
            step_bow()
        b1.record()
        torch.cuda.synchronize()
        bow_ms = b0.elapsed_time(b1) / args.steps
        next_rows = {"bow_transform": {"frames_per_s": NI / (bow_ms * 1e-3), "ms_per_step": bow_ms, "frames_per_step": NI,
                                       "vocabulary": "synthetic, ORBvoc shape k=10 L=6 (1 111 111 nodes), levelsup 4",
                                       "words_per_frame": float(o_cnt[:, 0].float().mean().item())}}
        if not args.no_cpu:
            sys.path.insert(0, os.path.join(ROOT, "oracle"))
            import oracle_py  # cpu_baseline leg only: the checker timed as the CPU port

            vo = oracle_py.Vocabulary(arrays=arr)
            dh = np.zeros((4, stride, 4), np.uint64)
            cudart.cudaMemcpy(C.c_void_p(dh.ctypes.data), C.c_void_p(desc_ptr), C.c_size_t(dh.nbytes), 2)
            t0 = time.time()
            reps = 0
            while time.time() - t0 < 1.0:
                for f in range(4):
                    vo.transform(dh[f, :cnt_host[f]], 4)
                reps += 4
            next_rows["bow_transform"]["cpu_port_frames_per_s_1_thread"] = reps / (time.time() - t0)
        del voc
        # C3: SearchByProjection2 of 20 000 local-map points into one KITTI-shaped frame (host structs in, host results
        # out, as Tracking calls it); C4: BoW-candidate matching of two TUM-shaped keyframes. Side measurements.
        try:
            sys.path.insert(0, os.path.join(ROOT, "tests"))
            import matchdata as MD

            rngn = np.random.default_rng(5)
            exk = snakeorb.ORBExtractor(2000, SCALE, NLEVELS, TH, THMIN, device=local_rank)
            kk, dk = exk.Detect(S.image(3000, 1241, 376))
            sc_k, _ = exk.getPyramid()
            Kk = (718.856, 718.856, 607.19, 185.22)
            fr = MD.make_frame(kk, dk, 1241, 376, rngn, S.quat_from_axis_angle(rngn.normal(size=3), 0.004), rngn.uniform(-0.05, 0.05, 3))
            mp = S.map_points_from_keypoints(fr.kp, fr.desc, Kk, rngn, 20000, sc_k)
            pts = MD.fine_points(mp, rngn)
            camk, pyrk = MD.camera(*Kk, 386.1), snakeorb.scale_pyramid_c(1.2, sc_k)
            ctxn = snakeorb.Context(device=local_rank)
            for _ in range(3):
                mres = ctxn.SearchByProjection2(fr.view, pts, camk, pyrk, 4.0, 0.8)
            t0 = time.perf_counter()
            for _ in range(20):
                ctxn.SearchByProjection2(fr.view, pts, camk, pyrk, 4.0, 0.8)
            gpu_ms = (time.perf_counter() - t0) / 20 * 1e3
            next_rows["search_projection_c3"] = {"ms_per_call": gpu_ms, "map_points": 20000, "frame_keypoints": int(fr.view.n),
                                                 "matches": int(mres[0]), "path": "host structs in, host results out"}
            d1 = dk[:1000]
            d2 = np.concatenate([S.flip_bits(d1[rngn.permutation(1000)[:600]], rngn, 12), dk[1000:1400]])
            fv1, fv2 = MD.featvec(d1, rngn), MD.featvec(d2, rngn)
            v1 = np.ones(len(d1), np.uint8)
            v2 = np.ones(len(d2), np.uint8)
            for _ in range(3):
                mb = ctxn.MatchBoW(d1, d2, fv1, fv2, v1, v2, 50, 0.75)
            t0 = time.perf_counter()
            for _ in range(50):
                ctxn.MatchBoW(d1, d2, fv1, fv2, v1, v2, 50, 0.75)
            next_rows["match_bow_c4"] = {"ms_per_call": (time.perf_counter() - t0) / 50 * 1e3, "features": [len(d1), len(d2)],
                                         "matches": int(mb[0]), "path": "host arrays in, host results out"}
            if not args.no_cpu:
                t0 = time.perf_counter()
                for _ in range(3):
                    oracle_py.search_projection_fine(fr.view, pts, camk, pyrk, 4.0, 0.8)
                next_rows["search_projection_c3"]["cpu_port_ms_per_call_1_thread"] = (time.perf_counter() - t0) / 3 * 1e3
                t0 = time.perf_counter()
                for _ in range(10):
                    oracle_py.match_bow(d1, d2, fv1, fv2, v1, v2, 50, 0.75)
                next_rows["match_bow_c4"]["cpu_port_ms_per_call_1_thread"] = (time.perf_counter() - t0) / 10 * 1e3
            # C1 latency: one 752x480 frame, host image in, host keypoints + descriptors out, stream-synchronous call
            # (median of 200 after 20 warm-ups) — what a live camera loop sees
            ex1 = snakeorb.ORBExtractor(1000, SCALE, NLEVELS, TH, THMIN, device=local_rank)
            img1 = S.image(1000, W, H)
            lat = []
            for i in range(220):
                t0 = time.perf_counter()
                k1, _ = ex1.Detect(img1)
                lat.append((time.perf_counter() - t0) * 1e3)
            lat = sorted(lat[20:])
            next_rows["single_frame_latency_c1"] = {"median_ms": lat[len(lat) // 2], "p95_ms": lat[int(len(lat) * 0.95)],
                                                    "keypoints": int(len(k1)), "path": "host image in, host results out, synchronous"}
            if not args.no_cpu:
                oe = oracle_py.Extractor(1000, SCALE, NLEVELS, TH, THMIN, 2)
                t0 = time.perf_counter()
                for _ in range(10):
                    oe.detect(img1)
                next_rows["single_frame_latency_c1"]["cpu_port_ms_2_threads"] = (time.perf_counter() - t0) / 10 * 1e3
        except Exception as exc:  # noqa: BLE001 — side measurements must never take the bench line down
            next_rows["side_measurement_error"] = repr(exc)[:200]

    # ---- end-to-end timing (host buffers) ---------------------------------------------------------------------------
    e2e = None
    if not args.no_e2e:
        n_warm = DEPTH * max(1, -(-len(e2e_lanes) // NCH))  # every lane and every result set used at least once
        for i in range(n_warm):
            step_e2e(i)
        barrier()
        tw0 = time.perf_counter()
        for i in range(n_warm, n_warm + args.steps):
            step_e2e(i)
        torch.cuda.synchronize()
        e2e = (time.perf_counter() - tw0) * 1e3  # host wall clock: the last fetch returns when the results are in host memory
        barrier()
        # the end-to-end path must deliver the same results as the device-resident one
        for par in range(DEPTH):
            assert int(h_cnt[par].sum()) == n_kp_step and int(h_nm[par].sum()) == n_match_step, "e2e results differ from the device path"

    t = torch.tensor([ms, e2e if e2e is not None else 0.0], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max, e2e_max = float(t[0]), float(t[1])
    value = world * P * args.steps / (ms_max * 1e-3)
    kp_per_s = world * n_kp_step * args.steps / (ms_max * 1e-3)

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        peak_src = "measured (MEASURED_PEAKS.json)" if "hbm_gbs" in peaks else "fallback (B200_PROFILING.md)"
        # dominant extraction kernel by event time inside the timed region
        names = {1: "pyramid", 2: "fast", 3: "distribute", 4: "blur", 5: "describe"}
        # raw FAST candidates per image: not fetched in the timed path; use the per-image mean of a probe image
        n_cand_img = sum(len(ex.candidates(l, 0)) for l in range(NLEVELS))
        ab = algorithmic_bytes(n_kp_step / NI, n_cand_img)
        ab["distribute"] = 8 * n_cand_img + 4 * n_kp_step / NI  # candidate records in, selected records out
        stage = {names[i]: float(stage_ms[i]) / args.steps for i in names}
        dom = max(stage, key=stage.get)
        launches = {names[i]: int(stage_launches[i]) // args.steps for i in names}
        bytes_launch = ab[dom] * NI / max(launches[dom], 1)
        dur = stage[dom] * 1e-3 / max(launches[dom], 1)
        achieved = bytes_launch / dur / 1e9
        e2e_bytes = sum(algorithmic_bytes(n_kp_step / NI, n_cand_img)[k] for k in ("pyramid", "fast", "blur", "describe"))
        extraction_ms = sum(stage.values())
        kname = {"fast": "k_fast_group", "pyramid": "k_resize"}.get(dom, "k_" + dom)
        # DRAM traffic of that kernel: dram__bytes_read + dram__bytes_write of the committed `ncu --set full` capture
        # (profiles/*_ncu_summary.csv, a 128-image launch of this same command), scaled to this launch's image count
        traffic, traffic_src, limiter = None, None, None
        try:
            import csv
            import glob

            newest = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_ncu_summary.csv")))[-1]
            rows_ = [r for r in csv.reader(open(newest)) if r and r[0] == kname]
            if rows_:
                per_img = (sum(float(r[2]) + float(r[3]) for r in rows_) * 1e6) / 128.0  # Mbyte columns, 128 images per capture
                traffic = per_img * NI / max(launches[dom], 1)
                traffic_src = os.path.relpath(newest, ROOT) + " (128-image capture, scaled per image)"
                # what the kernel is bound by when it is not the memory system: share of the issue slots it uses (same capture)
                limiter = {"what": "instruction issue (integer byte-SIMD, shared-memory lists), not DRAM",
                           "issue_active_pct": sum(float(r[6]) for r in rows_) / len(rows_),
                           "dram_pct_of_peak": sum(float(r[4]) for r in rows_) / len(rows_),
                           "warp_instructions_per_image": sum(float(r[9]) for r in rows_) / 128.0}
        except Exception:  # noqa: BLE001
            traffic = None
        roofline = {"bound": "hbm", "kernel": kname, "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                    "frac": achieved / hbm_peak, "traffic": traffic, "traffic_source": traffic_src, "limiter": limiter, "peak_source": peak_src,
                    "algorithmic_bytes_per_image": ab[dom], "images_per_launch": NI / max(launches[dom], 1),
                    "stage_ms_per_step": stage, "stage_timing_pass_ms_per_step": ms_profiled / args.steps,
                    "stage_timing": "separate pass of the same K steps with cudaEvents between the kernels (blur serialised "
                                    "on the main stream); `value` comes from the pass with the blur on a side stream",
                    "extraction_e2e": {"bytes_per_image": e2e_bytes, "achieved_gbs": e2e_bytes * NI / (extraction_ms * 1e-3) / 1e9,
                                       "frac": e2e_bytes * NI / (extraction_ms * 1e-3) / 1e9 / hbm_peak}}
        line = {"metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": ms_max / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic", "config": config,
                "keypoints_per_s": kp_per_s, "images_per_s": 2 * value,
                "keypoints_per_step": n_kp_step, "stereo_matches_per_step": n_match_step,
                "raw_fast_candidates_per_image": n_cand_img, "clocks": clocks,
                "gpu_launches": int(sum(int(x) for x in stage_launches[1:6])) + args.steps,  # + stereo kernel per step
                "roofline": roofline}
        if e2e is not None:
            h2d = NI * W * H
            d2h = NI * cap * (24 + 32) + NI * 4 + P * cap * 8 + P * 4 + 4
            line["e2e"] = {"value": world * P * args.steps / (e2e_max * 1e-3), "unit": "frames/s", "h2d_bytes_per_step": h2d,
                           "d2h_bytes_per_step": d2h, "ms_per_step": e2e_max / args.steps,
                           "host_affinity": numa, "pipeline": f"{NCH} chunks per step, the stream of chunks round robin over {len(e2e_lanes)} streams/handles (copy-compute overlap); the host waits for step s-1 while step s is in flight (two pinned result sets)"}
        if next_rows is not None:
            line["next_rows"] = next_rows
        if not args.no_cpu:
            try:
                popc = ctx.hamming_peak()
                line["popc_peak_pair_ops_per_s"] = popc
            except Exception as exc:  # noqa: BLE001
                line["popc_peak_pair_ops_per_s"] = None
            v, n, kps = cpu_pairs_per_s(distinct, cores, args.cpu_budget, None)
            line["cpu_baseline"] = {"value": v, "unit": "frames/s", "cores": cores, "kind": "port",
                                    "sample": f"{n} stereo pairs of the same workload, one oracle extractor per thread, "
                                              f"{cores} threads", "keypoints_per_s": kps}
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
