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


def upload_probe(torch, dist, dev, world):
    """GB/s of 8 x 64 MB pinned-host -> device copies on GPU `dev` while every other rank does the same on its GPU
    (the ranks start together). Returns the list of all ranks' rates."""
    with torch.cuda.device(dev):
        src = torch.empty(64 << 20, dtype=torch.uint8, pin_memory=True)
        dst = torch.empty(64 << 20, dtype=torch.uint8, device=f"cuda:{dev}")
        dst.copy_(src, non_blocking=True)
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.all_reduce(torch.zeros(1))  # gloo: a common starting line
        t0 = time.perf_counter()
        for _ in range(8):
            dst.copy_(src, non_blocking=True)
        torch.cuda.synchronize(dev)
        gbs = 8 * (64 << 20) / (time.perf_counter() - t0) / 1e9
    if world == 1:
        return [gbs]
    allr = [torch.zeros(1, dtype=torch.float64) for _ in range(world)]
    dist.all_gather(allr, torch.tensor([gbs], dtype=torch.float64))
    return [float(x) for x in allr]


def pick_device(torch, dist, local_rank, world):
    """Which GPU this rank drives, and how fast the ranks can upload when all of them do (the end-to-end ceiling of
    the box). With as many visible GPUs as ranks (or one rank) the GPU is LOCAL_RANK. With more visible GPUs than ranks
    the box is probed: on this pool's 8-GPU boxes GPUs 0-3 share one host path (about 100 GB/s in total: 51 GB/s each for
    two of them, 24 GB/s each for four) and GPUs 4-7 another, faster one (tools/scratch/numa8.py, profiles/r2c_*), so the
    ranks try three placements - the low block (rank r -> GPU r), the high block (rank r -> GPU visible - world + r) and
    one that alternates between the two halves of the box - with concurrent uploads, and keep the one with the highest
    total rate (the chunk queue of the end-to-end path balances unequal links, so the sum is what counts).
    Returns (device index, probe record)."""
    visible = torch.cuda.device_count()
    if world == 1 or visible <= world:
        dev = local_rank % max(visible, 1)
        return dev, {"visible_gpus": visible, "picked": "rank order", "h2d_gbs_per_gpu": upload_probe(torch, dist, dev, world)}
    half = visible // 2
    placements = {"low block": local_rank, "high block": visible - world + local_rank,
                  "alternating halves": (local_rank % 2) * half + local_rank // 2}
    rates = {name: upload_probe(torch, dist, dev, world) for name, dev in placements.items()}
    pick = max(rates, key=lambda k: sum(rates[k]))
    if sum(rates[pick]) < 1.05 * sum(rates["low block"]):
        pick = "low block"
    return placements[pick], {"visible_gpus": visible, "picked": pick, "h2d_gbs_per_gpu": rates[pick],
                              "h2d_gbs_per_gpu_of_every_placement": rates}


class HostArray:
    """Page-locked host array. One rank: plain pinned memory. Several ranks: one array in /dev/shm shared by all of
    them (the sequence and its results exist once on the box, as they would for a real dataset), registered with the
    CUDA driver in every process so copies to and from it are asynchronous."""

    def __init__(self, torch, name, shape, np_dtype, world, rank, barrier):
        import ctypes as C

        self.nbytes = int(np.prod(shape)) * np.dtype(np_dtype).itemsize
        self.path, self.rank, self._reg = None, rank, False
        if world == 1:
            self._t = torch.empty(self.nbytes, dtype=torch.uint8, pin_memory=True)
            self.np = self._t.numpy().view(np_dtype).reshape(shape)
        else:
            self.path = f"/dev/shm/snkbench_{os.environ.get('MASTER_PORT', '0')}_{name}"
            if rank == 0:
                import shutil

                if shutil.disk_usage("/dev/shm").free < self.nbytes + (1 << 28):
                    raise SystemExit(f"/dev/shm has no room for {self.nbytes} bytes ({name})")
                np.memmap(self.path, dtype=np.uint8, mode="w+", shape=(self.nbytes,)).flush()
            barrier()
            self.np = np.memmap(self.path, dtype=np.uint8, mode="r+", shape=(self.nbytes,)).view(np_dtype).reshape(shape)
            self._cudart = C.CDLL("libcudart.so")
            rc = self._cudart.cudaHostRegister(C.c_void_p(self.np.ctypes.data), C.c_size_t(self.nbytes), 1)  # portable
            if rc != 0:
                raise SystemExit(f"cudaHostRegister({name}, {self.nbytes} bytes) failed with {rc}")
            self._reg = True
        self.ptr = self.np.ctypes.data

    def close(self, barrier):
        import ctypes as C

        if self._reg:
            self._cudart.cudaHostUnregister(C.c_void_p(self.ptr))
        if self.path:
            barrier()
            if self.rank == 0 and os.path.exists(self.path):
                os.unlink(self.path)


def c5_frame(distinct, f, out):
    """Pair f of the C5 sequence: one of the distinct synthetic pairs, circularly shifted by a frame-dependent amount in
    both axes (the same shift for left and right, so the stereo geometry is kept): every frame has different bytes."""
    nd = len(distinct) // 2
    src, rep = f % nd, f // nd
    for e in range(2):
        out[e] = np.roll(distinct[2 * src + e], (7 * rep, 3 * rep), axis=(0, 1))


def run_c5(args, env):
    """BASELINE.json configs[4]: the whole sequence, sharded by frame over the ranks through the shared chunk queue."""
    import ctypes as C
    import hashlib

    torch, dist, snakeorb = env["torch"], env["dist"], env["snakeorb"]
    rank, world, store, barrier, Lane, cap = env["rank"], env["world"], env["store"], env["barrier"], env["Lane"], env["cap"]
    ChunkQueue, local_rank, numa, config = env["ChunkQueue"], env["local_rank"], env["numa"], dict(env["config"])
    F, CPc = args.frames, args.c5_chunk
    seq = HostArray(torch, "seq", (F, 2, H, W), np.uint8, world, rank, barrier)
    o_kps = HostArray(torch, "kps", (2 * F, cap, 24), np.uint8, world, rank, barrier)
    o_desc = HostArray(torch, "desc", (2 * F, cap, 4), np.uint64, world, rank, barrier)
    o_cnt = HostArray(torch, "cnt", (2 * F,), np.int32, world, rank, barrier)
    o_rp = HostArray(torch, "rp", (F, cap), np.float32, world, rank, barrier)
    o_dp = HostArray(torch, "dp", (F, cap), np.float32, world, rank, barrier)
    o_nm = HostArray(torch, "nm", (F,), np.int32, world, rank, barrier)
    outs = (o_kps, o_desc, o_cnt, o_rp, o_dp, o_nm)
    distinct = synth_pairs(16, seed0=10000)
    for f in range(rank, F, world):  # every rank writes its share of the frames
        c5_frame(distinct, f, seq.np[f])
    barrier()
    DCH = max(CPc, 256)  # the device-resident pass has no upload to hide: larger chunks
    lanes = [Lane(DCH) for _ in range(args.lanes)]
    dev_seq = torch.empty((F, 2, H, W), dtype=torch.uint8, device="cuda")  # 2.7 GB: the sequence resident in HBM
    rc = C.CDLL("libcudart.so").cudaMemcpy(C.c_void_p(dev_seq.data_ptr()), C.c_void_p(seq.ptr), C.c_size_t(seq.nbytes), 1)
    assert rc == 0

    def one_pass(key, on_device, chunk, nlanes):
        for o in outs:
            if rank == 0:
                o.np.reshape(-1).view(np.uint8)[:] = 0xCD  # poison: a frame nobody processed shows in the digest
        barrier()
        q = ChunkQueue(F, chunk, store=store, key=key)
        base = dev_seq.data_ptr() if on_device else seq.ptr
        barrier()
        t0 = time.perf_counter()
        i = mine = 0
        while True:
            lanes[i % nlanes].wait()
            c = q.pull()
            if c is None:
                break
            lo, hi = q.frames(c)
            lanes[i % nlanes].run(base + lo * 2 * W * H, on_device, hi - lo, o_kps.ptr + 2 * lo * cap * 24, o_desc.ptr + 2 * lo * cap * 32,
                                      o_cnt.ptr + 2 * lo * 4, o_rp.ptr + lo * cap * 4, o_dp.ptr + lo * cap * 4, o_nm.ptr + lo * 4)
            i += 1
            mine += 1
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        barrier()
        return dt, mine

    def digest():
        """sha256 over the per-frame results in frame order: keypoints and descriptors of both images trimmed to their
        counts, right_points / depth of the left image's keypoints, the match count."""
        h = hashlib.sha256()
        cnt = o_cnt.np
        for f in range(F):
            for e in range(2):
                n = int(cnt[2 * f + e])
                h.update(o_kps.np[2 * f + e, :n].tobytes())
                h.update(o_desc.np[2 * f + e, :n].tobytes())
            n = int(cnt[2 * f])
            h.update(o_rp.np[f, :n].tobytes())
            h.update(o_dp.np[f, :n].tobytes())
            h.update(o_nm.np[f:f + 1].tobytes())
        return h.hexdigest(), int(cnt.sum()), int(o_nm.np.sum())

    for w in range(max(args.warmup, 1)):
        one_pass(f"c5_warm_{w}", False, CPc, len(lanes))
        one_pass(f"c5_warmd_{w}", True, DCH, len(lanes))
    res = {}
    for mode, on_device in (("host", False), ("device", True)):
        times, chunks = [], None
        for k in range(max(args.steps, 1)):
            dt, mine = one_pass(f"c5_{mode}_{k}", on_device, DCH if on_device else CPc, len(lanes))
            t = torch.tensor([dt], dtype=torch.float64, device="cuda")
            g = [torch.zeros(1, dtype=torch.int64, device="cuda") for _ in range(world)]
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                dist.all_gather(g, torch.tensor([mine], dtype=torch.int64, device="cuda"))
                chunks = [int(x) for x in g]
            else:
                chunks = [mine]
            times.append(float(t))
        res[mode] = (times, chunks, digest() if rank == 0 else None)
    for ln in lanes:
        ln.ex.check()
        ln.cx.check()
    if rank == 0:
        (th, ch, (dg_h, nkp, nmatch)), (td, cd, (dg_d, _, _)) = res["host"], res["device"]
        assert dg_h == dg_d, "host-fed and device-resident passes returned different results"
        K = len(th)
        config.update({"workload": f"C5 {F}-pair synthetic EuRoC-length stereo sequence (distinct frames), ORB extract x2 (1200 kp, 8 "
                                   f"levels, 1.2, FAST 20/7) + stereo row search per pair; frames sharded over {world} GPU(s) in "
                                   f"chunks of {CPc} pairs from a shared queue; results in host memory in frame order",
                       "pairs_in_sequence": F, "chunk_pairs": CPc, "chunk_pairs_device_resident": DCH,
                       "l2": f"{seq.nbytes / 1e9:.2f} GB of distinct input per pass (>> 126 MB L2)"})
        config.pop("pairs_per_step_per_gpu", None), config.pop("images_per_step_per_gpu", None)
        line = {"metric": METRIC, "value": F * K / sum(td), "unit": "frames/s", "n_gpus": world, "steps": K, "warmup": max(args.warmup, 1),
                "ms_per_step": sum(td) / K * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u8",
                "data": "synthetic", "config": config, "sequence_wall_s_device_resident": td, "sequence_wall_s_host_fed": th,
                "keypoints_per_sequence": nkp, "stereo_matches_per_sequence": nmatch, "results_sha256": dg_h,
                "chunks_per_rank_device": cd,
                "e2e": {"value": F * K / sum(th), "unit": "frames/s", "h2d_bytes_per_step": seq.nbytes,
                        "d2h_bytes_per_step": sum(o.nbytes for o in outs), "ms_per_step": sum(th) / K * 1e3,
                        "chunks_per_rank": ch, "host_affinity": numa, "device_choice": env["h2d_probe"],
                        "upload_ceiling_frames_per_s": sum(env["h2d_probe"]["h2d_gbs_per_gpu"]) * 1e9 / (2 * W * H),
                        "pipeline": f"{len(lanes)} streams/handles per rank, chunks of {CPc} pairs from a queue shared by the ranks"},
                # per chunk: 7 resize levels, FAST, quadtree, blur, describe, stereo
                "gpu_launches": 12 * ((F + CPc - 1) // CPc) * 2 * K}
        print(json.dumps(line))
    for o in (seq,) + outs:
        o.close(barrier)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


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
    ap.add_argument("--workload", default="c2", choices=["c2", "c5"],
                    help="c2 (default, the contract's line): one resident batch per step, weak scaling. c5: BASELINE.json "
                         "configs[4] as written - a 3680-pair sequence of distinct frames in pinned host memory, sharded by "
                         "frame over the ranks, results back in host memory in frame order, wall clock for the whole "
                         "sequence (strong scaling); --steps / --warmup count passes over the sequence")
    ap.add_argument("--frames", type=int, default=3680, help="c5: stereo pairs in the sequence")
    ap.add_argument("--c5-chunk", type=int, default=128, help="c5: stereo pairs per chunk of the shared queue")
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
                "vs_baseline": None, "dtype": "u8", "data": "synthetic", "config": config,  # the same keys as the CUDA arm
                "cpu_baseline": {"value": v, "unit": "frames/s", "cores": threads, "kind": "port",
                                 "sample": f"a bounded sample of the step: {per_step} of its {args.pairs} stereo pairs per step "
                                           f"x {K} steps, one extractor per thread"},
                "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        print(json.dumps(line))
        return 0

    import torch
    import torch.distributed as dist

    import snakeorb

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (libsnakeorb has no CPU fallback)")
    if world > 1:
        # gloo carries the control plane (device choice, chunk queue statistics), NCCL the timing barrier / reduction
        dist.init_process_group("cpu:gloo,cuda:nccl")
    device_index, h2d_probe = pick_device(torch, dist, local_rank, world)
    torch.cuda.set_device(device_index)
    # keep this rank's host threads and its pinned buffers (first touch) on the NUMA node of its GPU: with several ranks
    # per box the host-to-device copies otherwise cross the socket interconnect
    numa = "unset"
    try:
        import pynvml

        pynvml.nvmlInit()
        pynvml.nvmlDeviceSetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(device_index))
        numa = "cpu affinity of GPU %d (%d cpus)" % (device_index, len(os.sched_getaffinity(0)))
    except Exception as exc:  # noqa: BLE001
        numa = "unset (%s)" % type(exc).__name__
    store = None
    if world > 1:
        from torch.distributed.distributed_c10d import _get_default_store

        store = _get_default_store()
    local_rank = device_index  # everything below addresses the chosen device

    _bar = torch.zeros(1, device="cuda")

    def barrier():
        if world > 1:
            dist.all_reduce(_bar)  # NCCL, on this rank's device
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

    # ---- end-to-end path: host (pinned) buffers in and out. The step's batch is cut into chunks; a chunk runs on one of
    # `--lanes` extractor handles / streams (H2D, all kernels, D2H on that stream, no host wait in between), so the upload
    # of one chunk, the kernels of another and the download of a third overlap; every byte still crosses PCIe inside
    # the timed region. Which rank takes which chunk is decided by a queue shared by all ranks (snakeorb.sharding
    # .ChunkQueue): a rank asks for the next chunk when one of its lanes is free, so GPUs behind a slower host link take
    # fewer chunks instead of holding the whole job back.
    import ctypes as C

    from snakeorb.sharding import ChunkQueue

    cudart = C.CDLL("libcudart.so")
    NCH = args.chunks if (NI % (2 * args.chunks) == 0) else 1
    CI, CP = NI // NCH, P // NCH
    DEPTH = max(2, args.depth)
    NSLOT = DEPTH * NCH  # pinned result slots of this rank, reused round robin (a streaming consumer reads slot i while i+1.. fill)
    h_kps = torch.empty((NSLOT, CI, cap, 24), dtype=torch.uint8, pin_memory=True)
    h_desc = torch.empty((NSLOT, CI, cap, 4), dtype=torch.int64, pin_memory=True)
    h_cnt = torch.empty((NSLOT, CI), dtype=torch.int32, pin_memory=True)
    h_rp = torch.empty((NSLOT, CP, cap), dtype=torch.float32, pin_memory=True)
    h_dp = torch.empty((NSLOT, CP, cap), dtype=torch.float32, pin_memory=True)
    h_nm = torch.empty((NSLOT, CP), dtype=torch.int32, pin_memory=True)

    class Lane:
        def __init__(self, max_pairs):
            self.st = torch.cuda.Stream()
            self.ex = snakeorb.ORBExtractor(NFEAT, SCALE, NLEVELS, TH, THMIN, max_batch=2 * max_pairs, device=local_rank)
            self.cx = snakeorb.Context(device=local_rank)
            self.ex.set_stream(self.st.cuda_stream)
            self.cx.set_stream(self.st.cuda_stream)
            self.rp = torch.empty((max_pairs, cap), dtype=torch.float32, device="cuda")
            self.dp = torch.empty((max_pairs, cap), dtype=torch.float32, device="cuda")
            self.nm = torch.empty((max_pairs,), dtype=torch.int32, device="cuda")
            self.ev = None

        def wait(self):
            if self.ev is not None:
                self.ev.synchronize()

        def run(self, src_ptr, on_device, n_pairs, o_kps, o_desc, o_cnt, o_rp, o_dp, o_nm):
            """One chunk: images at src_ptr -> keypoints, descriptors, counts, right_points, depth, match counts at
            the given host addresses. Everything is queued on the lane's stream; self.ev marks the end."""
            self.ex.enqueue_ptr(src_ptr, 2 * n_pairs, H, W, W, W * H, on_device)
            self.cx.stereo_orb_device(self.ex, n_pairs, sp, self.rp.data_ptr(), self.dp.data_ptr(), self.nm.data_ptr())
            snakeorb._check(snakeorb.lib().snk_orb_fetch_async(self.ex.h, C.c_void_p(o_kps), C.c_void_p(o_desc), cap, C.c_void_p(o_cnt)))
            s_ = C.c_void_p(self.st.cuda_stream)
            for dst, src, nb in ((o_rp, self.rp, 4 * n_pairs * cap), (o_dp, self.dp, 4 * n_pairs * cap), (o_nm, self.nm, 4 * n_pairs)):
                rc = cudart.cudaMemcpyAsync(C.c_void_p(dst), C.c_void_p(src.data_ptr()), C.c_size_t(nb), 2, s_)
                assert rc == 0, "cudaMemcpyAsync failed"
            self.ev = torch.cuda.Event()
            self.ev.record(self.st)

    e2e_lanes = [Lane(CP) for _ in range(args.lanes)] if not args.no_e2e else []
    slot_slice = [-1] * NSLOT  # which slice of the batch each result slot holds (for the check after the run)

    def e2e_chunk(i, k):
        """Local chunk number i (lane and result slot round robin) processes slice k of this rank's pinned batch."""
        lane, slot = e2e_lanes[i % len(e2e_lanes)], i % NSLOT
        slot_slice[slot] = k
        lane.run(host.data_ptr() + k * CI * W * H, False, CP, h_kps[slot].data_ptr(), h_desc[slot].data_ptr(), h_cnt[slot].data_ptr(),
                 h_rp[slot].data_ptr(), h_dp[slot].data_ptr(), h_nm[slot].data_ptr())

    if args.workload == "c5":
        return run_c5(args, locals())

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
        # the same step on a NOISY variant of the batch (Gaussian sensor noise, sigma = 5 grey levels: 2.3x the raw FAST
        # candidates, level 0 beyond the quadtree kernel's small shared-memory array): density regressions show up here
        try:
            rngz = np.random.default_rng(77)
            noisy = np.clip(batch.astype(np.int16) + np.rint(rngz.normal(0, 5.0, (32,) + batch.shape[1:])).astype(np.int16)[np.arange(NI) % 32],
                            0, 255).astype(np.uint8)
            dev_noisy = torch.from_numpy(noisy).cuda()
            for _ in range(3):
                ex.enqueue_ptr(dev_noisy.data_ptr(), NI, H, W, W, W * H, True)
                ctx.stereo_orb_device(ex, P, sp, rp.data_ptr(), dp.data_ptr(), nm.data_ptr())
            z0, z1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            z0.record()
            for _ in range(args.steps):
                ex.enqueue_ptr(dev_noisy.data_ptr(), NI, H, W, W, W * H, True)
                ctx.stereo_orb_device(ex, P, sp, rp.data_ptr(), dp.data_ptr(), nm.data_ptr())
            z1.record()
            torch.cuda.synchronize()
            zms = z0.elapsed_time(z1) / args.steps
            next_rows["noisy_sigma5_c2"] = {"frames_per_s": P / (zms * 1e-3), "ms_per_step": zms,
                                            "raw_fast_candidates_per_image": sum(len(ex.candidates(l, 0)) for l in range(NLEVELS)),
                                            "vs_clean_batch": (ms / args.steps) / zms}
            del dev_noisy
            ex.enqueue_ptr(dev.data_ptr(), NI, H, W, W, W * H, True)  # the clean batch's results are read again below
            ctx.stereo_orb_device(ex, P, sp, rp.data_ptr(), dp.data_ptr(), nm.data_ptr())
            torch.cuda.synchronize()
        except Exception as exc:  # noqa: BLE001
            next_rows["noisy_sigma5_c2"] = {"error": repr(exc)[:200]}
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
            # the same search with the frame and the local map resident on the device (snk_frame_dev / snk_localmap): what is
            # left per call is the kernels and the download of the per-point flags and the assignment
            fdv, lmv = ctxn.frame_dev(fr.view), ctxn.localmap(20000)
            lmv.upload(pts)
            for _ in range(3):
                ctxn.SearchByProjection2Resident(fdv, lmv, camk, pyrk, 4.0, 0.8, keep_assignment=False)
            lmv.upload(pts)
            t0 = time.perf_counter()
            for _ in range(20):
                mres2 = ctxn.SearchByProjection2Resident(fdv, lmv, camk, pyrk, 4.0, 0.8, keep_assignment=False)
            res_ms = (time.perf_counter() - t0) / 20 * 1e3
            next_rows["search_projection_c3"].update({"resident_ms_per_call": res_ms, "resident_matches": int(mres2[0]),
                                                      "resident_path": "frame and local map on the device, flags and assignment downloaded"})
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
            # the loop detector's candidate loop in one call (snk_match_bow_batch): 20 candidates like the one above
            cands20 = []
            for _ in range(20):
                dc = np.ascontiguousarray(np.concatenate([S.flip_bits(d1[rngn.permutation(1000)[:600]], rngn, 12), dk[1000:1400]]))
                cands20.append((dc, MD.featvec(dc, rngn), v2))
            for _ in range(3):
                ctxn.MatchBoWBatch(d1, fv1, v1, cands20, 50, 0.75)
            t0 = time.perf_counter()
            for _ in range(20):
                nmb, _m = ctxn.MatchBoWBatch(d1, fv1, v1, cands20, 50, 0.75)
            next_rows["match_bow_c4"].update({"batch20_ms_per_candidate": (time.perf_counter() - t0) / 20 / 20 * 1e3,
                                              "batch20_matches_mean": float(nmb.mean())})
            if not args.no_cpu:
                t0 = time.perf_counter()
                for _ in range(3):
                    oracle_py.search_projection_fine(fr.view, pts, camk, pyrk, 4.0, 0.8)
                next_rows["search_projection_c3"]["cpu_port_ms_per_call_1_thread"] = (time.perf_counter() - t0) / 3 * 1e3
                ops3 = oracle_py.search_projection_fine(fr.view, pts, camk, pyrk, 4.0, 0.8)[-1]  # distance() calls of the reference
                next_rows["search_projection_c3"].update({"pair_ops_per_call": int(ops3),
                                                          "pair_ops_per_s_resident": ops3 / (res_ms * 1e-3)})
                opsb = oracle_py.match_bow(d1, d2, fv1, fv2, v1, v2, 50, 0.75)[-1]
                next_rows["match_bow_c4"].update({"pair_ops_per_call": int(opsb),
                                                  "pair_ops_per_s": opsb / (next_rows["match_bow_c4"]["ms_per_call"] * 1e-3)})
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
    e2e, my_chunks = None, 0
    if not args.no_e2e:
        i = 0
        for _ in range(max(NSLOT, len(e2e_lanes), 3 * NCH)):  # warm-up: every lane and every result slot used at least once
            e2e_lanes[i % len(e2e_lanes)].wait()
            e2e_chunk(i, i % NCH)
            i += 1
        torch.cuda.synchronize()
        # the timed job: K steps' worth of chunks for every rank, handed out by the shared queue
        queue = ChunkQueue(world * args.steps * P, CP, store=store, key="snk_e2e_chunks")
        barrier()
        tw0 = time.perf_counter()
        while True:
            e2e_lanes[i % len(e2e_lanes)].wait()  # ask for work only when a lane is free
            c = queue.pull()
            if c is None:
                break
            e2e_chunk(i, c % NCH)
            i += 1
            my_chunks += 1
        torch.cuda.synchronize()
        e2e = (time.perf_counter() - tw0) * 1e3  # host wall clock: the results of this rank's last chunk are in host memory
        barrier()
        # the end-to-end path must deliver the same results as the device-resident one
        cnt_t, nm_t = torch.from_numpy(cnt_host), nm.cpu()
        for slot in range(NSLOT):
            k = slot_slice[slot]
            if k < 0:
                continue
            assert int(h_cnt[slot].sum()) == int(cnt_t[k * CI:(k + 1) * CI].sum()) and \
                int(h_nm[slot].sum()) == int(nm_t[k * CP:(k + 1) * CP].sum()), "e2e results differ from the device path"
        for ln in e2e_lanes:
            ln.ex.check()  # device-side fault flags of the asynchronous path
            ln.cx.check()

    t = torch.tensor([ms, e2e if e2e is not None else 0.0], dtype=torch.float64, device="cuda")
    chunks_per_rank = [my_chunks]
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        g = [torch.zeros(1, dtype=torch.int64, device="cuda") for _ in range(world)]
        dist.all_gather(g, torch.tensor([my_chunks], dtype=torch.int64, device="cuda"))
        chunks_per_rank = [int(x) for x in g]
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
        # DRAM traffic of that kernel: dram__bytes_read + dram__bytes_write of the committed `ncu --set full` capture of THIS
        # command at THIS launch shape (profiles/r2x_ncu_summary.csv: 512 images per launch, base units), per launch
        traffic, traffic_src, limiter = None, None, None
        try:
            import csv

            cap_csv = os.path.join(ROOT, "profiles", "r2x_ncu_summary.csv")
            cap_images = 512
            # (template kernels are listed as `void name<args>`; skip launches whose blocks return at once)
            rows_ = [r for r in csv.reader(open(cap_csv)) if r and any(tok.startswith(kname) for tok in r[0].split("<")[0].split()) and float(r[9]) > 5e6]
            if rows_:
                per_img = sum(float(r[2]) + float(r[3]) for r in rows_) / cap_images
                traffic = per_img * NI / len(rows_)
                traffic_src = os.path.relpath(cap_csv, ROOT) + f" ({cap_images}-image launches, the bench's own shape)"
                # what the kernel is bound by when it is not the memory system: share of the issue slots it uses (same capture)
                limiter = {"what": "instruction issue (integer byte-SIMD, shared-memory lists), not DRAM",
                           "issue_active_pct": sum(float(r[6]) for r in rows_) / len(rows_),
                           "dram_pct_of_peak": sum(float(r[4]) for r in rows_) / len(rows_),
                           "warp_instructions_per_image": sum(float(r[9]) for r in rows_) / cap_images}
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
                           "host_affinity": numa, "chunks_per_rank": chunks_per_rank, "device_choice": h2d_probe,
                           # what the box's host side allows when every rank uploads at once (probe at start-up): the
                           # step's input bytes at the sum of the measured per-GPU rates
                           "upload_ceiling_frames_per_s": sum(h2d_probe["h2d_gbs_per_gpu"]) * 1e9 / (2 * W * H),
                           "pipeline": f"chunks of {CP} pairs ({NCH} per step), handed out to the ranks by a shared queue; on a rank the chunks go round robin over {len(e2e_lanes)} streams/handles (copy-compute overlap), a rank asks for the next chunk when a lane is free; {NSLOT} pinned result slots per rank"}
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
            # two more CPU figures, informational: the port rebuilt -O3 -march=native on this host, and
            # the reference's own kernels (cv2 4.13 SIMD + its compiled Saiga pieces) under a Python driver
            anchors = {}
            for variant in ("native", "hybrid"):
                try:
                    out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "cpu_anchor.py"), "--variant", variant,
                                          "--threads", str(cores), "--budget", str(min(args.cpu_budget, 8.0))],
                                         capture_output=True, text=True, timeout=240)
                    anchors[variant] = json.loads(out.stdout.strip().splitlines()[-1]) if out.returncode == 0 else \
                        {"unavailable": (out.stderr.strip().splitlines() or ["failed"])[-1][:200]}
                except Exception as exc:  # noqa: BLE001
                    anchors[variant] = {"unavailable": repr(exc)[:200]}
            line["cpu_anchors"] = {"o3_march_native_port": anchors["native"], "cv2_simd_hybrid": anchors["hybrid"]}
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
