"""The N>1 path on CPU: two gloo ranks shard a short stereo sequence by frame (snakeorb.sharding), each rank runs its
block (here through the CPU oracle — the sharding logic is what is under test, the GPU kernels are covered by the
-m gpu tests), results are gathered in frame order and must equal the single-process run; the timing reduction
bench.py uses (max over ranks) is exercised too."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_range_covers_and_balances():
    from snakeorb.sharding import batches, shard_range

    for F in (0, 1, 7, 3680):
        for G in (1, 2, 4, 8):
            blocks = [shard_range(F, r, G) for r in range(G)]
            assert blocks[0][0] == 0 and blocks[-1][1] == F
            assert all(blocks[i][1] == blocks[i + 1][0] for i in range(G - 1))
            sizes = [b - a for a, b in blocks]
            assert max(sizes) - min(sizes) <= 1
    assert shard_range(3680, 3, 8) == (1380, 1840)
    assert batches(10, 25, 6) == [(10, 16), (16, 22), (22, 25)]
    with pytest.raises(ValueError):
        shard_range(10, 2, 2)


def _worker(rank, world, port, n_frames, out_dir):
    for p in (ROOT, os.path.join(ROOT, "snake-slam_b200"), os.path.join(ROOT, "oracle")):
        sys.path.insert(0, p)
    import torch
    import torch.distributed as dist

    import oracle_py as O
    from snakeorb import synthetic as S
    from snakeorb.sharding import shard_range

    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    lo, hi = shard_range(n_frames, rank, world)
    ex = O.Extractor(300, 1.2, 4, 20, 7, 1)
    counts = []
    digest = []
    for f in range(lo, hi):
        left = S.image(900 + f, 240, 180)
        k, d = ex.detect(left)
        counts.append(len(k))
        digest.append(int(np.bitwise_xor.reduce(d.reshape(-1)) & np.uint64(0x7fffffff)) if len(d) else 0)
    # gather variable-length per-frame results in rank order
    mine = torch.tensor([counts, digest], dtype=torch.int64).reshape(2, -1)
    sizes = [torch.zeros(1, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(sizes, torch.tensor([hi - lo], dtype=torch.int64))
    mx = int(max(s.item() for s in sizes))
    pad = torch.zeros((2, mx), dtype=torch.int64)
    pad[:, :hi - lo] = mine
    got = [torch.zeros((2, mx), dtype=torch.int64) for _ in range(world)]
    dist.all_gather(got, pad)
    # the timing reduction of bench.py: every rank contributes its elapsed time, the job time is the max
    t = torch.tensor([10.0 + rank], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        allc = np.concatenate([g[:, :int(s.item())].numpy() for g, s in zip(got, sizes)], axis=1)
        np.save(os.path.join(out_dir, "gathered.npy"), allc)
        np.save(os.path.join(out_dir, "tmax.npy"), t.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_two_gloo_ranks_equal_single_process(tmp_path, oracle):
    import torch.multiprocessing as mp

    from snakeorb import synthetic as S

    n_frames, world, port = 5, 2, 29000 + os.getpid() % 2000
    mp.spawn(_worker, args=(world, port, n_frames, str(tmp_path)), nprocs=world, join=True)
    got = np.load(tmp_path / "gathered.npy")
    ex = oracle.Extractor(300, 1.2, 4, 20, 7, 1)
    want_c, want_d = [], []
    for f in range(n_frames):
        k, d = ex.detect(S.image(900 + f, 240, 180))
        want_c.append(len(k))
        want_d.append(int(np.bitwise_xor.reduce(d.reshape(-1)) & np.uint64(0x7fffffff)) if len(d) else 0)
    assert got.shape == (2, n_frames)
    assert list(got[0]) == want_c and list(got[1]) == want_d
    assert float(np.load(tmp_path / "tmax.npy")[0]) == 11.0


def _queue_worker(rank, world, port, n_frames, chunk, out_dir):
    for p in (ROOT, os.path.join(ROOT, "snake-slam_b200")):
        sys.path.insert(0, p)
    import time

    import torch.distributed as dist
    from torch.distributed.distributed_c10d import _get_default_store

    from snakeorb.sharding import ChunkQueue

    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    q = ChunkQueue(n_frames, chunk, store=_get_default_store(), key="test_queue")
    mine = []
    while True:
        c = q.pull()
        if c is None:
            break
        mine.append(c)
        time.sleep(0.002 * (1 + 3 * rank))  # ranks of different speed: the faster one ends up with more chunks
    np.save(os.path.join(out_dir, f"chunks_{rank}.npy"), np.array(mine, np.int64))
    dist.barrier()
    dist.destroy_process_group()


def test_chunk_queue_hands_out_every_chunk_once_across_ranks(tmp_path):
    """The dynamic frame assignment of bench.py's end-to-end path: two gloo ranks of different speed pull chunks of a
    3680-frame sequence from one store-backed counter; every chunk is taken exactly once, the faster rank takes more,
    and the chunk -> frame range map covers the sequence in order."""
    import torch.multiprocessing as mp

    from snakeorb.sharding import ChunkQueue

    n_frames, chunk, world, port = 3680, 128, 2, 31000 + os.getpid() % 2000
    mp.spawn(_queue_worker, args=(world, port, n_frames, chunk, str(tmp_path)), nprocs=world, join=True)
    got = [np.load(tmp_path / f"chunks_{r}.npy") for r in range(world)]
    allc = np.sort(np.concatenate(got))
    q = ChunkQueue(n_frames, chunk)
    assert q.n_chunks == 29 and list(allc) == list(range(29))
    assert len(got[0]) > len(got[1]) > 0
    ranges = [q.frames(c) for c in range(q.n_chunks)]
    assert ranges[0] == (0, 128) and ranges[-1] == (3584, 3680) and all(ranges[i][1] == ranges[i + 1][0] for i in range(28))
    # one process, no store: a plain counter
    assert [q.pull() for _ in range(31)] == list(range(29)) + [None, None]
    with pytest.raises(ValueError):
        ChunkQueue(10, 0)


CONTRACT_KEYS = ["metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
                 "vs_baseline", "dtype", "data", "config", "e2e", "gpu_launches"]


def test_reference_arm_prints_the_contract_line():
    """bench.py --impl reference: the reference side of the driver's ratio runs without a GPU (it is the CPU path), prints
    ONE JSON line with the contract's keys, impl = reference, a cpu_baseline describing the run and an e2e equal to it."""
    import json
    import subprocess
    import sys

    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0",
                          "--pairs", "2"], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    for k in CONTRACT_KEYS + ["impl", "cpu_baseline"]:
        assert k in d, k
    assert d["impl"] == "reference" and d["gpu_launches"] == 0 and d["vs_baseline"] is None and d["value"] > 0
    assert d["cpu_baseline"]["kind"] in ("port", "reference") and d["cpu_baseline"]["value"] == d["value"] == d["e2e"]["value"]
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
    assert "workload" in d["config"] and "model" not in d["config"]


def test_committed_bench_line_has_every_contract_key():
    """The newest bench line under profiles/ (written on a B200) carries the contract's keys plus roofline, cpu_baseline,
    clocks; guards bench.py's JSON assembly against silently dropping one."""
    import glob
    import json

    newest = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_bench.json")))[-1]
    d = json.loads(open(newest).read().strip().splitlines()[-1])
    for k in CONTRACT_KEYS + ["roofline", "cpu_baseline", "clocks"]:
        assert k in d, (newest, k)
    for k in ("bound", "achieved", "peak", "unit", "frac", "traffic"):
        assert k in d["roofline"], k
    assert d["roofline"]["bound"] in ("hbm", "tensor") and abs(d["roofline"]["frac"] - d["roofline"]["achieved"] / d["roofline"]["peak"]) < 1e-9
    for k in ("value", "unit", "cores", "kind", "sample"):
        assert k in d["cpu_baseline"], k
    for k in ("value", "unit", "h2d_bytes_per_step", "d2h_bytes_per_step"):
        assert k in d["e2e"], k
    assert d["e2e"]["h2d_bytes_per_step"] > 0 and d["gpu_launches"] > 0 and d["e2e"]["value"] != d["value"]
    assert d["clocks"]["reasons"] == [] or all("thermal" not in r and "hw_slowdown" not in r for r in d["clocks"]["reasons"])
