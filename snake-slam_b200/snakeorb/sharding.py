"""Frame sharding for multi-GPU runs: every frame / stereo pair is an independent unit, so a sequence
of F frames is cut into contiguous blocks, one per rank; nothing is exchanged on the data path and the per-frame
outputs are concatenated in frame order on the host."""


def shard_range(n_frames, rank, world):
    """Contiguous block [lo, hi) of rank `rank`: blocks differ in size by at most one frame and cover [0, n_frames)."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    base, rem = divmod(n_frames, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def batches(lo, hi, batch):
    """Batches of at most `batch` frames inside [lo, hi)."""
    out = []
    while lo < hi:
        out.append((lo, min(lo + batch, hi)))
        lo = out[-1][1]
    return out


def gather_in_frame_order(per_rank_results):
    """Concatenate the ranks' per-frame result lists (rank order == frame order for contiguous blocks)."""
    out = []
    for r in per_rank_results:
        out.extend(r)
    return out
