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


class ChunkQueue:
    """Dynamic alternative to the contiguous blocks above for boxes whose GPUs are not fed equally fast (host-to-device
    bandwidth differs per PCIe root / NUMA node when every GPU uploads): the sequence is cut into chunks of `chunk`
    frames and every rank takes the next free chunk when one of its streams falls idle. The counter lives in a
    torch.distributed store (`store.add` is atomic across ranks); without a store (one process) it is a local integer.
    Nothing else is exchanged: a chunk id fixes the frames and the place of their results, so the outputs still land in
    frame order."""

    def __init__(self, n_frames, chunk, store=None, key="snk_chunk_queue"):
        if chunk <= 0 or n_frames < 0:
            raise ValueError("bad chunk / frame count")
        self.n_frames, self.chunk = n_frames, chunk
        self.n_chunks = (n_frames + chunk - 1) // chunk
        self.store, self.key, self._next = store, key, 0

    def frames(self, c):
        """Frame range [lo, hi) of chunk c."""
        lo = c * self.chunk
        return lo, min(lo + self.chunk, self.n_frames)

    def pull(self):
        """The next unclaimed chunk id, or None when the sequence is used up."""
        if self.store is None:
            c, self._next = self._next, self._next + 1
        else:
            c = self.store.add(self.key, 1) - 1
        return c if c < self.n_chunks else None
