"""Multi-GPU host logic: streams are independent (one PhyDecoder per stream, src/phy/decoder.rs:16-33), so a batch
shards over ranks with no data-path collective. The only exchange is the reduction of the decode counters (and of the
step time) after a batch: NCCL through the C ABI (tmk_comm_*, Engine.allreduce_sum / allreduce_max). No torch here:
`reduce_stats` takes any object with those two methods (the CPU tests pass a gloo-backed stand-in)."""
import numpy as np

# measured cost per sample relative to a 4B5B src/phy stream at 10 dB (profiles/, bench C2 legs): Manchester windows
# are longer but its frames rarer per sample; the chirp + BPSK path is an FFT filter + a 440-tap exact walk
CLASS_COST = {"4b5b": 1.0, "manchester": 0.9, "chirp": 5.0}


def shard_streams(n_samples, rank, world, cost=None):
    """Contiguous shard [lo, hi) of the streams for `rank`, balanced by work.

    n_samples: per-stream lengths; cost: optional per-stream cost per sample (CLASS_COST values, or a measured
    figure): the work of a stream is n_samples * cost. Boundaries depend only on (n_samples, cost, world), so every
    rank computes the same partition without talking to the others."""
    n = np.asarray(n_samples, dtype=np.float64)
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    if len(n) == 0:
        return 0, 0
    w = np.maximum(n, 1.0)  # empty streams still cost a slot
    if cost is not None:
        c = np.asarray(cost, dtype=np.float64)
        if c.shape != n.shape or (c <= 0).any():
            raise ValueError("cost: one positive value per stream")
        w = w * c
    cum = np.concatenate([[0.0], np.cumsum(w)])
    total = cum[-1]
    bounds = [int(np.searchsorted(cum, total * r / world, side="left")) for r in range(world)] + [len(n)]
    bounds[0] = 0
    for r in range(1, world + 1):
        bounds[r] = max(bounds[r], bounds[r - 1])
    return bounds[rank], bounds[rank + 1]


def reduce_stats(stats, ms, comm=None):
    """Sum the counters dict over ranks and take the max of `ms` (device time of the step).
    comm: an Engine with a communicator (tmk_comm_init), or any object with allreduce_sum(list[int]) and
    allreduce_max(list[float]); None = single rank. Returns (stats, ms) on every rank."""
    if comm is None:
        return dict(stats), float(ms)
    keys = sorted(stats)
    tot = comm.allreduce_sum([int(stats[k]) for k in keys])
    mx = comm.allreduce_max([float(ms)])
    return {k: int(v) for k, v in zip(keys, tot)}, float(mx[0])
