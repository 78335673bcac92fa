"""Multi-GPU host logic: streams are independent (one PhyDecoder per stream, src/phy/decoder.rs:16-33),
so a batch shards over ranks with no data-path collective. The only exchange is the reduction of the
decode counters (and of the step time) after a batch: torch.distributed all_reduce, NCCL on GPUs,
gloo in the CPU tests."""
import numpy as np


def shard_streams(n_samples, rank, world):
    """Contiguous shard [lo, hi) of the streams for `rank`, balanced by sample count.

    n_samples: per-stream lengths. Boundaries depend only on (n_samples, world), so every rank
    computes the same partition without talking to the others."""
    n = np.asarray(n_samples, dtype=np.int64)
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    if len(n) == 0:
        return 0, 0
    cum = np.concatenate([[0], np.cumsum(np.maximum(n, 1))])  # empty streams still cost a slot
    total = cum[-1]
    bounds = [int(np.searchsorted(cum, total * r / world, side="left")) for r in range(world)] + [len(n)]
    bounds[0] = 0
    for r in range(1, world + 1):
        bounds[r] = max(bounds[r], bounds[r - 1])
    return bounds[rank], bounds[rank + 1]


def reduce_stats(stats, ms, device=None, group=None):
    """Sum the counters dict over ranks and take the max of `ms` (device time of the step).
    Returns (stats, ms) on every rank; a no-op without an initialised process group."""
    try:
        import torch
        import torch.distributed as dist
    except ImportError:  # pragma: no cover
        return dict(stats), float(ms)
    if not (dist.is_available() and dist.is_initialized()):
        return dict(stats), float(ms)
    keys = sorted(stats)
    t = torch.tensor([int(stats[k]) for k in keys], dtype=torch.int64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    m = torch.tensor([float(ms)], dtype=torch.float64, device=device)
    dist.all_reduce(m, op=dist.ReduceOp.MAX, group=group)
    return {k: int(v) for k, v in zip(keys, t.cpu().tolist())}, float(m.item())
