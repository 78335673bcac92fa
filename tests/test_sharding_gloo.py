"""Multi-GPU host logic on CPU: world_size-2 gloo process group (streams shard with no
data-path collective; only the decode counters are reduced). The product reduces through NCCL behind the C ABI
(tmk_comm_*, tests/test_gpu_comm.py on >= 2 GPUs); here a gloo-backed stand-in with the same two methods drives the
same host logic (tmkphy.sharding has no torch in it)."""
import os
import socket

import numpy as np
import pytest


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


class GlooComm:
    """the two reductions of Engine (allreduce_sum / allreduce_max) over a torch.distributed gloo group"""

    def allreduce_sum(self, values):
        import torch
        import torch.distributed as dist
        t = torch.tensor([int(v) for v in values], dtype=torch.int64)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return t.tolist()

    def allreduce_max(self, values):
        import torch
        import torch.distributed as dist
        t = torch.tensor([float(v) for v in values], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.tolist()


def _worker(rank, world, port, lengths, out_dir):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, os.path.join(root, "trackmaker-rs_b200"))
    import torch.distributed as dist
    from tmkphy import sharding
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = sharding.shard_streams(lengths, rank, world)
    local = {"samples": int(np.sum(lengths[lo:hi])), "frames_ok": (hi - lo) * 3, "detections": 7 + rank}
    tot, ms = sharding.reduce_stats(local, 10.0 + rank, GlooComm())
    np.save(os.path.join(out_dir, f"r{rank}.npy"), np.array([lo, hi, tot["samples"], tot["frames_ok"], tot["detections"], ms]))
    dist.barrier()
    dist.destroy_process_group()


def test_shard_partition_properties():
    import sys
    from tmkphy import sharding
    rng = np.random.default_rng(3)
    for world in (1, 2, 3, 8):
        for n in (0, 1, 5, 64, 1000):
            lengths = rng.integers(0, 100000, n)
            cuts = [sharding.shard_streams(lengths, r, world) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == n
            assert all(cuts[r][1] == cuts[r + 1][0] for r in range(world - 1))  # disjoint cover, in order
            if n >= 64 * world:
                per = [lengths[a:b].sum() for a, b in cuts]
                assert max(per) <= 1.5 * (lengths.sum() / world) + lengths.max()
    with pytest.raises(ValueError):
        sharding.shard_streams([1, 2], 2, 2)


def test_shard_cost_weights():
    """mixed batches (BASELINE config 5): a chirp stream costs ~5 x a src/phy stream per sample; the weighted
    partition equalises work, not samples"""
    from tmkphy import sharding
    n = 3000
    lengths = np.full(n, 1 << 18)
    cost = np.where(np.arange(n) < 1000, sharding.CLASS_COST["chirp"], sharding.CLASS_COST["4b5b"])  # chirp streams first
    for world in (2, 4, 8):
        cuts = [sharding.shard_streams(lengths, r, world, cost) for r in range(world)]
        work = [float((lengths[a:b] * cost[a:b]).sum()) for a, b in cuts]
        assert cuts[0][0] == 0 and cuts[-1][1] == n and max(work) <= 1.02 * min(work) + lengths[0] * 5
        plain = [sharding.shard_streams(lengths, r, world) for r in range(world)]
        pwork = [float((lengths[a:b] * cost[a:b]).sum()) for a, b in plain]
        assert max(pwork) > 1.5 * min(pwork)  # the sample-count partition is what the weights fix
    assert sharding.reduce_stats({"a": 1}, 2.5) == ({"a": 1}, 2.5)
    with pytest.raises(ValueError):
        sharding.shard_streams([1, 2], 0, 2, cost=[1.0])


def test_gloo_world2_counters(tmp_path):
    import torch.multiprocessing as mp
    lengths = np.random.default_rng(5).integers(1000, 480000, 257)
    port = _free_port()
    mp.spawn(_worker, args=(2, port, lengths, str(tmp_path)), nprocs=2, join=True)
    r0, r1 = np.load(tmp_path / "r0.npy"), np.load(tmp_path / "r1.npy")
    assert r0[0] == 0 and r0[1] == r1[0] and r1[1] == len(lengths)
    for r in (r0, r1):
        assert r[2] == lengths.sum() and r[3] == 3 * len(lengths) and r[4] == 15 and r[5] == 11.0
