"""2+ rank check of the NCCL entry points of the C ABI (run under torchrun on a box with >= 2 GPUs):
torchrun --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/comm_check.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "trackmaker-rs_b200"))
import bench  # noqa: E402
import tmkphy  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
eng = tmkphy.Engine(local)
bench.comm_setup(eng, tmkphy, rank, world)
assert eng.comm_size() == world
s = eng.allreduce_sum([rank + 1, 10, 1 << 40])
assert s == [world * (world + 1) // 2, 10 * world, world << 40], s
m = eng.allreduce_max([float(rank), -float(rank)])
assert m == [float(world - 1), 0.0], m
from tmkphy import sharding  # noqa: E402
tot, ms = sharding.reduce_stats({"a": rank, "b": 3}, 5.0 + rank, eng)
assert tot == {"a": world * (world - 1) // 2, "b": 3 * world} and ms == 4.0 + world
g = eng.allgather([rank * 10 + 1, 1 << 41])
assert g == [[r * 10 + 1, 1 << 41] for r in range(world)], g
eng.comm_barrier()
if rank == 0:
    print(f"comm ok: {world} ranks")
eng.close()
