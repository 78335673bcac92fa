#!/usr/bin/env python
"""Bare concurrent pinned host -> device copies, one rank per GPU (is the e2e scaling limit the box or
the engine's staging?). Launch: python -m torch.distributed.run --nproc-per-node N tools/h2d_sweep.py  (or plain python for N = 1).
Every rank copies `GB` gigabytes `reps` times with tmk_memcpy_h2d (cudaMemcpy from pinned memory, no kernels) between two
barriers; rank 0 prints the per-rank and aggregate GB/s."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "trackmaker-rs_b200"))
import bench  # noqa: E402
import tmkphy  # noqa: E402

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
eng = tmkphy.Engine(local)
bench.comm_setup(eng, tmkphy, rank, world)
gb, reps = float(os.environ.get("GB", "2")), 4
n = int(gb * 1e9) // 4
h = eng.pinned_array((n,), np.float32)
h[:] = 1.0
d = eng.device_alloc(n * 4)
eng.h2d(d, h)  # warm-up
for bind in (False, True):
    if bind:
        eng.bind_host_thread()
    eng.synchronize()
    eng.comm_barrier()
    t0 = time.perf_counter()
    for _ in range(reps):
        eng.h2d(d, h)
    eng.synchronize()
    dt = time.perf_counter() - t0
    eng.comm_barrier()
    rate = reps * n * 4 / dt / 1e9
    rates = eng.allreduce_max([rate, -rate]) if world > 1 else [rate, -rate]
    tot = eng.allreduce_sum([int(rate * 1000)])[0] / 1000 if world > 1 else rate
    if rank == 0:
        print(f"H2D ranks={world} numa_bind={bind} per-rank max {rates[0]:.1f} min {-rates[1]:.1f} GB/s aggregate {tot:.1f} GB/s", flush=True)
eng.device_free(d)
eng.close()
