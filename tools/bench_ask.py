#!/usr/bin/env python
"""Timing of the chirp + BPSK path (examples/ask.rs) on one GPU: n streams x 480 000 samples of ask frames
with AWGN, resident in HBM; CUDA-event stage times; oracle on a few streams beside it."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "trackmaker-rs_b200"))
import tmkphy  # noqa: E402
from oracle import oracle as orc  # noqa: E402

n_streams = int(sys.argv[1]) if len(sys.argv) > 1 else 512
n_per = 480000
rng = np.random.default_rng(3)
eng = tmkphy.Engine(0)
base = []
for k in range(8):
    fr = rng.integers(0, 2, (92, 100)).astype(np.uint8)
    fr[:, :8] = [[((i % 100 + 1) >> (7 - j)) & 1 for j in range(8)] for i in range(92)]
    x = orc.ask_modulate(fr, rng.integers(0, 100, 92), rng.integers(0, 100, 92))[:n_per]
    x = np.pad(x, (0, n_per - len(x)))
    snr = (-3, 0, 3, 6, 10, 15, 20, 30)[k]
    base.append((x + rng.normal(0, np.sqrt(0.5) * 10 ** (-snr / 20), n_per)).astype(np.float32))
s = np.concatenate([base[i % 8] for i in range(n_streams)])
st, total = tmkphy.make_streams([n_per] * n_streams, tmkphy.make_config(1))
d = eng.device_alloc(total * 4)
eng.h2d(d, s)
for _ in range(2):
    recs = eng.ask_decode_batch(d, st, total_samples=total)
ms = [eng.ask_last_stage_ms() for _ in range(3) if eng.ask_decode_batch(d, st, total_samples=total) is not None]
t = time.time()
want = np.concatenate([orc.ask_demodulate(base[i], stream=i) for i in range(8)])
cpu_s = time.time() - t
got8 = recs[recs["stream"] < 8]
same = len(got8) == len(want) and all(np.array_equal(got8[f], want[f]) for f in ("stream", "start_index", "id", "bits", "accepted"))
scan = float(np.median([m["scan"] for m in ms])); walk = float(np.median([m["walk"] for m in ms]))
print(json.dumps({"path": "ask (chirp+BPSK)", "streams": n_streams, "samples": total, "scan_ms": scan, "walk_ms": walk,
                  "msamples_per_s": total / (scan + walk) / 1e3, "scan_gflops": 880.0 * total / scan / 1e6,
                  "records": int(len(recs)), "accepted": int(recs["accepted"].sum()),
                  "cpu_oracle_1core_msamples_per_s": 8 * n_per / cpu_s / 1e6, "parity_8_streams": "bit-exact" if same else "MISMATCH"}))
