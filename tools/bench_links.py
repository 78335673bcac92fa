#!/usr/bin/env python
"""Throughput of the carried-state streaming decode: n links fed with 25 ms polls (1200 samples, csma.rs:446) out of
captures resident in HBM. usage: bench_links.py [n_links] [n_polls]"""
import ctypes as C
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "trackmaker-rs_b200"))
import tmkphy  # noqa: E402

n_links = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
n_polls = int(sys.argv[2]) if len(sys.argv) > 2 else 100
poll = 1200
n_per = poll * n_polls
eng = tmkphy.Engine(0)
cfg = tmkphy.make_config(tmkphy.LINE_4B5B, local_addr=1)
d_x = eng.device_alloc(n_links * n_per * 4)
eng.synth_batch(cfg, 4242, n_links, n_per, 15.0, 0.7, device_ptr=d_x, want_truth=False)
bank = tmkphy.PhyDecoderBank(eng, [cfg] * n_links)
lib, h = eng._lib, bank._h
rec_cap, pay_cap = 8 * n_links + 1024, 300 * n_links + 65536
recs = np.zeros(rec_cap, tmkphy.REC_DTYPE)
pl = np.zeros(pay_cap, np.uint8)
nr, npl = C.c_uint64(0), C.c_uint64(0)
lens = np.full(n_links, poll, np.uint64)
base = np.arange(n_links, dtype=np.uint64) * np.uint64(n_per)
frames = 0
times = []
p_x, p_lens, p_recs, p_pl = C.c_void_p(d_x), C.c_void_p(lens.ctypes.data), C.c_void_p(recs.ctypes.data), C.c_void_p(pl.ctypes.data)
r_nr, r_npl = C.byref(nr), C.byref(npl)
for k in range(n_polls):
    offs = base + np.uint64(k * poll)
    p_offs = C.c_void_p(offs.ctypes.data)  # (the pointer objects are made outside the timed call: a Rust caller has none of this)
    t0 = time.perf_counter()
    rc = lib.tmk_links_process_samples(h, p_x, n_links * n_per, p_offs, p_lens, p_recs, rec_cap, r_nr, p_pl, pay_cap, r_npl)
    times.append(time.perf_counter() - t0)
    eng._check(rc)
    frames += int((recs[:nr.value]["status"] == 0).sum())
# the same captures in one shot
st, _ = tmkphy.make_streams([n_per] * n_links, cfg)
r1, _ = eng.decode_batch(d_x, st, total_samples=n_links * n_per)
ms = np.array(times[5:]) * 1e3
print({"links": n_links, "polls": n_polls, "ms_per_poll_median": round(float(np.median(ms)), 3),
       "ms_per_poll_p99": round(float(np.percentile(ms, 99)), 3),
       "Msamples_per_s": round(n_links * poll / float(np.median(ms)) / 1e3, 1),
       "realtime_factor": round(25.0 / float(np.median(ms)), 1),
       "frames_ok_streaming": frames, "frames_ok_one_shot": int((r1["status"] == 0).sum())})
