#!/usr/bin/env python
"""How much would the chirp walk gain from more warps per capture? The C2 captures decoded as whole streams and as
halves / quarters presented as separate streams (results differ at the cuts; the walk time is what is looked at)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "trackmaker-rs_b200"))
import tmkphy
eng = tmkphy.Engine(0)
n_s, n_per = 4096, 480000
snrs = [-3.0, 0.0, 3.0, 6.0, 10.0, 15.0, 20.0, 30.0]
d = eng.device_alloc(n_s * n_per * 4)
for g, snr in enumerate(snrs):
    eng.ask_synth_batch(0xA5C0 + g, 512, n_per, snr, device_ptr=d + g * 512 * n_per * 4)
for parts in (1, 2, 4, 8):
    st, total = tmkphy.make_streams([n_per // parts] * (n_s * parts), tmkphy.make_config(1))
    for _ in range(2):
        recs = eng.ask_decode_batch(d, st, total_samples=total)
    ms = []
    for _ in range(3):
        eng.ask_decode_batch(d, st, total_samples=total)
        ms.append(eng.ask_last_stage_ms())
    print(f"parts {parts}: filter {np.median([m['scan'] for m in ms]):.2f} ms walk {np.median([m['walk'] for m in ms]):.2f} ms records {len(recs)}", flush=True)
