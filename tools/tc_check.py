"""Focused check of the tensor-core scan: correlation numerators against float64 numpy, then the filter contract
numbers, on a few small captures (run under `timeout`)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "trackmaker-rs_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers  # noqa: E402
import tmkphy  # noqa: E402
from oracle import oracle as orc  # noqa: E402

eng = tmkphy.Engine(0)
rng = np.random.default_rng(1)
for kind in (1, 0):
    cfg = tmkphy.make_config(kind, local_addr=1)
    pre = orc.preamble(kind).astype(np.float64)
    P = len(pre)
    caps = [helpers.make_capture(orc, cfg, rng, n_frames=n, snr_db=snr) for n, snr in ((3, 10), (12, 6), (40, 20))]
    caps.append(rng.normal(0, 1, 8192 * 3 + 777).astype(np.float32))
    s, st = helpers.pack(caps, cfg, tmkphy)
    dot = eng.scan_dot_probe(s, st)
    worst = 0.0
    for i, c in enumerate(caps):
        x = c.astype(np.float64)
        D = np.correlate(x, pre, mode="valid")
        o = int(st[i]["offset"])
        got = dot[o:o + len(D)].astype(np.float64)
        cs = np.concatenate([[0.0], np.cumsum(x * x)])
        E = cs[P:] - cs[:-P]
        err = np.abs(got - D) / np.sqrt(np.maximum(E, 1e-30) * P)
        k = int(np.argmax(err))
        worst = max(worst, float(err.max()))
        print(f"kind {kind} capture {i}: n={len(c)} max normalised error {err.max():.3e} at window {k} (got {got[k]:.5f} want {D[k]:.5f})", flush=True)
    print(f"kind {kind}: worst {worst:.3e}  {'OK' if worst < 1.2e-3 else 'BAD'}", flush=True)
    recs, pl = eng.decode_batch(s, st)
    want, wpl, _ = orc.decode_batch(s, st)
    try:
        helpers.assert_records_equal(recs, pl, want, wpl, "tc", corr_tol=None)
        print(f"kind {kind}: decode bit-exact, {len(want)} records", flush=True)
    except AssertionError as ex:
        print(f"kind {kind}: decode MISMATCH {ex}", flush=True)
eng.close()
