"""BASELINE.json configs 4 and 5 as parity cases: hour-long sparse captures with clock drift, and a
mixed-modulation batch (stream_id % 3 -> 4B5B / Manchester / chirp + BPSK), bit for bit against the oracle."""
import numpy as np
import pytest

import helpers
from test_gpu_ask import assert_ask_equal

pytestmark = pytest.mark.gpu
MAN, FB = 0, 1


def hour_capture(orc, cfg, rng, n_total, ppm, amplitude, snr_db, every=24000):
    """sparse frames (one per ~0.5 s) at random offsets, resampled by (1 + ppm * 1e-6) with linear interpolation
    (SURVEY §8d clock drift), scaled, AWGN"""
    n_src = int(n_total * (1 + abs(ppm) * 1e-6)) + 16
    src = np.zeros(n_src, np.float32)
    pos, n_frames = int(rng.integers(0, every)), 0
    frames = helpers.random_frames(rng, 64, cfg.local_addr)
    waves = [orc.modulate(cfg, [f], 0) for f in frames]
    while pos + 14000 < n_src:
        w = waves[n_frames % len(waves)]
        src[pos:pos + len(w)] = w
        n_frames += 1
        pos += len(w) + int(rng.integers(every // 2, every * 3 // 2))
    out = np.empty(n_total, np.float32)
    step = 1 << 23
    sigma = amplitude * 10 ** (-snr_db / 20.0)
    for a in range(0, n_total, step):
        b = min(n_total, a + step)
        t = np.arange(a, b, dtype=np.float64) * (1.0 + ppm * 1e-6)
        i0 = t.astype(np.int64)
        fr = (t - i0).astype(np.float32)
        seg = src[i0] * (1 - fr) + src[i0 + 1] * fr
        out[a:b] = seg * np.float32(amplitude) + rng.standard_normal(b - a, dtype=np.float32) * np.float32(sigma)
    return out, n_frames


@pytest.mark.parametrize("hours", [1.0])
def test_config4_hour_captures_with_drift(engine, orc, hours):
    """C4 (scaled to two captures): 172.8 M samples each, one frame per ~0.5 s, +-100 ppm drift, amplitude
    U[0.3, 1], AWGN 10 dB: every record (offsets, verdicts, payload) bit-exact; long idle stretches, 45 000 tiles
    per capture, a single warp walking an hour of audio"""
    import tmkphy
    rng = np.random.default_rng(404)
    n_total = int(172_800_000 * hours)
    caps, cfgs, planted = [], [], 0
    for kind, ppm in ((FB, +93.0), (MAN, -71.0)):
        cfg = orc.make_config(kind, local_addr=1)
        x, nf = hour_capture(orc, cfg, rng, n_total, ppm, float(rng.uniform(0.3, 1.0)), 10.0)
        caps.append(x)
        cfgs.append(cfg)
        planted += nf
    samples, st = helpers.pack(caps, cfgs, tmkphy)
    del caps
    want, want_pl, _ = orc.decode_batch(samples, st)
    got, got_pl = engine.decode_batch(samples, st, rec_cap=len(want) + 64, payload_cap=len(want_pl) + 4096)
    helpers.assert_records_equal(got, got_pl, want, want_pl, "C4", corr_tol=None)
    ok = int((want["status"] == 0).sum()) + int((want["status"] == 4).sum())
    assert planted > 10000 and ok > 0.5 * planted, (planted, ok)
    # offsets span the whole hour
    assert int(want["preamble_start"].max()) > n_total - 200000


def test_config5_mixed_modulations(engine, orc):
    """C5 (scaled): stream_id % 3 -> 4B5B / Manchester / chirp + BPSK, 2^16 samples per stream, AWGN"""
    import tmkphy
    from test_gpu_ask import oracle_batch
    rng = np.random.default_rng(505)
    n_streams, n_per = 384, 1 << 16
    phy_caps, phy_cfgs, ask_caps = [], [], []
    for s in range(n_streams):
        k = s % 3
        if k < 2:
            cfg = orc.make_config(FB if k == 0 else MAN, local_addr=1)
            x = np.zeros(n_per, np.float32)
            c = helpers.make_capture(orc, cfg, rng, n_frames=14, snr_db=float(rng.choice([3, 6, 10, 20])),
                                     multipath=bool(s % 2), amplitude=float(rng.uniform(0.3, 1.0)))[:n_per]
            x[:len(c)] = c
            phy_caps.append(x)
            phy_cfgs.append(cfg)
        else:
            fr = rng.integers(0, 2, (10, 100)).astype(np.uint8)
            for i in range(10):
                fr[i, :8] = [((i + 1) >> (7 - j)) & 1 for j in range(8)]
            g1, g2 = rng.integers(0, 100, 10), rng.integers(0, 100, 10)
            w = orc.ask_modulate(fr, g1, g2)[:n_per]
            x = np.zeros(n_per, np.float32)
            x[:len(w)] = w
            x += rng.normal(0, 0.2, n_per).astype(np.float32)
            ask_caps.append(x)
    s1, st1 = helpers.pack(phy_caps, phy_cfgs, tmkphy)
    want, want_pl, _ = orc.decode_batch(s1, st1)
    got, got_pl = engine.decode_batch(s1, st1)
    helpers.assert_records_equal(got, got_pl, want, want_pl, "C5 src/phy", corr_tol=None)
    assert (want["status"] == 0).sum() > 1000
    s2, st2 = helpers.pack(ask_caps, [orc.make_config(FB)] * len(ask_caps), tmkphy)
    w2 = oracle_batch(orc, s2, st2)
    g2_ = engine.ask_decode_batch(s2, st2)
    assert_ask_equal(g2_, w2, "C5 chirp")
    assert (w2["accepted"] == 1).sum() > 600
