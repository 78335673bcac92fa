"""Segmented walk (long one-shot captures walked by several warps at once, walk.cu / tmk_internal.cuh): records must be
exactly the serial decoder's — compared with the oracle bit for bit — whatever the segment length, including frames that
straddle segment boundaries, false detections inside frame bodies, marginal SNR and corrupted frames."""
import os

import numpy as np
import pytest

import helpers

pytestmark = pytest.mark.gpu
FB, MAN = 1, 0


@pytest.fixture
def seg_env():
    saved = {k: os.environ.get(k) for k in ("TMK_WALK_SEG_MIN", "TMK_WALK_SEGMENTS")}
    yield
    for k, v in saved.items():
        if v is None:
            os.environ.pop(k, None)
        else:
            os.environ[k] = v


def long_capture(orc, cfg, rng, n_frames, snr_db, corrupt=0.0, **kw):
    c = helpers.make_capture(orc, cfg, rng, n_frames=n_frames, snr_db=snr_db, **kw)
    if corrupt:
        for _ in range(int(corrupt * n_frames)):
            a = int(rng.integers(0, len(c) - 400))
            c[a:a + int(rng.integers(20, 400))] *= -1.0  # flipped stretches: header / CRC failures, false preambles
    return c


@pytest.mark.parametrize("seg_min", [4096, 20000, 131072])
def test_segmented_walk_matches_oracle(engine, orc, seg_env, seg_min):
    import tmkphy
    os.environ["TMK_WALK_SEG_MIN"] = str(seg_min)
    os.environ.pop("TMK_WALK_SEGMENTS", None)
    rng = np.random.default_rng(1000 + seg_min)
    cf, cm = tmkphy.make_config(FB, local_addr=1), tmkphy.make_config(MAN, local_addr=1)
    caps = [long_capture(orc, cf, rng, 260, 12.0, corrupt=0.1, gap_range=(48, 600)),
            long_capture(orc, cm, rng, 170, 6.0, gap_range=(48, 3000)),          # 6 dB sits on the threshold
            long_capture(orc, cf, rng, 220, 3.0, gap_range=(48, 900)),           # mostly undetected: long searches
            rng.normal(0, 0.3, 700_000).astype(np.float32),                     # noise only
            np.zeros(300_000, np.float32),                                      # silence
            long_capture(orc, cm, rng, 40, 20.0, multipath=True, gap_range=(20000, 60000)),  # sparse frames
            np.zeros(0, np.float32), long_capture(orc, cf, rng, 3, 15.0)]
    cfgs = [cf, cm, cf, cf, cm, cm, cf, cf]
    s, st = helpers.pack(caps, cfgs, tmkphy)
    engine.stats_reset()
    recs, pl = engine.decode_batch(s, st)
    stats = engine.stats()
    want, wpl = [], []
    base = 0
    for i, (c, cfg) in enumerate(zip(caps, cfgs)):
        r, p = orc.decode_stream(cfg, c, stream_id=i)
        r["payload_off"] += base
        base += len(p)
        want.append(r)
        wpl.append(p)
    want, wpl = np.concatenate(want), np.concatenate(wpl)
    helpers.assert_records_equal(recs, pl, want, wpl, f"seg_min {seg_min}")
    assert len(recs) > 300
    # the verdict tallies are those of the records that count
    assert stats["detections"] == len(want) and stats["frames_ok"] == int((want["status"] == 0).sum())
    assert stats["crc_fail"] == int((want["status"] == 5).sum()) and stats["not_for_us"] == int((want["status"] == 4).sum())
    assert stats["samples"] == sum(len(c) for c in caps)
    # and the plain walk gives the same records
    os.environ["TMK_WALK_SEGMENTS"] = "0"
    recs0, pl0 = engine.decode_batch(s, st)
    helpers.assert_records_equal(recs0, pl0, want, wpl, "plain walk")


def test_segmented_walk_with_chunked_host_staging(engine, orc, seg_env):
    """host-resident captures staged in chunks (copy of chunk g + 1 under the kernels of chunk g) AND segmented walks: every
    chunk walks its own range of segment walkers in two passes"""
    import tmkphy
    saved = os.environ.get("TMK_CHUNK_MSAMPLES")
    os.environ["TMK_WALK_SEG_MIN"] = "16384"
    os.environ["TMK_CHUNK_MSAMPLES"] = "1"  # chunks of about one capture
    try:
        rng = np.random.default_rng(77)
        cf, cm = tmkphy.make_config(FB, local_addr=1), tmkphy.make_config(MAN, local_addr=1)
        caps = [long_capture(orc, cf, rng, 230, 10.0, gap_range=(48, 900)), long_capture(orc, cm, rng, 150, 7.0, gap_range=(48, 2000)),
                long_capture(orc, cf, rng, 200, 5.0, corrupt=0.1, gap_range=(48, 700)), long_capture(orc, cm, rng, 5, 20.0)]
        cfgs = [cf, cm, cf, cm]
        s, st = helpers.pack(caps, cfgs, tmkphy)
        assert len(s) > (3 << 19)  # more than one chunk of 2^20 samples
        recs, pl = engine.decode_batch(s, st)
        want, wpl, base = [], [], 0
        for i, (c, cfg) in enumerate(zip(caps, cfgs)):
            r, p = orc.decode_stream(cfg, c, stream_id=i)
            r["payload_off"] += base
            base += len(p)
            want.append(r)
            wpl.append(p)
        helpers.assert_records_equal(recs, pl, np.concatenate(want), np.concatenate(wpl), "chunked host staging + segments")
    finally:
        if saved is None:
            os.environ.pop("TMK_CHUNK_MSAMPLES", None)
        else:
            os.environ["TMK_CHUNK_MSAMPLES"] = saved
