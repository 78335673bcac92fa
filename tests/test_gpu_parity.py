"""Parity of the CUDA path (through the C ABI) against the oracle on the same seeded inputs.
Bit-exact: verdicts, preamble/frame start offsets, header fields, payload bytes, modulated
waveforms. Within 1e-5*max(|ref|,1): correlation values."""
import os

import numpy as np
import pytest

import helpers

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
MAN, FB = 0, 1


def oracle_batch(orc, samples, streams):
    recs, pl, _ = orc.decode_batch(samples, streams, n_threads=8)
    return recs, pl


def check_batch(engine, orc, captures, cfgs, ctx="", corr_tol=1e-5):
    import tmkphy
    s, st = helpers.pack(captures, cfgs, tmkphy)
    got, gpl = engine.decode_batch(s, st)
    want, wpl = oracle_batch(orc, s, st)
    helpers.assert_records_equal(got, gpl, want, wpl, ctx, corr_tol)
    return want


@pytest.mark.parametrize("kind", [FB, MAN])
def test_modulate_bit_exact(engine, orc, kind):
    import tmkphy
    rng = np.random.default_rng(1)
    cfg = tmkphy.make_config(kind)
    frames = helpers.random_frames(rng, 9) + [(1, 3, 4, 5, bytes(range(256)) * 2)]
    got = engine.modulate(cfg, [tmkphy.Frame(tmkphy.FrameType(t), q, s, d, p) for t, q, s, d, p in frames], 48)
    want = orc.modulate(cfg, frames, 48)
    assert got.dtype == np.float32 and np.array_equal(got, want)
    assert len(engine.modulate(cfg, [], 48)) == 0


@pytest.mark.parametrize("kind,spl,pb", [(MAN, 2, 2), (FB, 4, 2), (FB, 1, 1), (MAN, 5, 3)])
def test_modulate_other_shapes(engine, orc, kind, spl, pb):
    import tmkphy
    cfg = tmkphy.make_config(kind, spl=spl, preamble_bytes=pb)
    frames = helpers.random_frames(np.random.default_rng(2), 4)
    got = engine.modulate(cfg, [tmkphy.Frame(tmkphy.FrameType(t), q, s, d, p) for t, q, s, d, p in frames], 7)
    assert np.array_equal(got, orc.modulate(cfg, frames, 7))


@pytest.mark.parametrize("kind", [FB, MAN])
@pytest.mark.parametrize("chunk", [0, 1200])
def test_loopback_think_different(engine, orc, kind, chunk):
    """BASELINE config 1: modulate assets/think-different.txt, demodulate, compare (src/main.rs:480-589)."""
    import tmkphy
    text = open(os.path.join(GOLDEN, "think-different.txt"), "rb").read()
    frames = [tmkphy.Frame.new_data(i, 0, 1, text[o:o + 128]) for i, o in enumerate(range(0, len(text), 128))]
    enc = tmkphy.PhyEncoder(engine, 3, 2, kind)
    wave = enc.encode_frames(frames, 48)
    assert len(wave) == (32136 if kind == FB else 51216)
    out, recs, pl = tmkphy.decode_capture(engine, tmkphy.make_config(kind, local_addr=1, chunk=chunk), wave)
    assert b"".join(f.data for f in out) == text and [f.sequence for f in out] == list(range(8))
    stride, P = (4158, 60) if kind == FB else (6624, 96)
    assert list(recs["preamble_start"]) == [i * stride for i in range(8)]
    assert list(recs["frame_start"]) == [i * stride + P for i in range(8)]
    cfg = orc.make_config(kind, local_addr=1, chunk=chunk)
    want, wpl = orc.decode_stream(cfg, wave)
    helpers.assert_records_equal(recs, pl, want, wpl, "loopback")
    # the reference's own call: ONE PhyDecoder::process_samples over the whole capture (src/main.rs:540)
    dec = tmkphy.PhyDecoder(engine, 3, 2, kind, local_addr=1)
    out2 = dec.process_samples(wave)
    assert [f.data for f in out2] == [f.data for f in out]
    dec.close()


@pytest.mark.parametrize("kind", [FB, MAN])
@pytest.mark.parametrize("chunk", [0, 1200, 257])
def test_decode_snr_sweep(engine, orc, kind, chunk):
    """SURVEY §8d AWGN sweep; 6 dB sits on the 0.9 threshold, where exactness is hardest."""
    import tmkphy
    rng = np.random.default_rng(100 + kind + chunk)
    caps, cfgs = [], []
    for snr in (-3, 0, 3, 5, 6, 7, 8, 10, 15, 20, 30, None):
        for rep in range(3):
            cfg = tmkphy.make_config(kind, local_addr=1, chunk=chunk)
            caps.append(helpers.make_capture(orc, cfg, rng, n_frames=8, snr_db=snr,
                                             amplitude=float(rng.uniform(0.3, 1.0))))
            cfgs.append(cfg)
    want = check_batch(engine, orc, caps, cfgs, f"snr kind={kind} chunk={chunk}")
    assert (want["status"] == 0).sum() > 100


@pytest.mark.parametrize("kind", [FB, MAN])
def test_decode_multipath(engine, orc, kind):
    import tmkphy
    rng = np.random.default_rng(300 + kind)
    caps, cfgs = [], []
    for i in range(24):
        cfg = tmkphy.make_config(kind, local_addr=1, chunk=(0, 1200)[i % 2])
        caps.append(helpers.make_capture(orc, cfg, rng, n_frames=6, snr_db=float(rng.choice([8, 12, 20])),
                                         multipath=True))
        cfgs.append(cfg)
    check_batch(engine, orc, caps, cfgs, "multipath")


def test_mixed_batch_ragged_and_edge_lengths(engine, orc):
    """mixed line codes / chunk sizes / local addresses in one batch; empty and too-short streams;
    captures cut in the middle of a frame (TRUNCATED) and of a preamble."""
    import tmkphy
    rng = np.random.default_rng(5)
    caps, cfgs = [], []
    for i in range(40):
        kind = i % 2
        cfg = tmkphy.make_config(kind, local_addr=(1, 2, -1)[i % 3], chunk=(0, 1200, 48, 5000)[i % 4])
        c = helpers.make_capture(orc, cfg, rng, n_frames=4, snr_db=(None, 12)[i % 2 == 0])
        if i % 5 == 0:
            c = c[:int(rng.integers(1, len(c)))]
        caps.append(c)
        cfgs.append(cfg)
    for n in (0, 1, 59, 60, 61, 95, 96, 97, 300):
        caps.append(np.ones(n, np.float32))
        cfgs.append(tmkphy.make_config(n % 2))
    want = check_batch(engine, orc, caps, cfgs, "mixed")
    assert (want["status"] == 6).any() and (want["status"] == 4).any()


@pytest.mark.parametrize("kind", [FB, MAN])
def test_corrupted_frames(engine, orc, kind):
    """sample flips inside header and payload: HDR_PARSE / BAD_LEN / SHORT_DECODE / CRC_FAIL verdicts
    and their consume rules (decoder.rs:266-355)."""
    import tmkphy
    rng = np.random.default_rng(17 + kind)
    caps, cfgs = [], []
    P = 60 if kind == FB else 96
    for i in range(60):
        cfg = tmkphy.make_config(kind, local_addr=1, chunk=(0, 1200)[i % 2])
        c = helpers.make_capture(orc, cfg, rng, n_frames=5, snr_db=None, lead=100)
        for _ in range(int(rng.integers(1, 6))):
            pos = int(rng.integers(100 + P, len(c) - 400))
            ln = int(rng.integers(3, 40))
            c[pos:pos + ln] = -c[pos:pos + ln] if rng.random() < 0.7 else 0.0
        caps.append(c)
        cfgs.append(cfg)
    want = check_batch(engine, orc, caps, cfgs, "corrupt")
    assert len(set(want["status"])) >= 4


def test_amplitude_steps_and_silence(engine, orc):
    """ill-conditioned rolling energy: loud bursts followed by quiet frames, digital silence, DC offset.
    The reference's rolling window energy drifts here; the engine must still agree bit for bit."""
    import tmkphy
    rng = np.random.default_rng(23)
    caps, cfgs = [], []
    for i in range(32):
        kind = i % 2
        cfg = tmkphy.make_config(kind, local_addr=1, chunk=(0, 1200)[(i // 2) % 2])
        quiet = helpers.make_capture(orc, cfg, rng, n_frames=4, snr_db=(None, 25, 40)[i % 3], amplitude=1e-3 * (1 + i))
        loud = rng.normal(0, 30.0, int(rng.integers(200, 3000))).astype(np.float32)
        c = np.concatenate([loud, quiet, loud[:500] * 0.01, quiet * 7.0])
        if i % 4 == 3:
            c = c + 0.5  # DC offset
        caps.append(c.astype(np.float32))
        cfgs.append(cfg)
    check_batch(engine, orc, caps, cfgs, "amplitude steps", corr_tol=None)


def test_non_finite_samples(engine, orc):
    import tmkphy
    rng = np.random.default_rng(29)
    caps, cfgs = [], []
    for i in range(8):
        cfg = tmkphy.make_config(i % 2, local_addr=1, chunk=(0, 1200)[(i // 2) % 2])
        c = helpers.make_capture(orc, cfg, rng, n_frames=5, snr_db=15)
        c[int(rng.integers(0, len(c)))] = (np.nan, np.inf, -np.inf, 3e38)[i % 4]
        caps.append(c)
        cfgs.append(cfg)
    check_batch(engine, orc, caps, cfgs, "non-finite", corr_tol=None)


@pytest.mark.parametrize("kind,spl,pb", [(MAN, 2, 2), (FB, 4, 2), (FB, 3, 1), (MAN, 3, 4), (FB, 2, 3)])
def test_other_shapes_decode(engine, orc, kind, spl, pb):
    """the general-shape scan kernel: spl / preamble_bytes other than the shipped constants"""
    import tmkphy
    rng = np.random.default_rng(31 + spl + pb)
    caps, cfgs = [], []
    for i in range(10):
        cfg = tmkphy.make_config(kind, local_addr=1, spl=spl, preamble_bytes=pb, chunk=(0, 1200)[i % 2])
        caps.append(helpers.make_capture(orc, cfg, rng, n_frames=4, snr_db=(None, 14)[i % 2], max_len=40))
        cfgs.append(cfg)
    check_batch(engine, orc, caps, cfgs, f"shape {kind}/{spl}/{pb}")


@pytest.mark.parametrize("kind", [FB, MAN])
def test_correlation_values(engine, orc, kind):
    """north_star asks for correlation values within 1e-5 relative; the probe replays the reference's
    rolling energy and AVX-order dot literally, so they are equal bit for bit"""
    import tmkphy
    rng = np.random.default_rng(41 + kind)
    cfg = tmkphy.make_config(kind, local_addr=1)
    c = helpers.make_capture(orc, cfg, rng, n_frames=5, snr_db=10)
    s, st = helpers.pack([c], cfg, tmkphy)
    got = engine.correlate(s, st)[:len(c) - (60 if kind == FB else 96) + 1]
    want = orc.correlate(cfg, c)
    assert len(got) == len(want)
    assert np.array_equal(got, want)
    assert want.max() > 0.9


def test_channel_busy(engine, orc):
    import tmkphy
    rng = np.random.default_rng(43)
    caps = [np.zeros(19, np.float32), np.zeros(20, np.float32), rng.normal(0, 0.1, 500).astype(np.float32),
            rng.normal(0, 0.4, 500).astype(np.float32), np.full(33, 0.5, np.float32), np.full(33, -0.50001, np.float32)]
    s, st = helpers.pack(caps, tmkphy.make_config(FB), tmkphy)
    got = engine.channel_busy(s, st)
    assert list(got) == [orc.channel_busy(c) for c in caps]


def test_capacity_and_errors(engine, orc):
    import ctypes as C
    import tmkphy
    from tmkphy import _abi
    rng = np.random.default_rng(47)
    cfg = tmkphy.make_config(FB, local_addr=1)
    c = helpers.make_capture(orc, cfg, rng, n_frames=6, snr_db=None)
    s, st = helpers.pack([c], cfg, tmkphy)
    recs = np.zeros(2, _abi.REC_DTYPE)
    pl = np.zeros(16, np.uint8)
    nr, npl = C.c_uint64(0), C.c_uint64(0)
    rc = engine._lib.tmk_decode_batch(engine._h, s.ctypes.data_as(C.c_void_p), s.size, st.ctypes.data_as(C.c_void_p),
                                      1, recs.ctypes.data_as(C.c_void_p), 2, C.byref(nr), pl.ctypes.data_as(C.c_void_p),
                                      16, C.byref(npl))
    assert rc == _abi.E_CAPACITY and nr.value == 6  # counts still returned
    bad = st.copy()
    bad["offset"] = 2
    with pytest.raises(tmkphy.TmkError):
        engine.decode_batch(s, bad)
    bad = st.copy()
    bad["cfg"]["threshold"] = 0.0
    with pytest.raises(tmkphy.TmkError):
        engine.decode_batch(s, bad)
    bad = st.copy()
    bad["n_samples"] = s.size + 1
    with pytest.raises(tmkphy.TmkError):
        engine.decode_batch(s, bad)


def test_synth_batch_and_device_resident_decode(engine, orc):
    """the bench's generator: GPU-made captures decode identically on the GPU and in the oracle,
    and at 30 dB every frame the generator planted comes back."""
    import tmkphy
    for kind in (FB, MAN):
        cfg = tmkphy.make_config(kind, local_addr=1)
        n_streams, n_per = 48, 48000
        dptr = engine.device_alloc(n_streams * n_per * 4)
        try:
            _, truth = engine.synth_batch(cfg, 1234 + kind, n_streams, n_per, 30.0, 0.8, device_ptr=dptr)
            st, total = tmkphy.make_streams([n_per] * n_streams, cfg)
            assert total == n_streams * n_per
            got, gpl = engine.decode_batch(dptr, st, total_samples=total)
            host = np.zeros(total, np.float32)
            engine.d2h(host, dptr)
        finally:
            engine.device_free(dptr)
        want, wpl = oracle_batch(orc, host, st)
        helpers.assert_records_equal(got, gpl, want, wpl, "synth")
        assert len(truth) > 300
        assert np.array_equal(got["preamble_start"], truth["preamble_start"])
        assert np.array_equal(got["stream"], truth["stream"]) and np.array_equal(got["len"], truth["len"])
        ok = truth["dst"] == 1
        assert (got["status"][ok] == 0).all() and (got["status"][~ok] == 4).all()
        assert np.array_equal(got["crc"], truth["crc"])


def test_stats_and_launch_count(engine, orc):
    import tmkphy
    rng = np.random.default_rng(53)
    cfg = tmkphy.make_config(FB, local_addr=1)
    c = helpers.make_capture(orc, cfg, rng, n_frames=6, snr_db=20)
    s, st = helpers.pack([c], cfg, tmkphy)
    engine.stats_reset()
    n0 = engine.launch_count()
    recs, _ = engine.decode_batch(s, st)
    stt = engine.stats()
    assert engine.launch_count() >= n0 + 2
    assert stt["samples"] == len(c) and stt["detections"] == len(recs)
    assert stt["frames_ok"] == int((recs["status"] == 0).sum())
    assert stt["not_for_us"] == int((recs["status"] == 4).sum())


@pytest.mark.parametrize("kind,spl,pb", [(FB, 3, 2), (MAN, 3, 2), (FB, 4, 2), (MAN, 2, 1)])
def test_scan_filter_contract(engine, orc, kind, spl, pb):
    """the scan kernels (fast for the shipped constants, general otherwise) against float64 numpy:
    every window with corr >= 0.85/0.9*thr is marked, none clearly below is; block summaries agree"""
    import tmkphy
    rng = np.random.default_rng(61 + kind + spl)
    cfg = tmkphy.make_config(kind, local_addr=1, spl=spl, preamble_bytes=pb)
    caps = [helpers.make_capture(orc, cfg, rng, n_frames=12, snr_db=snr, amplitude=amp, max_len=60)
            for snr, amp in ((6, 1.0), (10, 0.3), (None, 1.0), (3, 50.0), (20, 1e-3))]
    caps.append(caps[0][:4000])
    caps.append(caps[1][:481 + len(orc.preamble(kind, spl, pb))])
    s, st = helpers.pack(caps, cfg, tmkphy)
    info, bits, base = engine.scan_probe(s, st)
    B = engine.scan_block()
    pre = orc.preamble(kind, spl, pb).astype(np.float64)
    P = len(pre)
    thr_c = np.float32(0.9) * np.float32(0.9444444)
    for i, c in enumerate(caps):
        x = c.astype(np.float64)
        nwin = len(x) - P + 1
        D = np.correlate(x, pre, mode="valid")
        cs = np.concatenate([[0.0], np.cumsum(x * x)])
        E = cs[P:] - cs[:-P]
        corr = np.where(E > 0, D / np.sqrt(np.maximum(E, 1e-300) * P), 0.0)
        nb = (nwin + B - 1) // B
        b0 = int(base[i])
        got_bits = np.unpackbits(bits[b0:b0 + nb].view(np.uint8), bitorder="little").reshape(nb, B)[:, :B].reshape(-1)[:nwin]
        must = corr >= thr_c * 1.002
        must_not = corr < thr_c * 0.998
        assert got_bits[must].all(), f"capture {i}: {(must & (got_bits == 0)).sum()} windows above the filter level are unmarked"
        assert not got_bits[must_not].any(), f"capture {i}: windows below the filter level are marked"
        for b in range(nb):
            e = E[b * B:(b + 1) * B]
            cc = corr[b * B:(b + 1) * B]
            emin, emax, rmax, ncand = info[b0 + b]
            tol = 1e-3 * e.max() + 1e-30
            assert abs(emin - e.min()) <= tol and abs(emax - e.max()) <= tol, (i, b, emin, e.min(), emax, e.max())
            want_r = max(0.0, cc.max()) ** 2
            # max corr^2 is an upper bound the walk relies on; tight when the block is well conditioned
            assert rmax >= want_r * (1 - 2e-3) - 1e-6, (i, b, rmax, want_r)
            if e.min() >= 1e-3 * e.max():
                assert rmax <= want_r * (1 + 2e-3) + 1e-6, (i, b, rmax, want_r)
            assert ncand == got_bits[b * B:(b + 1) * B].sum()


def test_exactly_sized_caller_buffers(engine, orc):
    """record and payload buffers sized to what the oracle returns must be enough: the first pass counts bytes of
    speculative decodes no walk links (false preambles inside frame bodies) and of segment walkers before their meeting
    point; the engine recounts with the plain walk before it calls a buffer too small"""
    import os
    import tmkphy
    rng = np.random.default_rng(2025)
    for kind in (FB, MAN):
        cfg = tmkphy.make_config(kind, local_addr=1)
        caps = [helpers.make_capture(orc, cfg, rng, n_frames=40, snr_db=s, gap_range=(48, 300)) for s in (None, 15, 8, 6, 4)]
        for c in caps:  # flipped stretches: in-body false preambles, header failures
            for _ in range(6):
                a = int(rng.integers(0, len(c) - 500))
                c[a:a + int(rng.integers(30, 400))] *= -1.0
        s, st = helpers.pack(caps, cfg, tmkphy)
        want, wpl = [], []
        base = 0
        for i, c in enumerate(caps):
            r, p = orc.decode_stream(cfg, c, stream_id=i)
            r["payload_off"] += base
            base += len(p)
            want.append(r)
            wpl.append(p)
        want, wpl = np.concatenate(want), np.concatenate(wpl)
        saved = os.environ.get("TMK_WALK_SEG_MIN")
        try:
            for seg_min in (None, "8192"):
                if seg_min:
                    os.environ["TMK_WALK_SEG_MIN"] = seg_min
                got, gpl = engine.decode_batch(s, st, rec_cap=len(want), payload_cap=max(1, len(wpl)))
                helpers.assert_records_equal(got, gpl, want, wpl, f"exact caps kind={kind} seg_min={seg_min}")
                if len(want) > 1:  # one record short: refused, with the exact count (the Python wrapper would grow and retry)
                    import ctypes as C
                    from tmkphy import _abi
                    recs = np.zeros(len(want) - 1, _abi.REC_DTYPE)
                    pl = np.zeros(len(wpl), np.uint8)
                    nr, npl = C.c_uint64(0), C.c_uint64(0)
                    rc = engine._lib.tmk_decode_batch(engine._h, s.ctypes.data_as(C.c_void_p), s.size, st.ctypes.data_as(C.c_void_p),
                                                      len(st), recs.ctypes.data_as(C.c_void_p), len(recs), C.byref(nr),
                                                      pl.ctypes.data_as(C.c_void_p), len(pl), C.byref(npl))
                    assert rc == _abi.E_CAPACITY and nr.value == len(want) and npl.value == len(wpl)
        finally:
            if saved is None:
                os.environ.pop("TMK_WALK_SEG_MIN", None)
            else:
                os.environ["TMK_WALK_SEG_MIN"] = saved


@pytest.mark.parametrize("kind", [FB, MAN])
def test_frame_length_limits(engine, orc, kind):
    """payload lengths on the limits of the length rule (decoder.rs:277-286): ACK with 0, DATA with 0 (refused), 1, 128
    (what the encoder's callers send at most), 255, 256 (max_frame_bytes: still accepted), 257 and 300 (refused: consume 1
    sample, the body is then searched for preambles), and a decoder configured with a smaller max_frame_bytes"""
    import tmkphy
    rng = np.random.default_rng(31 + kind)
    lens = [(2, 0), (1, 0), (1, 1), (1, 128), (1, 255), (1, 256), (1, 257), (1, 300), (1, 7), (2, 5)]
    frames = [(t, i, 9, 1, bytes(rng.integers(0, 256, n, dtype=np.uint8))) for i, (t, n) in enumerate(lens)]
    for max_bytes, chunk in ((256, 0), (256, 1200), (64, 0), (300, 0)):
        cfg = tmkphy.make_config(kind, local_addr=1, chunk=chunk)
        cfg.max_frame_bytes = max_bytes
        parts = [np.zeros(300, np.float32)]
        for f in frames:
            parts += [orc.modulate(cfg, [f], 0), np.zeros(int(rng.integers(60, 400)), np.float32)]
        x = np.concatenate(parts)
        caps = [x, (x + rng.normal(0, 0.15, len(x))).astype(np.float32)]
        want = check_batch(engine, orc, caps, cfg, f"length limits kind={kind} max={max_bytes} chunk={chunk}")
        ok = set(int(v) for v in want[(want["status"] == 0) & (want["stream"] == 0)]["len"])
        assert {0, 1, 7, 5} <= ok and (128 in ok) == (max_bytes >= 128) and (256 in ok) == (max_bytes >= 256)
        assert (300 in ok) == (max_bytes >= 300) and (257 in ok) == (max_bytes >= 257)
        assert (want["status"] == 2).sum() >= (1 if max_bytes >= 300 else 3)  # BAD_LEN: DATA with len 0, frames beyond max_frame_bytes
