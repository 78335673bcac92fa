"""GPU side of the round-2 boundary additions, all through the C ABI and against the oracle:
the impaired capture generator (multipath / clock drift / amplitude), the chirp + BPSK generator,
WAV container ingest (dump.rs:23-46 written, codec.rs:114-148 read), and the NCCL counter reduction."""
import struct

import numpy as np
import pytest

import helpers
from test_ingest import pcm_fmt, riff

pytestmark = pytest.mark.gpu
MAN, FB = 0, 1


@pytest.mark.parametrize("kind", [FB, MAN])
def test_synth_impairments_decode_bit_exact(engine, orc, kind):
    """captures with the 3-tap multipath FIR, +-100 ppm drift and per-stream amplitude: whatever they contain, the
    engine and the oracle decode them identically; most planted frames survive at 15 dB"""
    import tmkphy
    cfg = tmkphy.make_config(kind, local_addr=1)
    n_s, n_per = 24, 96000
    for imp in ({"multipath": True}, {"drift_ppm": 100.0, "amp_min": 0.3}, {"multipath": True, "drift_ppm": 60.0, "gap": (2000, 9000)}):
        x, truth = engine.synth_batch_ex(cfg, 4242 + kind, n_s, n_per, 15.0, 0.8, **imp)
        assert len(truth) > 100 and np.isfinite(x).all()
        st, _ = tmkphy.make_streams([n_per] * n_s, cfg)
        want, wpl, _ = orc.decode_batch(x, st)
        got, gpl = engine.decode_batch(x, st)
        helpers.assert_records_equal(got, gpl, want, wpl, f"synth {imp}", corr_tol=None)
        ok = int(np.isin(want["status"], (0, 4)).sum())
        # (the echoes put the preamble correlation near h0 = 0.93 +- its autocorrelation at the echo lags: with
        #  multipath a good part of the frames stays below the 0.9 threshold, for the oracle and the engine alike)
        assert ok > (0.25 if imp.get("multipath") else 0.6) * len(truth), (imp, ok, len(truth))
        if "drift_ppm" in imp:  # truth positions are rounded capture positions: detections sit within a sample of them
            t0 = truth[truth["stream"] == 0]["preamble_start"].astype(np.int64)
            d0 = want[(want["stream"] == 0) & np.isin(want["status"], (0, 4))]["preamble_start"].astype(np.int64)
            assert all(np.abs(t0 - d).min() <= 2 for d in d0)
    # no impairment requested: the _ex entry is the plain generator, bit for bit
    a, ta = engine.synth_batch(cfg, 99, 4, 48000, 10.0, 0.7)
    b, tb = engine.synth_batch_ex(cfg, 99, 4, 48000, 10.0, 0.7)
    assert np.array_equal(a, b) and np.array_equal(ta, tb)


def test_ask_synth_decodes_bit_exact(engine, orc):
    import tmkphy
    from test_gpu_ask import assert_ask_equal, oracle_batch
    n_s, n_per = 12, 1 << 16
    x, planted = engine.ask_synth_batch(7, n_s, n_per, 10.0)
    assert planted >= n_s * 11
    st, _ = tmkphy.make_streams([n_per] * n_s, tmkphy.make_config(FB))
    want = oracle_batch(orc, x, st)
    got = engine.ask_decode_batch(x, st)
    assert_ask_equal(got, want, "ask synth")
    # (frames follow each other after 0..200 samples like the reference's TX loop: the correlator restarts empty behind
    #  a frame, ask.rs:717-723, so the receiver — the reference's included — loses about every other one)
    assert int(want["accepted"].sum()) > 0.4 * planted
    ids = want[(want["stream"] == 3) & (want["accepted"] == 1)]["id"].astype(int)
    assert len(ids) > 3 and (np.diff(ids) > 0).all() and ids[0] <= 2  # ids count up along the capture


def test_wav_ingest_formats(engine, orc):
    """every sample type of convert_audio_buffer_to_f32 (codec.rs:114-148), channel 0 of interleaved files, chunk
    walk with extra chunks: the device unpack equals the oracle's numpy reader bit for bit"""
    import tmkphy
    rng = np.random.default_rng(11)
    files = []
    files.append(tmkphy.wav_bytes(rng.integers(-32768, 32768, 1001).astype(np.int16)))                      # S16 mono, odd length
    files.append(riff(pcm_fmt(2, 44100, 16), rng.integers(0, 256, 4 * 777, dtype=np.uint8).tobytes()))      # S16 stereo
    files.append(riff(pcm_fmt(1, 8000, 8), rng.integers(0, 256, 333, dtype=np.uint8).tobytes()))            # U8
    ext = pcm_fmt(3, 48000, 24, tag=0xFFFE) + struct.pack("<HHI", 22, 24, 7) + struct.pack("<H", 1) + b"\0" * 14
    files.append(riff(ext, rng.integers(0, 256, 9 * 210, dtype=np.uint8).tobytes()))                        # S24, 3 channels, extensible
    files.append(riff(pcm_fmt(1, 48000, 32), rng.integers(0, 256, 4 * 500, dtype=np.uint8).tobytes(),
                      extra=b"LIST" + struct.pack("<I", 3) + b"abc\0"))                                     # S32, odd LIST chunk
    f32 = rng.normal(0, 1, 256).astype(np.float32)
    f32[:4] = [np.inf, -0.0, 1e-42, np.float32(3.4e38)]
    files.append(riff(pcm_fmt(1, 48000, 32, tag=3), f32.tobytes()))                                         # F32
    files.append(tmkphy.wav_bytes(np.zeros(0, np.int16)))                                                   # empty data chunk
    got = engine.wav_to_f32(files)
    for i, f in enumerate(files):
        want = orc.wav_to_f32(f)
        assert got[i].dtype == np.float32 and np.array_equal(got[i].view(np.uint32), want.view(np.uint32)), i
    with pytest.raises(tmkphy.TmkError):
        engine.wav_to_f32([files[0], b"RIFFxxxxWAVEjunk"])


@pytest.mark.parametrize("kind", [FB, MAN])
def test_decode_batch_wav(engine, orc, kind):
    """recordings as the reference stores them (dump_to_wav: (sample * 32767) as i16 into a 16-bit WAV) decoded from
    the file bytes: records equal the oracle's decode of the samples its reader returns"""
    import tmkphy
    rng = np.random.default_rng(21 + kind)
    cfg = tmkphy.make_config(kind, local_addr=1)
    files, caps = [], []
    for i in range(6):
        x = helpers.make_capture(orc, cfg, rng, n_frames=5, snr_db=float(rng.choice([10, 20])), amplitude=0.6)
        pcm = engine.pcm16_encode(x)
        assert np.array_equal(pcm, orc.pcm16_encode(x))
        if i % 2:  # a stereo recording: channel 1 is another capture, the reader keeps channel 0
            other = rng.integers(-3000, 3000, len(pcm)).astype(np.int16)
            files.append(riff(pcm_fmt(2, 48000, 16), np.stack([pcm, other], 1).tobytes()))
        else:
            files.append(tmkphy.wav_bytes(pcm))
        caps.append(orc.wav_to_f32(files[-1]))
        assert np.array_equal(caps[-1], orc.pcm16_to_f32(pcm))
    got, gpl, st = engine.decode_batch_wav(files, cfg)
    assert [int(n) for n in st["n_samples"]] == [len(c) for c in caps]
    s, st2 = helpers.pack(caps, [cfg] * len(caps), tmkphy)
    assert np.array_equal(st["offset"], st2["offset"])
    want, wpl, _ = orc.decode_batch(s, st2)
    helpers.assert_records_equal(got, gpl, want, wpl, "wav", corr_tol=None)
    assert (want["status"] == 0).sum() >= 15


def test_comm_single_rank_and_errors(engine):
    """libnccl is opened at run time; a one-rank communicator reduces to the identity (the 2+ rank path runs under
    torchrun in bench.py --gpus N and in tools/comm_check.py)"""
    import tmkphy
    assert engine.comm_size() == 1
    assert engine.allreduce_sum([1, 2, 3]) == [1, 2, 3] and engine.allreduce_max([2.5]) == [2.5]
    assert engine.allgather([7, 1 << 50]) == [[7, 1 << 50]]
    uid = tmkphy.comm_unique_id()
    assert len(uid) == 128 and any(uid)
    engine.comm_init(uid, 0, 1)
    try:
        assert engine.comm_size() == 1
        assert engine.allreduce_sum([5, 1 << 40]) == [5, 1 << 40]
        assert engine.allreduce_max([-1.0, 3.25]) == [-1.0, 3.25]
        assert engine.allgather([9, 8, 7]) == [[9, 8, 7]]  # ncclAllGather over one rank
        st = engine.stats_allreduce()
        assert st == engine.stats()
        engine.comm_barrier()
        with pytest.raises(tmkphy.TmkError):
            engine.comm_init(uid, 0, 1)  # one communicator per engine
    finally:
        engine.comm_destroy()
    with pytest.raises(tmkphy.TmkError):
        engine.comm_init(uid, 2, 2)


@pytest.mark.parametrize("kind", [1, 0])
def test_host_generator_matches_device(engine, orc, kind):
    """oracle/synth_host.py (what `bench.py --impl reference` decodes) against tmk_synth_batch: the same frames at the same
    offsets with the same bytes, the noise equal up to the rounding of the device's fast log / sin / cos"""
    import tmkphy
    from oracle import synth_host
    cfg = tmkphy.make_config(kind, local_addr=1)
    seed, n_s, n_per = 0x7A4D + 3, 6, 48000
    for snr in (10.0, 0.0, 100.0):
        x, truth = engine.synth_batch(cfg, seed, n_s, n_per, snr, 0.7)
        x = x.reshape(n_s, n_per)
        for s in (0, 3, 5):
            h, frames = synth_host.capture(orc, cfg, seed, s, n_per, snr, 0.7)
            t = truth[truth["stream"] == s]
            assert [int(v) for v in t["preamble_start"]] == [f[0] for f in frames] and len(frames) > 3
            assert [(int(r["type"]), int(r["seq"]), int(r["src"]), int(r["dst"]), int(r["len"])) for r in t] == [f[1:] for f in frames]
            sigma = 0.0 if snr >= 100 else 0.7 * 10 ** (-snr / 20)
            assert np.abs(h - x[s]).max() <= 1e-5 * max(1.0, 8 * sigma), (snr, s, np.abs(h - x[s]).max())
            if snr >= 100:
                assert np.array_equal(h, x[s])  # no noise: the frames alone, exactly
            # and the decoder sees the same frames in both
            a, pa = orc.decode_stream(cfg, h)
            b, pb = orc.decode_stream(cfg, x[s])
            if snr >= 10:
                assert np.array_equal(a["frame_start"], b["frame_start"]) and bytes(pa) == bytes(pb)
