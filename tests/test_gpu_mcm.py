"""Parity of the multi-carrier modem (FSK / OFDM demodulation on the batched FFT, csrc/mcm.cu) through the C ABI against
its self-oracle (oracle/mcm_modem.hpp). The specification is build-defined — the reference has no FFT / FSK / OFDM code —
so these results are "parity unpinned by the reference"; what is compared: spectra, modulated tracks, sync offsets,
sync power, verdicts, header bytes and payload bytes, all bit for bit."""
import ctypes as C
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
FSK, OFDM = 0, 1
CAP = {FSK: 25, OFDM: 100}


def oracle_batch(orc, mode, s, st):
    recs, pays, base = [], [], 0
    for i, t in enumerate(st):
        r, p = orc.mcm_demodulate(mode, s[int(t["offset"]):int(t["offset"]) + int(t["n_samples"])], stream=i)
        r["payload_off"] += base
        base += len(p)
        recs.append(r)
        pays.append(p)
    return (np.concatenate(recs) if recs else np.zeros(0)), (np.concatenate(pays) if pays else np.zeros(0, np.uint8))


def assert_mcm_equal(got, gpl, want, wpl, ctx=""):
    assert len(got) == len(want), f"{ctx}: {len(got)} records, oracle {len(want)}"
    for f in ("stream", "status", "start_index", "seq", "len", "crc", "mode"):
        assert np.array_equal(got[f], want[f]), f"{ctx}: field {f} differs"
    assert np.array_equal(got["sync_power"].view(np.uint32), want["sync_power"].view(np.uint32)), f"{ctx}: sync_power"
    for g, w in zip(got, want):
        if w["status"] != 2:
            a = bytes(gpl[int(g["payload_off"]):int(g["payload_off"]) + int(g["len"])])
            b = bytes(wpl[int(w["payload_off"]):int(w["payload_off"]) + int(w["len"])])
            assert a == b, f"{ctx}: payload differs at start_index {w['start_index']}"


def multipath(x, rng):
    h = np.zeros(41)
    h[0], h[int(rng.integers(2, 13))], h[int(rng.integers(13, 41))] = 1.0, 0.35, -0.2
    h /= np.sqrt((h * h).sum())
    return np.convolve(x.astype(np.float64), h)[:len(x)]


@pytest.mark.parametrize("n", [64, 256])
def test_fft_bit_exact(engine, orc, n):
    rng = np.random.default_rng(100 + n)
    rows = rng.normal(0, 1, (257, n)).astype(np.float32)
    rows[0] = 0
    rows[1, :] = 1e-30   # subnormal products
    rows[2, 3] = np.inf
    rows[3, 5] = np.nan
    rows[4] = 3e18       # squares overflow later, sums do not
    got, want = engine.mcm_fft(rows), orc.mcm_fft(rows)
    g = np.stack([got.real, got.imag]).astype(np.float32).view(np.uint32)
    w = np.stack([want.real, want.imag]).astype(np.float32).view(np.uint32)
    nan = np.isnan(np.stack([want.real, want.imag]))
    assert np.array_equal(g[~nan], w[~nan]) and np.isnan(np.stack([got.real, got.imag])[nan]).all()
    ref = np.fft.fft(rows[5:].astype(np.float64), axis=1)
    assert (np.abs(got[5:] - ref).max(axis=1) / np.abs(ref).max(axis=1) < 2e-6).all()


@pytest.mark.parametrize("mode", [FSK, OFDM])
def test_modulate_bit_exact_and_loopback(engine, orc, mode):
    import tmkphy
    text = open(os.path.join(GOLDEN, "think-different.txt"), "rb").read()
    pays = [text[i:i + CAP[mode]] for i in range(0, len(text), CAP[mode])] + [b""]
    rng = np.random.default_rng(5)
    g1, g2 = rng.integers(0, 300, len(pays)), rng.integers(60, 300, len(pays))
    seqs = np.arange(len(pays)) & 0xFF
    got, want = engine.mcm_modulate(mode, pays, seqs, g1, g2), orc.mcm_modulate(mode, pays, seqs, g1, g2)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    x = np.concatenate([got, np.zeros(600, np.float32)])
    st, _ = tmkphy.make_streams([len(x)], tmkphy.make_config(1))
    recs, pl = engine.mcm_decode_batch(mode, x, st)
    wrecs, wpl = oracle_batch(orc, mode, x, st)
    assert_mcm_equal(recs, pl, wrecs, wpl, "loopback")
    assert len(recs) == len(pays) and (recs["status"] == 0).all()
    # (payload bytes sit wherever the warps' atomic reservations put them: gather in record order)
    assert b"".join(bytes(pl[int(r["payload_off"]):int(r["payload_off"]) + int(r["len"])]) for r in recs) == text
    assert len(engine.mcm_modulate(mode, [], [], [], [])) == 0


@pytest.mark.parametrize("mode", [FSK, OFDM])
def test_decode_awgn_multipath_batch(engine, orc, mode):
    """ragged batch: clean, AWGN from 25 dB down to -6 dB (CRC failures, broken length bytes), multipath, an empty
    capture, a capture that ends inside a body, frames of the OTHER mode"""
    import tmkphy
    rng = np.random.default_rng(40 + mode)
    caps = []
    for k, snr in enumerate([None, 25, 15, 10, 6, 3, 0, -3, -6, 12, 12]):
        n_f = int(rng.integers(3, 9))
        pays = [bytes(rng.integers(0, 256, int(rng.integers(0, CAP[mode] + 1)), dtype=np.uint8)) for _ in range(n_f)]
        x = orc.mcm_modulate(mode, pays, rng.integers(0, 256, n_f), rng.integers(0, 500, n_f), rng.integers(100, 500, n_f))
        x = np.concatenate([x, np.zeros(int(rng.integers(300, 900)), np.float32)]).astype(np.float64)
        if k >= 9:
            x = multipath(x, rng)
        if snr is not None:
            x = x + rng.normal(0, 0.5 * 10 ** (-snr / 20), len(x))
        caps.append(x.astype(np.float32))
    caps.append(np.zeros(0, np.float32))
    caps.append(caps[1][:300 + 440 + 2500].copy())
    other = orc.mcm_modulate(1 - mode, [b"abc", b"defg"], [1, 2], [100, 100], [300, 300])
    caps.append(np.concatenate([other, np.zeros(500, np.float32)]))
    st, total = tmkphy.make_streams([len(c) for c in caps], tmkphy.make_config(1))
    s = np.zeros(total, np.float32)
    for c, t in zip(caps, st):
        s[int(t["offset"]):int(t["offset"]) + len(c)] = c
    recs, pl = engine.mcm_decode_batch(mode, s, st)
    wrecs, wpl = oracle_batch(orc, mode, s, st)
    assert_mcm_equal(recs, pl, wrecs, wpl, "batch")
    assert (recs["status"] == 0).sum() >= 20 and (recs["status"] != 0).sum() >= 5
    ms = engine.mcm_last_stage_ms()
    assert ms["filter"] > 0 and ms["demod"] > 0


@pytest.mark.parametrize("mode", [FSK, OFDM])
def test_synth_batch_parity(engine, orc, mode):
    """the bench's own generator (multipath + AWGN sweep), device-resident samples"""
    import tmkphy
    n_s, n_per = 24, 1 << 16
    d_x = engine.device_alloc(n_s * n_per * 4)
    try:
        planted = 0
        for g, snr in enumerate([20.0, 8.0, 2.0]):
            _, nf = engine.mcm_synth_batch(mode, 0x51 + g, 8, n_per, snr, multipath=(g != 0), device_ptr=d_x + g * 8 * n_per * 4)
            planted += nf
        st, total = tmkphy.make_streams([n_per] * n_s, tmkphy.make_config(1))
        recs, pl = engine.mcm_decode_batch(mode, d_x, st, total_samples=total)
        s = np.zeros(total, np.float32)
        engine.d2h(s, d_x)
    finally:
        engine.device_free(d_x)
    wrecs, wpl = oracle_batch(orc, mode, s, st)
    assert_mcm_equal(recs, pl, wrecs, wpl, "synth")
    # (the reference's lock rule needs sync_power > 2 power: with the chirp's own power at 0.5 it holds only at high SNR,
    #  ask.rs:711-714 — the noisy multipath groups lock rarely)
    first = recs[recs["stream"] < 8]
    assert planted >= n_s * 10 and len(first) >= 80 and (first["status"] == 0).mean() > 0.95


def test_errors(engine):
    import tmkphy
    x = np.zeros(8192, np.float32)
    st, _ = tmkphy.make_streams([len(x)], tmkphy.make_config(1))
    with pytest.raises(tmkphy.TmkError):
        engine.mcm_decode_batch(2, x, st)
    with pytest.raises(tmkphy.TmkError):
        engine.mcm_modulate(FSK, [b"x" * 26], [0], [0], [0])
    with pytest.raises(tmkphy.TmkError):
        engine.mcm_fft(np.zeros((2, 128), np.float32))
    y = engine.mcm_modulate(OFDM, [b"a" * 50, b"b" * 50], [1, 2], [100, 100], [400, 400])
    st, _ = tmkphy.make_streams([len(y)], tmkphy.make_config(1))
    with pytest.raises(tmkphy.TmkError) as ei:
        engine.mcm_decode_batch(OFDM, y, st, rec_cap=1)
    assert ei.value.code == tmkphy._abi.E_CAPACITY
    recs = np.zeros(4, tmkphy._abi.MCM_REC_DTYPE)
    pay = np.zeros(60, np.uint8)
    n, pl = C.c_uint64(0), C.c_uint64(0)
    rc = engine._lib.tmk_mcm_decode_batch(engine._h, OFDM, y.ctypes.data_as(C.c_void_p), len(y), st.ctypes.data_as(C.c_void_p), 1,
                                          recs.ctypes.data_as(C.c_void_p), 4, C.byref(n), pay.ctypes.data_as(C.c_void_p), 60, C.byref(pl))
    assert rc == tmkphy._abi.E_CAPACITY and n.value == 2 and pl.value == 100
