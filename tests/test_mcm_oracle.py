"""The self-oracle of the multi-carrier modem (oracle/mcm_modem.hpp; build-defined FSK / OFDM specification — the
reference has no such code, parity unpinned by the reference). CPU only: the specification's FFT against numpy.fft,
loopback of the reference's own text asset, AWGN / multipath behaviour, verdicts."""
import os

import numpy as np
import pytest

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
FSK, OFDM = 0, 1
CAP = {FSK: 25, OFDM: 100}


def chunks(text, n):
    return [text[i:i + n] for i in range(0, len(text), n)]


def multipath(x, rng):
    h = np.zeros(41)
    h[0], h[int(rng.integers(2, 13))], h[int(rng.integers(13, 41))] = 1.0, 0.35, -0.2
    h /= np.sqrt((h * h).sum())
    return np.convolve(x.astype(np.float64), h)[:len(x)]


@pytest.mark.parametrize("n", [64, 256])
def test_fft_matches_numpy(orc, n):
    rng = np.random.default_rng(n)
    rows = rng.normal(0, 1, (50, n)).astype(np.float32)
    rows[0] = 0
    rows[1] = 1
    rows[2] = np.cos(2 * np.pi * 5 * np.arange(n) / n)
    got = orc.mcm_fft(rows)
    want = np.fft.fft(rows.astype(np.float64), axis=1)
    err = np.abs(got - want).max(axis=1)
    scale = np.maximum(np.abs(want).max(axis=1), 1e-30)
    assert (err[1:] / scale[1:] < 2e-6).all()
    assert (got[0] == 0).all()
    assert abs(got[2][5] - n / 2) < 1e-3 * n and abs(got[2][n - 5] - n / 2) < 1e-3 * n


@pytest.mark.parametrize("mode", [FSK, OFDM])
def test_loopback_think_different(orc, mode):
    """every byte of the reference's text asset through modulate -> demodulate, clean channel"""
    text = open(os.path.join(GOLDEN, "think-different.txt"), "rb").read()
    pays = chunks(text, CAP[mode])
    rng = np.random.default_rng(3)
    g1, g2 = rng.integers(50, 300, len(pays)), rng.integers(50, 300, len(pays))
    x = orc.mcm_modulate(mode, pays, np.arange(len(pays)) & 0xFF, g1, g2)
    assert len(x) == len(pays) * (440 + 4752) + g1.sum() + g2.sum()
    recs, pl = orc.mcm_demodulate(mode, np.concatenate([x, np.zeros(600, np.float32)]))
    assert len(recs) == len(pays) and (recs["status"] == 0).all()
    assert [int(v) for v in recs["seq"]] == [i & 0xFF for i in range(len(pays))]
    assert bytes(pl) == text
    # the sync peak is the last sample of the chirp
    starts = np.cumsum(np.concatenate([[0], (g1 + g2 + 440 + 4752)[:-1]])) + g1 + 439
    assert np.array_equal(recs["start_index"], starts.astype(np.uint64))


@pytest.mark.parametrize("mode", [FSK, OFDM])
def test_noise_and_multipath(orc, mode):
    rng = np.random.default_rng(11 + mode)
    pays = [bytes(rng.integers(0, 256, int(rng.integers(1, CAP[mode] + 1)), dtype=np.uint8)) for _ in range(12)]
    x = orc.mcm_modulate(mode, pays, np.arange(12), rng.integers(100, 400, 12), rng.integers(100, 400, 12))
    x = np.concatenate([x, np.zeros(600, np.float32)])
    y = multipath(x, rng) + rng.normal(0, 0.5 * 10 ** (-20 / 20), len(x))  # 20 dB against the body power 0.25
    recs, pl = orc.mcm_demodulate(mode, y.astype(np.float32))
    assert len(recs) == 12 and (recs["status"] == 0).sum() >= 11
    ok = [bytes(pl[int(r["payload_off"]):int(r["payload_off"]) + int(r["len"])]) for r in recs if r["status"] == 0]
    assert all(p in pays for p in ok)
    # at -6 dB the frames still lock (the chirp is 3 dB above the body) but the payloads break: CRC failures, never a wrong OK
    y = x + rng.normal(0, 0.5 * 10 ** (6 / 20), len(x))
    recs, pl = orc.mcm_demodulate(mode, y.astype(np.float32))
    bad = [bytes(pl[int(r["payload_off"]):int(r["payload_off"]) + int(r["len"])]) for r in recs if r["status"] == 0]
    assert set(np.unique(recs["status"])) <= {0, 2, 5}
    assert sum(p not in pays for p in bad) <= 1  # an 8-bit checksum lets 1 in 256 broken frames through


def test_verdicts(orc):
    pays = [b"", b"x" * 100, b"hello"]
    x = orc.mcm_modulate(OFDM, pays, [7, 8, 9], [200] * 3, [200] * 3)
    x = np.concatenate([x, np.zeros(600, np.float32)])
    recs, pl = orc.mcm_demodulate(OFDM, x)
    assert list(recs["status"]) == [0, 0, 0] and list(recs["len"]) == [0, 100, 5] and list(recs["seq"]) == [7, 8, 9]
    assert recs["crc"][0] == 0 and bytes(pl) == b"x" * 100 + b"hello"
    # an FSK frame read as OFDM and the other way round: garbage bytes -> BAD_LEN or CRC_FAIL, records still one per lock
    recs, _ = orc.mcm_demodulate(FSK, x)
    assert len(recs) == 3 and set(recs["status"]) <= {2, 5, 0}
    # a capture that ends inside a body: the frame is never gathered, no record
    recs, _ = orc.mcm_demodulate(OFDM, x[:200 + 440 + 3000])
    assert len(recs) == 0
    assert len(orc.mcm_demodulate(OFDM, np.zeros(0, np.float32))[0]) == 0
