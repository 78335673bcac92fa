"""Seeded synthetic captures for the parity tests: frames modulated by the ORACLE
encoder, random gaps, AWGN, 3-tap multipath, amplitude scaling."""
import numpy as np


def random_frames(rng, n, local_addr=1, max_len=128, p_ack=0.1, p_foreign=0.1):
    frames = []
    for i in range(n):
        ack = rng.random() < p_ack
        ln = 0 if ack else int(rng.integers(1, max_len + 1))
        dst = local_addr if rng.random() >= p_foreign else (local_addr + 1 + int(rng.integers(0, 200))) % 256
        frames.append((2 if ack else 1, i & 0xFF, int(rng.integers(0, 256)), dst,
                       bytes(rng.integers(0, 256, ln, dtype=np.uint8))))
    return frames


def make_capture(orc, cfg, rng, n_frames=6, snr_db=None, multipath=False, amplitude=1.0, lead=None,
                 gap_range=(48, 1500), tail=200, max_len=128):
    """-> float32 capture"""
    parts = [np.zeros(int(rng.integers(0, 900)) if lead is None else lead, np.float32)]
    for f in random_frames(rng, n_frames, cfg.local_addr if cfg.local_addr >= 0 else 1, max_len):
        parts.append(orc.modulate(cfg, [f], 0))
        parts.append(np.zeros(int(rng.integers(*gap_range)), np.float32))
    parts.append(np.zeros(tail, np.float32))
    x = np.concatenate(parts).astype(np.float64) * amplitude
    if multipath:
        h = np.zeros(41)
        h[0] = 1.0
        h[int(rng.integers(2, 13))] = 0.35
        h[int(rng.integers(13, 41))] = -0.2
        h /= np.sqrt((h * h).sum())
        x = np.convolve(x, h)[:len(x)]
    if snr_db is not None:
        x = x + rng.normal(0.0, amplitude * 10 ** (-snr_db / 20.0), len(x))
    return x.astype(np.float32)


def pack(captures, cfgs, tmkphy):
    """ragged batch -> (samples, streams)"""
    st, total = tmkphy.make_streams([len(c) for c in captures], cfgs)
    s = np.zeros(total, np.float32)
    for i, c in enumerate(captures):
        o = int(st[i]["offset"])
        s[o:o + len(c)] = c
    return s, st


def assert_records_equal(got, got_pl, want, want_pl, ctx="", corr_tol=1e-5):
    """verdicts, offsets, header fields and payload bytes bit-exact; corr within corr_tol
    (None: skip — the record's corr uses an accurately summed window energy, the reference a rolling one
    that drifts on ill-conditioned captures; tmk_correlate is the bit-exact probe)"""
    assert len(got) == len(want), f"{ctx}: {len(got)} records, oracle {len(want)}"
    for f in ("stream", "status", "preamble_start", "frame_start", "consumed", "len", "crc", "type", "seq", "src",
              "dst", "hdr_valid"):
        if not np.array_equal(got[f], want[f]):
            i = int(np.nonzero(got[f] != want[f])[0][0])
            raise AssertionError(f"{ctx}: field {f} differs at record {i}: {got[i]} vs oracle {want[i]}")
    if corr_tol is not None:
        tol = corr_tol * np.maximum(np.abs(want["corr"]), 1.0)
        assert (np.abs(got["corr"] - want["corr"]) <= tol).all(), f"{ctx}: corr differs"
    for g, w in zip(got, want):
        if w["status"] in (0, 4, 5):
            a = bytes(got_pl[int(g["payload_off"]):int(g["payload_off"]) + int(g["len"])])
            b = bytes(want_pl[int(w["payload_off"]):int(w["payload_off"]) + int(w["len"])])
            assert a == b, f"{ctx}: payload differs at preamble_start {w['preamble_start']}"
