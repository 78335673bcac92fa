"""Two independent restatements of the reference's decoder against each other: the C++ oracle
(oracle/phy_decoder.hpp) and a numpy-f32 one written from the Rust (oracle/second_restatement.py) decode the same random
captures — one-shot and fed in random chunks — and must agree on every verdict, offset, header byte and payload byte.
Plus the reference's other stored fixture (ref/checkutils/proj2/INPUT.bin, 6250 B) as a loopback through every modem, and
the frame-length literal of docs/proj/report.md:166-169. None of this pins parity against the Rust binary itself (no
rustc here or on the GPU box); it removes single-reader error from the restatement."""
import os

import numpy as np
import pytest

import helpers

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
FB, MAN = 1, 0


def events_of(orc_recs, orc_pl):
    out = []
    for r in orc_recs:
        pay = bytes(orc_pl[int(r["payload_off"]):int(r["payload_off"]) + int(r["len"])]) if r["status"] in (0, 4, 5) else b""
        hdr = ([int(r["len"]) >> 8, int(r["len"]) & 255, int(r["crc"]), int(r["type"]), int(r["seq"]), int(r["src"]), int(r["dst"])]
               if r["hdr_valid"] else None)
        out.append((int(r["status"]), int(r["preamble_start"]), int(r["frame_start"]), int(r["consumed"]), hdr, pay))
    return out


@pytest.mark.parametrize("kind", [FB, MAN])
def test_two_restatements_agree(orc, kind):
    import tmkphy
    from oracle.second_restatement import PhyDecoderPy
    rng = np.random.default_rng(77 + kind)
    cfg = tmkphy.make_config(kind, local_addr=1)
    n_caps, n_events, statuses = 300, 0, set()
    for k in range(n_caps):
        snr = [None, 20.0, 10.0, 7.0, 6.0, 5.0, 3.0][k % 7]
        c = helpers.make_capture(orc, cfg, rng, n_frames=int(rng.integers(1, 4)), snr_db=snr, multipath=(k % 5 == 4),
                                 gap_range=(48, 400), tail=int(rng.integers(0, 300)), max_len=20,
                                 amplitude=float(rng.uniform(0.2, 1.0)))
        if k % 3 == 0 and len(c) > 600:  # flipped / zeroed stretches: every verdict and consume rule
            a = int(rng.integers(0, len(c) - 300))
            c[a:a + int(rng.integers(10, 300))] *= np.float32(-1.0 if k % 2 else 0.0)
        if k % 4 == 0:
            feeds = [len(c)]
        else:
            feeds, left = [], len(c)
            while left:
                f = min(left, int(rng.choice([1, 7, 48, 257, 1200, 3000])))
                feeds.append(f)
                left -= f
        recs, pl, _ = orc.decode_feeds(cfg, c, feeds)
        want = [e for e in events_of(recs, pl) if e[0] != 6]  # a frame still being gathered is state, not an event
        dec = PhyDecoderPy(kind, local_addr=1)
        pos = 0
        for f in feeds:
            dec.process_samples(c[pos:pos + f])
            pos += f
        got = [(s, p, fs, cons, (list(h) if h is not None else None), d) for s, p, fs, cons, h, d in dec.events]
        assert got == want, f"capture {k} (snr {snr}, feeds {feeds[:6]}...): {got} != {want}"
        n_events += len(want)
        statuses |= {e[0] for e in want}
    assert n_events > 300 and {0, 1, 5} <= statuses, (n_events, statuses)


def test_input_bin_loopback_all_modems(orc):
    """ref/checkutils/proj2/INPUT.bin (= OUTPUT.bin there): 6250 bytes through src/phy at both line codes, the chirp + BPSK
    modem and the build-defined FSK / OFDM modem, decoded back to the same bytes"""
    import tmkphy
    data = open(os.path.join(GOLDEN, "INPUT.bin"), "rb").read()
    assert len(data) == 6250
    chunks = [data[i:i + 128] for i in range(0, len(data), 128)]
    assert len(chunks) == 49 and len(chunks[-1]) == 106
    for kind in (FB, MAN):
        cfg = tmkphy.make_config(kind, local_addr=2)
        x = orc.modulate(cfg, [(1, i & 255, 1, 2, c) for i, c in enumerate(chunks)], 48)
        recs, pl = orc.decode_stream(cfg, x)
        assert len(recs) == 49 and (recs["status"] == 0).all() and bytes(pl) == data
        assert list(recs["len"]) == [128] * 48 + [106] and list(recs["seq"]) == list(range(49))
    bits = np.unpackbits(np.frombuffer(data, np.uint8))
    n_fr = (len(bits) + 91) // 92
    fr = np.zeros((n_fr, 100), np.uint8)
    for i in range(n_fr):
        fr[i, :8] = [((i % 100 + 1) >> (7 - j)) & 1 for j in range(8)]
        seg = bits[92 * i:92 * (i + 1)]
        fr[i, 8:8 + len(seg)] = seg
    x = np.concatenate([orc.ask_modulate(fr, [60] * n_fr, [60] * n_fr), np.zeros(400, np.float32)])
    recs = orc.ask_demodulate(x)
    assert len(recs) == n_fr and recs["accepted"].all()
    got = np.concatenate([np.unpackbits(r["bits"])[8:100] for r in recs])[:len(bits)]
    assert np.array_equal(got, bits)
    for mode, cap in ((0, 25), (1, 100)):
        pays = [data[i:i + cap] for i in range(0, len(data), cap)]
        x = np.concatenate([orc.mcm_modulate(mode, pays, np.arange(len(pays)) & 255, [80] * len(pays), [80] * len(pays)),
                            np.zeros(400, np.float32)])
        recs, pl = orc.mcm_demodulate(mode, x)
        assert len(recs) == len(pays) and (recs["status"] == 0).all() and bytes(pl) == data


def test_report_frame_length_literal(orc):
    """docs/proj/report.md:166-169: 64 B payload + 7 B header = 568 bits; Manchester at 3 samples per level = 3408 samples"""
    import ctypes as C
    import tmkphy
    from tmkphy import _abi
    assert orc.samples_for_bits(MAN, 3, (64 + 7) * 8) == 3408
    cfg = tmkphy.make_config(MAN)
    assert _abi.load_library().tmk_samples_for_bits(C.byref(cfg), 568) == 3408
    x = orc.modulate(tmkphy.make_config(MAN, local_addr=2), [(1, 0, 1, 2, bytes(64))], 0)
    assert len(x) == 96 + 3408


def test_two_restatements_agree_chirp_modem(orc):
    """Rows a14-a16 (examples/ask.rs:574-796): the C++ oracle (oracle/ask_modem.hpp) against the numpy-f32 restatement
    written from the Rust (oracle/second_restatement_ask.py). Transmitter sample for sample; receiver record for record
    (lock position, the bits of the locking sync power, id, verdict, the 108 decided bits) on noisy, scaled, clipped,
    truncated and frame-free captures. The sine tables agree within one ulp: glibc's sinf, which the oracle (and the
    Rust binary on linux-gnu) calls, is not correctly rounded, the second restatement's sine is — the receivers are
    compared on the oracle's tables, and once more each on its own tables through a clean loopback."""
    from oracle import second_restatement_ask as sra
    pre2, car2 = sra.tables(4752)
    pre, car = orc.ask_tables(4752)
    for mine, theirs in ((pre2, pre), (car2, car)):
        ulp = np.abs(mine.view(np.int32).astype(np.int64) - theirs.view(np.int32).astype(np.int64))
        assert ulp.max() <= 1 and (ulp != 0).mean() < 0.03, (int(ulp.max()), float((ulp != 0).mean()))
    assert pre2[0] == 0.0 and pre[0] == 0.0

    rng = np.random.default_rng(2026)
    n_recs = n_acc = n_rej = 0
    for k in range(48):
        nf = int(rng.integers(0, 5))
        bits = rng.integers(0, 2, (nf, 100)).astype(np.uint8)
        for f in range(nf):  # ids inside and outside 1..=100 (ask.rs:771)
            ident = int(rng.integers(1, 101)) if rng.random() < 0.7 else int(rng.choice([0, 101, 200, 255]))
            bits[f, :8] = [(ident >> (7 - j)) & 1 for j in range(8)]
        g1 = rng.integers(0, 100, nf) if k % 3 else rng.integers(0, 3000, nf)
        g2 = rng.integers(0, 100, nf)
        x = orc.ask_modulate(bits, g1, g2) if nf else np.zeros(int(rng.integers(1, 5000)), np.float32)
        assert np.array_equal(x, sra.modulate(bits, g1, g2, pre, car)) or nf == 0
        x = x * np.float32(rng.uniform(0.05, 1.0))
        snr = [None, 30.0, 15.0, 8.0, 3.0, 0.0, -3.0][k % 7]
        if snr is not None:
            sigma = np.sqrt(np.mean(x.astype(np.float64) ** 2) + 1e-12) * 10 ** (-snr / 20) if nf else 0.3
            x = (x + rng.normal(0, sigma, len(x))).astype(np.float32)
        if k % 5 == 1 and len(x) > 3000:   # capture ends inside a frame: the decode never completes
            x = x[:len(x) - int(rng.integers(1, 3000))]
        if k % 5 == 2 and len(x) > 3000:   # a zeroed / flipped stretch: lost locks, wrong ids
            a = int(rng.integers(0, len(x) - 2000))
            x[a:a + int(rng.integers(50, 2000))] *= np.float32(0.0 if k % 2 else -1.0)
        if k % 5 == 3:                     # clipping
            x = np.clip(x, -0.3, 0.3)
        want = orc.ask_demodulate(x)
        got = sra.demodulate(x, pre, car)
        assert len(got) == len(want), (k, len(got), len(want))
        for g, w in zip(got, want):
            assert g[0] == int(w["start_index"]) and np.float32(g[1]).view(np.uint32) == w["sync_power"].view(np.uint32), (k, g[:2], w)
            assert g[2] == int(w["id"]) and bool(g[3]) == bool(w["accepted"]), (k, g[2:4], w)
            assert np.array_equal(np.unpackbits(w["bits"])[:108], g[4].astype(np.uint8)), k
        n_recs += len(want)
        n_acc += int(sum(bool(w["accepted"]) for w in want))
        n_rej += int(sum(not bool(w["accepted"]) for w in want))
    assert n_recs > 60 and n_acc > 30 and n_rej > 5, (n_recs, n_acc, n_rej)

    # each restatement end to end on ITS OWN tables: a clean loopback returns the frames it was given
    bits = rng.integers(0, 2, (6, 100)).astype(np.uint8)
    for f in range(6):
        bits[f, :8] = [((f + 1) >> (7 - j)) & 1 for j in range(8)]
    g1, g2 = rng.integers(1, 100, 6), rng.integers(0, 100, 6)
    got = sra.demodulate(sra.modulate(bits, g1, g2, pre2, car2), pre2, car2)
    assert [g[2] for g in got] == [1, 2, 3, 4, 5, 6] and all(g[3] for g in got)
    assert all(np.array_equal(g[4][:100].astype(np.uint8), bits[f]) and not g[4][100:].any() for f, g in enumerate(got))


@pytest.mark.parametrize("kind", [FB, MAN])
def test_two_restatements_agree_transmitter(orc, kind):
    """Rows a9 / a10 (encoder.rs:41-85, frame.rs:74-107): the C++ oracle's modulate against the numpy restatement's, sample
    for sample, over data / ack frames of 0..300 bytes, several shapes and gaps; plus the reference's own encoder test
    (encoder.rs:106-128: 4B5B, 2 samples per level, 2 preamble bytes, 3 payload bytes -> 40 + 200 samples)."""
    import tmkphy
    from oracle.second_restatement import encode_frames
    rng = np.random.default_rng(5 + kind)
    for spl, pb, gap in ((3, 2, 48), (2, 2, 0), (1, 1, 7), (5, 3, 100)):
        frames = helpers.random_frames(rng, 7, max_len=300) + [(1, 9, 8, 7, b""), (2, 255, 0, 255, bytes(range(256)))]
        cfg = tmkphy.make_config(kind, spl=spl, preamble_bytes=pb)
        want = orc.modulate(cfg, frames, gap)
        got = encode_frames(kind, frames, gap, spl, pb)
        assert got.dtype == np.float32 and np.array_equal(got, want), (spl, pb, gap)
    assert len(encode_frames(kind, [], 48)) == 0 and len(orc.modulate(tmkphy.make_config(kind), [], 48)) == 0
    if kind == FB:
        x = encode_frames(FB, [(1, 1, 0, 1, bytes([0x12, 0x34, 0x56]))], 0, spl=2, preamble_bytes=2)
        assert len(x) == 40 + 200  # 16 preamble bits and 80 frame bits: 5 levels per 4 bits, 2 samples per level
