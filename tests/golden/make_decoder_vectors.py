"""Generates tests/golden/decoder_vectors.npz — frozen input / output vectors of the hot path.

Run from the repo root:  python tests/golden/make_decoder_vectors.py

WHAT THE VECTORS ARE (and are not): the reference is Rust and no Rust toolchain exists in this image or on the GPU box,
so these are NOT outputs of the reference binary. They are outputs of the C++ oracle (oracle/, the line-cited
restatement of src/phy/decoder.rs and examples/ask.rs) on fixed captures, frozen at the commit that added them, after
the second, independent numpy restatements (oracle/second_restatement*.py) had produced the same records for every
capture below (the script asserts that before it writes). They pin three things that a live oracle cannot: the oracle
against silent drift, the GPU path against a FIXED answer instead of a moving one, and the numpy restatements against
both. Parity against the reference itself stays "unpinned" (DESIGN.md §5).

Captures are stored as 16-bit PCM (the form the reference stores recordings in, src/utils/dump.rs:37-40) and enter every
decoder as pcm / 32768 in float32, which is exact, so the fixture does not depend on any random generator or libm.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "trackmaker-rs_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))

FB, MAN = 1, 0
OUT = os.path.join(HERE, "decoder_vectors.npz")


def quantise(x, scale=16384.0):
    """float capture -> int16 PCM with 6 dB of headroom for the noise peaks"""
    return np.clip(np.rint(np.asarray(x, np.float64) * scale), -32768, 32767).astype(np.int16)


def main():
    from oracle import oracle as orc
    from oracle import second_restatement_ask as sra
    from oracle.second_restatement import PhyDecoderPy
    import helpers
    orc.build()
    rng = np.random.default_rng(20261105)

    # ---- src/phy decoder (decoder.rs:68-356): 16 captures, both line codes, clean .. 6.5 dB, multipath, damage, chunked feeds
    pcm, cap_len, cap_kind, cap_chunk = [], [], [], []
    recs_all, pl_all = [], []
    pl_off = 0
    for kind in (FB, MAN):
        for v in range(8):
            snr = [None, 20.0, 12.0, 9.0, 8.0, 7.0, 6.5, None][v]  # below ~6 dB the 0.9 threshold is never reached
            chunk = [0, 1200, 257, 0, 1200, 0, 257, 1200][v]
            cfg = orc.make_config(kind, local_addr=1, chunk=chunk)
            c = helpers.make_capture(orc, cfg, rng, n_frames=5, snr_db=snr, multipath=(v in (2, 5)), gap_range=(48, 600),
                                     tail=int(rng.integers(0, 300)), max_len=40, amplitude=1.0)
            if v == 7:  # a flipped and a zeroed stretch: header / CRC / length verdicts and their consume rules
                a = int(rng.integers(200, len(c) // 2))
                c[a:a + 150] *= np.float32(-1.0)
                b = int(rng.integers(len(c) // 2, len(c) - 400))
                c[b:b + 90] = 0.0
            if v == 1:  # short flipped runs inside frames: CRC failures, header damage
                for _ in range(6):
                    a = int(rng.integers(100, len(c) - 100))
                    c[a:a + int(rng.integers(3, 9))] *= np.float32(-1.0)
            if v == 3:  # the capture ends inside a frame
                c = c[:len(c) - int(rng.integers(600, 1200))]
            q = quantise(c)
            x = q.astype(np.float32) / np.float32(32768.0)
            recs, pl = orc.decode_stream(cfg, x, stream_id=len(pcm))
            # the second restatement (numpy, written from the Rust) must tell the same story before the vector is frozen
            dec = PhyDecoderPy(kind, local_addr=1)
            pos = 0
            while pos < len(x):
                n = len(x) - pos if chunk == 0 else min(chunk, len(x) - pos)
                dec.process_samples(x[pos:pos + n])
                pos += n
            want = [(int(r["status"]), int(r["preamble_start"]), int(r["frame_start"])) for r in recs if r["status"] != 6]
            got = [(e[0], e[1], e[2]) for e in dec.events]
            assert got == want, (kind, v, got, want)
            recs = recs.copy()
            out_pl = []
            for r in recs:  # payload offsets re-based onto the fixture's own payload array
                n = int(r["len"]) if r["status"] in (0, 4, 5) else 0
                out_pl.append(pl[int(r["payload_off"]):int(r["payload_off"]) + n])
                r["payload_off"] = pl_off
                pl_off += n
            pcm.append(q)
            cap_len.append(len(q))
            cap_kind.append(kind)
            cap_chunk.append(chunk)
            recs_all.append(recs)
            pl_all += out_pl
    recs_all = np.concatenate(recs_all)
    pl_all = np.concatenate(pl_all) if pl_all else np.zeros(0, np.uint8)
    assert len(recs_all) >= 40 and len({int(s) for s in recs_all["status"]}) >= 3, (len(recs_all), set(recs_all["status"]))

    # ---- chirp + BPSK modem (examples/ask.rs:668-796): 4 captures
    pre, car = orc.ask_tables(4752)
    ask_pcm, ask_len, ask_recs = [], [], []
    for v in range(4):
        nf = 3
        bits = rng.integers(0, 2, (nf, 100)).astype(np.uint8)
        for f in range(nf):
            ident = [int(rng.integers(1, 101)), int(rng.integers(1, 101)), 200][f] if v != 2 else int(rng.integers(1, 101))
            bits[f, :8] = [(ident >> (7 - j)) & 1 for j in range(8)]
        g1, g2 = rng.integers(0, 100, nf), rng.integers(0, 100, nf)
        x = orc.ask_modulate(bits, g1, g2).astype(np.float64)
        snr = [None, 10.0, 3.0, -3.0][v]
        if snr is not None:
            x = x + rng.normal(0, np.sqrt(0.5) * 10 ** (-snr / 20), len(x))
        q = quantise(x, 8192.0)
        xq = q.astype(np.float32) / np.float32(32768.0)
        r = orc.ask_demodulate(xq, stream=v)
        g = sra.demodulate(xq, pre, car)
        assert [(a[0], int(a[2]), bool(a[3])) for a in g] == [(int(b["start_index"]), int(b["id"]), bool(b["accepted"])) for b in r], v
        ask_pcm.append(q)
        ask_len.append(len(q))
        ask_recs.append(r)
    ask_recs = np.concatenate(ask_recs)
    assert len(ask_recs) >= 8

    np.savez_compressed(OUT, pcm=np.concatenate(pcm), cap_len=np.array(cap_len, np.int64), cap_kind=np.array(cap_kind, np.int32),
                        cap_chunk=np.array(cap_chunk, np.int32), records=recs_all, payload=pl_all.astype(np.uint8),
                        ask_pcm=np.concatenate(ask_pcm), ask_len=np.array(ask_len, np.int64), ask_records=ask_recs)
    print(OUT, os.path.getsize(OUT), "bytes;", len(recs_all), "records (statuses", sorted({int(s) for s in recs_all["status"]}), "),",
          len(pl_all), "payload bytes;", len(ask_recs), "chirp records,", int(ask_recs["accepted"].sum()), "accepted")


if __name__ == "__main__":
    main()
