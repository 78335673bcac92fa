"""The frozen vectors of tests/golden/decoder_vectors.npz (made by tests/golden/make_decoder_vectors.py, which says what
they are: oracle outputs on fixed 16-bit PCM captures, frozen after the independent numpy restatements agreed — NOT
outputs of the Rust binary, which cannot be built here) against everything that decodes:

  CPU (`-m "not gpu"`): the C++ oracle, one-shot per capture and as a threaded batch, bit for bit incl. `corr`; the numpy
                        restatements of the decoder and of the chirp modem.
  GPU (`-m gpu`, tests/test_zy_gpu_golden.py): the CUDA path through the C ABI.

A change of any of them that moves a verdict, an offset, a header byte, a payload byte or a decided bit fails here
against a FIXED answer rather than against a checker that may have moved with it."""
import os

import numpy as np
import pytest

import helpers

VEC = os.path.join(os.path.dirname(__file__), "golden", "decoder_vectors.npz")


def load():
    z = np.load(VEC)
    caps, pos = [], 0
    for n in z["cap_len"]:
        caps.append(z["pcm"][pos:pos + int(n)].astype(np.float32) / np.float32(32768.0))  # exact in float32
        pos += int(n)
    ask, pos = [], 0
    for n in z["ask_len"]:
        ask.append(z["ask_pcm"][pos:pos + int(n)].astype(np.float32) / np.float32(32768.0))
        pos += int(n)
    return z, caps, ask


def batch_of(tmkphy, z, caps):
    cfgs = [tmkphy.make_config(int(k), local_addr=1, chunk=int(c)) for k, c in zip(z["cap_kind"], z["cap_chunk"])]
    return helpers.pack(caps, cfgs, tmkphy)


def assert_ask_equal(got, want, ctx=""):
    assert len(got) == len(want), f"{ctx}: {len(got)} records, golden {len(want)}"
    for f in ("stream", "accepted", "start_index", "id"):
        assert np.array_equal(got[f], want[f]), f"{ctx}: field {f} differs"
    assert np.array_equal(got["bits"], want["bits"]), f"{ctx}: decided bits differ"
    assert np.array_equal(got["sync_power"].view(np.uint32), want["sync_power"].view(np.uint32)), f"{ctx}: sync_power"


def test_fixture_is_what_the_script_says():
    z, caps, ask = load()
    assert len(caps) == 16 and len(ask) == 4 and z["records"].dtype.itemsize == 48 and z["ask_records"].dtype.itemsize == 40
    assert {0, 1, 3, 4, 5, 6} <= {int(s) for s in z["records"]["status"]}  # ok, header, short decode, not for us, CRC, truncated
    assert len(z["records"]) == 63 and len(z["ask_records"]) == 12 and 0 < z["ask_records"]["accepted"].sum() < 12
    assert os.path.getsize(VEC) < 400_000


def test_oracle_against_golden(orc):
    import tmkphy
    z, caps, ask = load()
    want, wpl = z["records"], z["payload"]
    got_all = []
    for i, c in enumerate(caps):  # one decoder per capture, the way the vectors were made
        cfg = orc.make_config(int(z["cap_kind"][i]), local_addr=1, chunk=int(z["cap_chunk"][i]))
        recs, pl = orc.decode_stream(cfg, c, stream_id=i)
        w = want[want["stream"] == i]
        helpers.assert_records_equal(recs, pl, w, wpl, f"capture {i}", corr_tol=None)
        assert np.array_equal(recs["corr"].view(np.uint32), w["corr"].view(np.uint32)), f"capture {i}: corr bits"
        got_all.append(len(recs))
    assert sum(got_all) == len(want)
    s, st = batch_of(tmkphy, z, caps)  # the threaded batch entry the GPU tests and the bench compare with
    recs, pl, _ = orc.decode_batch(s, st, n_threads=3)
    helpers.assert_records_equal(recs, pl, want, wpl, "batch")
    r = np.concatenate([orc.ask_demodulate(a, stream=i) for i, a in enumerate(ask)])
    assert_ask_equal(r, z["ask_records"], "chirp")


def test_numpy_restatements_against_golden(orc):
    from oracle import second_restatement_ask as sra
    from oracle.second_restatement import PhyDecoderPy
    z, caps, ask = load()
    want, wpl = z["records"], z["payload"]
    for i, c in enumerate(caps):
        dec = PhyDecoderPy(int(z["cap_kind"][i]), local_addr=1)
        chunk, pos = int(z["cap_chunk"][i]), 0
        while pos < len(c):
            n = len(c) - pos if chunk == 0 else min(chunk, len(c) - pos)
            dec.process_samples(c[pos:pos + n])
            pos += n
        w = [r for r in want[want["stream"] == i] if r["status"] != 6]  # a frame still being gathered is state, not an event
        assert len(dec.events) == len(w), i
        for (status, p0, fs, cons, hdr, data), r in zip(dec.events, w):
            assert (status, p0, fs, cons) == (int(r["status"]), int(r["preamble_start"]), int(r["frame_start"]), int(r["consumed"])), (i, r)
            if r["hdr_valid"]:
                assert list(hdr) == [int(r["len"]) >> 8, int(r["len"]) & 255, int(r["crc"]), int(r["type"]), int(r["seq"]), int(r["src"]), int(r["dst"])]
            if r["status"] in (0, 4, 5):
                assert bytes(data) == bytes(wpl[int(r["payload_off"]):int(r["payload_off"]) + int(r["len"])]), (i, r)
    pre, car = orc.ask_tables(4752)  # (the tables are libm's: see test_two_restatements_agree_chirp_modem)
    for i, a in enumerate(ask):
        w = z["ask_records"][z["ask_records"]["stream"] == i]
        got = sra.demodulate(a, pre, car)
        assert len(got) == len(w), i
        for g, r in zip(got, w):
            assert (g[0], g[2], bool(g[3])) == (int(r["start_index"]), int(r["id"]), bool(r["accepted"]))
            assert np.float32(g[1]).view(np.uint32) == r["sync_power"].view(np.uint32)
            assert np.array_equal(np.unpackbits(r["bits"])[:108], g[4].astype(np.uint8))
