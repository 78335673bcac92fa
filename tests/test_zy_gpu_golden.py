"""The CUDA path (through the C ABI) against the FROZEN vectors of tests/golden/decoder_vectors.npz — a fixed answer,
not a live checker (tests/test_golden_vectors.py holds the oracle and the numpy restatements to the same file on the
CPU). Verdicts, offsets, header bytes, payload bytes, chirp lock positions, sync-power bits and decided bits: bit-exact;
`corr` within 1e-5 (tmkphy.h: not part of the bit-exact contract). Collected late on purpose."""
import numpy as np
import pytest

import helpers
from test_golden_vectors import assert_ask_equal, batch_of, load

pytestmark = pytest.mark.gpu


def test_decoder_batch_against_golden(engine):
    import tmkphy
    z, caps, _ = load()
    s, st = batch_of(tmkphy, z, caps)
    got, gpl = engine.decode_batch(s, st)  # 16 captures, both line codes, chunk 0 / 257 / 1200 mixed in one batch
    helpers.assert_records_equal(got, gpl, z["records"], z["payload"], "golden batch")


def test_decoder_single_captures_against_golden(engine):
    import tmkphy
    z, caps, _ = load()
    for i, c in enumerate(caps):  # one capture per call: the small-batch path of a live receiver
        cfg = tmkphy.make_config(int(z["cap_kind"][i]), local_addr=1, chunk=int(z["cap_chunk"][i]))
        _, recs, pl = tmkphy.decode_capture(engine, cfg, c)
        want = z["records"][z["records"]["stream"] == i].copy()
        want["stream"] = 0
        helpers.assert_records_equal(recs, pl, want, z["payload"], f"golden capture {i}")


def test_chirp_modem_against_golden(engine):
    import tmkphy
    z, _, ask = load()
    st, total = tmkphy.make_streams([len(a) for a in ask], tmkphy.make_config(1))
    s = np.zeros(total, np.float32)
    for a, d in zip(ask, st):
        s[int(d["offset"]):int(d["offset"]) + len(a)] = a
    assert_ask_equal(engine.ask_decode_batch(s, st), z["ask_records"], "golden chirp")
