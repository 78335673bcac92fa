"""Carried-state streaming decode (tmk_links_*) against ONE oracle PhyDecoder per link fed with
the same process_samples calls: the records every call returns, bit for bit, and the positions counted from the
link's first sample."""
import numpy as np
import pytest

import helpers

pytestmark = pytest.mark.gpu
MAN, FB = 0, 1


def feed_sizes(rng, n, kind):
    """a list of call sizes covering n samples"""
    if kind == "poll":      # the 25 ms receiver poll, src/mac/csma.rs:446
        sizes = [1200] * (n // 1200 + 1)
    elif kind == "odd":
        sizes = [257] * (n // 257 + 1)
    elif kind == "big":
        sizes = [20000] * (n // 20000 + 1)
    else:                   # ragged: empty calls, tiny calls, calls longer than a frame
        sizes = []
        tot = 0
        while tot < n:
            s = int(rng.choice([0, 1, 7, 59, 60, 61, 300, 1200, 4800, 9000, 25000]))
            sizes.append(s)
            tot += s
    return sizes


def run_links(engine, orc, captures, cfgs, sizes_per_link, ctx):
    import tmkphy
    n = len(captures)
    bank = tmkphy.PhyDecoderBank(engine, cfgs)
    n_calls = max(len(s) for s in sizes_per_link)
    pos = [0] * n
    got_per_call = []
    for k in range(n_calls):
        chunks = []
        for i in range(n):
            ln = sizes_per_link[i][k] if k < len(sizes_per_link[i]) else 0
            ln = min(ln, len(captures[i]) - pos[i])
            chunks.append(captures[i][pos[i]:pos[i] + ln])
            pos[i] += ln
        got_per_call.append(bank.process_samples(chunks))
    # the oracle: one decoder per link, the same calls
    for i in range(n):
        want, wpl, feed = orc.decode_feeds(cfgs[i], captures[i], sizes_per_link[i], stream_id=i)
        for k in range(n_calls):
            recs, pl = got_per_call[k]
            g = recs[recs["stream"] == i]
            w = want[feed == k]
            helpers.assert_records_equal(g, pl, w, wpl, f"{ctx} link {i} call {k}")
    st = [bank.state(i) for i in range(n)]
    bank.close()
    return st


@pytest.mark.parametrize("kind", [FB, MAN])
@pytest.mark.parametrize("cadence", ["poll", "odd", "big", "ragged"])
def test_links_match_fed_decoders(engine, orc, kind, cadence):
    import tmkphy
    rng = np.random.default_rng(100 + kind)
    cfgs = [tmkphy.make_config(kind, local_addr=1) for _ in range(6)]
    snrs = [None, 30.0, 12.0, 6.0, 3.0, None]
    caps = [helpers.make_capture(orc, cfgs[i], rng, n_frames=5 + i, snr_db=snrs[i], multipath=(i == 2),
                                 gap_range=(0, 2500) if i == 5 else (48, 1500)) for i in range(6)]
    sizes = [feed_sizes(rng, len(c), cadence) for c in caps]
    run_links(engine, orc, caps, cfgs, sizes, f"{cadence}/{kind}")


def test_links_mixed_configs_and_empty_links(engine, orc):
    """links with different line codes / addresses in one bank, one link that never receives a sample, one that
    only receives noise"""
    import tmkphy
    rng = np.random.default_rng(7)
    cfgs = [tmkphy.make_config(FB, local_addr=1), tmkphy.make_config(MAN, local_addr=2),
            tmkphy.make_config(FB, local_addr=-1), tmkphy.make_config(MAN, local_addr=1),
            tmkphy.make_config(FB, local_addr=1, spl=4, preamble_bytes=3)]
    caps = [helpers.make_capture(orc, cfgs[0], rng, n_frames=4, snr_db=20.0),
            helpers.make_capture(orc, cfgs[1], rng, n_frames=4, snr_db=None),
            np.zeros(0, np.float32),
            rng.normal(0, 0.3, 30000).astype(np.float32),
            helpers.make_capture(orc, cfgs[4], rng, n_frames=3, snr_db=15.0)]
    sizes = [feed_sizes(rng, len(c), "ragged") if len(c) else [0, 0, 0] for c in caps]
    st = run_links(engine, orc, caps, cfgs, sizes, "mixed")
    assert st[2] == {"buffered": 0, "buffer_offset": 0, "decoding": False}


def test_links_frame_split_across_many_calls_and_reset(engine, orc):
    """a 256-byte frame arriving in 100-sample calls stays in Decoding state until it is complete; reset drops it"""
    import tmkphy
    cfg = tmkphy.make_config(FB, local_addr=1)
    wave = orc.modulate(cfg, [(1, 9, 2, 1, bytes(range(256)))], 0)
    x = np.concatenate([np.zeros(333, np.float32), wave, np.zeros(500, np.float32)])
    bank = tmkphy.PhyDecoderBank(engine, [cfg])
    seen = []
    decoding_calls = 0
    for p in range(0, len(x), 100):
        recs, pl = bank.process_samples([x[p:p + 100]])
        decoding_calls += bank.state(0)["decoding"]
        seen.append((recs, pl))
    allrecs = np.concatenate([r for r, _ in seen])
    assert len(allrecs) == 1 and allrecs[0]["status"] == 0 and allrecs[0]["preamble_start"] == 333
    assert decoding_calls > 50
    want, wpl, feed = orc.decode_feeds(cfg, x, [100] * (len(x) // 100 + 1))
    k = int(feed[0])
    helpers.assert_records_equal(seen[k][0], seen[k][1], want, wpl, "split frame")
    # reset in the middle of a frame: the frame is gone, positions restart
    bank.reset()
    for p in range(0, 333 + 3000, 100):
        bank.process_samples([x[p:p + 100]])
    assert bank.state(0)["decoding"]
    bank.reset(0)
    assert bank.state(0) == {"buffered": 0, "buffer_offset": 0, "decoding": False}
    recs, _ = bank.process_samples([x])
    assert len(recs) == 1 and recs[0]["preamble_start"] == 333
    bank.close()


def test_links_argument_errors(engine):
    import tmkphy
    cfg = tmkphy.make_config(FB)
    bank = tmkphy.PhyDecoderBank(engine, [cfg, cfg])
    with pytest.raises(AssertionError):
        bank.process_samples([np.zeros(4, np.float32)])
    with pytest.raises(tmkphy.TmkError):
        bank.reset(5)
    recs, pl = bank.process_samples([np.zeros(0, np.float32), np.zeros(0, np.float32)])
    assert len(recs) == 0 and len(pl) == 0
    bank.close()


@pytest.mark.parametrize("kind", [FB, MAN])
def test_phy_decoder_is_streaming(engine, orc, kind):
    """tmkphy.PhyDecoder carries its state like the reference's (decoder.rs:68-112): fed with the 1200-sample polls of
    src/mac/csma.rs:446-465 it returns, call by call, the frames one oracle decoder returns for the same calls; reset()
    starts over (decoder.rs:114-119)."""
    import tmkphy
    rng = np.random.default_rng(77 + kind)
    cfg = tmkphy.make_config(kind, local_addr=1)
    cap = helpers.make_capture(orc, cfg, rng, n_frames=10, snr_db=15.0)
    sizes = [1200] * (len(cap) // 1200 + 1)
    want, wpl, feed = orc.decode_feeds(cfg, cap, sizes)
    dec = tmkphy.PhyDecoder(engine, 3, 2, kind, local_addr=1)
    for rnd in range(2):
        pos, n_frames = 0, 0
        for k, ln in enumerate(sizes):
            frames = dec.process_samples(cap[pos:pos + ln])
            pos += ln
            w = want[feed == k]
            helpers.assert_records_equal(dec.last_records, dec.last_payload, w, wpl, f"PhyDecoder call {k}")
            ok = w[w["status"] == 0]
            assert [f.data for f in frames] == [bytes(wpl[int(r["payload_off"]):int(r["payload_off"]) + int(r["len"])]) for r in ok]
            n_frames += len(frames)
        assert n_frames == int((want["status"] == 0).sum()) and n_frames >= 5
        dec.reset()
    dec.close()
