"""Parity of the chirp + BPSK modem (examples/ask.rs:526-858) through the C ABI against the oracle:
tables, modulated track and every record field bit-exact (sync peaks, sync power, decided bits, ids)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def ask_frames(text, n=100):
    """examples/ask.rs:546-569: bits 0-7 = frame id 1.., bits 8-99 = the next 92 payload bits, wrapping"""
    tb = np.unpackbits(np.frombuffer(text, np.uint8))
    fr = np.zeros((n, 100), np.uint8)
    bi = 0
    for i in range(n):
        fr[i, :8] = [((i + 1) >> (7 - j)) & 1 for j in range(8)]
        for j in range(8, 100):
            if bi >= len(tb):
                bi = 0
            fr[i, j] = tb[bi]
            bi += 1
    return fr


def assert_ask_equal(got, want, ctx=""):
    assert len(got) == len(want), f"{ctx}: {len(got)} records, oracle {len(want)}"
    for f in ("stream", "accepted", "start_index", "id"):
        assert np.array_equal(got[f], want[f]), f"{ctx}: field {f} differs"
    assert np.array_equal(got["bits"], want["bits"]), f"{ctx}: decided bits differ"
    assert np.array_equal(got["sync_power"].view(np.uint32), want["sync_power"].view(np.uint32)), f"{ctx}: sync_power"


def oracle_batch(orc, s, st):
    out = [orc.ask_demodulate(s[int(t["offset"]):int(t["offset"]) + int(t["n_samples"])], stream=i) for i, t in enumerate(st)]
    return np.concatenate(out) if out else np.zeros(0)


def test_ask_tables_and_modulate_bit_exact(engine, orc):
    pre, car = engine.ask_tables(48000)
    opre, ocar = orc.ask_tables(48000)
    assert np.array_equal(pre.view(np.uint32), opre.view(np.uint32)) and np.array_equal(car.view(np.uint32), ocar.view(np.uint32))
    assert pre[0] == 0.0 and abs(pre).max() <= 1.0
    rng = np.random.default_rng(1)
    fr = rng.integers(0, 2, (7, 100)).astype(np.uint8)
    g1, g2 = rng.integers(0, 100, 7), rng.integers(0, 100, 7)
    got, want = engine.ask_modulate(fr, g1, g2), orc.ask_modulate(fr, g1, g2)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    assert len(engine.ask_modulate(np.zeros((0, 100), np.uint8), [], [])) == 0


def test_ask_loopback_think_different(engine, orc):
    """the reference's own offline test: 100 frames of assets/think-different.txt, decoded text matches"""
    import tmkphy
    text = open(os.path.join(GOLDEN, "think-different.txt"), "rb").read()
    fr = ask_frames(text)
    rng = np.random.default_rng(7)
    g1, g2 = rng.integers(0, 100, 100), rng.integers(0, 100, 100)
    x = engine.ask_modulate(fr, g1, g2)
    st, total = tmkphy.make_streams([len(x)], tmkphy.make_config(1))
    got = engine.ask_decode_batch(x, st)
    assert_ask_equal(got, orc.ask_demodulate(x), "loopback")
    assert len(got) == 100 and got["accepted"].all() and list(got["id"]) == list(range(1, 101))
    bits = np.unpackbits(got["bits"], axis=1)[:, 8:100].reshape(-1)
    decoded = np.packbits(bits[:len(bits) // 8 * 8]).tobytes()
    n = min(len(decoded), len(text))
    assert decoded[:n] == text[:n]
    assert got["start_index"][0] == g1[0] + 439


def impaired(rng, x, snr_db=None, amp=1.0, multipath=False):
    y = x.astype(np.float64) * amp
    if multipath:
        h = np.zeros(41)
        h[0], h[int(rng.integers(2, 13))], h[int(rng.integers(13, 41))] = 1.0, 0.35, -0.2
        y = np.convolve(y, h / np.sqrt((h * h).sum()))[:len(y)]
    if snr_db is not None:
        y = y + rng.normal(0, amp * np.sqrt(0.5) * 10 ** (-snr_db / 20), len(y))
    return y.astype(np.float32)


def test_ask_batch_impairments(engine, orc):
    """ragged batch: AWGN sweep, amplitude scaling, multipath, long gaps, truncation inside a frame and inside a
    preamble, silence, streams shorter than the preamble, a loud burst before quiet frames"""
    import tmkphy
    rng = np.random.default_rng(11)
    caps = []
    for i, snr in enumerate((None, 30, 20, 12, 8, 5, 3, 0, -3)):
        fr = rng.integers(0, 2, (6, 100)).astype(np.uint8)
        fr[:, :8] = [[((k + 1) >> (7 - j)) & 1 for j in range(8)] for k in range(6)]
        if i % 3 == 2:
            fr[2, :8] = 1  # id 255: decoded but not accepted
        x = orc.ask_modulate(fr, rng.integers(0, 3000, 6), rng.integers(0, 100, 6))
        caps.append(impaired(rng, x, snr, amp=float(rng.uniform(0.2, 1.0)), multipath=(i % 2 == 1)))
    base = orc.ask_modulate(rng.integers(0, 2, (3, 100)).astype(np.uint8), [50, 0, 10], [0, 30, 500])
    caps += [base[:6000], base[:250], base[:5192 + 50 + 300], np.zeros(3000, np.float32), np.zeros(100, np.float32),
             np.zeros(0, np.float32), impaired(rng, base, 10)[:20000]]
    loud = rng.normal(0, 20.0, 4000).astype(np.float32)
    caps.append(np.concatenate([loud, base * 0.05, loud[:700], base * 0.3]).astype(np.float32))
    st, total = tmkphy.make_streams([len(c) for c in caps], tmkphy.make_config(1))
    s = np.zeros(total, np.float32)
    for i, c in enumerate(caps):
        s[int(st[i]["offset"]):int(st[i]["offset"]) + len(c)] = c
    got = engine.ask_decode_batch(s, st)
    want = oracle_batch(orc, s, st)
    assert_ask_equal(got, want, "impairments")
    assert got["accepted"].sum() > 30 and (got["accepted"] == 0).any()


@pytest.mark.parametrize("seed", [0, 1])
def test_ask_marginal_detection_many_streams(engine, orc, seed):
    """256 captures whose sync peaks sit ON the `sync_power > 2 power` rule (ask.rs:711-714: amplitude and noise drawn
    so that 1.1 a ~ 2 (a^2 / 2 + sigma^2)), plus DC offsets, bursts and silence: the FFT filter in front of the walk may
    only ever add candidates, so every record must still be the oracle's"""
    import tmkphy
    rng = np.random.default_rng(900 + seed)
    caps = []
    for s in range(256):
        nf = int(rng.integers(3, 9))
        fr = rng.integers(0, 2, (nf, 100)).astype(np.uint8)
        for i in range(nf):
            fr[i, :8] = [((i + 1) >> (7 - j)) & 1 for j in range(8)]
        x = orc.ask_modulate(fr, rng.integers(0, 700, nf), rng.integers(0, 700, nf))
        a = float(rng.uniform(0.25, 1.4))
        # marginal: 1.1 a = 2 (a^2 / 2 + sigma^2) (1 + d), d in +-8 %
        d = float(rng.uniform(-0.08, 0.08))
        s2 = max(1.1 * a / (2 * (1 + d)) - a * a / 2, 1e-6)
        x = x * np.float32(a) + rng.normal(0, np.sqrt(s2), len(x)).astype(np.float32)
        kind = s % 8
        if kind == 1:
            x += np.float32(rng.uniform(-0.2, 0.2))                 # DC offset
        elif kind == 2:
            p = int(rng.integers(0, len(x) - 400))
            x[p:p + 300] += rng.normal(0, 3.0, 300).astype(np.float32)  # a loud burst, then quiet again
        elif kind == 3:
            p = int(rng.integers(0, len(x) - 3000))
            x[p:p + 2500] = 0.0                                     # a stretch of exact silence
        elif kind == 4:
            x = np.concatenate([np.zeros(int(rng.integers(0, 3000)), np.float32), x])
        caps.append(x.astype(np.float32))
    st, total = tmkphy.make_streams([len(c) for c in caps], tmkphy.make_config(1))
    buf = np.zeros(total, np.float32)
    for c, t in zip(caps, st):
        buf[int(t["offset"]):int(t["offset"]) + len(c)] = c
    want = oracle_batch(orc, buf, st)
    got = engine.ask_decode_batch(buf, st)
    assert_ask_equal(got, want, f"marginal seed={seed}")
    assert 300 < len(want) < 256 * 8 and (want["accepted"] == 1).sum() > 200


@pytest.mark.parametrize("seg_min", [8192, 20000, 65536])
def test_ask_segmented_walk(engine, orc, seg_min):
    """captures cut into segments walked at once (ask.cu: segmented walk): whatever the segment length — frames that
    straddle a cut, peaks pending at a cut, marginal locks, bursts, silence — the records are the serial receiver's"""
    import os
    import tmkphy
    rng = np.random.default_rng(4000 + seg_min)
    caps = []
    for s in range(12):
        nf = int(rng.integers(25, 60))
        fr = rng.integers(0, 2, (nf, 100)).astype(np.uint8)
        for i in range(nf):
            fr[i, :8] = [((i % 100 + 1) >> (7 - j)) & 1 for j in range(8)]
        x = orc.ask_modulate(fr, rng.integers(0, 900 if s % 3 else 100, nf), rng.integers(0, 300, nf))
        a = float(rng.uniform(0.3, 1.2))
        if s % 4 == 0:   # marginal locks: 1.1 a ~ 2 (a^2 / 2 + sigma^2)
            s2 = max(1.1 * a / (2 * (1 + float(rng.uniform(-0.05, 0.05)))) - a * a / 2, 1e-6)
            x = x * np.float32(a) + rng.normal(0, np.sqrt(s2), len(x)).astype(np.float32)
        else:
            x = impaired(rng, x, float(rng.choice([30, 15, 11])), amp=a, multipath=(s % 4 == 1))
        if s % 5 == 2:
            p = int(rng.integers(0, len(x) - 5000))
            x[p:p + 3000] = 0.0
        caps.append(x.astype(np.float32))
    caps += [np.zeros(100000, np.float32), rng.normal(0, 0.3, 150000).astype(np.float32)]
    st, total = tmkphy.make_streams([len(c) for c in caps], tmkphy.make_config(1))
    buf = np.zeros(total, np.float32)
    for c, t in zip(caps, st):
        buf[int(t["offset"]):int(t["offset"]) + len(c)] = c
    want = oracle_batch(orc, buf, st)
    saved = {k: os.environ.get(k) for k in ("TMK_ASK_SEG_MIN", "TMK_ASK_SEGMENTS")}
    try:
        os.environ["TMK_ASK_SEG_MIN"] = str(seg_min)
        os.environ["TMK_ASK_SEGMENTS"] = "20000"
        got = engine.ask_decode_batch(buf, st)
        assert_ask_equal(got, want, f"segmented seg_min={seg_min}")
        os.environ["TMK_ASK_SEGMENTS"] = "0"
        assert_ask_equal(engine.ask_decode_batch(buf, st), want, "one walker per capture")
    finally:
        for k, v in saved.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
    assert len(want) > 250


def test_ask_misaligned_device_buffer_is_refused(engine):
    """the filter and the power chain read 16 bytes at a time: a device pointer that is not 16-byte aligned is an argument
    error, not a misaligned access"""
    import tmkphy
    d = engine.device_alloc(4 * 20000)
    try:
        st, total = tmkphy.make_streams([8000], tmkphy.make_config(1))
        with pytest.raises(tmkphy.TmkError) as ei:
            engine.ask_decode_batch(d + 4, st, total_samples=total)
        assert ei.value.code == tmkphy._abi.E_INVALID
        with pytest.raises(tmkphy.TmkError):
            engine.mcm_decode_batch(1, d + 8, st, total_samples=total)
    finally:
        engine.device_free(d)


def test_ask_power_brackets_withdrawn(engine, orc):
    """the exact power chain is computed in parallel from brackets that start at a LOCAL bound (ask_power_kernel) and are
    verified afterwards: captures built to break the local bound — a burst hundreds of times louder than what follows, a
    NaN, an inf, a sample whose square overflows, long silence — must still give the serial receiver's records"""
    import tmkphy
    rng = np.random.default_rng(77)
    fr = rng.integers(0, 2, (9, 100)).astype(np.uint8)
    fr[:, :8] = [[((k + 1) >> (7 - j)) & 1 for j in range(8)] for k in range(9)]
    base = orc.ask_modulate(fr, rng.integers(0, 600, 9), rng.integers(0, 600, 9))
    noisy = impaired(rng, base, 20)
    caps = []
    for kind in range(8):
        x = noisy.copy()
        if kind == 0:
            x[3000:3400] += rng.normal(0, 300.0, 400).astype(np.float32)      # a burst 300 x louder than the rest
        elif kind == 1:
            x = np.concatenate([rng.normal(0, 50.0, 9000).astype(np.float32), noisy * np.float32(0.02)])
        elif kind == 2:
            x[100] = np.nan
        elif kind == 3:
            x[20000] = np.inf
        elif kind == 4:
            x[7000] = np.float32(3e19)                                       # finite, its square is not
        elif kind == 5:
            x = np.concatenate([noisy, np.zeros(30000, np.float32), noisy])    # digital silence between two bursts of frames
        elif kind == 6:
            x[9000] = np.float32(-2e19)
            x = np.concatenate([x, noisy])
        caps.append(x.astype(np.float32))
    st, total = tmkphy.make_streams([len(c) for c in caps], tmkphy.make_config(1))
    buf = np.zeros(total, np.float32)
    for c, t in zip(caps, st):
        buf[int(t["offset"]):int(t["offset"]) + len(c)] = c
    with np.errstate(all="ignore"):
        want = oracle_batch(orc, buf, st)
    got = engine.ask_decode_batch(buf, st)
    assert_ask_equal(got, want, "withdrawn brackets")
    per = [int((want["stream"] == i).sum()) for i in range(8)]
    assert per[7] == 9 and per[5] == 18 and per[2] == 0 and per[0] >= 5, per
