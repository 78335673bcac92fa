"""tmk_channel_busy_noalloc (static __device__ scratch, no cudaMalloc) against tmk_channel_busy through the C ABI.

Collected last on purpose. The results are int8 verdicts: the comparison is exact,
like every other test of this suite. Two calls in a row share the static array; the second one is built so that
stale flags or descriptors of the first would show."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _batch(rng, lens, loud):
    """captures of the given lengths: quiet noise, plus one sample above 0.5 somewhere in the `loud` ones"""
    caps = []
    for i, n in enumerate(lens):
        c = rng.uniform(-0.4, 0.4, n).astype(np.float32)
        if n and i in loud:
            c[int(rng.integers(0, n))] = np.float32(0.75 if i % 2 else -0.75)
        caps.append(c)
    return caps


def _run_both(engine, tmkphy, caps):
    cfg = tmkphy.make_config(tmkphy.LINE_4B5B)
    st, total = tmkphy.make_streams([len(c) for c in caps], cfg)
    s = np.zeros(max(total, 1), np.float32)
    for c, d in zip(caps, st):
        s[int(d["offset"]):int(d["offset"]) + len(c)] = c
    want = engine.channel_busy(s, st)  # the existing entry point (host buffer, engine scratch)
    d_x = engine.device_alloc(s.nbytes)  # the caller's own input buffer; the call under test allocates nothing
    try:
        engine.h2d(d_x, s)
        got = engine.channel_busy_noalloc(d_x, total, st)
        # device-resident output as well
        d_out = engine.device_alloc(max(1, len(st)))
        got_dev = np.full(len(st), 99, np.int8)
        engine._check(engine._lib.tmk_channel_busy_noalloc(engine._h, C.c_void_p(d_x), total, st.ctypes.data_as(C.c_void_p),
                                                           len(st), C.c_void_p(d_out)))
        engine.d2h(got_dev, d_out)
        engine.device_free(d_out)
    finally:
        engine.device_free(d_x)
    return want, got, got_dev


def test_noalloc_matches_channel_busy_two_calls(engine):
    import tmkphy
    rng = np.random.default_rng(20261105)
    lens = [19, 20, 500, 70000, 200001, 33, 0, 131072, 65536, 65537, 1000]
    loud1 = {1, 3, 4, 7, 9}
    want, got, got_dev = _run_both(engine, tmkphy, _batch(rng, lens, loud1))
    assert list(want) == [-1, 1, 0, 1, 1, 0, -1, 1, 0, 1, 0]
    assert np.array_equal(got, want) and np.array_equal(got_dev, want)
    # second call in a row: the streams that were loud are quiet now, and the batch is shorter than the first one,
    # so flags or descriptors left in the static array by the first call would change the answer
    lens2 = [70000, 21, 200001, 500, 65536, 18]
    loud2 = {1, 4}
    want2, got2, got_dev2 = _run_both(engine, tmkphy, _batch(rng, lens2, loud2))
    assert list(want2) == [0, 1, 0, 0, 1, -1]
    assert np.array_equal(got2, want2) and np.array_equal(got_dev2, want2)


def test_noalloc_limits(engine):
    import tmkphy
    cfg = tmkphy.make_config(tmkphy.LINE_4B5B)
    n = tmkphy._abi.NOALLOC_MAX_STREAMS
    d_x = engine.device_alloc(4 * 32 * (n + 1))
    try:
        engine.h2d(d_x, np.full(32 * (n + 1), 0.6, np.float32))
        st, total = tmkphy.make_streams([32] * (n + 1), cfg)
        with pytest.raises(tmkphy.TmkError) as ei:  # one stream beyond the static scratch: refused, with a clear message
            engine.channel_busy_noalloc(d_x, total, st)
        assert ei.value.code == tmkphy._abi.E_CAPACITY and "TMK_NOALLOC_MAX_STREAMS" in str(ei.value)
        st, total = tmkphy.make_streams([32] * n, cfg)  # exactly the maximum works
        assert (engine.channel_busy_noalloc(d_x, total, st) == 1).all()
        host = np.zeros(64, np.float32)
        st, total = tmkphy.make_streams([64], cfg)
        with pytest.raises(tmkphy.TmkError):  # host samples would need staging memory
            engine._check(engine._lib.tmk_channel_busy_noalloc(engine._h, host.ctypes.data_as(C.c_void_p), total,
                                                               st.ctypes.data_as(C.c_void_p), 1,
                                                               np.zeros(1, np.int8).ctypes.data_as(C.c_void_p)))
    finally:
        engine.device_free(d_x)
