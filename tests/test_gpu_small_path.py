"""Small-batch path of tmk_decode_batch (engine.cu: one block with counters, exit states, ordered records and payload; one
copy of the usual amount, one synchronisation): a busy batch that has more records / payload than the first copy brings
over, the same batch through the general path, and caller buffers on the limit."""
import os

import numpy as np
import pytest

import helpers

pytestmark = pytest.mark.gpu
FB, MAN = 1, 0


@pytest.fixture
def path_env():
    saved = os.environ.get("TMK_NO_SMALL_PATH")
    yield
    if saved is None:
        os.environ.pop("TMK_NO_SMALL_PATH", None)
    else:
        os.environ["TMK_NO_SMALL_PATH"] = saved


def oracle_records(orc, caps, cfgs):
    want, wpl, base = [], [], 0
    for i, (c, cfg) in enumerate(zip(caps, cfgs)):
        r, p = orc.decode_stream(cfg, c, stream_id=i)
        r["payload_off"] += base
        base += len(p)
        want.append(r)
        wpl.append(p)
    return np.concatenate(want), np.concatenate(wpl)


def test_busy_small_batch(engine, orc, path_env):
    """three dense captures: ~1500 records and ~90 KB of payload — more than the first copy of the small path holds
    (512 + n / 2 records, 32 KB + 40 n bytes), fewer than its block (so the second copy runs, not the general path)"""
    import tmkphy
    rng = np.random.default_rng(77)
    cf, cm = tmkphy.make_config(FB, local_addr=1), tmkphy.make_config(MAN, local_addr=1)
    caps = [helpers.make_capture(orc, cf, rng, n_frames=620, snr_db=15.0, gap_range=(48, 300)),
            helpers.make_capture(orc, cm, rng, n_frames=520, snr_db=12.0, gap_range=(48, 300)),
            helpers.make_capture(orc, cf, rng, n_frames=400, snr_db=20.0, gap_range=(48, 200), max_len=128)]
    cfgs = [cf, cm, cf]
    s, st = helpers.pack(caps, cfgs, tmkphy)
    want, wpl = oracle_records(orc, caps, cfgs)
    assert len(want) > 1200 and len(wpl) > 40000
    os.environ.pop("TMK_NO_SMALL_PATH", None)
    recs, pl = engine.decode_batch(s, st)
    helpers.assert_records_equal(recs, pl, want, wpl, "small path, busy")
    os.environ["TMK_NO_SMALL_PATH"] = "1"
    recs0, pl0 = engine.decode_batch(s, st)
    helpers.assert_records_equal(recs0, pl0, want, wpl, "general path")
    assert np.array_equal(recs["corr"], recs0["corr"])


def test_small_batch_buffers_on_the_limit(engine, orc, path_env):
    """caller buffers of exactly the needed size, one record / one byte less (TMK_E_CAPACITY, nothing returned), through
    the raw C ABI so that nothing grows them"""
    import ctypes as C
    import tmkphy
    os.environ.pop("TMK_NO_SMALL_PATH", None)
    rng = np.random.default_rng(78)
    cf = tmkphy.make_config(FB, local_addr=1)
    caps = [helpers.make_capture(orc, cf, rng, n_frames=30, snr_db=18.0), helpers.make_capture(orc, cf, rng, n_frames=25, snr_db=18.0)]
    s, st = helpers.pack(caps, [cf, cf], tmkphy)
    want, wpl = oracle_records(orc, caps, [cf, cf])
    lib, h = engine._lib, engine._h

    def call(rec_cap, pay_cap):
        recs = np.zeros(max(rec_cap, 1), tmkphy.REC_DTYPE)
        pl = np.zeros(max(pay_cap, 1), np.uint8)
        nr, npl = C.c_uint64(0), C.c_uint64(0)
        rc = lib.tmk_decode_batch(h, s.ctypes.data_as(C.c_void_p), len(s), st.ctypes.data_as(C.c_void_p), len(st),
                                  recs.ctypes.data_as(C.c_void_p), rec_cap, C.byref(nr), pl.ctypes.data_as(C.c_void_p), pay_cap,
                                  C.byref(npl))
        return rc, recs[:min(nr.value, rec_cap)], pl[:min(npl.value, pay_cap)]

    rc, recs, pl = call(len(want), len(wpl))
    assert rc == 0
    helpers.assert_records_equal(recs, pl, want, wpl, "exact buffers")
    rc, _, _ = call(len(want) - 1, len(wpl))
    assert rc == tmkphy._abi.E_CAPACITY
    rc, _, _ = call(len(want), len(wpl) - 1)
    assert rc == tmkphy._abi.E_CAPACITY
    rc, recs, pl = call(len(want) + 7, len(wpl) + 100)
    assert rc == 0
    helpers.assert_records_equal(recs, pl, want, wpl, "roomy buffers")


@pytest.mark.parametrize("small", [True, False])
def test_fresh_engine_without_cluster_pass(orc, path_env, small):
    """a NEW engine whose first decode runs without the cluster pass (TMK_NO_CLUSTER): the event that times the cluster pass
    has never been recorded, and asking for its elapsed time used to leave a CUDA error behind for the next call"""
    import tmkphy
    saved = os.environ.get("TMK_NO_CLUSTER")
    os.environ["TMK_NO_CLUSTER"] = "1"
    if small:
        os.environ.pop("TMK_NO_SMALL_PATH", None)
    else:
        os.environ["TMK_NO_SMALL_PATH"] = "1"
    eng = tmkphy.Engine(0)
    try:
        rng = np.random.default_rng(79)
        cf = tmkphy.make_config(FB, local_addr=1)
        caps = [helpers.make_capture(orc, cf, rng, n_frames=12, snr_db=15.0)]
        s, st = helpers.pack(caps, [cf], tmkphy)
        want, wpl = oracle_records(orc, caps, [cf])
        for _ in range(2):
            recs, pl = eng.decode_batch(s, st)
            helpers.assert_records_equal(recs, pl, want, wpl, "no cluster pass")
            eng.last_stage_ms()
            eng.debug_counters()
        bank = tmkphy.PhyDecoderBank(eng, [cf])
        got = bank.process_samples([caps[0][:5000]])
        bank.close()
        assert len(got[0]) >= 0
    finally:
        eng.close()
        if saved is None:
            os.environ.pop("TMK_NO_CLUSTER", None)
        else:
            os.environ["TMK_NO_CLUSTER"] = saved
