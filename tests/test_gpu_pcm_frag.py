"""16-bit PCM ingest and IP fragmentation / reassembly above the PHY (§8 f3) through the C ABI,
bit for bit against the oracle restatements of src/utils/dump.rs, src/audio/codec.rs and src/net/fragmentation.rs."""
import numpy as np
import pytest

import helpers

pytestmark = pytest.mark.gpu
MAN, FB = 0, 1


# ------------------------------------------------------------------ f2: 16-bit PCM
@pytest.mark.parametrize("n", [0, 1, 3, 4, 5, 1023, 48000, 1 << 20])
def test_pcm16_conversions(engine, orc, n):
    rng = np.random.default_rng(n + 5)
    x = (rng.normal(0, 0.6, n)).astype(np.float32)
    special = np.array([0.0, -0.0, 1.0, -1.0, 0.5, -0.5, 1.00001, -1.00001, 2.0, -3.0, np.nan, np.inf, -np.inf,
                        1e-9, -1e-9, 32766.6 / 32767, -32767.9 / 32767, 1e30, -1e30, 3.0518e-5], np.float32)
    x[:min(n, len(special))] = special[:min(n, len(special))]
    pcm = engine.pcm16_encode(x)
    assert pcm.dtype == np.int16 and np.array_equal(pcm, orc.pcm16_encode(x))
    back = engine.pcm16_to_f32(pcm)
    assert np.array_equal(back, orc.pcm16_to_f32(pcm))
    allv = np.arange(-32768, 32768, dtype=np.int16)[:max(n, 1)]
    assert np.array_equal(engine.pcm16_to_f32(allv), orc.pcm16_to_f32(allv))


@pytest.mark.parametrize("where", ["host", "device"])
@pytest.mark.parametrize("snr", [None, 12.0, 8.0])
def test_decode_batch_pcm16(engine, orc, where, snr):
    """int16 captures decode exactly as the oracle decodes the widened samples (codec.rs:121-125)"""
    import tmkphy
    rng = np.random.default_rng(41 if snr is None else int(snr))
    caps, cfgs = [], []
    for i in range(24):
        cfg = orc.make_config(FB if i % 2 else MAN, local_addr=1, chunk=[0, 1200, 257][i % 3])
        c = helpers.make_capture(orc, cfg, rng, n_frames=5, snr_db=snr, amplitude=float(rng.uniform(0.2, 0.9)))
        caps.append(c)
        cfgs.append(cfg)
    caps.append(np.zeros(0, np.float32))
    cfgs.append(orc.make_config(FB))
    caps.append(np.zeros(37, np.float32))
    cfgs.append(orc.make_config(MAN))
    samples, st = helpers.pack(caps, cfgs, tmkphy)
    pcm = orc.pcm16_encode(samples)          # what dump_to_wav would have stored
    widened = orc.pcm16_to_f32(pcm)
    want, want_pl, _ = orc.decode_batch(widened, st)
    if where == "host":
        got, got_pl = engine.decode_batch_pcm16(pcm, st)
    else:
        d = engine.device_alloc(pcm.nbytes)
        try:
            engine.h2d(d, pcm)
            got, got_pl = engine.decode_batch_pcm16(d, st, total_samples=pcm.size)
        finally:
            engine.device_free(d)
    assert (want["status"] == 0).sum() >= 20
    helpers.assert_records_equal(got, got_pl, want, want_pl, f"pcm16 {where} snr={snr}")
    # and the f32 entry point on the widened samples gives the very same records
    got2, got2_pl = engine.decode_batch(widened, st)
    helpers.assert_records_equal(got2, got2_pl, want, want_pl, "f32 of widened")


def test_pcm16_multichunk_host(engine, orc, monkeypatch):
    """host int16 captures staged in several chunks (copy overlapped with the widen + scan of the chunk before)"""
    import tmkphy
    monkeypatch.setenv("TMK_CHUNK_MSAMPLES", "1")
    rng = np.random.default_rng(9)
    cfg = orc.make_config(FB, local_addr=1)
    caps = [np.concatenate([helpers.make_capture(orc, cfg, rng, n_frames=8, snr_db=15.0) for _ in range(10)]) for _ in range(12)]
    samples, st = helpers.pack(caps, [cfg] * len(caps), tmkphy)
    assert samples.size > 2 << 20
    pcm = orc.pcm16_encode(samples * np.float32(0.7))
    want, want_pl, _ = orc.decode_batch(orc.pcm16_to_f32(pcm), st)
    got, got_pl = engine.decode_batch_pcm16(pcm, st)
    helpers.assert_records_equal(got, got_pl, want, want_pl, "pcm16 chunks")


# ------------------------------------------------------------------ f3: fragments
def ip_packet(rng, n_data, ident=None, src=None, options=0, proto=17):
    """an IPv4 packet with `options` option bytes (multiple of 4) and n_data payload bytes"""
    ihl = 20 + options
    h = bytearray(ihl)
    h[0] = 0x40 | (ihl // 4)
    tl = ihl + n_data
    h[2:4] = tl.to_bytes(2, "big")
    h[4:6] = (int(rng.integers(0, 65536)) if ident is None else ident).to_bytes(2, "big")
    h[8], h[9] = 64, proto
    h[10:12] = bytes(rng.integers(0, 256, 2, dtype=np.uint8))
    h[12:16] = bytes(src) if src is not None else bytes(rng.integers(0, 256, 4, dtype=np.uint8))
    h[16:20] = bytes(rng.integers(0, 256, 4, dtype=np.uint8))
    if options:
        h[20:ihl] = bytes(rng.integers(1, 256, options, dtype=np.uint8))
    return bytes(h) + bytes(rng.integers(0, 256, n_data, dtype=np.uint8))


def assert_pkts_equal(got, gb, want, wb, ctx):
    assert len(got) == len(want), f"{ctx}: {len(got)} records, oracle {len(want)}"
    for f in ("link", "status", "frame", "len", "data_off", "identification", "n_fragments"):
        if not np.array_equal(got[f], want[f]):
            i = int(np.nonzero(got[f] != want[f])[0][0])
            raise AssertionError(f"{ctx}: field {f} differs at record {i}: {got[i]} vs oracle {want[i]}")
    assert np.array_equal(gb, wb), f"{ctx}: bytes differ"


@pytest.mark.parametrize("mtu", [10, 28, 60, 100, 128, 576, 1500])
def test_fragment_batch(engine, orc, mtu):
    rng = np.random.default_rng(mtu)
    pkts = []
    for i in range(300):
        opt = int(rng.choice([0, 0, 0, 4, 8, 40]))
        n = int(rng.choice([0, 1, 7, 8, 9, 40, 100, 108, 300, 1000, 1480, 3000, 9000]))
        pkts.append(ip_packet(rng, n, options=opt))
    pkts += [b"", b"\x45" * 5, b"\x45" * 19, b"\x41" + bytes(199), b"\x4f" + bytes(39), b"\x4f" + bytes(700), bytes(2000),
             b"\x45" * 20, b"\x45" * 21, b"\x46" * 24 + bytes(64)]
    first = 65500 if mtu == 100 else 7
    want, wb, wid = orc.fragment_batch(pkts, mtu, first)
    got, gb, gid = engine.fragment_batch(pkts, mtu, first)
    assert_pkts_equal(got, gb, want, wb, f"fragment mtu={mtu}")
    assert gid == wid
    ok = got[got["status"] == 0]
    assert (ok["len"] <= mtu).all()
    if mtu == 100:
        assert wid < 65500  # the identification counter wrapped


def test_fragment_reference_vectors(engine):
    """the reference's own unit tests (src/net/fragmentation.rs:426-515) through the C ABI"""
    import tmkphy
    fr = tmkphy.IpFragmenter(engine, 500)
    assert fr.fragment_packet(bytes(100)) == [bytes(100)]                       # test_no_fragmentation_needed
    frags = tmkphy.IpFragmenter(engine, 100).fragment_packet(b"\x45" * 20 + bytes(300))   # test_fragmentation_basic
    assert len(frags) > 1 and all(len(f) <= 100 for f in frags)
    hdr = bytes([0x45, 0, 0, 0, 0, 0, 0, 0, 0x40, 0x11, 0, 0, 192, 168, 1, 1, 192, 168, 1, 2])
    payload = bytes(range(100))
    pkt = bytearray(hdr + payload)
    pkt[2:4] = len(pkt).to_bytes(2, "big")
    frags = tmkphy.IpFragmenter(engine, 60).fragment_packet(bytes(pkt))         # test_fragment_assemble
    assert len(frags) > 1
    recs, pl = recs_of([(0, f) for f in frags])
    pk, by = engine.reassemble_batch(recs, pl, 1)
    out = tmkphy.packets_of(pk, by)
    assert len(out) == 1 and out[0][1][20:] == payload and int(pk[-1]["frame"]) == len(frags) - 1


def recs_of(items, types=None, statuses=None):
    """frame records + payload buffer as tmk_decode_batch would return them: items = [(link, data bytes)] sorted by link"""
    from tmkphy._abi import REC_DTYPE
    recs = np.zeros(len(items), REC_DTYPE)
    blob = bytearray()
    for i, (link, data) in enumerate(items):
        recs[i]["stream"] = link
        recs[i]["status"] = 0 if statuses is None else statuses[i]
        recs[i]["type"] = 1 if types is None else types[i]
        recs[i]["len"] = len(data)
        recs[i]["payload_off"] = len(blob)
        recs[i]["hdr_valid"] = 1
        blob += data
        blob += bytes(int(i % 3))  # payload offsets need not be dense
    return recs, np.frombuffer(bytes(blob) + b"\0", np.uint8).copy()


@pytest.mark.parametrize("seed", range(6))
def test_reassemble_records(engine, orc, seed):
    """random arrival orders, duplicates, losses, interleaved datagrams, key reuse, malformed and filtered frames"""
    rng = np.random.default_rng(100 + seed)
    n_links = 70
    items, types, statuses = [], [], []
    for link in range(n_links):
        if link % 17 == 5:
            continue  # a link without frames
        srcs = [bytes(rng.integers(0, 256, 4, dtype=np.uint8)) for _ in range(2)]
        arrivals = []
        for d in range(int(rng.integers(1, 9))):
            mtu = int(rng.choice([44, 60, 128]))
            ident = int(rng.integers(0, 3))           # few identifications: keys collide and get reused
            pkt = ip_packet(rng, int(rng.choice([0, 10, 100, 333, 1000])), ident=ident, src=srcs[int(rng.integers(0, 2))],
                            options=int(rng.choice([0, 0, 8])))
            fr, by, _ = orc.fragment_batch([pkt], mtu, ident)
            frs = [bytes(by[int(r["data_off"]):int(r["data_off"]) + int(r["len"])]) for r in fr]
            mode = rng.integers(0, 6)
            if mode == 1:
                rng.shuffle(frs)
            elif mode == 2 and len(frs) > 1:
                frs.pop(int(rng.integers(0, len(frs))))           # a loss: never completes (or completes later)
            elif mode == 3:
                frs.insert(int(rng.integers(0, len(frs) + 1)), frs[int(rng.integers(0, len(frs)))])  # a duplicate
            elif mode == 4:
                frs = frs[::-1]
            arrivals.append(frs)
        # interleave the datagrams of the link
        while any(arrivals):
            k = int(rng.integers(0, len(arrivals)))
            if arrivals[k]:
                f = arrivals[k].pop(0)
                t, s = 1, 0
                u = rng.random()
                if u < 0.05:
                    t = 2                                     # Ack carrying data: accepted
                elif u < 0.08:
                    s = int(rng.choice([4, 5, 1]))            # NOT_FOR_US / CRC_FAIL / HDR_PARSE: not a Frame
                items.append((link, f)); types.append(t); statuses.append(s)
            if rng.random() < 0.06:
                junk = [bytes(rng.integers(0, 256, int(rng.integers(1, 20)), dtype=np.uint8)),   # < 20 bytes
                        b"\x43" + bytes(40), b"\x4f" + bytes(30), b"", ip_packet(rng, 50)][int(rng.integers(0, 5))]
                items.append((link, junk)); types.append(2 if junk == b"" else 1); statuses.append(0)
    recs, pl = recs_of(items, types, statuses)
    want, wb = orc.reassemble_batch(recs, pl, n_links)
    got, gb = engine.reassemble_batch(recs, pl, n_links)
    assert (want["status"] == 0).sum() > 50 and (want["n_fragments"] > 1).sum() > 20
    assert_pkts_equal(got, gb, want, wb, f"reassemble seed={seed}")
    # device-resident records / payload give the same
    dr, dp = engine.device_alloc(recs.nbytes), engine.device_alloc(pl.nbytes)
    try:
        engine.h2d(dr, recs)
        engine.h2d(dp, pl)
        import ctypes as C
        from tmkphy import _abi
        pk = np.zeros(len(recs), _abi.PKT_DTYPE)
        by = np.zeros(pl.size + 64, np.uint8)
        n, nb = C.c_uint64(0), C.c_uint64(0)
        rc = engine._lib.tmk_reassemble_batch(engine._h, C.c_void_p(dr), len(recs), C.c_void_p(dp), pl.size, n_links,
                                              pk.ctypes.data_as(C.c_void_p), len(pk), C.byref(n),
                                              by.ctypes.data_as(C.c_void_p), by.size, C.byref(nb))
        assert rc == 0
        assert_pkts_equal(pk[:n.value], by[:nb.value], want, wb, "device records")
    finally:
        engine.device_free(dr)
        engine.device_free(dp)


def test_reassemble_limits_and_capacity(engine, orc):
    from tmkphy import _abi
    rng = np.random.default_rng(3)
    # 512 fragments under one key reassemble; 600 do not (engine limit, reported, never silently wrong)
    def frags_of(n_frag, ident):
        pkt = ip_packet(rng, 8 * n_frag, ident=ident, src=b"\x0a\x00\x00\x01")
        fr, by, _ = orc.fragment_batch([pkt], 28, ident)
        assert len(fr) == n_frag
        return pkt, [bytes(by[int(r["data_off"]):int(r["data_off"]) + int(r["len"])]) for r in fr]
    pkt_a, fa = frags_of(512, 1)
    pkt_b, fb = frags_of(600, 2)
    order = rng.permutation(512)
    items = [(0, fa[i]) for i in order] + [(1, f) for f in fb] + [(1, ip_packet(rng, 30))]
    recs, pl = recs_of(items)
    got, gb = engine.reassemble_batch(recs, pl, 2)
    want, wb = orc.reassemble_batch(recs, pl, 2)
    g0 = got[got["link"] == 0]
    assert len(g0) == 1 and g0[0]["status"] == 0 and g0[0]["n_fragments"] == 512
    a = bytearray(pkt_a)
    assert bytes(gb[int(g0[0]["data_off"]):int(g0[0]["data_off"]) + int(g0[0]["len"])])[20:] == bytes(a[20:])
    assert_pkts_equal(g0, gb[:int(g0[0]["len"])], want[want["link"] == 0], wb[:int(g0[0]["len"])], "512 fragments")
    g1 = got[got["link"] == 1]
    assert (g1["status"] == _abi.PKT_E_TOO_MANY).sum() >= 1 and g1[-1]["status"] == 0 and g1[-1]["n_fragments"] == 1
    # capacity: counts are returned, nothing is written past the caps
    import tmkphy
    with pytest.raises(tmkphy.TmkError):
        engine.reassemble_batch(recs, pl, 2, pkt_cap=1, byte_cap=16)
    # empty inputs
    e_recs, e_pl = recs_of([])
    pk, by = engine.reassemble_batch(e_recs, e_pl, 4)
    assert len(pk) == 0 and len(by) == 0


def test_phy_to_packets_end_to_end(engine, orc):
    """packets -> fragments (GPU) -> frames -> waveform (GPU) -> noisy capture -> decode (GPU) -> reassemble (GPU)"""
    import tmkphy
    rng = np.random.default_rng(77)
    n_links = 12
    caps, cfgs, sent = [], [], []
    for link in range(n_links):
        kind = FB if link % 2 else MAN
        cfg = orc.make_config(kind, local_addr=2)
        fragger = tmkphy.IpFragmenter(engine, 128)     # MAX_FRAME_DATA_SIZE, src/utils/consts.rs:37
        pkts = [ip_packet(rng, int(rng.choice([20, 100, 400, 900])), src=b"\xc0\xa8\x01\x01") for _ in range(4)]
        sent.append(pkts)
        frames, seq = [], 0
        for frs in fragger.fragment_packets(pkts):
            for f in frs:
                frames.append(tmkphy.Frame.new_data(seq & 0xff, 1, 2, f))
                seq += 1
        wave = tmkphy.PhyEncoder(engine, 3, 2, kind).encode_frames(frames, 48)
        x = np.concatenate([np.zeros(100, np.float32), wave, np.zeros(300, np.float32)])
        x = (x * 0.6 + rng.normal(0, 0.03, len(x))).astype(np.float32)
        caps.append(x)
        cfgs.append(cfg)
    samples, st = helpers.pack(caps, cfgs, tmkphy)
    recs, pl = engine.decode_batch(samples, st)
    want_r, want_pl, _ = orc.decode_batch(samples, st)
    helpers.assert_records_equal(recs, pl, want_r, want_pl, "e2e decode")
    pk, by = engine.reassemble_batch(recs, pl, n_links)
    wpk, wby = orc.reassemble_batch(want_r, want_pl, n_links)
    assert_pkts_equal(pk, by, wpk, wby, "e2e reassemble")
    out = tmkphy.packets_of(pk, by)
    for link in range(n_links):
        mine = [p for l, p in out if l == link]
        assert len(mine) == 4
        for a, b in zip(mine, sent[link]):
            # reassembly keeps everything but identification (fragmented packets) / flags
            assert a[20:] == b[20:] and a[:4] == b[:4] and a[8:20] == b[8:20]
