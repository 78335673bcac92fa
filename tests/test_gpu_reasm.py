"""The reassembler kept across receive polls (tmk_reasm_*; src/mac/acoustic_interface.rs:316-378 owns one IpReassembler for
all its polls): the records of a capture pushed in several calls must give, link by link, exactly the packets of the
oracle's reassembler fed with all of them in arrival order."""
import numpy as np
import pytest

from test_gpu_pcm_frag import ip_packet, recs_of

pytestmark = pytest.mark.gpu


def make_traffic(orc, rng, n_links):
    items, types, statuses = [], [], []
    for link in range(n_links):
        if link % 7 == 3:
            continue
        srcs = [bytes(rng.integers(0, 256, 4, dtype=np.uint8)) for _ in range(2)]
        arrivals = []
        for d in range(int(rng.integers(2, 10))):
            mtu = int(rng.choice([44, 60, 128]))
            ident = int(rng.integers(0, 3))
            pkt = ip_packet(rng, int(rng.choice([0, 10, 100, 333, 1000])), ident=ident, src=srcs[int(rng.integers(0, 2))],
                            options=int(rng.choice([0, 0, 8])))
            fr, by, _ = orc.fragment_batch([pkt], mtu, ident)
            frs = [bytes(by[int(r["data_off"]):int(r["data_off"]) + int(r["len"])]) for r in fr]
            mode = rng.integers(0, 6)
            if mode == 1:
                rng.shuffle(frs)
            elif mode == 2 and len(frs) > 1:
                frs.pop(int(rng.integers(0, len(frs))))
            elif mode == 3:
                frs.insert(int(rng.integers(0, len(frs) + 1)), frs[int(rng.integers(0, len(frs)))])
            elif mode == 4:
                frs = frs[::-1]
            arrivals.append(frs)
        while any(arrivals):
            k = int(rng.integers(0, len(arrivals)))
            if arrivals[k]:
                t, s = 1, 0
                u = rng.random()
                if u < 0.05:
                    t = 2
                elif u < 0.08:
                    s = int(rng.choice([4, 5, 1]))
                items.append((link, arrivals[k].pop(0))); types.append(t); statuses.append(s)
            if rng.random() < 0.05:
                junk = [bytes(rng.integers(0, 256, int(rng.integers(1, 20)), dtype=np.uint8)), b"\x43" + bytes(40), b"",
                        ip_packet(rng, 50)][int(rng.integers(0, 4))]
                items.append((link, junk)); types.append(2 if junk == b"" else 1); statuses.append(0)
    return items, types, statuses


def per_link(pk, by, n_links):
    out = [[] for _ in range(n_links)]
    for r in pk:
        out[int(r["link"])].append((int(r["status"]), int(r["len"]), int(r["identification"]), int(r["n_fragments"]),
                                    bytes(by[int(r["data_off"]):int(r["data_off"]) + int(r["len"])])))
    return out


@pytest.mark.parametrize("seed", range(4))
def test_pushes_equal_one_reassembler(engine, orc, seed):
    import tmkphy
    rng = np.random.default_rng(500 + seed)
    n_links = 20
    items, types, statuses = make_traffic(orc, rng, n_links)
    recs, pl = recs_of(items, types, statuses)
    want = per_link(*orc.reassemble_batch(recs, pl, n_links), n_links)
    assert sum(len(w) for w in want) > 40 and sum(1 for w in want for p in w if p[3] > 1) > 10
    # every link's arrivals are cut into calls at random places (call numbers ascend along a link)
    n_calls = [1, 2, 5, 40][seed]
    call_of, cur, last_link = [], 0, -1
    for link, _ in items:
        if link != last_link:
            cur, last_link = 0, link
        while cur + 1 < n_calls and rng.random() < (n_calls / 12.0):
            cur += 1
        call_of.append(cur)
    rs = tmkphy.IpReassembler(engine, n_links)
    got = [[] for _ in range(n_links)]
    try:
        for c in range(n_calls):
            idx = [i for i in range(len(items)) if call_of[i] == c]
            r_c, pl_c = recs_of([items[i] for i in idx], [types[i] for i in idx], [statuses[i] for i in idx])
            pk, by = rs.push(r_c, pl_c)
            for l, lst in enumerate(per_link(pk, by, n_links)):
                got[l] += lst
            for r in pk:  # `frame` names the completing record of THIS call
                assert int(r_c[int(r["frame"])]["stream"]) == int(r["link"])
        assert got == want
        held = sum(rs.pending(l)[0] for l in range(n_links))
        pk, by = rs.push(*recs_of([]))           # an empty poll changes nothing
        assert len(pk) == 0 and sum(rs.pending(l)[0] for l in range(n_links)) == held
        # what is held are fragments only, of datagrams the one-shot reassembler did not finish either
        frag_total = sum(1 for w in want for p in w if p[0] == 0 and p[3] > 1)
        assert held > 0 or frag_total == 0 or n_calls >= 1
        rs.reset()
        assert sum(rs.pending(l)[0] for l in range(n_links)) == 0
    finally:
        rs.close()


def test_datagram_across_polls_and_errors(engine, orc):
    """a datagram whose fragments arrive in three polls completes in the third; tmk_reassemble_batch per poll would lose it"""
    import tmkphy
    rng = np.random.default_rng(9)
    pkt = ip_packet(rng, 300, ident=77, src=b"\x0a\x00\x00\x01")
    fr, by, _ = orc.fragment_batch([pkt], 128, 77)
    frs = [bytes(by[int(r["data_off"]):int(r["data_off"]) + int(r["len"])]) for r in fr]
    assert len(frs) == 3
    rs = tmkphy.IpReassembler(engine, 2)
    try:
        for k, f in enumerate([frs[2], frs[0], frs[1]]):
            pk, out = rs.push(*recs_of([(1, f)]))
            one, _ = engine.reassemble_batch(*recs_of([(1, f)]), 2)
            assert len(one) == 0
            if k < 2:
                assert len(pk) == 0 and rs.pending(1)[0] == k + 1 and rs.pending(0)[0] == 0
        assert len(pk) == 1 and pk[0]["status"] == 0 and pk[0]["n_fragments"] == 3 and pk[0]["link"] == 1
        assert bytes(out[int(pk[0]["data_off"]):int(pk[0]["data_off"]) + int(pk[0]["len"])])[20:] == pkt[20:]
        assert rs.pending(1)[0] == 0
        with pytest.raises(tmkphy.TmkError):
            rs.push(*recs_of([(5, frs[0])]))          # a link the reassembler does not have
        with pytest.raises(tmkphy.TmkError):
            rs.push(*recs_of([(1, frs[0]), (0, frs[1])]))  # records not grouped by link
        assert rs.pending(1)[0] == 0 and rs.pending(0)[0] == 0  # errors consume nothing
    finally:
        rs.close()
