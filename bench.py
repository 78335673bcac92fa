#!/usr/bin/env python
"""bench.py — Msamples/s of the fused preamble-sync + demod path (BASELINE.json metric).

A "step" is one pass of the hot path (tmk_decode_batch: scan -> cluster -> walk -> record ordering) over one batch
of synthetic captures. The HEADLINE workload is BASELINE config 2: 4096 streams x 10 s x 48 kHz per GPU (480 000
samples each, 7.86 GB of fp32), src/phy 4B5B frames at random offsets, AWGN SNR sweep {-3..30 dB}, weak scaling
(every rank owns its own 4096 streams; no data-path collective, NCCL only sums the decode counters afterwards).

  value        whole-job Msamples/s, captures already resident in HBM, CUDA events on the engine's stream
  e2e          the same through the C ABI with HOST (pinned) sample buffers: H2D + kernels + D2H of records
  roofline     the PIPELINE: 4 B/sample algorithmic bytes / the whole step time vs the measured HBM peak, with
               per-kernel entries (scan / cluster / walk: ms, algorithmic bytes, ncu DRAM traffic)
  cpu_baseline the oracle (C++ restatement of the reference's src/phy, AVX dot) on the box's host cores,
               records compared bit for bit with the GPU's
  configs      every other BASELINE config at its stated size, each with Msamples/s, stage times, the fraction of
               the HBM roofline of the whole step, and bit-exact parity against the oracle on a bounded sample:
               C2 Manchester, C2 fed in 1200-sample process_samples calls, C3 (16 384 x 2^20, multipath) for the
               src/phy line codes AND for the FSK / OFDM demodulation on the batched FFT (build-defined spec),
               C4 (8 hour-long captures with clock drift), C5 (131 072 mixed-modulation streams x 2^18, sharded
               over the ranks: the strong-scaling leg)

--impl reference times the CPU implementation alone, on a bounded sample of the SAME captures (the GPU arm's generator restated
on the host, oracle/synth_host.py) (the reference is Rust; there is no rustc here or on the GPU
box — probed, gpurun_out/r2_probe.txt — so oracle/_ref cannot exist: kind = "port").
No torch anywhere: under torchrun the ranks find each other through RANK / WORLD_SIZE / LOCAL_RANK and exchange
the NCCL id through a file; the reduction is tmk_comm_* of the C ABI.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "trackmaker-rs_b200"))

N_STREAMS = 4096
N_PER_STREAM = 480000
SNRS = [-3.0, 0.0, 3.0, 6.0, 10.0, 15.0, 20.0, 30.0]
METRIC = "Msamples/s sync+demod (48 kHz streams)"
WORKLOAD = "C2: 4096 src/phy 4B5B streams x 10 s @ 48 kHz per GPU, AWGN SNR sweep -3..30 dB"
# dram__bytes_read.sum + dram__bytes_write.sum per launch at the headline workload, from the ncu --set full capture
# named below; null in the JSON line for any other workload
NCU_TRAFFIC_FILE = os.path.join(ROOT, "profiles", "ncu_traffic.json")
FIELDS = ("stream", "status", "preamble_start", "frame_start", "consumed", "len", "crc", "type", "seq", "src", "dst",
          "hdr_valid")


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler(threading.Thread):
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False

    def run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                self.rows.append([c.strip() for c in out.strip().split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        rows = [r for r in self.rows if len(r) == 6 and r[0].isdigit()]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[2 + i] == "Active" for r in rows)]
        return {"sm_mhz": float(np.median([int(r[0]) for r in rows])), "sm_max_mhz": float(rows[0][1]),
                "reasons": reasons, "samples": len(rows)}


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def payload_in_record_order(recs, pl):
    keep = np.isin(recs["status"], (0, 4, 5))
    return b"".join(pl[int(o):int(o) + int(n)].tobytes() for o, n in zip(recs["payload_off"][keep], recs["len"][keep]))


def records_equal(grecs, gpl, crecs, cpl):
    return (len(grecs) == len(crecs) and all(np.array_equal(grecs[f], crecs[f]) for f in FIELDS)
            and payload_in_record_order(grecs, gpl) == payload_in_record_order(crecs, cpl))


# ------------------------------------------------------------------------------------------------ rendezvous
def comm_setup(eng, tmkphy, rank, world):
    """one NCCL communicator per engine; rank 0 hands the id to the others through a file (all ranks of a launch are
    processes of one node and children of one launcher)"""
    if world <= 1:
        return
    tag = f"{os.environ.get('MASTER_PORT', '0')}_{os.environ.get('TORCHELASTIC_RUN_ID', 'x')}_{os.getppid()}"
    path = os.path.join(os.environ.get("TMPDIR", "/tmp"), f"tmkphy_ncclid_{tag}")
    if rank == 0:
        uid = tmkphy.comm_unique_id()
        with open(path + ".tmp", "wb") as f:
            f.write(uid)
        os.replace(path + ".tmp", path)
    else:
        t0 = time.time()
        while not os.path.exists(path):
            if time.time() - t0 > 180:
                raise RuntimeError("rank 0 never published the NCCL id")
            time.sleep(0.01)
        uid = open(path, "rb").read()
    eng.comm_init(uid, rank, world)
    eng.comm_barrier()
    if rank == 0:
        try:
            os.remove(path)
        except OSError:
            pass


# ------------------------------------------------------------------------------------------------ legs
def synth_groups(eng, cfg, seed, n_s, n_per, d_x, snrs, **imp):
    """n_s captures at d_x (device), SNR group g = stream // (n_s / len(snrs)); returns frames planted"""
    per_g, planted = n_s // len(snrs), 0
    for g, snr in enumerate(snrs):
        cnt = per_g if g + 1 < len(snrs) else n_s - per_g * (len(snrs) - 1)
        if cnt <= 0:
            continue
        if imp:
            _, nt = eng.synth_batch_ex(cfg, seed + g, cnt, n_per, snr, 0.7, device_ptr=d_x + g * per_g * n_per * 4,
                                       want_truth=False, **imp)
        else:
            _, nt = eng.synth_batch(cfg, seed + g, cnt, n_per, snr, 0.7, device_ptr=d_x + g * per_g * n_per * 4,
                                    want_truth=False)
        planted += nt
    return planted


class Decode:
    """resident-data decode steps through the C ABI (device sample / record / payload buffers)"""

    def __init__(self, eng, d_x, total, st, rec_cap=None, pay_cap=None):
        self.eng, self.d_x, self.total, self.st = eng, d_x, total, st
        self.rec_cap = rec_cap or (total // 256 + 4 * len(st))
        self.pay_cap = pay_cap or (total // 24 + 65536)
        self.d_recs = eng.device_alloc(self.rec_cap * 48)
        self.d_pay = eng.device_alloc(self.pay_cap)
        self.nr, self.npl = C.c_uint64(0), C.c_uint64(0)
        self.stp = st.ctypes.data_as(C.c_void_p)

    def step(self):
        e = self.eng
        e._check(e._lib.tmk_decode_batch(e._h, C.c_void_p(self.d_x), self.total, self.stp, len(self.st),
                                         C.c_void_p(self.d_recs), self.rec_cap, C.byref(self.nr),
                                         C.c_void_p(self.d_pay), self.pay_cap, C.byref(self.npl)))

    def timed(self, steps, warmup, barrier=None):
        """-> (ms for `steps` steps, stage times per step from extra untimed steps, launches per step)"""
        for _ in range(warmup):
            self.step()
        if barrier:
            barrier()
        l0 = self.eng.launch_count()
        self.eng.timer_start()
        for _ in range(steps):
            self.step()
        ms = self.eng.timer_stop()
        launches = self.eng.launch_count() - l0
        if barrier:
            barrier()
        stage = {"scan": 0.0, "cluster": 0.0, "walk": 0.0, "order": 0.0}
        n_extra = 2
        for _ in range(n_extra):  # stage breakdown outside the timed region
            self.step()
            s = self.eng.last_stage_ms()
            cl = self.eng.debug_counters()["cluster_us"] * 1e-3
            stage["scan"] += s["scan"] / n_extra
            stage["cluster"] += cl / n_extra
            stage["walk"] += (s["walk"] - cl) / n_extra
            stage["order"] += s["order"] / n_extra
        return ms, stage, launches

    def free(self):
        self.eng.device_free(self.d_recs)
        self.eng.device_free(self.d_pay)


def parity_sample(eng, orc, tmkphy, d_x, st, idx, threads):
    """the captures st[idx] copied back from HBM, decoded by the oracle (all host threads) and by the engine through
    the host-buffer path; -> (bit-exact?, records, oracle Msamples/s)"""
    lens = [int(st[i]["n_samples"]) for i in idx]
    cfgs = [tmkphy.PhyConfig(*[st[i]["cfg"][k].item() for k in ("line_code", "spl", "preamble_bytes", "local_addr", "threshold",
                                                         "max_frame_bytes", "chunk")]) for i in idx]
    sub_st, total = tmkphy.make_streams(lens, cfgs)
    sub = np.zeros(total, np.float32)
    for k, i in enumerate(idx):
        o = int(sub_st[k]["offset"])
        eng.d2h(sub[o:o + lens[k]], d_x + int(st[i]["offset"]) * 4)
    crecs, cpl, sec = orc.decode_batch(sub, sub_st, n_threads=threads)
    grecs, gpl = eng.decode_batch(sub, sub_st, rec_cap=len(crecs) + 64, payload_cap=len(cpl) + 4096)
    return records_equal(grecs, gpl, crecs, cpl), int(len(crecs)), total / sec / 1e6


def phy_leg(eng, orc, tmkphy, name, d_x, st, total, steps, warmup, hbm, n_parity, threads, extra=None):
    dec = Decode(eng, d_x, total, st)
    ms, stage, _ = dec.timed(steps, warmup)
    res = {"workload": name, "msamples_per_s": total * steps / (ms * 1e-3) / 1e6, "ms_per_step": ms / steps,
           "stage_ms": {k: round(v, 4) for k, v in stage.items()},
           "frac_hbm": 4.0 * total / (ms / steps * 1e-3) / 1e9 / hbm, "records": int(dec.nr.value),
           "walk_counters": eng.debug_counters()}
    dec.free()
    if orc is not None and n_parity:
        idx = np.unique(np.linspace(0, len(st) - 1, n_parity).astype(np.int64))
        same, nrec, cpu = parity_sample(eng, orc, tmkphy, d_x, st, idx, threads)
        res["parity"] = "bit-exact" if same else "MISMATCH"
        res["parity_sample"] = f"{len(idx)} captures, {nrec} records, oracle {cpu:.0f} Msamples/s on {threads} threads"
    if extra:
        res.update(extra)
    return res


def ask_leg(eng, tmkphy, orc, d_x, n_s, n_per, hbm, check_streams=8):
    """BASELINE config 2 read literally ("PSK streams"): the chirp + BPSK modem of examples/ask.rs over the same
    n_s x n_per samples resident in HBM; captures from tmk_ask_synth_batch (distinct frames per stream, AWGN sweep)."""
    per_g = n_s // len(SNRS)
    planted = 0
    for g, snr in enumerate(SNRS):
        cnt = per_g if g + 1 < len(SNRS) else n_s - per_g * (len(SNRS) - 1)
        _, nf = eng.ask_synth_batch(0xA5C0 + g, cnt, n_per, snr, device_ptr=d_x + g * per_g * n_per * 4)
        planted += nf
    st, total = tmkphy.make_streams([n_per] * n_s, tmkphy.make_config(1))
    for _ in range(2):
        recs = eng.ask_decode_batch(d_x, st, total_samples=total)
    ms = []
    for _ in range(3):
        eng.ask_decode_batch(d_x, st, total_samples=total)
        ms.append(eng.ask_last_stage_ms())
    scan = float(np.median([m["scan"] for m in ms]))
    walk = float(np.median([m["walk"] for m in ms]))
    # the direct-form correlation is 2 * 440 + 6 flop per sample: % of the fp32 peak it would need
    fp32_peak = 148 * 128 * 2 * 1.965e9
    res = {"workload": f"{n_s} chirp+BPSK streams x {n_per} samples (examples/ask.rs frames, AWGN sweep -3..30 dB)",
           "value": total / (scan + walk) / 1e3, "unit": "Msamples/s", "filter_ms": scan, "walk_ms": walk,
           "frames_planted": planted, "frames": int(len(recs)), "accepted": int(recs["accepted"].sum()),
           "frac_hbm": 4.0 * total / ((scan + walk) * 1e-3) / 1e9 / hbm,
           "filter_frac_hbm": 4.0 * total / (scan * 1e-3) / 1e9 / hbm,
           "filter_frac_fp32_direct_form": 886.0 * total / (scan * 1e-3) / fp32_peak}
    if orc is not None:
        idx = np.unique(np.linspace(0, n_s - 1, check_streams).astype(np.int64))
        t = time.time()
        same = True
        for i in idx:
            x = np.zeros(n_per, np.float32)
            eng.d2h(x, d_x + int(i) * n_per * 4)
            want = orc.ask_demodulate(x, stream=int(i))
            got = recs[recs["stream"] == i]
            same = same and len(got) == len(want) and all(np.array_equal(got[f], want[f]) for f in
                                                          ("stream", "start_index", "id", "bits", "accepted"))
        res["cpu_oracle_1core_msamples_per_s"] = len(idx) * n_per / (time.time() - t) / 1e6
        res["parity"] = "bit-exact" if same else "MISMATCH"
        res["parity_sample"] = f"{len(idx)} captures"
    return res


def links_leg(eng, orc, tmkphy, n_links=256, n_polls=100, poll=1200, check_links=8):
    """Carried-state streaming decode (PhyDecoder::process_samples, decoder.rs:68-112, at the cadence of
    csma.rs:446 — 25 ms polls of 1200 samples): n_links decoders fed poll by poll out of captures resident in HBM through
    tmk_links_process_samples; a poll is one call, timed on the host from call to return (records and payload in host
    memory). Parity: every record of `check_links` links, poll by poll, against the oracle's decoder fed with the same calls."""
    import ctypes as C
    n_per = poll * n_polls
    cfg = tmkphy.make_config(tmkphy.LINE_4B5B, local_addr=1)
    d_x = eng.device_alloc(n_links * n_per * 4)
    eng.synth_batch(cfg, 4242, n_links, n_per, 15.0, 0.7, device_ptr=d_x, want_truth=False)
    bank = tmkphy.PhyDecoderBank(eng, [cfg] * n_links)
    lib, h = eng._lib, bank._h
    rec_cap, pay_cap = 8 * n_links + 1024, 300 * n_links + 65536
    recs = np.zeros(rec_cap, tmkphy.REC_DTYPE)
    pl = np.zeros(pay_cap, np.uint8)
    nr, npl = C.c_uint64(0), C.c_uint64(0)
    lens = np.full(n_links, poll, np.uint64)
    base = np.arange(n_links, dtype=np.uint64) * np.uint64(n_per)
    p_x, p_lens, p_recs, p_pl = C.c_void_p(d_x), C.c_void_p(lens.ctypes.data), C.c_void_p(recs.ctypes.data), C.c_void_p(pl.ctypes.data)
    times, got = [], []
    for k in range(n_polls):
        offs = base + np.uint64(k * poll)
        p_offs = C.c_void_p(offs.ctypes.data)
        t0 = time.perf_counter()
        rc = lib.tmk_links_process_samples(h, p_x, n_links * n_per, p_offs, p_lens, p_recs, rec_cap, C.byref(nr), p_pl, pay_cap, C.byref(npl))
        times.append(time.perf_counter() - t0)
        eng._check(rc)
        got.append((recs[:nr.value].copy(), pl[:npl.value].copy()))
    ms = np.array(times[5:]) * 1e3
    res = {"workload": f"{n_links} links x {n_polls} polls of {poll} samples (25 ms at 48 kHz), 4B5B frames, AWGN 15 dB, captures resident in HBM",
           "ms_per_poll": float(np.median(ms)), "ms_per_poll_p99": float(np.percentile(ms, 99)),
           "realtime_factor": 25.0 / float(np.median(ms)), "Msamples_per_s": n_links * poll / float(np.median(ms)) / 1e3,
           "frames_ok": int(sum(int((r["status"] == 0).sum()) for r, _ in got))}
    if orc is not None:
        same, n_cmp = True, 0
        for i in np.unique(np.linspace(0, n_links - 1, check_links).astype(np.int64)):
            x = np.zeros(n_per, np.float32)
            eng.d2h(x, d_x + int(i) * n_per * 4)
            want, wpl, feed = orc.decode_feeds(cfg, x, [poll] * n_polls, stream_id=int(i))
            for k in range(n_polls):
                r, p = got[k]
                g, w = r[r["stream"] == i], want[feed == k]
                n_cmp += len(w)
                same = same and records_equal(g, p, w, wpl)
        res["parity"] = "bit-exact" if same else "MISMATCH"
        res["parity_sample"] = f"{check_links} links, {n_cmp} records, poll by poll"
    bank.close()
    eng.device_free(d_x)
    return res


def mcm_leg(eng, orc, tmkphy, hbm, n_s, n_per, check_streams=4):
    """BASELINE config 3 as worded: FSK / OFDM demodulation with the batched FFT over n_s captures x n_per samples
    (half FSK, half OFDM), 3-tap multipath + AWGN sweep. Build-defined specification (the reference has no FSK / OFDM /
    FFT code): parity is against the build's own oracle and is UNPINNED by the reference."""
    try:
        d_x = eng.device_alloc(n_s * n_per * 4)
    except tmkphy.TmkError:
        return {"skipped": f"cannot allocate {n_s * n_per * 4 / 1e9:.1f} GB"}
    half = n_s // 2
    per_g = max(1, half // len(SNRS))
    planted = [0, 0]
    for m in (0, 1):
        g0 = 0
        for g, snr in enumerate(SNRS):
            cnt = per_g if g + 1 < len(SNRS) else half - g0
            cnt = min(cnt, half - g0)
            if cnt <= 0:
                break
            _, nf = eng.mcm_synth_batch(m, 0xC3F0 + 16 * m + g, cnt, n_per, snr, multipath=True,
                                        device_ptr=d_x + (m * half + g0) * n_per * 4)
            planted[m] += nf
            g0 += cnt
    st, tot = tmkphy.make_streams([n_per] * half, tmkphy.make_config(1))
    cap = tot // 5000 + 2 * half
    res = {"workload": f"C3 (FSK/OFDM wording): {n_s} captures x {n_per} samples, {half} FSK + {half} OFDM frames streams, "
                       "multipath h = [1, 0.35 @ t1, -0.2 @ t2] + AWGN sweep -3..30 dB; chirp sync (examples/ask.rs) + "
                       "batched shared-memory FFT demodulation, symbol decision, CRC-8",
           "parity_pinned_by_reference": False, "modes": {}}
    total_ms, ok_all = 0.0, True
    for m, name in ((0, "fsk"), (1, "ofdm")):
        base = d_x + m * half * n_per * 4
        recs, pl = eng.mcm_decode_batch(m, base, st, rec_cap=cap, total_samples=tot)  # warm-up
        ms = []
        for _ in range(2):
            recs, pl = eng.mcm_decode_batch(m, base, st, rec_cap=cap, total_samples=tot)
            ms.append(eng.mcm_last_stage_ms())
        stage = {k: float(np.median([x[k] for x in ms])) for k in ("filter", "walk", "demod")}
        t = sum(stage.values())
        total_ms += t
        r = {"msamples_per_s": tot / t / 1e3, "stage_ms": {k: round(v, 3) for k, v in stage.items()},
             "frames_planted": planted[m], "records": int(len(recs)), "frames_ok": int((recs["status"] == 0).sum()),
             "payload_bytes": int(len(pl)),
             "demod_ffts": int(len(recs)) * (56 if m == 0 else 14), "fft_points": 64 if m == 0 else 256}
        if orc is not None:
            same = True
            for i in np.unique(np.linspace(0, half - 1, check_streams).astype(np.int64)):
                x = np.zeros(n_per, np.float32)
                eng.d2h(x, base + int(i) * n_per * 4)
                want, wpl = orc.mcm_demodulate(m, x, stream=int(i))
                got = recs[recs["stream"] == i]
                same = same and len(got) == len(want) and all(
                    np.array_equal(got[f], want[f]) for f in ("stream", "status", "start_index", "seq", "len", "crc"))
                if same:
                    gp = b"".join(bytes(pl[int(a["payload_off"]):int(a["payload_off"]) + int(a["len"])]) for a in got if a["status"] != 2)
                    same = gp == bytes(wpl)
            r["parity"] = "bit-exact vs the build's own oracle" if same else "MISMATCH"
            r["parity_sample"] = f"{check_streams} captures"
            ok_all = ok_all and same
        res["modes"][name] = r
    res["msamples_per_s"] = n_s * n_per / total_ms / 1e3
    res["ms_per_step"] = total_ms
    res["frac_hbm"] = 4.0 * n_s * n_per / (total_ms * 1e-3) / 1e9 / hbm
    if orc is not None:
        res["parity"] = "bit-exact" if ok_all else "MISMATCH"
        res["parity_note"] = "against oracle/mcm_modem.hpp (build-defined spec); unpinned by the reference"
    eng.device_free(d_x)
    return res


def c5_leg(eng, orc, tmkphy, rank, world, hbm, steps, warmup, n_total, n_per, threads, barrier):
    """BASELINE config 5: n_total mixed-modulation streams x n_per samples, stream_id % 3 -> 4B5B / Manchester /
    chirp+BPSK, sharded over the ranks in contiguous id ranges (every range holds the same class and SNR mix, so the
    shards cost the same); every rank lays its shard out class-major. Strong scaling: the job is fixed."""
    from tmkphy import sharding
    lo, hi = sharding.shard_streams([n_per] * n_total, rank, world)
    ids = np.arange(lo, hi)
    cls = [ids[ids % 3 == k] for k in range(3)]
    n_loc = len(ids)
    d_x = None
    while d_x is None:
        try:
            d_x = eng.device_alloc(n_loc * n_per * 4)
        except tmkphy.TmkError:
            return {"workload": "C5", "skipped": f"cannot allocate {n_loc * n_per * 4 / 1e9:.1f} GB on this GPU"}
    base = [0, len(cls[0]), len(cls[0]) + len(cls[1])]
    planted = 0
    cfgs = [tmkphy.make_config(tmkphy.LINE_4B5B, local_addr=1), tmkphy.make_config(tmkphy.LINE_MANCHESTER, local_addr=1)]
    for k in range(2):
        if len(cls[k]):
            planted += synth_groups(eng, cfgs[k], 0xC500 + 97 * rank + 10 * k, len(cls[k]), n_per,
                                    d_x + base[k] * n_per * 4, SNRS)
    n_ask = len(cls[2])
    if n_ask:
        per_g = max(1, n_ask // len(SNRS))
        g0 = 0
        for g, snr in enumerate(SNRS):
            cnt = per_g if g + 1 < len(SNRS) else n_ask - g0
            cnt = min(cnt, n_ask - g0)
            if cnt <= 0:
                break
            _, nf = eng.ask_synth_batch(0xC5A0 + 13 * rank + g, cnt, n_per, snr, device_ptr=d_x + (base[2] + g0) * n_per * 4)
            planted += nf
            g0 += cnt
    n_phy = len(cls[0]) + len(cls[1])
    st_phy, _ = tmkphy.make_streams([n_per] * n_phy, [cfgs[0]] * len(cls[0]) + [cfgs[1]] * len(cls[1]))
    st_ask, _ = tmkphy.make_streams([n_per] * n_ask, cfgs[0])
    d_ask = d_x + n_phy * n_per * 4
    tot_phy, tot_ask = n_phy * n_per, n_ask * n_per
    dec = Decode(eng, d_x, tot_phy, st_phy, rec_cap=tot_phy // 1024 + 4 * n_phy + 1024, pay_cap=tot_phy // 48 + 65536)
    ask_cap = max(64, tot_ask // 4096 + 2 * n_ask)

    # the chirp records come back to the host: one pinned buffer for all steps (a pageable one allocated per step cost ~15 ms)
    h_arecs = eng.pinned_array((ask_cap,), tmkphy._abi.ASK_REC_DTYPE) if n_ask else None
    n_arecs = C.c_uint64(0)
    stp_ask = st_ask.ctypes.data_as(C.c_void_p)

    def step():
        if n_phy:
            dec.step()
        if n_ask:
            eng._check(eng._lib.tmk_ask_decode_batch(eng._h, C.c_void_p(d_ask), tot_ask, stp_ask, n_ask,
                                                     h_arecs.ctypes.data_as(C.c_void_p), ask_cap, C.byref(n_arecs)))
            return h_arecs[:int(n_arecs.value)]
        return np.zeros(0, tmkphy._abi.ASK_REC_DTYPE)

    for _ in range(warmup):
        arecs = step()
    barrier()
    eng.timer_start()
    phy_ms = ask_ms = 0.0
    for _ in range(steps):
        arecs = step()
    ms = eng.timer_stop()
    barrier()
    step()  # stage split outside the timed region
    phy_ms = sum(eng.last_stage_ms()[k] for k in ("scan", "walk", "order")) if n_phy else 0.0
    am = eng.ask_last_stage_ms() if n_ask else {"scan": 0.0, "walk": 0.0}
    ms_all = eng.allreduce_max([ms])[0]
    rank_ms = eng.allreduce_max([ms, -ms])
    counts = eng.allreduce_sum([int(dec.nr.value), len(arecs), planted, n_loc])
    per_rank = eng.allgather([int(ms / steps * 1000), int(dec.nr.value) + len(arecs)])  # ncclAllGather: step time (us), records
    res = {"workload": f"C5: {n_total} mixed streams x {n_per} samples (stream_id % 3 -> 4B5B / Manchester / chirp+BPSK, "
                       f"AWGN sweep), sharded over {world} rank(s) in contiguous id ranges",
           "scaling": "strong", "msamples_per_s": n_total * n_per * steps / (ms_all * 1e-3) / 1e6,
           "ms_per_step": ms_all / steps, "rank_ms_per_step_max_min": [rank_ms[0] / steps, -rank_ms[1] / steps],
           "frac_hbm": 4.0 * n_total * n_per / world / (ms_all / steps * 1e-3) / 1e9 / hbm,
           "stage_ms_rank0": {"src_phy": round(phy_ms, 3), "chirp_filter": round(am["scan"], 3), "chirp_walk": round(am["walk"], 3)},
           "streams": counts[3], "frames_planted": counts[2], "records_src_phy": counts[0], "records_chirp": counts[1],
           "per_rank_ms": [r[0] / 1000.0 for r in per_rank], "per_rank_records": [r[1] for r in per_rank]}
    dec.free()
    if orc is not None and rank == 0:
        ok = True
        if n_phy:
            idx = np.unique(np.linspace(0, n_phy - 1, 48).astype(np.int64))
            ok, nrec, _ = parity_sample(eng, orc, tmkphy, d_x, st_phy, idx, threads)
        n_chk = 0
        if n_ask:
            for i in np.unique(np.linspace(0, n_ask - 1, 12).astype(np.int64)):
                x = np.zeros(n_per, np.float32)
                eng.d2h(x, d_ask + int(i) * n_per * 4)
                want = orc.ask_demodulate(x, stream=int(i))
                got = arecs[arecs["stream"] == i]
                ok = ok and len(got) == len(want) and all(np.array_equal(got[f], want[f]) for f in
                                                          ("stream", "start_index", "id", "bits", "accepted"))
                n_chk += 1
        res["parity"] = "bit-exact" if ok else "MISMATCH"
        res["parity_sample"] = f"rank 0: {min(48, n_phy)} src/phy captures + {n_chk} chirp captures against the oracle"
    if h_arecs is not None:
        eng.pinned_free(h_arecs.ctypes.data)
    eng.device_free(d_x)
    return res


def run_reference(args):
    """--impl reference: the CPU implementation alone, on rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    from oracle import oracle as orc
    import tmkphy
    orc.build()
    threads = host_threads()
    cfg = tmkphy.make_config(tmkphy.LINE_4B5B, local_addr=1)
    # a bounded sample of the SAME workload: the captures of the GPU arm's generator (tmk_synth_batch, seeds 0x7A4D + group),
    # restated on the host (oracle/synth_host.py: same frames, offsets and bytes; noise equal up to the rounding of the
    # device's fast log / sin — checked in tests/test_gpu_ingest.py), so this arm needs no GPU and none of the product's kernels
    from oracle import synth_host
    n_s = max(256, 32 * threads) // len(SNRS) * len(SNRS)
    n_per = N_PER_STREAM  # whole 10 s streams, 1/8 of the captures of the workload per step on a 16-core host
    per_g = n_s // len(SNRS)
    caps = []
    for g, snr in enumerate(SNRS):  # GPU stream g * 512 + s  <->  call seed 0x7A4D + g, local stream s
        for sidx in range(per_g):
            caps.append(synth_host.capture(orc, cfg, 0x7A4D + g, sidx, n_per, snr, 0.7)[0])
    samples = np.concatenate(caps).astype(np.float32)
    st, total = tmkphy.make_streams([n_per] * n_s, cfg)
    secs = []
    for it in range(args.warmup + args.steps):
        _, _, sec = orc.decode_batch(samples, st, n_threads=threads)
        if it >= args.warmup:
            secs.append(sec)
    t = float(np.sum(secs))
    v = total * args.steps / t / 1e6
    sample = (f"{n_s} of the {N_STREAMS} captures of the workload per step (streams 0..{per_g - 1} of each of the 8 SNR groups, "
              f"the GPU arm's generator restated on the host), {n_per} samples each")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "Msamples/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": t / args.steps * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "streams_per_gpu": N_STREAMS, "samples_per_stream": N_PER_STREAM, "line_code": "4b5b",
                   "chunk": 0, "sample": sample},
        "cpu_baseline": {"value": v, "unit": "Msamples/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="tmkphy")
    ap.add_argument("--streams", type=int, default=N_STREAMS)
    ap.add_argument("--samples", type=int, default=N_PER_STREAM)
    ap.add_argument("--line-code", default="4b5b")
    ap.add_argument("--chunk", type=int, default=0)
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--cpu-max-captures", type=int, default=0, help="bound the cpu_baseline sample (0: by time)")
    ap.add_argument("--no-numa-bind", action="store_true", help="do not bind the rank to the GPU's NUMA node")
    ap.add_argument("--no-ask", action="store_true", help="skip the chirp+BPSK (examples/ask.rs) leg")
    ap.add_argument("--no-links", action="store_true", help="skip the carried-state streaming (links) leg")
    ap.add_argument("--no-pcm16", action="store_true", help="skip the 16-bit PCM end-to-end leg")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the C2-Manchester / chunk / C3 / C4 / C5 legs")
    ap.add_argument("--configs", default="", help="comma-separated subset of man,chunk,c3,mcm,c4,c5")
    ap.add_argument("--mcm-streams", type=int, default=16384)
    ap.add_argument("--c5-streams", type=int, default=131072)
    ap.add_argument("--snrs", default="", help="comma-separated SNRs in dB replacing the C2 sweep (diagnostics)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    if args.snrs:
        global SNRS
        SNRS = [float(v) for v in args.snrs.split(",")]
    if args.impl == "reference":
        return run_reference(args)

    # stdout carries exactly ONE JSON line: anything a library prints there (NCCL's version banner) goes to stderr
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    import tmkphy
    eng = tmkphy.Engine(local_rank)  # raises without a B200: no CPU fallback
    comm_setup(eng, tmkphy, rank, world)
    kind = tmkphy.LINE_4B5B if args.line_code.lower() == "4b5b" else tmkphy.LINE_MANCHESTER
    cfg = tmkphy.make_config(kind, local_addr=1, chunk=args.chunk)
    n_s, n_per = args.streams, args.samples
    headline = (n_s, n_per, kind, args.chunk) == (N_STREAMS, N_PER_STREAM, tmkphy.LINE_4B5B, 0) and not args.snrs
    st, total = tmkphy.make_streams([n_per] * n_s, cfg)
    hbm, which = peaks()
    threads = host_threads()
    orc = None
    if not args.no_cpu:
        from oracle import oracle as orc
        orc.build()

    def barrier():
        eng.synchronize()
        eng.comm_barrier()

    # ---- synthetic captures, generated on the GPU straight into HBM (SNR group g = stream // (n_s/8))
    d_x = eng.device_alloc(total * 4)
    planted = synth_groups(eng, cfg, 0x7A4D + 1000 * rank, n_s, n_per, d_x, SNRS)
    dec = Decode(eng, d_x, total, st)
    sampler = ClockSampler(local_rank)
    sampler.start()
    ms_total, stage, launches = dec.timed(args.steps, args.warmup, barrier)
    sampler.stop_flag = True
    sampler.join()
    stats_recs, stats_pay = int(dec.nr.value), int(dec.npl.value)
    walk_dbg = eng.debug_counters()
    rec_cap, pay_cap = dec.rec_cap, dec.pay_cap
    lib, h, stp, nr, npl = eng._lib, eng._h, dec.stp, dec.nr, dec.npl

    # ---- e2e: pinned host buffers through the same C-ABI call (H2D + kernels + D2H of records/payload)
    e2e_ms = None
    prev_affinity = eng.bind_host_thread() if not args.no_numa_bind else None
    if args.e2e_steps > 0:
        h_x = eng.pinned_array((total,), np.float32)
        eng.d2h(h_x, d_x)
        h_recs = eng.pinned_array((rec_cap,), tmkphy.REC_DTYPE)
        h_pay = eng.pinned_array((pay_cap,), np.uint8)
        xp, rp, pp = h_x.ctypes.data_as(C.c_void_p), h_recs.ctypes.data_as(C.c_void_p), h_pay.ctypes.data_as(C.c_void_p)

        def step_e2e():
            eng._check(lib.tmk_decode_batch(h, xp, total, stp, n_s, rp, rec_cap, C.byref(nr), pp, pay_cap, C.byref(npl)))

        step_e2e()
        barrier()
        eng.timer_start()
        for _ in range(args.e2e_steps):
            step_e2e()
        e2e_ms = eng.timer_stop()
        barrier()
        e2e_d2h = int(nr.value) * 48 + int(npl.value)
        eng.pinned_free(h_x.ctypes.data)
        h_x = None

    # ---- the same captures as 16-bit PCM (what dump_to_wav stores): host int16 -> tmk_decode_batch_pcm16
    pcm_ms, pcm_recs = None, 0
    if args.e2e_steps > 0 and not args.no_pcm16:
        d_p = eng.device_alloc(total * 2)
        eng._check(lib.tmk_pcm16_encode(h, C.c_void_p(d_x), total, C.c_void_p(d_p)))
        h_p = eng.pinned_array((total,), np.int16)
        eng.d2h(h_p, d_p)
        eng.device_free(d_p)
        pcp = h_p.ctypes.data_as(C.c_void_p)

        def step_pcm():
            eng._check(lib.tmk_decode_batch_pcm16(h, pcp, total, stp, n_s, rp, rec_cap, C.byref(nr), pp, pay_cap, C.byref(npl)))

        step_pcm()
        barrier()
        eng.timer_start()
        for _ in range(args.e2e_steps):
            step_pcm()
        pcm_ms = eng.timer_stop()
        barrier()
        pcm_recs = int(nr.value)
        pcm_d2h = int(nr.value) * 48 + int(npl.value)
        eng.pinned_free(h_p.ctypes.data)
    if args.e2e_steps > 0:
        eng.pinned_free(h_recs.ctypes.data)
        eng.pinned_free(h_pay.ctypes.data)
    if prev_affinity is not None:
        os.sched_setaffinity(0, prev_affinity)  # the cpu_baseline leg uses every host core again

    # ---- max over ranks of the device times, sum of the decode counters: the only collective (NCCL, tmk_comm_*)
    names = ["records", "payload_bytes", "frames_planted"] + list(tmkphy.STAT_NAMES)
    loc_stats = eng.stats()
    counters = eng.allreduce_sum([stats_recs, stats_pay, planted] + [loc_stats[k] for k in tmkphy.STAT_NAMES])
    ms_all = eng.allreduce_max([ms_total])[0]
    e2e_all = eng.allreduce_max([e2e_ms])[0] if e2e_ms is not None else None
    pcm_all = eng.allreduce_max([pcm_ms])[0] if pcm_ms is not None else None

    out = None
    if rank == 0:
        step_ms = ms_all / args.steps
        achieved = 4.0 * total / (step_ms * 1e-3) / 1e9
        traffic = None
        kern = {}
        tr = {}
        if headline and os.path.exists(NCU_TRAFFIC_FILE):
            tr = json.load(open(NCU_TRAFFIC_FILE))
        for k in ("scan", "cluster", "walk"):
            kern[k] = {"ms": stage[k], "frac_of_step": stage[k] / max(1e-9, sum(stage.values())),
                       "traffic": tr.get(k)}
        kern["scan"].update({"algorithmic_bytes": 4 * total, "achieved": 4.0 * total / (stage["scan"] * 1e-3) / 1e9,
                             "frac": 4.0 * total / (stage["scan"] * 1e-3) / 1e9 / hbm, "bound": "hbm (in practice: instruction issue)"})
        kern["cluster"].update({"algorithmic_bytes": "frame spans + candidate blocks (data dependent)", "bound": "latency / issue"})
        kern["walk"].update({"algorithmic_bytes": "block summaries + candidate entries (data dependent)", "bound": "latency"})
        if tr and all(tr.get(k) is not None for k in ("scan", "cluster", "walk")):
            traffic = sum(tr[k] for k in ("scan", "cluster", "walk"))
        out = {
            "metric": METRIC, "value": world * total * args.steps / (ms_all * 1e-3) / 1e6, "unit": "Msamples/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": step_ms,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD if headline else f"{n_s} streams x {n_per} samples, line code {args.line_code}, chunk {args.chunk}",
                       "streams_per_gpu": n_s, "samples_per_stream": n_per, "line_code": args.line_code,
                       "chunk": args.chunk, "l2": "inputs (7.9 GB/GPU) far larger than the 126 MB L2",
                       "sharding": f"streams x{world}, no data-path collective; counters reduced by ncclAllReduce (tmk_comm_*)"},
            "gpu_launches": launches,
            "stage_ms_per_step": stage,
            "roofline": {"bound": "hbm", "scope": "pipeline: whole step (scan + cluster + walk + order), 4 B per sample read once",
                         "achieved": achieved, "peak": hbm, "unit": "GB/s", "frac": achieved / hbm,
                         "traffic": traffic, "traffic_source": tr.get("source"), "peak_source": which,
                         "algorithmic_bytes_per_step": 4 * total, "kernels": kern},
            "clocks": sampler.summary(),
            "decode": {**{k: int(v) for k, v in zip(names, counters)}, "walk_last_step": walk_dbg},
        }
        if e2e_all is not None:
            out["e2e"] = {"value": world * total * args.e2e_steps / (e2e_all * 1e-3) / 1e6, "unit": "Msamples/s",
                          "h2d_bytes_per_step": 4 * total + st.nbytes, "d2h_bytes_per_step": e2e_d2h,
                          "steps": args.e2e_steps}
        if pcm_all is not None:
            out["e2e_pcm16"] = {"value": world * total * args.e2e_steps / (pcm_all * 1e-3) / 1e6, "unit": "Msamples/s",
                                "h2d_bytes_per_step": 2 * total + st.nbytes, "d2h_bytes_per_step": pcm_d2h,
                                "records": pcm_recs,
                                "note": "same captures stored as 16-bit PCM (dump.rs:37-40), widened on the GPU"}
        if orc is not None:
            def gather(n_cpu):
                """n_cpu of the captures, spread evenly over the SNR groups, copied back from HBM"""
                idx = np.linspace(0, n_s - 1, n_cpu).astype(np.int64)
                sub = np.empty(n_cpu * n_per, np.float32)
                for k, i in enumerate(idx):
                    eng.d2h(sub[k * n_per:(k + 1) * n_per], d_x + int(i) * n_per * 4)
                return sub

            def cpu_leg(sub, n_cpu):
                st_c, tot_c = tmkphy.make_streams([n_per] * n_cpu, cfg)
                recs, pl, sec = orc.decode_batch(sub[:tot_c], st_c, n_threads=threads)
                return tot_c / sec / 1e6, recs, pl, sec

            # size the sample for about 10 s of host-core work (bounded by the whole workload)
            n_cpu = min(n_s, max(64, 2 * threads))
            sub = gather(n_cpu)
            v, crecs, cpl, sec = cpu_leg(sub, n_cpu)
            want = int(min(n_s, args.cpu_max_captures or n_s, max(n_cpu, 10.0 * v * 1e6 / n_per)))
            if want > n_cpu:
                n_cpu = want
                sub = gather(n_cpu)
                v, crecs, cpl, sec = cpu_leg(sub, n_cpu)
            st_sub, _ = tmkphy.make_streams([n_per] * n_cpu, cfg)
            grecs, gpl = eng.decode_batch(sub, st_sub, rec_cap=len(crecs) + 16, payload_cap=len(cpl) + 4096)
            same = records_equal(grecs, gpl, crecs, cpl)
            corr_ok = bool(len(grecs) == len(crecs) and
                           (np.abs(grecs["corr"] - crecs["corr"]) <= 1e-5 * np.maximum(np.abs(crecs["corr"]), 1.0)).all())
            out["cpu_baseline"] = {"value": v, "unit": "Msamples/s", "cores": threads, "kind": "port",
                                   "sample": f"{n_cpu} of the {n_s} captures x {n_per} samples, {sec:.2f} s",
                                   "parity_vs_gpu": "bit-exact" if same else "MISMATCH", "records": int(len(crecs)),
                                   "corr_within_1e-5": corr_ok}
            del sub

    # ---- the other BASELINE configs (1 GPU: all of them; several GPUs: only the sharded C5 leg)
    want_cfg = set(c for c in args.configs.split(",") if c) or {"man", "chunk", "c3", "mcm", "c4", "c5"}
    run_cfgs = headline and not args.no_configs
    configs = {}
    if run_cfgs and world == 1:
        if "chunk" in want_cfg:
            cfg_c = tmkphy.make_config(kind, local_addr=1, chunk=1200)
            st_c, _ = tmkphy.make_streams([n_per] * n_s, cfg_c)
            configs["C2_4b5b_chunk1200"] = phy_leg(
                eng, orc, tmkphy, "C2 captures (4B5B) fed as 1200-sample process_samples calls (the live receiver poll, csma.rs:446)",
                d_x, st_c, total, 3, 2, hbm, 64, threads)
        if "man" in want_cfg:
            cfg_m = tmkphy.make_config(tmkphy.LINE_MANCHESTER, local_addr=1)
            st_m, _ = tmkphy.make_streams([n_per] * n_s, cfg_m)
            pl_m = synth_groups(eng, cfg_m, 0x7B4D, n_s, n_per, d_x, SNRS)
            configs["C2_manchester"] = phy_leg(eng, orc, tmkphy, "C2 with Manchester frames: 4096 x 480 000 samples, AWGN sweep",
                                               d_x, st_m, total, 5, 3, hbm, 128, threads, {"frames_planted": pl_m})
    if rank == 0 and world == 1 and not args.no_ask and (n_s, n_per) == (N_STREAMS, N_PER_STREAM):
        out["chirp_bpsk"] = ask_leg(eng, tmkphy, orc, d_x, n_s, n_per, hbm)
    dec.free()
    eng.device_free(d_x)
    if rank == 0 and world == 1 and not args.no_links:
        out["links"] = links_leg(eng, orc, tmkphy)
    if run_cfgs and world == 1:
        if "c3" in want_cfg:
            # C3: 16 384 captures x 2^20 samples (68.7 GB resident), 3-tap multipath + AWGN sweep; half 4B5B, half Manchester
            n3, per3 = 16384, 1 << 20
            try:
                d3 = eng.device_alloc(n3 * per3 * 4)
            except tmkphy.TmkError:
                d3 = None
                configs["C3_multipath"] = {"skipped": "cannot allocate 68.7 GB"}
            if d3:
                c_f = tmkphy.make_config(tmkphy.LINE_4B5B, local_addr=1)
                c_m = tmkphy.make_config(tmkphy.LINE_MANCHESTER, local_addr=1)
                pl3 = synth_groups(eng, c_f, 0xC300, n3 // 2, per3, d3, SNRS, multipath=True)
                pl3 += synth_groups(eng, c_m, 0xC380, n3 // 2, per3, d3 + (n3 // 2) * per3 * 4, SNRS, multipath=True)
                st3, tot3 = tmkphy.make_streams([per3] * n3, [c_f] * (n3 // 2) + [c_m] * (n3 // 2))
                configs["C3_multipath"] = phy_leg(
                    eng, orc, tmkphy, "C3: 16 384 captures x 2^20 samples, multipath h = [1, 0.35 @ t1, -0.2 @ t2] + AWGN sweep, "
                    "8192 4B5B + 8192 Manchester (the FSK/OFDM demod of this config is the mcm leg)", d3, st3, tot3, 3, 2,
                    hbm, 96, threads, {"frames_planted": pl3})
                eng.device_free(d3)
        if "mcm" in want_cfg:
            configs["C3_fsk_ofdm"] = mcm_leg(eng, orc, tmkphy, hbm, args.mcm_streams, 1 << 20)
        if "c4" in want_cfg:
            # C4: 8 hour-long captures (172.8 M samples each), one frame per ~0.5 s, +-100 ppm drift, amplitude U[0.3, 1], 10 dB
            n4, per4 = 8, 172_800_000
            d4 = eng.device_alloc(n4 * per4 * 4)
            cfgs4 = [tmkphy.make_config(tmkphy.LINE_4B5B if i % 2 == 0 else tmkphy.LINE_MANCHESTER, local_addr=1) for i in range(n4)]
            pl4 = 0
            for i in range(n4):
                _, nt = eng.synth_batch_ex(cfgs4[i], 0xC400 + i, 1, per4, 10.0, 1.0, drift_ppm=100.0, gap=(12000, 36000),
                                           amp_min=0.3, device_ptr=d4 + i * per4 * 4, want_truth=False)
                pl4 += nt
            st4, tot4 = tmkphy.make_streams([per4] * n4, cfgs4)
            configs["C4_hour_captures"] = phy_leg(
                eng, orc, tmkphy, "C4: 8 captures x 172.8 M samples (1 h @ 48 kHz), one frame per ~0.5 s, +-100 ppm clock drift, "
                "amplitude U[0.3, 1], AWGN 10 dB; each capture walked by ~440 warps at once (segmented walk)", d4, st4, tot4, 3, 2, hbm, 2, threads,
                {"frames_planted": pl4})
            eng.device_free(d4)
    if run_cfgs and "c5" in want_cfg:
        r5 = c5_leg(eng, orc, tmkphy, rank, world, hbm, 3, 2, args.c5_streams, 1 << 18, threads, barrier)
        if rank == 0:
            configs["C5_mixed"] = r5
    if rank == 0:
        if configs:
            out["configs"] = configs
        sys.stdout.flush()
        os.write(real_stdout, (json.dumps(out) + "\n").encode())
    eng.comm_barrier()
    eng.close()


if __name__ == "__main__":
    main()
