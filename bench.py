#!/usr/bin/env python
"""bench.py — Msamples/s of the fused preamble-sync + demod path (BASELINE.json metric).

A "step" is one pass of the hot path (tmk_decode_batch: scan -> walk -> record ordering) over one
batch of synthetic captures: BASELINE config 2, 4096 streams x 10 s x 48 kHz (480 000 samples each,
7.86 GB of fp32 per GPU), src/phy frames at random offsets, AWGN SNR sweep {-3..30 dB}.
Streams shard across GPUs with no data-path collective (weak scaling: every rank owns its own 4096
streams); NCCL only all-reduces the decode counters after the timed region.

  value        whole-job Msamples/s, captures already resident in HBM, CUDA events on the engine's stream
  e2e          the same through the C ABI with HOST (pinned) sample buffers: H2D + kernels + D2H of records
  roofline     the scan kernel: 4 B/sample algorithmic bytes / its CUDA-event duration vs measured HBM peak
  cpu_baseline the oracle (C++ restatement of the reference's src/phy, AVX dot) on the box's host cores,
               on a bounded sample of the same captures, records compared bit for bit with the GPU's

--impl reference times that CPU implementation alone (the reference is Rust; no rustc in this image, so
oracle/_ref cannot exist: kind = "port").
"""
import argparse
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
NCU_SCAN_TRAFFIC = 7.890631e9 + 79.098112e6  # bytes per scan launch at C2, ncu --set full (profiles/r01_ncu_full_summary_final2.txt)
N_PER_STREAM = 480000
SNRS = [-3.0, 0.0, 3.0, 6.0, 10.0, 15.0, 20.0, 30.0]
METRIC = "Msamples/s sync+demod (48 kHz streams)"
WORKLOAD = "C2: 4096 src/phy 4B5B streams x 10 s @ 48 kHz per GPU, AWGN SNR sweep -3..30 dB"


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


def make_streams(tmkphy, cfg, n_streams, n_per):
    st, total = tmkphy.make_streams([n_per] * n_streams, cfg)
    return st, total


def cpu_leg(orc, tmkphy, host_samples, n_streams, n_per, cfg, threads):
    """oracle over n_streams captures with all host threads -> (Msamples/s, records, payload)"""
    st, total = make_streams(tmkphy, cfg, n_streams, n_per)
    recs, pl, sec = orc.decode_batch(host_samples[:total], st, n_threads=threads)
    return total / sec / 1e6, recs, pl, sec


def ask_leg(eng, tmkphy, d_x, n_s, n_per, check=True):
    """BASELINE config 2 read literally ("PSK streams"): the chirp + BPSK modem of examples/ask.rs over the same
    n_s x n_per samples, resident in HBM (frames from the engine's own TX kernel, AWGN sweep on the host, 8 distinct
    captures tiled over the streams). Extra evidence beside the headline src/phy workload."""
    rng = np.random.default_rng(3)
    base = []
    for k, snr in enumerate(SNRS[:8]):
        fr = rng.integers(0, 2, (92, 100)).astype(np.uint8)
        fr[:, :8] = [[((i % 100 + 1) >> (7 - j)) & 1 for j in range(8)] for i in range(92)]
        x = eng.ask_modulate(fr, rng.integers(0, 100, 92), rng.integers(0, 100, 92))[:n_per]
        x = np.pad(x, (0, n_per - len(x)))
        base.append((x + rng.normal(0, np.sqrt(0.5) * 10 ** (-snr / 20), n_per)).astype(np.float32))
    nb = len(base)
    for i in range(n_s):
        eng.h2d(d_x + i * n_per * 4, base[i % nb])
    st, total = tmkphy.make_streams([n_per] * n_s, tmkphy.make_config(1))
    for _ in range(2):
        recs = eng.ask_decode_batch(d_x, st, total_samples=total)
    ms = []
    for _ in range(3):
        eng.ask_decode_batch(d_x, st, total_samples=total)
        ms.append(eng.ask_last_stage_ms())
    scan = float(np.median([m["scan"] for m in ms]))
    walk = float(np.median([m["walk"] for m in ms]))
    res = {"workload": f"{n_s} chirp+BPSK streams x {n_per} samples (examples/ask.rs frames, AWGN sweep)",
           "value": total / (scan + walk) / 1e3, "unit": "Msamples/s", "filter_ms": scan, "walk_ms": walk,
           "frames": int(len(recs)), "accepted": int(recs["accepted"].sum())}
    if check:
        from oracle import oracle as orc
        import time
        t = time.time()
        want = np.concatenate([orc.ask_demodulate(base[i], stream=i) for i in range(nb)])
        sec = time.time() - t
        got = recs[recs["stream"] < nb]
        same = len(got) == len(want) and all(np.array_equal(got[f], want[f]) for f in
                                              ("stream", "start_index", "id", "bits", "accepted"))
        res["cpu_oracle_1core_msamples_per_s"] = nb * n_per / sec / 1e6
        res["parity_vs_gpu"] = "bit-exact" if same else "MISMATCH"
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
    # a bounded sample of the same workload: captures come from the oracle's own modulator here,
    # so this arm needs no GPU and none of the product's kernels
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import helpers
    rng = np.random.default_rng(7)
    n_s = max(256, 32 * threads)
    n_per = N_PER_STREAM  # whole 10 s streams, 1/8 of the captures of the workload per step on a 16-core host
    caps = []
    base = [helpers.make_capture(orc, cfg, np.random.default_rng(100 + i), n_frames=150, snr_db=SNRS[i % len(SNRS)],
                                 gap_range=(48, 4800)) for i in range(16)]
    for i in range(n_s):
        b = base[i % 16]
        caps.append(np.resize(b, n_per) if len(b) < n_per else b[:n_per])
    samples = np.concatenate(caps).astype(np.float32)
    st, total = make_streams(tmkphy, cfg, n_s, n_per)
    secs = []
    for it in range(args.warmup + args.steps):
        _, _, sec = orc.decode_batch(samples, st, n_threads=threads)
        if it >= args.warmup:
            secs.append(sec)
    t = float(np.sum(secs))
    v = total * args.steps / t / 1e6
    sample = f"{n_s} streams x {n_per} samples per step (oracle-modulated 4B5B frames, AWGN sweep)"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "Msamples/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": t / args.steps * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "sample": sample},
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
    ap.add_argument("--no-pcm16", action="store_true", help="skip the 16-bit PCM end-to-end leg")
    ap.add_argument("--no-cpu", action="store_true")
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
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    import tmkphy
    eng = tmkphy.Engine(local_rank)  # raises without a B200: no CPU fallback
    kind = tmkphy.LINE_4B5B if args.line_code.lower() == "4b5b" else tmkphy.LINE_MANCHESTER
    cfg = tmkphy.make_config(kind, local_addr=1, chunk=args.chunk)
    n_s, n_per = args.streams, args.samples
    total = n_s * n_per
    st, _ = make_streams(tmkphy, cfg, n_s, n_per)

    # ---- synthetic captures, generated on the GPU straight into HBM (SNR group g = stream // (n_s/8))
    d_x = eng.device_alloc(total * 4)
    per_g = n_s // len(SNRS)
    planted = 0
    for g, snr in enumerate(SNRS):
        cnt = per_g if g + 1 < len(SNRS) else n_s - per_g * (len(SNRS) - 1)
        if cnt <= 0:
            continue
        _, nt = eng.synth_batch(cfg, 0x7A4D + 1000 * rank + g, cnt, n_per, snr, 0.7,
                                device_ptr=d_x + g * per_g * n_per * 4, want_truth=False)
        planted += nt
    rec_cap = total // 256 + 4 * n_s
    pay_cap = total // 24 + 65536
    d_recs = eng.device_alloc(rec_cap * 48)
    d_pay = eng.device_alloc(pay_cap)

    import ctypes as C
    lib, h = eng._lib, eng._h
    nr, npl = C.c_uint64(0), C.c_uint64(0)
    stp = st.ctypes.data_as(C.c_void_p)

    def step_resident():
        rc = lib.tmk_decode_batch(h, C.c_void_p(d_x), total, stp, n_s, C.c_void_p(d_recs), rec_cap, C.byref(nr),
                                  C.c_void_p(d_pay), pay_cap, C.byref(npl))
        eng._check(rc)

    def barrier():
        eng.synchronize()
        if dist is not None:
            dist.barrier()
        eng.synchronize()

    for _ in range(args.warmup):
        step_resident()
    sampler = ClockSampler(local_rank)
    sampler.start()
    barrier()
    l0 = eng.launch_count()
    stage = {"scan": 0.0, "walk": 0.0, "order": 0.0}
    cluster_ms = 0.0
    eng.timer_start()
    for _ in range(args.steps):
        step_resident()
        s = eng.last_stage_ms()
        for k in stage:
            stage[k] += s[k]
        cluster_ms += eng.debug_counters()["cluster_us"] * 1e-3  # part of "walk": the per-block cluster pass
    ms_total = eng.timer_stop()
    launches = eng.launch_count() - l0
    barrier()
    sampler.stop_flag = True
    sampler.join()
    stats_recs, stats_pay = int(nr.value), int(npl.value)

    # ---- e2e: pinned host buffers through the same C-ABI call (H2D + kernels + D2H of records/payload)
    e2e_ms, h_x = None, None
    # pinned host buffers live on the GPU's own NUMA node (matters from 4 ranks up: two sockets feed eight PCIe links)
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
        e2e_recs = int(nr.value)
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
        eng.pinned_free(h_p.ctypes.data)

    if prev_affinity is not None:
        os.sched_setaffinity(0, prev_affinity)  # the cpu_baseline leg uses every host core again

    # ---- max over ranks of the device times, sum of the decode counters (the only collective)
    from tmkphy import sharding
    names = ["records", "payload_bytes", "frames_planted"] + list(tmkphy.STAT_NAMES)
    local = dict(zip(names, [stats_recs, stats_pay, planted] + [eng.stats()[k] for k in tmkphy.STAT_NAMES]))
    dev = None
    if dist is not None:
        import torch
        dev = torch.device("cuda", local_rank)
    summed, ms_all = sharding.reduce_stats(local, ms_total, device=dev)
    e2e_all = e2e_ms
    if e2e_ms is not None:
        _, e2e_all = sharding.reduce_stats({}, e2e_ms, device=dev)
    counters = np.array([summed[k] for k in names], dtype=np.int64)

    if rank == 0:
        hbm, which = peaks()
        scan_ms = stage["scan"] / args.steps
        achieved = 4.0 * total / (scan_ms * 1e-3) / 1e9
        out = {
            "metric": METRIC, "value": world * total * args.steps / (ms_all * 1e-3) / 1e6, "unit": "Msamples/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_all / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD if (n_s, n_per, kind) == (N_STREAMS, N_PER_STREAM, tmkphy.LINE_4B5B) and not args.snrs
                       else f"{n_s} streams x {n_per} samples, line code {args.line_code}",
                       "streams_per_gpu": n_s, "samples_per_stream": n_per, "line_code": args.line_code,
                       "chunk": args.chunk, "l2": "inputs (7.9 GB/GPU) far larger than the 126 MB L2",
                       "sharding": f"streams x{world}, no data-path collective"},
            "gpu_launches": launches,
            "stage_ms_per_step": {"scan": stage["scan"] / args.steps, "cluster": cluster_ms / args.steps,
                                  "walk": (stage["walk"] - cluster_ms) / args.steps, "order": stage["order"] / args.steps},
            "roofline": {"bound": "hbm", "kernel": "preamble scan", "achieved": achieved, "peak": hbm, "unit": "GB/s",
                         "frac": achieved / hbm,
                         # dram__bytes_read.sum + dram__bytes_write.sum of one launch at this workload
                         # (profiles/r01_ncu_full_summary_final.txt); null for any other workload
                         "traffic": NCU_SCAN_TRAFFIC if (n_s, n_per, kind) == (N_STREAMS, N_PER_STREAM, tmkphy.LINE_4B5B)
                         and not args.snrs else None,
                         "peak_source": which,
                         "algorithmic_bytes_per_launch": 4 * total},
            "clocks": sampler.summary(),
            "decode": {"records": int(counters[0]), "payload_bytes": int(counters[1]), "frames_planted": int(counters[2]),
                       **{k: int(v) for k, v in zip(tmkphy.STAT_NAMES, counters[3:])},
                       "walk_last_step": eng.debug_counters()},
        }
        if e2e_all is not None:
            out["e2e"] = {"value": world * total * args.e2e_steps / (e2e_all * 1e-3) / 1e6, "unit": "Msamples/s",
                          "h2d_bytes_per_step": 4 * total + st.nbytes, "d2h_bytes_per_step": e2e_d2h,
                          "steps": args.e2e_steps}
        if pcm_ms is not None:
            # rank 0's own time (every rank runs the same step on its own PCIe link)
            out["e2e_pcm16"] = {"value": world * total * args.e2e_steps / (pcm_ms * 1e-3) / 1e6, "unit": "Msamples/s",
                                "h2d_bytes_per_step": 2 * total + st.nbytes, "records": pcm_recs,
                                "note": "same captures stored as 16-bit PCM (dump.rs:37-40), widened on the GPU"}
        if not args.no_cpu:
            from oracle import oracle as orc
            orc.build()
            threads = host_threads()

            def gather(n_cpu):
                """n_cpu of the captures, spread evenly over the SNR groups, copied back from HBM"""
                idx = np.linspace(0, n_s - 1, n_cpu).astype(np.int64)
                sub = np.empty(n_cpu * n_per, np.float32)
                for k, i in enumerate(idx):
                    eng.d2h(sub[k * n_per:(k + 1) * n_per], d_x + int(i) * n_per * 4)
                return sub

            # size the sample for about 10 s of host-core work (bounded by the whole workload)
            n_cpu = min(n_s, max(64, 2 * threads))
            sub = gather(n_cpu)
            v, crecs, cpl, sec = cpu_leg(orc, tmkphy, sub, n_cpu, n_per, cfg, threads)
            want = int(min(n_s, args.cpu_max_captures or n_s, max(n_cpu, 10.0 * v * 1e6 / n_per)))
            if want > n_cpu:
                n_cpu = want
                sub = gather(n_cpu)
                v, crecs, cpl, sec = cpu_leg(orc, tmkphy, sub, n_cpu, n_per, cfg, threads)
            st_sub, _ = make_streams(tmkphy, cfg, n_cpu, n_per)
            grecs, gpl = eng.decode_batch(sub, st_sub, rec_cap=len(crecs) + 16, payload_cap=len(cpl) + 16)
            fields = ("stream", "status", "preamble_start", "frame_start", "consumed", "len", "crc", "type", "seq",
                      "src", "dst", "hdr_valid")

            def payload_in_record_order(recs, pl):
                keep = np.isin(recs["status"], (0, 4, 5))
                return b"".join(pl[int(o):int(o) + int(n)].tobytes() for o, n in
                                zip(recs["payload_off"][keep], recs["len"][keep]))

            same = (len(grecs) == len(crecs) and all(np.array_equal(grecs[f], crecs[f]) for f in fields)
                    and payload_in_record_order(grecs, gpl) == payload_in_record_order(crecs, cpl))
            out["cpu_baseline"] = {"value": v, "unit": "Msamples/s", "cores": threads, "kind": "port",
                                   "sample": f"{n_cpu} of the {n_s} captures x {n_per} samples, {sec:.2f} s",
                                   "parity_vs_gpu": "bit-exact" if same else "MISMATCH", "records": int(len(crecs))}
        if world == 1 and not args.no_ask and (n_s, n_per) == (N_STREAMS, N_PER_STREAM):
            out["chirp_bpsk"] = ask_leg(eng, tmkphy, d_x, n_s, n_per, check=not args.no_cpu)
        sys.stdout.flush()
        os.write(real_stdout, (json.dumps(out) + "\n").encode())
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    eng.close()


if __name__ == "__main__":
    main()
