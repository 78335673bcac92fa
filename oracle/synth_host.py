"""ORACLE — TEST INFRASTRUCTURE ONLY. Host restatement of the capture generator of the measurement harness
(trackmaker-rs_b200/csrc/synth.cu: counter-hash Gaussian noise, per-stream frame plan, hashed payload bytes), so that
`bench.py --impl reference` — which may load nothing but oracle/ — decodes THE SAME captures as the GPU arm: same seeds,
same frames at the same offsets, same payload bytes; the noise differs from the device's only by the rounding of its fast
logarithm / sine (|difference| < 1e-5, tests/test_gpu_ingest.py::test_host_generator_matches_device). Only the
unimpaired generator (no multipath, no drift, amplitude scale 1) is restated: that is the headline workload.
The frames are modulated by the oracle's encoder (src/phy/encoder.rs restated), not by anything of the product."""
import numpy as np

U64 = np.uint64
_M1, _M2, _GOLD = U64(0xBF58476D1CE4E5B9), U64(0x94D049BB133111EB), U64(0x9E3779B97F4A7C15)


def mix64(z):
    """splitmix64 finaliser on uint64 arrays / scalars (wrapping arithmetic)"""
    with np.errstate(over="ignore"):
        z = (np.asarray(z, dtype=U64) + _GOLD).astype(U64)
        z = ((z ^ (z >> U64(30))) * _M1).astype(U64)
        z = ((z ^ (z >> U64(27))) * _M2).astype(U64)
        return z ^ (z >> U64(31))


def _m(z):
    return int(mix64(U64(z & 0xFFFFFFFFFFFFFFFF)))


def noise(seed, first_sample, n, sigma):
    """samples [first_sample, first_sample + n) of one tmk_synth_batch call (first_sample and n even)"""
    assert first_sample % 2 == 0 and n % 2 == 0
    k = np.arange(first_sample // 2, (first_sample + n) // 2, dtype=U64)
    h = mix64(U64(seed) ^ mix64(k))
    u1 = ((h >> U64(40)).astype(np.float32) + np.float32(1.0)) * np.float32(1.0 / 16777216.0)
    u2 = (h & U64(0xFFFFFF)).astype(np.float32) * np.float32(1.0 / 16777216.0)
    r = np.float32(sigma) * np.sqrt(np.float32(-2.0) * np.log(u1))
    a = np.float32(6.2831853) * u2
    out = np.empty(n, np.float32)
    out[0::2] = r * np.cos(a)
    out[1::2] = r * np.sin(a)
    return out


def plan(kind, seed, stream, n_per_stream, local_addr=1, spl=3, preamble_bytes=2, gap_min=48, gap_span=4752):
    """frame plan of one stream: list of (pos, type, seq, src, dst, len)   (synth.cu: plan_kernel)"""
    sfb = (lambda nb: nb * spl * 2) if kind == 0 else (lambda nb: ((nb + 3) // 4) * 5 * spl)  # LineCode::samples_for_bits
    P = sfb(preamble_bytes * 8)
    st = _m(seed ^ (stream << 20) ^ 0x5151)
    st = _m(st)
    pos = st % 4800
    me = local_addr if local_addr >= 0 else 1
    frames, seq = [], 0
    while True:
        st = _m(st)
        r = st
        ack = (r % 10) == 0
        ln = 0 if ack else 1 + ((r >> 8) % 128)
        foreign = ((r >> 24) % 10) == 0
        dst = (me + 1 + ((r >> 32) % 200)) & 0xFF if foreign else me
        src = (r >> 48) & 0xFF
        flen = P + sfb(8 * (7 + ln))
        if pos + flen > n_per_stream:
            break
        frames.append((pos, 2 if ack else 1, seq & 0xFF, src, dst, ln))
        seq += 1
        st = _m(st)
        pos += flen + gap_min + st % gap_span
    return frames


def payload(seed, stream, pos, ln):
    i = np.arange(ln, dtype=U64)
    h = mix64(U64(seed) ^ U64((stream << 40) & 0xFFFFFFFFFFFFFFFF) ^ U64((pos << 8) & 0xFFFFFFFFFFFFFFFF) ^ U64(0xABCD) ^ (i << U64(52)))
    return ((h >> U64(17)) & U64(0xFF)).astype(np.uint8).tobytes()


def capture(orc, cfg, seed, stream, n_per_stream, snr_db, amplitude=0.7):
    """stream `stream` of tmk_synth_batch(cfg, seed, n_streams > stream, n_per_stream, snr_db, amplitude) -> (samples, frames)"""
    kind = int(cfg.line_code)
    p10 = np.float32(np.power(np.float32(10.0), np.float32(-snr_db) / np.float32(20.0), dtype=np.float32))
    sigma = np.float32(0.0) if snr_db >= 100.0 else np.float32(amplitude) * p10
    x = noise(seed, stream * n_per_stream, n_per_stream, sigma)
    frames = plan(kind, seed, stream, n_per_stream, int(cfg.local_addr), int(cfg.spl), int(cfg.preamble_bytes))
    amp = np.float32(amplitude)
    for pos, typ, seq, src, dst, ln in frames:
        w = orc.modulate(cfg, [(typ, seq, src, dst, payload(seed, stream, pos, ln))], 0)
        x[pos:pos + len(w)] += amp * w
    return x, frames
