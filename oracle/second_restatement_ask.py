"""ORACLE — TEST INFRASTRUCTURE ONLY. A SECOND, independent restatement of the reference's chirp + BPSK modem
(examples/ask.rs:526-796, `test_sender_receiver`: tables :574-605, transmitter :607-637, receiver loop :668-796), written
from the Rust, not from oracle/ask_modem.hpp: numpy float32, no shared code with the C++ oracle.
tests/test_oracle_cross.py runs both on the same random captures and compares every record (lock position, the bits of
the locking sync power, id, verdict, all 108 decided bits). Like oracle/second_restatement.py it does NOT pin parity
against the reference itself (no Rust toolchain here or on the GPU box); it removes single-reader error for rows
a14-a16 of SURVEY §8.

f32 semantics: every arithmetic result is rounded to float32 once (Rust does not contract a*b+c); `iter().sum::<f32>()`
folds from 0.0 left to right, which `np.add.accumulate` over a float32 row reproduces (0.0 + x = x); `f32::sin` is taken
as the correctly rounded sine (math.sin of the exact f32 argument, rounded once) — glibc's sinf agrees with that except
for rare last-bit cases, which the cross test counts instead of hiding."""
import math

import numpy as np

F = np.float32
PRE_LEN, SPB, N_BITS = 440, 44, 108          # ask.rs:583-592 (220 + 220), :615 (44 samples per bit), :609-610 (100 + 8)
FRAME_SAMPLES = SPB * N_BITS                 # ask.rs:725
PI = F(math.pi)                              # std::f32::consts::PI


def _sin32(x):
    return F(math.sin(float(x)))


def tables(n_carrier=FRAME_SAMPLES):
    """(preamble[440], carrier[n_carrier]) as ask.rs:574-605 builds them."""
    fs = F(48000.0)
    t = [F(i) / fs for i in range(max(n_carrier, PRE_LEN))]
    two_pi_fc = F(2.0) * PI * F(10000.0)                                        # :580 — evaluated left to right
    carrier = np.array([_sin32(two_pi_fc * t[i]) for i in range(n_carrier)], F)
    f_p = [F(2000.0) + F(8000.0) * F(i) / F(219.0) for i in range(220)]          # :586-588
    f_p += [F(10000.0) - F(8000.0) * F(i) / F(219.0) for i in range(220)]        # :590-592
    pre = [_sin32(F(2.0) * PI * f_p[0] * t[0])]                                   # :597
    omega = F(0.0)
    for i in range(1, PRE_LEN):                                                   # :599-603
        dt = t[i] - t[i - 1]
        omega = omega + PI * (f_p[i] + f_p[i - 1]) * dt
        pre.append(_sin32(omega))
    return np.array(pre, F), carrier


def modulate(bits, gap1, gap2, pre, carrier):
    """ask.rs:607-637: per frame  zeros(gap1) | preamble | 108 bits x 44 samples of +-carrier | zeros(gap2); the 8 bits
    behind the 100 data bits are the zero "CRC" placeholder (:609-610); the carrier index restarts with every frame."""
    out = []
    for fr, g1, g2 in zip(bits, gap1, gap2):
        b = list(fr) + [0] * 8
        amp = np.repeat(np.where(np.array(b) == 1, F(1.0), F(-1.0)).astype(F), SPB)
        out += [np.zeros(int(g1), F), pre, carrier[:FRAME_SAMPLES] * amp, np.zeros(int(g2), F)]
    return np.concatenate(out) if out else np.zeros(0, F)


def _window_sums(x, lo, hi, fifo_from, pre):
    """sync_power numerators for i in [lo, hi): the fifo holds x[i-439 .. i], with zeros where the index is below
    `fifo_from` (the fifo was cleared when the receiver last left the decode state, :719) or below 0."""
    idx = np.arange(lo, hi)[:, None] + np.arange(-PRE_LEN + 1, 1)[None, :]
    w = np.where(idx >= max(fifo_from, 0), x[np.clip(idx, 0, len(x) - 1)], F(0.0)).astype(F)
    prod = w * pre[None, :]
    return np.add.accumulate(prod, axis=1, dtype=F)[:, -1]


def _decode(fifo, carrier):
    """ask.rs:726-768: mix with the frame-local carrier, 11-tap moving average clipped at the ends, integrate samples
    10..30 of every bit, decide by sign."""
    L = len(fifo)
    prod = fifo * carrier[:L]
    acc = np.zeros(L, F)
    j = np.arange(L)
    for o in range(-5, 6):                                   # k runs start..end upwards: the same order per j
        k = j + o
        ok = (k >= 0) & (k < L)
        acc = np.where(ok, acc + prod[np.clip(k, 0, L - 1)], acc).astype(F)
    width = (np.minimum(j + 6, L) - np.maximum(j - 5, 0)).astype(F)
    rc = (acc / width).astype(F)
    bits = np.zeros(N_BITS, bool)
    for b in range(N_BITS):
        s = F(0.0)
        for v in rc[10 + b * SPB:30 + b * SPB]:
            s = s + v
        bits[b] = s > F(0.0)
    return bits


def demodulate(x, pre, carrier, block=2048):
    """The receiver loop of ask.rs:668-796, literally: one EMA power chain over every sample, sync state with the
    440-sample fifo, lock 201 samples after the last rising peak, the sample at start_index + 201 dropped (:713-722 sets
    decode_fifo = rx[start_index+1 .. i) and the sample i itself is never pushed), 4752-sample decode.
    Returns a list of (start_index, sync_power_at_lock, id, accepted, bits[108])."""
    x = np.ascontiguousarray(x, F)
    n = len(x)
    c_keep = F(1.0) - F(1.0) / F(64.0)
    power = F(0.0)
    start_index, local_max, lock_power = 0, F(0.0), F(0.0)
    state, fifo_from = 0, 0
    decode_fifo = []
    recs = []
    sums, sums_lo = None, 0
    i = 0
    while i < n:
        cs = x[i]
        power = power * c_keep + cs * cs / F(64.0)                                 # :690-691
        if state == 0:
            if sums is None or not (sums_lo <= i < sums_lo + len(sums)):
                sums_lo = i
                sums = _window_sums(x, i, min(n, i + block), fifo_from, pre)
            sync_power = sums[i - sums_lo] / F(200.0)                               # :698-703
            if sync_power > power * F(2.0) and sync_power > local_max and sync_power > F(0.05):  # :705-710
                local_max = sync_power
                start_index = i
                lock_power = sync_power
            elif i > start_index + 200 and start_index != 0:                       # :711-721
                local_max = F(0.0)
                fifo_from = n + PRE_LEN      # cleared, and not written again before the state returns to 0
                state = 1
                decode_fifo = list(x[start_index + 1:i])
                sums = None
        else:
            decode_fifo.append(cs)                                                  # :723
            if len(decode_fifo) == FRAME_SAMPLES:
                bits = _decode(np.array(decode_fifo, F), carrier)
                ident = 0
                for k in range(8):                                                   # :763-768
                    if bits[k]:
                        ident += 1 << (7 - k)
                recs.append((start_index, lock_power, ident, 1 <= ident <= 100, bits))
                start_index = 0                                                      # :790-792
                decode_fifo = []
                state = 0
                fifo_from = i + 1            # the fifo is all zeros and takes its next sample at i + 1
        i += 1
    return recs
