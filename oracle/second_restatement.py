"""ORACLE — TEST INFRASTRUCTURE ONLY. A SECOND, independent restatement of the reference's streaming decoder, written
from the Rust (src/phy/decoder.rs:36-417, line_coding.rs:6-333, frame.rs:109-183, crc.rs:7-75), not from the C++
oracle: numpy float32 scalars / arrays, no shared code with oracle/*.hpp. tests/test_oracle_cross.py decodes the same
random captures with both and compares every verdict, offset and payload byte: a slip in either restatement (operation
order, consume rule, drain arithmetic) shows up as a difference. It does NOT pin parity against the reference itself —
no Rust toolchain exists here or on the GPU box — it removes single-reader error.

f32 semantics: every arithmetic result is a numpy float32 (each operation rounded once, no fusion), sums run in the
reference's order (`iter().sum::<f32>()` folds from 0.0 left to right; the AVX dot product keeps 8 lane accumulators,
adds the high half to the low half, sums the 4 lanes left to right, then the tail)."""
import numpy as np

F = np.float32
HEADER_BYTES = 7                      # frame.rs / consts.rs: [len_hi, len_lo, crc, type, seq, src, dst]
FIVE_B = [0b11110, 0b01001, 0b10100, 0b10101, 0b01010, 0b01011, 0b01110, 0b01111,
          0b10010, 0b10011, 0b10110, 0b10111, 0b11010, 0b11011, 0b11100, 0b11101]  # line_coding.rs:138-155
FIVE_B_INV = {s: n for n, s in enumerate(FIVE_B)}
OK, HDR_PARSE, BAD_LEN, SHORT_DECODE, NOT_FOR_US, CRC_FAIL = 0, 1, 2, 3, 4, 5


def crc8(data):                       # crc.rs:7-22
    crc = 0
    for b in data:
        crc ^= b
        for _ in range(8):
            crc = ((crc << 1) ^ 0x07) & 0xFF if crc & 0x80 else (crc << 1) & 0xFF
    return crc


def bits_to_bytes(bits):              # crc.rs:62-75 (MSB first, the last partial byte is padded with zeros)
    out = []
    for i in range(0, len(bits), 8):
        v = 0
        for j, b in enumerate(bits[i:i + 8]):
            if b:
                v |= 1 << (7 - j)
        out.append(v)
    return out


def seq_sum(v):
    """`iter().sum::<f32>()`"""
    s = F(0.0)
    for x in v:
        s = F(s + x)
    return s


class LineCode:
    def __init__(self, kind, spl):
        self.kind, self.spl = kind, spl   # kind 0 Manchester, 1 4B5B + NRZI

    def samples_for_bits(self, n):    # line_coding.rs:125-127 / 295-300
        return n * self.spl * 2 if self.kind == 0 else ((n + 3) // 4) * 5 * self.spl

    def encode(self, bits):           # line_coding.rs:74-91 / 202-240
        out = []
        if self.kind == 0:
            for b in bits:
                out += ([1.0] * self.spl + [-1.0] * self.spl) if b == 0 else ([-1.0] * self.spl + [1.0] * self.spl)
        else:
            enc = []
            for i in range(0, len(bits), 4):
                nib = 0
                for j, b in enumerate(bits[i:i + 4]):
                    if b:
                        nib |= 1 << (3 - j)
                enc += [(FIVE_B[nib] >> (4 - j)) & 1 for j in range(5)]
            level = 1.0
            for b in enc:
                if b:
                    level = -level
                out += [level] * self.spl
        return np.array(out, F)

    def preamble(self, nbytes):       # line_coding.rs:13-21
        return self.encode([0, 0, 1, 1, 0, 0, 1, 1] * (nbytes - 1) + [0, 1, 0, 1, 1, 0, 1, 0])

    def decode(self, x):              # line_coding.rs:93-123 / 243-293
        bits, spl = [], self.spl
        if self.kind == 0:
            for i in range(len(x) // (2 * spl)):
                a = F(seq_sum(x[2 * spl * i:2 * spl * i + spl]) / F(spl))
                b = F(seq_sum(x[2 * spl * i + spl:2 * spl * (i + 1)]) / F(spl))
                bits.append(0 if a > b else 1)
            return bits
        if len(x) == 0:
            return bits
        last, sym, cnt = F(1.0), 0, 0
        for i in range(len(x) // spl):
            cur = F(seq_sum(x[spl * i:spl * (i + 1)]) / F(spl))
            bit = 1 if F(last * cur) < 0 else 0
            if abs(cur) > F(1e-6):
                last = cur
            sym, cnt = (sym << 1) | bit, cnt + 1
            if cnt == 5:
                if sym not in FIVE_B_INV:
                    break
                bits += [(FIVE_B_INV[sym] >> (3 - j)) & 1 for j in range(4)]
                sym, cnt = 0, 0
        return bits


class PhyDecoderPy:
    """decoder.rs:16-356. `events` collects one tuple per detection: (status, preamble_start, frame_start, consumed,
    header bytes or None, payload bytes), positions counted from the first sample ever fed."""

    def __init__(self, kind, spl=3, preamble_bytes=2, local_addr=1, threshold=0.9, max_frame_bytes=256):
        self.lc = LineCode(kind, spl)
        self.pre = self.lc.preamble(preamble_bytes)
        self.pre_energy = F(np.sqrt(seq_sum(self.pre * self.pre)))      # decoder.rs:47-51
        self.thr, self.max_frame_bytes, self.local_addr = F(threshold), max_frame_bytes, local_addr
        self.buf = np.zeros(0, F)
        self.off, self.state, self.drained = 0, None, 0                 # state: None = Searching, int = Decoding(start)
        self.events, self.frames = [], []

    # compute_dot_product_avx for every window of `area` at once (decoder.rs:381-417)
    def _dots(self, area):
        P = len(self.pre)
        nw = len(area) - P + 1
        lanes = [np.zeros(nw, F) for _ in range(8)]
        i = 0
        while i + 8 <= P:
            for l in range(8):
                lanes[l] = lanes[l] + area[i + l:i + l + nw] * self.pre[i + l]
            i += 8
        q = [lanes[l] + lanes[l + 4] for l in range(4)]
        s = np.zeros(nw, F)
        for l in range(4):
            s = s + q[l]
        while i < P:
            s = s + area[i:i + nw] * self.pre[i]
            i += 1
        return s

    def _search(self):                # decoder.rs:123-238
        area = self.buf[self.off:]
        P = len(self.pre)
        if len(area) < P:
            return None
        nw = len(area) - P + 1
        dots = self._dots(area)
        sq = area * area
        e = seq_sum(sq[:P])
        for i in range(nw):
            corr = F(0.0) if e < F(1e-6) else F(dots[i] / F(np.sqrt(e) * self.pre_energy))
            if corr >= self.thr:
                sync_len = self.lc.samples_for_bits(8)
                sp = self.pre[P - sync_len:]
                sync_energy = F(np.sqrt(seq_sum(sp * sp)))
                margin = self.lc.samples_for_bits(1)
                expected = i + P - sync_len
                lo, hi = max(0, expected - margin), min(expected + margin, len(area) - sync_len)
                best, best_off = F(-1.0), expected
                for j in range(lo, hi + 1):
                    w = area[j:j + sync_len]
                    dot, we = F(0.0), F(0.0)
                    for a, b in zip(w, sp):
                        dot = F(dot + F(a * b))
                        we = F(we + F(a * a))
                    c = F(dot / F(np.sqrt(we) * sync_energy)) if (we > F(1e-6) and sync_energy > F(1e-6)) else F(0.0)
                    if c > best:
                        best, best_off = c, j
                self.state = self.off + best_off + sync_len
                self._det = self.drained + self.off + i  # where the preamble was detected, counted from the first sample fed
                return i
            if i + 1 < nw:
                e = F(F(e - sq[i]) + sq[i + P])
                if e < 0:
                    e = F(0.0)
        return nw

    def _decode(self, fs):            # decoder.rs:242-356
        P = len(self.pre)
        hs = self.lc.samples_for_bits(8 * HEADER_BYTES)
        if len(self.buf) < fs + hs:
            return None
        hb = bits_to_bytes(self.lc.decode(self.buf[fs:fs + hs]))
        ev = lambda st, consumed, hdr=None, data=b"": self.events.append(
            (st, self._det, self.drained + fs, consumed, hdr, bytes(data)))
        if len(hb) < HEADER_BYTES or hb[3] not in (1, 2):
            self.state = None
            ev(HDR_PARSE, hs, hb[:HEADER_BYTES] if len(hb) >= HEADER_BYTES else None)
            return hs
        ln = (hb[0] << 8) | hb[1]
        if (hb[3] == 1 and ln == 0) or ln > self.max_frame_bytes:
            self.state = None
            ev(BAD_LEN, 1, hb[:HEADER_BYTES])
            return 1
        total_bits = (HEADER_BYTES + ln) * 8
        total = self.lc.samples_for_bits(total_bits)
        if len(self.buf) < fs + total:
            return None
        bits = self.lc.decode(self.buf[fs:fs + total])
        consumed = P + self.lc.samples_for_bits(len(bits))
        self.state = None
        hdr = hb[:HEADER_BYTES]
        if len(bits) < total_bits:
            ev(SHORT_DECODE, consumed, hdr)
            return consumed
        by = bits_to_bytes(bits)
        data = by[HEADER_BYTES:HEADER_BYTES + ln]
        if hb[6] != self.local_addr:
            ev(NOT_FOR_US, consumed, hdr, data)
            return consumed
        if crc8(data) != hb[2]:       # frame.rs:155-159
            ev(CRC_FAIL, consumed, hdr, data)
            return consumed
        ev(OK, consumed, hdr, data)
        self.frames.append((hb[3], hb[4], hb[5], hb[6], bytes(data)))
        return consumed

    def process_samples(self, x):     # decoder.rs:68-112
        n0 = len(self.frames)
        self.buf = np.concatenate([self.buf, np.asarray(x, F)])
        while True:
            r = self._search() if self.state is None else self._decode(self.state)
            if r is None:
                break
            self.off += r
        if self.off > 0:
            drain = max(0, self.off - max(0, len(self.pre) - 1))
            if drain > 0:
                self.buf = self.buf[drain:]
                self.off -= drain
                self.drained += drain
                if self.state is not None:
                    self.state = max(0, self.state - drain)
        return self.frames[n0:]


def frame_bits(ftype, seq, src, dst, data):
    """frame.rs:74-107 (`to_bytes` + `to_bits`): [len_hi, len_lo, crc8(data), type, seq, src, dst, data...], MSB first."""
    by = [(len(data) >> 8) & 255, len(data) & 255, crc8(data), ftype, seq, src, dst] + list(data)
    return [(b >> (7 - j)) & 1 for b in by for j in range(8)]


def encode_frames(kind, frames, gap, spl=3, preamble_bytes=2):
    """encoder.rs:41-85: per frame the preamble followed by the line-coded frame bits; `gap` zeros BETWEEN frames only."""
    lc = LineCode(kind, spl)
    pre = lc.preamble(preamble_bytes)
    out = []
    for i, f in enumerate(frames):
        out += [pre, lc.encode(frame_bits(*f))]
        if i + 1 < len(frames):
            out.append(np.zeros(gap, F))
    return np.concatenate(out) if out else np.zeros(0, F)
