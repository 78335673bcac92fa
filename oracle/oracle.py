"""ORACLE — TEST INFRASTRUCTURE ONLY.

ctypes loader for oracle/liboracle.so, the CPU restatement of the reference's src/phy
(see phy_core.hpp / phy_decoder.hpp for the file:line citations). Imported only by tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
"""
import ctypes as C
import os
import subprocess
import sys

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(os.path.dirname(_HERE), "trackmaker-rs_b200"))
from tmkphy._abi import (CFG_DTYPE, DESC_DTYPE, REC_DTYPE, STREAM_DTYPE, FrameDesc, FrameRec,  # noqa: E402
                         PhyConfig, Stream)

LIB = os.path.join(_HERE, "liboracle.so")


def build(force=False):
    srcs = [os.path.join(_HERE, f) for f in ("oracle_capi.cpp", "phy_core.hpp", "phy_decoder.hpp", "ask_modem.hpp", "mcm_modem.hpp", "ip_fragments.hpp")]
    if force or not os.path.exists(LIB) or any(os.path.getmtime(s) > os.path.getmtime(LIB) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-B", "liboracle.so"], stdout=subprocess.DEVNULL)


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB):
            build()
        L = C.CDLL(LIB)
        vp, u64, u32 = C.c_void_p, C.c_uint64, C.c_uint32
        pcfg = C.POINTER(PhyConfig)
        sig = {
            "orc_crc8": (C.c_uint8, [vp, u64]),
            "orc_bytes_to_bits": (u64, [vp, u64, vp]),
            "orc_bits_to_bytes": (u64, [vp, u64, vp]),
            "orc_samples_for_bits": (u64, [C.c_int, u32, u64]),
            "orc_line_encode": (u64, [C.c_int, u32, vp, u64, vp, u64]),
            "orc_line_decode": (u64, [C.c_int, u32, vp, u64, vp, u64]),
            "orc_preamble": (u64, [C.c_int, u32, u32, vp, u64]),
            "orc_frame_to_bytes": (u64, [vp, vp, vp]),
            "orc_modulate": (u64, [pcfg, vp, u32, vp, u32, vp, u64]),
            "orc_decode_stream": (u64, [pcfg, vp, u64, u32, vp, u64, vp, u64, C.POINTER(u64)]),
            "orc_decode_feeds": (u64, [pcfg, vp, u64, vp, u32, u32, vp, vp, u64, vp, u64, C.POINTER(u64)]),
            "orc_decode_batch_mt": (u64, [vp, vp, u32, C.c_int, vp, u64, vp, u64, C.POINTER(u64),
                                          C.POINTER(C.c_double)]),
            "orc_correlate": (u64, [pcfg, vp, u64, vp]),
            "orc_rolling_energy": (None, [vp, u64, u32, u64, vp, u32, vp]),
            "orc_channel_busy": (C.c_int, [vp, u64]),
            "orc_hardware_threads": (C.c_int, []),
            "orc_pcm16_encode": (None, [vp, u64, vp]),
            "orc_pcm16_to_f32": (None, [vp, u64, vp]),
            "orc_fraginfo_to_u16": (C.c_uint16, [C.c_uint16, C.c_int, C.c_uint16]),
            "orc_fraginfo_from_u16": (None, [C.c_uint16, C.POINTER(C.c_int), C.POINTER(C.c_uint16)]),
            "orc_fragment_batch": (u64, [vp, vp, u32, u32, C.c_uint16, vp, u64, vp, u64, C.POINTER(u64),
                                         C.POINTER(C.c_uint16)]),
            "orc_reassemble_batch": (u64, [vp, u64, vp, u32, vp, u64, vp, u64, C.POINTER(u64)]),
            "orc_ask_tables": (None, [vp, vp, u64]),
            "orc_ask_modulate": (u64, [vp, u32, vp, vp, vp, u64]),
            "orc_ask_demodulate": (u64, [vp, u64, u32, vp, u64]),
            "orc_mcm_modulate": (u64, [C.c_int, vp, vp, vp, u32, vp, vp, vp, u64]),
            "orc_mcm_demodulate": (u64, [C.c_int, vp, u64, u32, vp, u64, vp, u64, C.POINTER(u64)]),
            "orc_mcm_fft": (None, [C.c_int, vp, u32, vp]),
        }
        for n, (r, a) in sig.items():
            f = getattr(L, n)
            f.restype, f.argtypes = r, a
        _lib = L
    return _lib


def make_config(line_code, local_addr=1, spl=3, preamble_bytes=2, threshold=0.9, max_frame_bytes=256, chunk=0):
    return PhyConfig(line_code, spl, preamble_bytes, local_addr, threshold, max_frame_bytes, chunk)


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def crc8(data: bytes) -> int:
    b = np.frombuffer(bytes(data), dtype=np.uint8)
    return lib().orc_crc8(_ptr(b) if len(b) else None, len(b))


def bytes_to_bits(data: bytes) -> np.ndarray:
    b = np.frombuffer(bytes(data), dtype=np.uint8)
    out = np.zeros(len(b) * 8, np.uint8)
    lib().orc_bytes_to_bits(_ptr(b), len(b), _ptr(out))
    return out


def bits_to_bytes(bits) -> bytes:
    b = np.ascontiguousarray(bits, dtype=np.uint8)
    out = np.zeros((len(b) + 7) // 8, np.uint8)
    lib().orc_bits_to_bytes(_ptr(b), len(b), _ptr(out))
    return out.tobytes()


def samples_for_bits(kind, spl, nbits):
    return lib().orc_samples_for_bits(kind, spl, nbits)


def line_encode(kind, spl, bits) -> np.ndarray:
    b = np.ascontiguousarray(bits, dtype=np.uint8)
    n = lib().orc_line_encode(kind, spl, _ptr(b), len(b), None, 0)
    out = np.zeros(n, np.float32)
    lib().orc_line_encode(kind, spl, _ptr(b), len(b), _ptr(out), n)
    return out


def line_decode(kind, spl, samples) -> np.ndarray:
    s = np.ascontiguousarray(samples, dtype=np.float32)
    out = np.zeros(len(s) + 8, np.uint8)
    n = lib().orc_line_decode(kind, spl, _ptr(s), len(s), _ptr(out), len(out))
    return out[:n]


def preamble(kind, spl=3, pbytes=2) -> np.ndarray:
    n = lib().orc_preamble(kind, spl, pbytes, None, 0)
    out = np.zeros(n, np.float32)
    lib().orc_preamble(kind, spl, pbytes, _ptr(out), n)
    return out


def make_descs(frames):
    """frames: list of (type, seq, src, dst, payload bytes) -> (desc array, payload array)"""
    d = np.zeros(len(frames), DESC_DTYPE)
    blob = bytearray()
    for i, (t, seq, src, dst, data) in enumerate(frames):
        d[i] = (len(blob), len(data), t, seq, src, dst)
        blob += bytes(data)
    pl = np.frombuffer(bytes(blob) + b"\0", dtype=np.uint8).copy()
    return d, pl


def frame_to_bytes(t, seq, src, dst, data) -> bytes:
    d, pl = make_descs([(t, seq, src, dst, data)])
    out = np.zeros(7 + len(data), np.uint8)
    n = lib().orc_frame_to_bytes(_ptr(d), _ptr(pl), _ptr(out))
    return out[:n].tobytes()


def modulate(cfg, frames, gap=48) -> np.ndarray:
    d, pl = make_descs(frames)
    n = lib().orc_modulate(C.byref(cfg), _ptr(d), len(d), _ptr(pl), gap, None, 0)
    out = np.zeros(n, np.float32)
    lib().orc_modulate(C.byref(cfg), _ptr(d), len(d), _ptr(pl), gap, _ptr(out), n)
    return out


def decode_stream(cfg, samples, stream_id=0):
    """-> (records ndarray[REC_DTYPE], payload ndarray[u8])"""
    s = np.ascontiguousarray(samples, dtype=np.float32)
    cap = max(16, len(s) // 8 + 16)
    recs = np.zeros(cap, REC_DTYPE)
    pl = np.zeros(max(1024, len(s) // 4), np.uint8)
    plen = C.c_uint64(0)
    n = lib().orc_decode_stream(C.byref(cfg), _ptr(s), len(s), stream_id, _ptr(recs), cap, _ptr(pl), len(pl),
                                C.byref(plen))
    assert n <= cap and plen.value <= len(pl)
    return recs[:n].copy(), pl[:plen.value].copy()


def decode_feeds(cfg, samples, feed_sizes, stream_id=0):
    """One PhyDecoder fed with process_samples calls of the given sizes.
    -> (records, payload, feed_of_rec): feed_of_rec[i] is the call that returned record i."""
    s = np.ascontiguousarray(samples, dtype=np.float32)
    fs = np.ascontiguousarray(feed_sizes, dtype=np.uint64)
    cap = max(16, len(s) // 8 + 16)
    recs = np.zeros(cap, REC_DTYPE)
    fo = np.zeros(cap, np.uint32)
    pl = np.zeros(max(1024, len(s) // 4), np.uint8)
    plen = C.c_uint64(0)
    n = lib().orc_decode_feeds(C.byref(cfg), _ptr(s), len(s), _ptr(fs), len(fs), stream_id, _ptr(recs), _ptr(fo), cap,
                               _ptr(pl), len(pl), C.byref(plen))
    assert n <= cap and plen.value <= len(pl)
    return recs[:n].copy(), pl[:plen.value].copy(), fo[:n].copy()


def decode_batch(samples, streams, n_threads=0, rec_cap=None, payload_cap=None):
    """-> (records, payload, seconds). streams: ndarray[STREAM_DTYPE]."""
    s = np.ascontiguousarray(samples, dtype=np.float32)
    st = np.ascontiguousarray(streams, dtype=STREAM_DTYPE)
    if n_threads <= 0:
        n_threads = os.cpu_count() or 1
    total = int(st["n_samples"].sum())
    rec_cap = rec_cap or max(1024, total // 64)
    payload_cap = payload_cap or max(4096, total // 8)
    recs = np.zeros(rec_cap, REC_DTYPE)
    pl = np.zeros(payload_cap, np.uint8)
    plen = C.c_uint64(0)
    sec = C.c_double(0)
    n = lib().orc_decode_batch_mt(_ptr(s), _ptr(st), len(st), n_threads, _ptr(recs), rec_cap, _ptr(pl),
                                  payload_cap, C.byref(plen), C.byref(sec))
    assert n <= rec_cap and plen.value <= payload_cap, "oracle output capacity too small"
    return recs[:n].copy(), pl[:plen.value].copy(), sec.value


def rolling_energy(samples, P, seed, windows) -> np.ndarray:
    """the reference's rolling window energy for a search entered at `seed`, at the ascending `windows`"""
    s = np.ascontiguousarray(samples, dtype=np.float32)
    w = np.ascontiguousarray(windows, dtype=np.uint64)
    out = np.zeros(len(w), np.float32)
    lib().orc_rolling_energy(_ptr(s), len(s), int(P), int(seed), _ptr(w), len(w), _ptr(out))
    return out


def correlate(cfg, samples) -> np.ndarray:
    s = np.ascontiguousarray(samples, dtype=np.float32)
    out = np.zeros(len(s), np.float32)
    n = lib().orc_correlate(C.byref(cfg), _ptr(s), len(s), _ptr(out))
    return out[:n]


def channel_busy(samples) -> int:
    s = np.ascontiguousarray(samples, dtype=np.float32)
    return lib().orc_channel_busy(_ptr(s), len(s))


def hardware_threads() -> int:
    return lib().orc_hardware_threads()


# ---- chirp + BPSK modem (examples/ask.rs)
def ask_tables(n_carrier=4752):
    pre, car = np.zeros(440, np.float32), np.zeros(n_carrier, np.float32)
    lib().orc_ask_tables(_ptr(pre), _ptr(car), n_carrier)
    return pre, car


def ask_modulate(bits, gap1, gap2) -> np.ndarray:
    """bits: (n_frames, 100) array of 0/1; gap1/gap2: zeros before / after each frame"""
    b = np.ascontiguousarray(bits, dtype=np.uint8).reshape(-1, 100)
    g1 = np.ascontiguousarray(gap1, dtype=np.uint32)
    g2 = np.ascontiguousarray(gap2, dtype=np.uint32)
    n = int(len(b) * (440 + 4752) + g1.sum() + g2.sum())
    out = np.zeros(max(n, 1), np.float32)
    got = lib().orc_ask_modulate(_ptr(b), len(b), _ptr(g1), _ptr(g2), _ptr(out), n)
    assert got == n
    return out[:n]


def ask_demodulate(samples, stream=0, cap=None) -> np.ndarray:
    from tmkphy._abi import ASK_REC_DTYPE
    s = np.ascontiguousarray(samples, dtype=np.float32)
    cap = cap or max(16, len(s) // 4000 + 4)
    recs = np.zeros(cap, ASK_REC_DTYPE)
    n = lib().orc_ask_demodulate(_ptr(s), len(s), stream, _ptr(recs), cap)
    assert n <= cap
    return recs[:n].copy()


# ---- multi-carrier modem (build-defined FSK / OFDM spec; parity unpinned by the reference)
def mcm_modulate(mode, payloads, seqs, gap1, gap2) -> np.ndarray:
    """payloads: list of bytes (<= 25 FSK / <= 100 OFDM each); seqs: one byte per frame"""
    off = np.zeros(len(payloads) + 1, np.uint32)
    off[1:] = np.cumsum([len(p) for p in payloads])
    blob = np.frombuffer(b"".join(payloads) + b"\0", np.uint8).copy()
    sq = np.ascontiguousarray(seqs, dtype=np.uint8)
    g1 = np.ascontiguousarray(gap1, dtype=np.uint32)
    g2 = np.ascontiguousarray(gap2, dtype=np.uint32)
    n = int(len(payloads) * (440 + 4752) + g1.sum() + g2.sum())
    out = np.zeros(max(n, 1), np.float32)
    got = lib().orc_mcm_modulate(mode, _ptr(blob), _ptr(off), _ptr(sq), len(payloads), _ptr(g1), _ptr(g2), _ptr(out), n)
    assert got == n
    return out[:n]


def mcm_demodulate(mode, samples, stream=0, cap=None):
    """-> (records in tmk_mcm_rec layout, payload bytes)"""
    from tmkphy._abi import MCM_REC_DTYPE
    s = np.ascontiguousarray(samples, dtype=np.float32)
    cap = cap or max(16, len(s) // 4000 + 4)
    recs = np.zeros(cap, MCM_REC_DTYPE)
    pay = np.zeros(cap * 100 + 16, np.uint8)
    pl = C.c_uint64(0)
    n = lib().orc_mcm_demodulate(mode, _ptr(s), len(s), stream, _ptr(recs), cap, _ptr(pay), len(pay), C.byref(pl))
    assert n <= cap and pl.value <= len(pay)
    return recs[:n].copy(), pay[:pl.value].copy()


def mcm_fft(rows) -> np.ndarray:
    """rows: (n_ffts, 64 | 256) real f32 -> complex64 spectra in the specification's operation order"""
    r = np.ascontiguousarray(rows, dtype=np.float32)
    out = np.zeros(r.shape + (2,), np.float32)
    lib().orc_mcm_fft(r.shape[1], _ptr(r), r.shape[0], _ptr(out))
    return out.view(np.complex64)[..., 0]


# ---- 16-bit PCM (dump.rs:37-40, codec.rs:121-125)
def pcm16_encode(samples) -> np.ndarray:
    s = np.ascontiguousarray(samples, dtype=np.float32)
    out = np.zeros(s.size, np.int16)
    lib().orc_pcm16_encode(_ptr(s), s.size, _ptr(out))
    return out


def pcm16_to_f32(pcm) -> np.ndarray:
    p = np.ascontiguousarray(pcm, dtype=np.int16)
    out = np.zeros(p.size, np.float32)
    lib().orc_pcm16_to_f32(_ptr(p), p.size, _ptr(out))
    return out


# ---- IP fragments (fragmentation.rs)
def fraginfo_to_u16(identification, more, offset) -> int:
    return lib().orc_fraginfo_to_u16(identification, int(more), offset)


def fraginfo_from_u16(v):
    more, off = C.c_int(0), C.c_uint16(0)
    lib().orc_fraginfo_from_u16(v, C.byref(more), C.byref(off))
    return bool(more.value), int(off.value)


def fragment_batch(packets, mtu, first_identification=0):
    """One IpFragmenter over a list of packets: (records[PKT_DTYPE], bytes, next_identification)."""
    from tmkphy._abi import PKT_DTYPE
    off = np.zeros(len(packets) + 1, np.uint64)
    if packets:
        off[1:] = np.cumsum([len(p) for p in packets], dtype=np.uint64)
    data = np.frombuffer(b"".join(bytes(p) for p in packets) + b"\0", dtype=np.uint8).copy()
    cap = len(packets) + int(off[-1]) // 8 + 8
    bcap = int(off[-1]) + 60 * cap + 64
    fr, by = np.zeros(cap, PKT_DTYPE), np.zeros(bcap, np.uint8)
    nb, nid = C.c_uint64(0), C.c_uint16(0)
    n = lib().orc_fragment_batch(_ptr(data), _ptr(off), len(packets), int(mtu), int(first_identification) & 0xffff,
                                 _ptr(fr), cap, _ptr(by), bcap, C.byref(nb), C.byref(nid))
    assert n <= cap and nb.value <= bcap
    return fr[:n].copy(), by[:nb.value].copy(), int(nid.value)


def reassemble_batch(recs, payload, n_links):
    from tmkphy._abi import PKT_DTYPE
    r = np.ascontiguousarray(recs, dtype=REC_DTYPE)
    pl = np.ascontiguousarray(payload, dtype=np.uint8)
    if pl.size == 0:
        pl = np.zeros(1, np.uint8)
    cap, bcap = max(16, len(r)), max(4096, pl.size + 64 * len(r))
    pk, by = np.zeros(cap, PKT_DTYPE), np.zeros(bcap, np.uint8)
    nb = C.c_uint64(0)
    n = lib().orc_reassemble_batch(_ptr(r), len(r), _ptr(pl), int(n_links), _ptr(pk), cap, _ptr(by), bcap, C.byref(nb))
    assert n <= cap and nb.value <= bcap
    return pk[:n].copy(), by[:nb.value].copy()


# ---- WAV container (dump.rs:23-46 writes, codec.rs:114-148 reads): numpy restatement, test infrastructure
def wav_to_f32(data: bytes) -> np.ndarray:
    """channel 0 of a RIFF/WAVE file as the reference's reader converts it (convert_audio_buffer_to_f32,
    src/audio/codec.rs:114-148): U8 (s - 128) / 128, S16 s / 32768, S24 s / 8388608, S32 s / 2147483648, F32 as is."""
    import struct
    assert data[:4] == b"RIFF" and data[8:12] == b"WAVE"
    pos, fmt = 12, None
    while pos + 8 <= len(data):
        cid, sz = data[pos:pos + 4], struct.unpack("<I", data[pos + 4:pos + 8])[0]
        body = pos + 8
        if cid == b"fmt ":
            tag, ch, rate, _, align, bits = struct.unpack("<HHIIHH", data[body:body + 16])
            if tag == 0xFFFE:
                tag = struct.unpack("<H", data[body + 24:body + 26])[0]
            fmt = (tag, ch, align, bits)
        elif cid == b"data":
            tag, ch, align, bits = fmt
            raw = np.frombuffer(data[body:body + sz], dtype=np.uint8)
            n = len(raw) // align
            fr = raw[:n * align].reshape(n, align)[:, :bits // 8].astype(np.uint32)  # channel 0
            if tag == 3:
                v = fr[:, 0] | (fr[:, 1] << 8) | (fr[:, 2] << 16) | (fr[:, 3] << 24)
                return v.astype(np.uint32).view(np.float32).copy()
            if bits == 8:
                return ((fr[:, 0].astype(np.float32) - np.float32(128.0)) / np.float32(128.0)).astype(np.float32)
            if bits == 16:
                v = (fr[:, 0] | (fr[:, 1] << 8)).astype(np.uint16).view(np.int16)
                return (v.astype(np.float32) / np.float32(32768.0)).astype(np.float32)
            if bits == 24:
                v = ((fr[:, 0] << 8) | (fr[:, 1] << 16) | (fr[:, 2] << 24)).astype(np.uint32).view(np.int32) >> 8
                return (v.astype(np.float32) / np.float32(8388608.0)).astype(np.float32)
            v = (fr[:, 0] | (fr[:, 1] << 8) | (fr[:, 2] << 16) | (fr[:, 3] << 24)).astype(np.uint32).view(np.int32)
            return (v.astype(np.float32) / np.float32(2147483648.0)).astype(np.float32)
        pos = body + sz + (sz & 1)
    raise ValueError("no data chunk")
