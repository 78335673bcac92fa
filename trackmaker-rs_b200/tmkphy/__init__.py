"""tmkphy — host-side mirror of the reference's PHY interface over libtmkphy's C ABI.

Names, argument meaning and error behaviour follow src/phy/mod.rs:4-13 of the reference:
  Frame / FrameType          src/phy/frame.rs:8-72
  LineCodingKind             src/phy/line_coding.rs:26-49
  PhyEncoder                 src/phy/encoder.rs:5-97
  PhyDecoder                 src/phy/decoder.rs:16-119
plus the batched entry point the reference does not have (decode_batch: many captures at once).

Everything that computes goes through libtmkphy.so (CUDA, sm_100a). There is no CPU path: if the
library is missing or no B200 is visible, constructing an Engine raises.
"""
import ctypes as C
from dataclasses import dataclass
from enum import IntEnum

import numpy as np

from . import _abi
from ._abi import (DESC_DTYPE, REC_DTYPE, STREAM_DTYPE, FRAME_OK, LINE_4B5B, LINE_MANCHESTER,  # noqa: F401
                   STATUS_NAMES, STAT_NAMES, PhyConfig)

# src/utils/consts.rs:24-48
SAMPLE_RATE = 48000
SAMPLES_PER_LEVEL = 3
PREAMBLE_PATTERN_BYTES = 2
MAX_FRAME_DATA_SIZE = 128
INTER_FRAME_GAP_SAMPLES = 48
PHY_HEADER_BYTES = 7


class TmkError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"libtmkphy error {code}: {msg}")
        self.code = code


class LineCodingKind(IntEnum):  # line_coding.rs:26-38
    Manchester = LINE_MANCHESTER
    FourBFiveB = LINE_4B5B

    def display_name(self):
        return "Manchester" if self == LineCodingKind.Manchester else "4B5B"


class FrameType(IntEnum):  # frame.rs:12-32
    Data = 0x01
    Ack = 0x02


@dataclass
class Frame:  # frame.rs:34-41
    frame_type: FrameType
    sequence: int
    src: int
    dst: int
    data: bytes = b""

    @staticmethod
    def new_data(sequence, src, dst, data):  # frame.rs:60-62
        return Frame(FrameType.Data, sequence & 0xFF, src & 0xFF, dst & 0xFF, bytes(data))

    @staticmethod
    def new_ack(sequence, frm, to):  # frame.rs:64-66
        return Frame(FrameType.Ack, sequence & 0xFF, frm & 0xFF, to & 0xFF, b"")


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None and a.size else None


class Engine:
    """One engine per GPU (tmk_engine). Not thread-safe: one host thread drives it."""

    def __init__(self, device=0):
        self._lib = _abi.load_library()
        h = C.c_void_p()
        rc = self._lib.tmk_engine_create(int(device), C.byref(h))
        if rc != 0:
            raise TmkError(rc, "tmk_engine_create failed: no usable sm_100 CUDA device (there is no CPU fallback)")
        self._h = h
        self.device = device

    def close(self):
        if getattr(self, "_h", None):
            self._lib.tmk_engine_destroy(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != 0:
            raise TmkError(rc, self._lib.tmk_last_error(self._h).decode())

    # ---- plumbing
    def device_alloc(self, nbytes):
        p = C.c_void_p()
        self._check(self._lib.tmk_device_alloc(self._h, int(nbytes), C.byref(p)))
        return p.value

    def device_free(self, p):
        self._check(self._lib.tmk_device_free(self._h, C.c_void_p(p)))

    def pinned_alloc(self, nbytes):
        p = C.c_void_p()
        self._check(self._lib.tmk_host_alloc_pinned(self._h, int(nbytes), C.byref(p)))
        return p.value

    def pinned_free(self, p):
        self._check(self._lib.tmk_host_free_pinned(self._h, C.c_void_p(p)))

    def pinned_array(self, shape, dtype):
        """numpy array over pinned host memory (freed with pinned_free(arr.ctypes.data))"""
        dtype = np.dtype(dtype)
        n = int(np.prod(shape))
        p = self.pinned_alloc(max(1, n * dtype.itemsize))
        buf = (C.c_char * (n * dtype.itemsize)).from_address(p)
        return np.frombuffer(buf, dtype=dtype, count=n).reshape(shape)

    def h2d(self, dptr, arr):
        arr = np.ascontiguousarray(arr)
        self._check(self._lib.tmk_memcpy_h2d(self._h, C.c_void_p(dptr), _ptr(arr), arr.nbytes))

    def d2h(self, arr, dptr):
        self._check(self._lib.tmk_memcpy_d2h(self._h, _ptr(arr), C.c_void_p(dptr), arr.nbytes))

    def synchronize(self):
        self._check(self._lib.tmk_synchronize(self._h))

    def stream_handle(self):
        return self._lib.tmk_engine_stream(self._h)

    def timer_start(self):
        self._check(self._lib.tmk_timer_start(self._h))

    def timer_stop(self):
        ms = C.c_float(0)
        self._check(self._lib.tmk_timer_stop(self._h, C.byref(ms)))
        return ms.value

    def pci_bus_id(self):
        buf = C.create_string_buffer(32)
        self._check(self._lib.tmk_device_pci_bus_id(self._h, buf, 32))
        return buf.value.decode()

    def bind_host_thread(self):
        """Pin the calling thread to the CPUs of the GPU's NUMA node (sysfs local_cpulist), so that pinned capture
        buffers allocated next are node-local and H2D copies do not cross sockets. Returns the previous affinity
        (restore with os.sched_setaffinity(0, prev)), or None when the topology is not exposed."""
        import os
        try:
            path = f"/sys/bus/pci/devices/{self.pci_bus_id()}/local_cpulist"
            cpus = set()
            for part in open(path).read().strip().split(","):
                if part:
                    a, _, b = part.partition("-")
                    cpus.update(range(int(a), int(b or a) + 1))
            prev = os.sched_getaffinity(0)
            cpus &= prev
            if not cpus or cpus == prev:
                return None
            os.sched_setaffinity(0, cpus)
            return prev
        except (OSError, ValueError, TmkError):
            return None

    def launch_count(self):
        return int(self._lib.tmk_launch_count(self._h))

    def last_kernel_ms(self):
        ms = C.c_float(0)
        self._check(self._lib.tmk_last_kernel_ms(self._h, C.byref(ms)))
        return ms.value

    def last_stage_ms(self):
        ms = (C.c_float * 4)()
        self._check(self._lib.tmk_last_stage_ms(self._h, ms))
        return {"scan": ms[0], "walk": ms[1], "order": ms[2], "h2d": ms[3]}

    def stats(self):
        c = (C.c_uint64 * _abi.STAT_COUNT)()
        self._check(self._lib.tmk_stats(self._h, c))
        return dict(zip(STAT_NAMES, [int(v) for v in c]))

    def stats_reset(self):
        self._check(self._lib.tmk_stats_reset(self._h))

    def debug_counters(self):
        """walk diagnostics of the last decode_batch call (tmk_debug_counters)"""
        c = (C.c_uint64 * 8)()
        self._check(self._lib.tmk_debug_counters(self._h, c))
        return dict(zip(("blocks_examined", "reanchors", "replay_steps", "summary_groups", "cluster_us",
                         "parallel_replay_steps", "parallel_replay_doubts"), [int(v) for v in c[:7]]))

    # ---- chirp + BPSK modem of examples/ask.rs
    def ask_tables(self, n_carrier=4752):
        pre, car = np.zeros(440, np.float32), np.zeros(n_carrier, np.float32)
        self._check(self._lib.tmk_ask_tables(self._h, _ptr(pre), _ptr(car), n_carrier))
        return pre, car

    def ask_modulate(self, bits, gap1, gap2):
        """TX of examples/ask.rs:608-637. bits: (n_frames, 100) of 0/1; gaps: zeros before / after each frame."""
        b = np.ascontiguousarray(bits, dtype=np.uint8).reshape(-1, 100)
        g1 = np.ascontiguousarray(gap1, dtype=np.uint32)
        g2 = np.ascontiguousarray(gap2, dtype=np.uint32)
        n = int(self._lib.tmk_ask_modulated_len(len(b), _ptr(g1), _ptr(g2)))
        out = np.zeros(max(n, 1), np.float32)
        ln = C.c_uint64(0)
        self._check(self._lib.tmk_ask_modulate(self._h, _ptr(b), len(b), _ptr(g1), _ptr(g2), _ptr(out), n, C.byref(ln)))
        return out[:int(ln.value)]

    def ask_decode_batch(self, samples, streams, rec_cap=None, total_samples=None):
        """RX of examples/ask.rs:668-796 over a batch of captures -> ndarray[ASK_REC_DTYPE]"""
        st = np.ascontiguousarray(streams, dtype=STREAM_DTYPE)
        if isinstance(samples, np.ndarray):
            s = np.ascontiguousarray(samples, dtype=np.float32)
            sp, total = _ptr(s), s.size
        else:
            sp, total = C.c_void_p(int(samples)), int(total_samples)
        tot = int(st["n_samples"].sum()) if len(st) else 0
        rec_cap = int(rec_cap) if rec_cap is not None else max(64, tot // 4096 + 2 * len(st))
        recs = np.zeros(rec_cap, _abi.ASK_REC_DTYPE)
        n = C.c_uint64(0)
        self._check(self._lib.tmk_ask_decode_batch(self._h, sp, total, _ptr(st), len(st), _ptr(recs), rec_cap, C.byref(n)))
        return recs[:int(n.value)].copy()

    def ask_last_stage_ms(self):
        ms = (C.c_float * 2)()
        self._check(self._lib.tmk_ask_last_stage_ms(self._h, ms))
        return {"scan": ms[0], "walk": ms[1]}

    # ---- multi-carrier modem (build-defined FSK / OFDM spec on the batched FFT; parity unpinned by the reference)
    def mcm_modulate(self, mode, payloads, seqs, gap1, gap2):
        """payloads: list of bytes (<= 25 FSK / <= 100 OFDM each) -> float32 track"""
        off = np.zeros(len(payloads) + 1, np.uint32)
        off[1:] = np.cumsum([len(p) for p in payloads])
        blob = np.frombuffer(b"".join(payloads) + b"\0", np.uint8).copy()
        sq = np.ascontiguousarray(seqs, dtype=np.uint8)
        g1 = np.ascontiguousarray(gap1, dtype=np.uint32)
        g2 = np.ascontiguousarray(gap2, dtype=np.uint32)
        n = int(self._lib.tmk_mcm_modulated_len(len(payloads), _ptr(g1), _ptr(g2)))
        out = np.zeros(max(n, 1), np.float32)
        ln = C.c_uint64(0)
        self._check(self._lib.tmk_mcm_modulate(self._h, int(mode), _ptr(blob), _ptr(off), _ptr(sq), len(payloads), _ptr(g1),
                                               _ptr(g2), _ptr(out), n, C.byref(ln)))
        return out[:int(ln.value)]

    def mcm_decode_batch(self, mode, samples, streams, rec_cap=None, total_samples=None):
        """-> (ndarray[MCM_REC_DTYPE] sorted by (stream, start_index), payload bytes)"""
        st = np.ascontiguousarray(streams, dtype=STREAM_DTYPE)
        if isinstance(samples, np.ndarray):
            s = np.ascontiguousarray(samples, dtype=np.float32)
            sp, total = _ptr(s), s.size
        else:
            sp, total = C.c_void_p(int(samples)), int(total_samples)
        tot = int(st["n_samples"].sum()) if len(st) else 0
        rec_cap = int(rec_cap) if rec_cap is not None else max(64, tot // 4096 + 2 * len(st))
        recs = np.zeros(rec_cap, _abi.MCM_REC_DTYPE)
        pay = np.zeros(rec_cap * _abi.MCM_MAX_PAYLOAD.get(int(mode), 100) + 16, np.uint8)
        n, pl = C.c_uint64(0), C.c_uint64(0)
        self._check(self._lib.tmk_mcm_decode_batch(self._h, int(mode), sp, total, _ptr(st), len(st), _ptr(recs), rec_cap,
                                                   C.byref(n), _ptr(pay), len(pay), C.byref(pl)))
        return recs[:int(n.value)].copy(), pay[:int(pl.value)].copy()

    def mcm_synth_batch(self, mode, seed, n_streams, n_per_stream, snr_db, multipath=False, device_ptr=None):
        """-> (samples or None when device_ptr is given, frames planted)"""
        nf = C.c_uint64(0)
        if device_ptr is None:
            out = np.zeros(n_streams * n_per_stream, np.float32)
            sp = _ptr(out)
        else:
            out, sp = None, C.c_void_p(device_ptr)
        self._check(self._lib.tmk_mcm_synth_batch(self._h, int(mode), seed, n_streams, n_per_stream, C.c_float(snr_db),
                                                  1 if multipath else 0, sp, C.byref(nf)))
        return out, int(nf.value)

    def mcm_fft(self, rows):
        """rows: (n_ffts, 64 | 256) real f32 -> complex64 spectra (the kernel's own FFT; parity probe)"""
        r = np.ascontiguousarray(rows, dtype=np.float32)
        out = np.zeros(r.shape + (2,), np.float32)
        self._check(self._lib.tmk_mcm_fft_probe(self._h, r.shape[1], _ptr(r), r.shape[0], _ptr(out)))
        return out.view(np.complex64)[..., 0]

    def mcm_last_stage_ms(self):
        ms = (C.c_float * 3)()
        self._check(self._lib.tmk_mcm_last_stage_ms(self._h, ms))
        return {"filter": ms[0], "walk": ms[1], "demod": ms[2]}

    # ---- the path
    def modulate(self, cfg, frames, gap=INTER_FRAME_GAP_SAMPLES, out_device_ptr=None):
        """PhyEncoder::encode_frames. frames: list of Frame. Returns float32 samples (host)."""
        d = np.zeros(len(frames), DESC_DTYPE)
        blob = bytearray()
        for i, f in enumerate(frames):
            d[i] = (len(blob), len(f.data), int(f.frame_type), f.sequence, f.src, f.dst)
            blob += bytes(f.data)
        pl = np.frombuffer(bytes(blob) + b"\0", dtype=np.uint8).copy()
        n = self._lib.tmk_modulated_len(C.byref(cfg), _ptr(d), len(d), gap)
        out_len = C.c_uint64(0)
        if out_device_ptr is not None:
            self._check(self._lib.tmk_modulate(self._h, C.byref(cfg), _ptr(d), len(d), _ptr(pl), len(blob), gap,
                                               C.c_void_p(out_device_ptr), n, C.byref(out_len)))
            return int(out_len.value)
        out = np.zeros(n, np.float32)
        self._check(self._lib.tmk_modulate(self._h, C.byref(cfg), _ptr(d), len(d), _ptr(pl), len(blob), gap,
                                           _ptr(out), n, C.byref(out_len)))
        assert out_len.value == n
        return out

    def decode_batch(self, samples, streams, rec_cap=None, payload_cap=None, total_samples=None):
        """tmk_decode_batch. samples: float32 ndarray (host) or an int device pointer;
        streams: ndarray[STREAM_DTYPE]. Returns (records ndarray[REC_DTYPE], payload ndarray[u8])."""
        st = np.ascontiguousarray(streams, dtype=STREAM_DTYPE)
        if isinstance(samples, np.ndarray):
            s = np.ascontiguousarray(samples, dtype=np.float32)
            sp, total = _ptr(s), s.size
        else:
            sp, total = C.c_void_p(int(samples)), int(total_samples)
        tot = int(st["n_samples"].sum()) if len(st) else 0
        rec_cap = int(rec_cap) if rec_cap is not None else max(1024, tot // 96 + 2 * len(st))
        payload_cap = int(payload_cap) if payload_cap is not None else max(65536, tot // 12)
        recs = np.zeros(rec_cap, REC_DTYPE)
        pl = np.zeros(payload_cap, np.uint8)
        nr, npl = C.c_uint64(0), C.c_uint64(0)
        rc = self._lib.tmk_decode_batch(self._h, sp, total, _ptr(st), len(st), _ptr(recs), rec_cap, C.byref(nr),
                                        _ptr(pl), payload_cap, C.byref(npl))
        if rc == _abi.E_CAPACITY and (nr.value > rec_cap or npl.value > payload_cap):
            return self.decode_batch(samples, streams, max(rec_cap, nr.value), max(payload_cap, npl.value),
                                     total_samples)
        self._check(rc)
        return recs[:nr.value].copy(), pl[:npl.value].copy()

    # ---- 16-bit PCM captures (src/utils/dump.rs:37-40, src/audio/codec.rs:121-125)
    def pcm16_encode(self, samples):
        s = np.ascontiguousarray(samples, dtype=np.float32)
        out = np.zeros(s.size, np.int16)
        self._check(self._lib.tmk_pcm16_encode(self._h, _ptr(s), s.size, _ptr(out)))
        return out

    def pcm16_to_f32(self, pcm):
        p = np.ascontiguousarray(pcm, dtype=np.int16)
        out = np.zeros(p.size, np.float32)
        self._check(self._lib.tmk_pcm16_to_f32(self._h, _ptr(p), p.size, _ptr(out)))
        return out

    def decode_batch_pcm16(self, pcm, streams, rec_cap=None, payload_cap=None, total_samples=None):
        """tmk_decode_batch_pcm16. pcm: int16 ndarray (host) or an int device pointer."""
        st = np.ascontiguousarray(streams, dtype=STREAM_DTYPE)
        if isinstance(pcm, np.ndarray):
            s = np.ascontiguousarray(pcm, dtype=np.int16)
            sp, total = _ptr(s), s.size
        else:
            sp, total = C.c_void_p(int(pcm)), int(total_samples)
        tot = int(st["n_samples"].sum()) if len(st) else 0
        rec_cap = int(rec_cap) if rec_cap is not None else max(1024, tot // 96 + 2 * len(st))
        payload_cap = int(payload_cap) if payload_cap is not None else max(65536, tot // 12)
        recs = np.zeros(rec_cap, REC_DTYPE)
        pl = np.zeros(payload_cap, np.uint8)
        nr, npl = C.c_uint64(0), C.c_uint64(0)
        self._check(self._lib.tmk_decode_batch_pcm16(self._h, sp, total, _ptr(st), len(st), _ptr(recs), rec_cap,
                                                     C.byref(nr), _ptr(pl), payload_cap, C.byref(npl)))
        return recs[:nr.value].copy(), pl[:npl.value].copy()

    # ---- IP fragments above the PHY (src/net/fragmentation.rs)
    def reassemble_batch(self, recs, payload, n_links, pkt_cap=None, byte_cap=None):
        """tmk_reassemble_batch: (packet records ndarray[PKT_DTYPE], bytes ndarray[u8])."""
        r = np.ascontiguousarray(recs, dtype=REC_DTYPE)
        pl = np.ascontiguousarray(payload, dtype=np.uint8)
        pkt_cap = int(pkt_cap) if pkt_cap is not None else max(16, len(r))
        byte_cap = int(byte_cap) if byte_cap is not None else max(4096, pl.size + 64 * len(r))
        pk = np.zeros(pkt_cap, _abi.PKT_DTYPE)
        by = np.zeros(byte_cap, np.uint8)
        n, nb = C.c_uint64(0), C.c_uint64(0)
        self._check(self._lib.tmk_reassemble_batch(self._h, _ptr(r), len(r), _ptr(pl), pl.size, int(n_links), _ptr(pk),
                                                   pkt_cap, C.byref(n), _ptr(by), byte_cap, C.byref(nb)))
        return pk[:n.value].copy(), by[:nb.value].copy()

    def fragment_batch(self, packets, mtu, first_identification=0, frag_cap=None, byte_cap=None):
        """tmk_fragment_batch over a list of bytes objects sent through one IpFragmenter.
        Returns (fragment records, bytes, next_identification)."""
        off = np.zeros(len(packets) + 1, np.uint64)
        off[1:] = np.cumsum([len(p) for p in packets], dtype=np.uint64) if packets else 0
        data = np.frombuffer(b"".join(bytes(p) for p in packets) + b"\0", dtype=np.uint8).copy()
        if frag_cap is None:
            frag_cap = len(packets) + sum(len(p) // max(8, (int(mtu) - 60) // 8 * 8) + 1 for p in packets)
        if byte_cap is None:
            byte_cap = int(off[-1]) + 60 * int(frag_cap) + 64
        fr = np.zeros(int(frag_cap), _abi.PKT_DTYPE)
        by = np.zeros(int(byte_cap), np.uint8)
        n, nb, nid = C.c_uint64(0), C.c_uint64(0), C.c_uint16(0)
        self._check(self._lib.tmk_fragment_batch(self._h, _ptr(data), _ptr(off), len(packets), int(mtu),
                                                 int(first_identification) & 0xffff, _ptr(fr), int(frag_cap), C.byref(n),
                                                 _ptr(by), int(byte_cap), C.byref(nb), C.byref(nid)))
        return fr[:n.value].copy(), by[:nb.value].copy(), int(nid.value)

    def correlate(self, samples, streams):
        s = np.ascontiguousarray(samples, dtype=np.float32)
        st = np.ascontiguousarray(streams, dtype=STREAM_DTYPE)
        out = np.zeros(s.size, np.float32)
        self._check(self._lib.tmk_correlate(self._h, _ptr(s), s.size, _ptr(st), len(st), _ptr(out)))
        return out

    def scan_probe(self, samples, streams):
        """raw scan-kernel output: (blk_info[n_blocks,4], cand_bits[n_blocks, block / 32], blk_base[n_streams]);
        block = scan_block() windows"""
        s = np.ascontiguousarray(samples, dtype=np.float32)
        st = np.ascontiguousarray(streams, dtype=STREAM_DTYPE)
        blk = self.scan_block()
        cap = int(sum((int(n) + blk - 1) // blk for n in st["n_samples"])) + 16 * len(st) + 16
        info = np.zeros((cap, 4), np.float32)
        bits = np.zeros((cap, blk // 32), np.uint32)
        base = np.zeros(len(st), np.uint64)
        nb = C.c_uint64(0)
        self._check(self._lib.tmk_scan_probe(self._h, _ptr(s), s.size, _ptr(st), len(st), _ptr(info), _ptr(bits), cap,
                                             C.byref(nb), _ptr(base)))
        return info[:nb.value], bits[:nb.value], base

    def scan_block(self):
        return int(self._lib.tmk_scan_block())

    def scan_dot_probe(self, samples, streams):
        """the filter's correlation numerators per window (tmk_scan_dot_probe)"""
        s = np.ascontiguousarray(samples, dtype=np.float32)
        st = np.ascontiguousarray(streams, dtype=STREAM_DTYPE)
        out = np.zeros(s.size, np.float32)
        self._check(self._lib.tmk_scan_dot_probe(self._h, _ptr(s), s.size, _ptr(st), len(st), _ptr(out)))
        return out

    def replay_probe(self, cfg, samples, seed, windows):
        """tmk_replay_probe -> (E_ref of the windows for a search entered at `seed`, counters dict)"""
        s = np.ascontiguousarray(samples, dtype=np.float32)
        w = np.ascontiguousarray(windows, dtype=np.uint64)
        out = np.zeros(len(w), np.float32)
        cnt = (C.c_uint64 * 3)()
        self._check(self._lib.tmk_replay_probe(self._h, _ptr(s), s.size, C.byref(cfg), int(seed), _ptr(w), len(w),
                                               _ptr(out), cnt))
        return out, {"literal_steps": int(cnt[0]), "parallel_steps": int(cnt[1]), "parallel_doubts": int(cnt[2])}

    def channel_busy(self, samples, streams):
        s = np.ascontiguousarray(samples, dtype=np.float32)
        st = np.ascontiguousarray(streams, dtype=STREAM_DTYPE)
        out = np.zeros(len(st), np.int8)
        self._check(self._lib.tmk_channel_busy(self._h, _ptr(s), s.size, _ptr(st), len(st), _ptr(out)))
        return out

    def channel_busy_noalloc(self, device_ptr, total_samples, streams):
        """tmk_channel_busy_noalloc: samples resident at `device_ptr`; all device scratch is the library's static array"""
        st = np.ascontiguousarray(streams, dtype=STREAM_DTYPE)
        out = np.zeros(len(st), np.int8)
        self._check(self._lib.tmk_channel_busy_noalloc(self._h, C.c_void_p(device_ptr), int(total_samples), _ptr(st), len(st),
                                                       _ptr(out)))
        return out

    def synth_batch(self, cfg, seed, n_streams, n_per_stream, snr_db, amplitude=1.0, device_ptr=None,
                    want_truth=True):
        """Synthetic captures. Returns (samples or None when device_ptr is given, truth records)."""
        total = n_streams * n_per_stream
        cap = (total // 300 + n_streams + 16) if want_truth else 0
        truth = np.zeros(cap, REC_DTYPE) if want_truth else None
        nt = C.c_uint64(0)
        if device_ptr is None:
            out = np.zeros(total, np.float32)
            sp = _ptr(out)
        else:
            out, sp = None, C.c_void_p(int(device_ptr))
        self._check(self._lib.tmk_synth_batch(self._h, C.byref(cfg), seed, n_streams, n_per_stream, snr_db, amplitude,
                                              sp, _ptr(truth) if want_truth else None, cap, C.byref(nt)))
        return out, (truth[:nt.value].copy() if want_truth else int(nt.value))


    def synth_batch_ex(self, cfg, seed, n_streams, n_per_stream, snr_db, amplitude=1.0, multipath=False, drift_ppm=0.0,
                       gap=(0, 0), amp_min=0.0, device_ptr=None, want_truth=True):
        """tmk_synth_batch_ex: the generator with the impairments of SURVEY §8d (multipath FIR, clock drift,
        per-stream amplitude, sparse frame plans)."""
        total = n_streams * n_per_stream
        cap = (total // 300 + n_streams + 16) if want_truth else 0
        truth = np.zeros(cap, REC_DTYPE) if want_truth else None
        nt = C.c_uint64(0)
        prm = _abi.SynthParams(snr_db, amplitude, 1 if multipath else 0, drift_ppm, int(gap[0]), int(gap[1]), amp_min, 0)
        if device_ptr is None:
            out = np.zeros(total, np.float32)
            sp = _ptr(out)
        else:
            out, sp = None, C.c_void_p(int(device_ptr))
        self._check(self._lib.tmk_synth_batch_ex(self._h, C.byref(cfg), seed, n_streams, n_per_stream, C.byref(prm), sp,
                                                 _ptr(truth) if want_truth else None, cap, C.byref(nt)))
        return out, (truth[:nt.value].copy() if want_truth else int(nt.value))

    def ask_synth_batch(self, seed, n_streams, n_per_stream, snr_db, device_ptr=None):
        """tmk_ask_synth_batch: chirp + BPSK captures. -> (samples or None, frames planted)"""
        total = n_streams * n_per_stream
        nf = C.c_uint64(0)
        if device_ptr is None:
            out = np.zeros(total, np.float32)
            sp = _ptr(out)
        else:
            out, sp = None, C.c_void_p(int(device_ptr))
        self._check(self._lib.tmk_ask_synth_batch(self._h, seed, n_streams, n_per_stream, snr_db, sp, C.byref(nf)))
        return out, int(nf.value)

    # ---- WAV container ingest (src/utils/dump.rs:23-46 writes them, src/audio/codec.rs:114-148 reads them)
    def wav_to_f32(self, files):
        """channel 0 of every WAVE file (list of bytes) as f32 -> list of float32 arrays"""
        blob, off = _blob(files)
        offs, lens = np.zeros(len(files), np.uint64), np.zeros(len(files), np.uint64)
        total = C.c_uint64(0)
        rc = self._lib.tmk_wav_to_f32(self._h, _ptr(blob), blob.size - 1, _ptr(off), len(files), None, 0, _ptr(offs),
                                      _ptr(lens), C.byref(total))
        if rc not in (0, _abi.E_CAPACITY):
            self._check(rc)
        out = np.zeros(max(1, total.value), np.float32)
        self._check(self._lib.tmk_wav_to_f32(self._h, _ptr(blob), blob.size - 1, _ptr(off), len(files), _ptr(out),
                                             out.size, _ptr(offs), _ptr(lens), C.byref(total)))
        return [out[int(o):int(o) + int(n)].copy() for o, n in zip(offs, lens)]

    def decode_batch_wav(self, files, cfgs, rec_cap=None, payload_cap=None):
        """tmk_decode_batch_wav: WAVE files (list of bytes) in, (records, payload, streams) out; record.stream = file"""
        blob, off = _blob(files)
        cf = (PhyConfig * len(files))(*(cfgs if isinstance(cfgs, (list, tuple)) else [cfgs] * len(files)))
        st = np.zeros(len(files), STREAM_DTYPE)
        tot = blob.size // 2
        rec_cap = int(rec_cap) if rec_cap is not None else max(1024, tot // 96 + 2 * len(files))
        payload_cap = int(payload_cap) if payload_cap is not None else max(65536, tot // 12)
        recs = np.zeros(rec_cap, REC_DTYPE)
        pl = np.zeros(payload_cap, np.uint8)
        nr, npl = C.c_uint64(0), C.c_uint64(0)
        self._check(self._lib.tmk_decode_batch_wav(self._h, _ptr(blob), blob.size - 1, _ptr(off), len(files), cf,
                                                   _ptr(st), _ptr(recs), rec_cap, C.byref(nr), _ptr(pl), payload_cap,
                                                   C.byref(npl)))
        return recs[:nr.value].copy(), pl[:npl.value].copy(), st

    # ---- counter reduction over the GPUs of a box (NCCL, one process per GPU)
    def comm_init(self, unique_id, rank, world):
        buf = C.create_string_buffer(bytes(unique_id), 128)
        self._check(self._lib.tmk_comm_init(self._h, buf, int(rank), int(world)))

    def comm_size(self):
        return int(self._lib.tmk_comm_size(self._h))

    def stats_allreduce(self):
        c = (C.c_uint64 * _abi.STAT_COUNT)()
        self._check(self._lib.tmk_stats_allreduce(self._h, c))
        return dict(zip(STAT_NAMES, [int(v) for v in c]))

    def allreduce_sum(self, values):
        c = (C.c_uint64 * len(values))(*[int(v) for v in values])
        self._check(self._lib.tmk_comm_allreduce_sum_u64(self._h, c, len(values)))
        return [int(v) for v in c]

    def allreduce_max(self, values):
        c = (C.c_double * len(values))(*[float(v) for v in values])
        self._check(self._lib.tmk_comm_allreduce_max_f64(self._h, c, len(values)))
        return [float(v) for v in c]

    def allgather(self, values):
        """-> list (one per rank) of lists: every rank's values (tmk_comm_allgather_u64; at most 8 values)"""
        n, world = len(values), self._lib.tmk_comm_size(self._h)
        mine = (C.c_uint64 * n)(*[int(v) for v in values])
        out = (C.c_uint64 * (n * world))()
        self._check(self._lib.tmk_comm_allgather_u64(self._h, mine, n, out))
        return [[int(out[r * n + i]) for i in range(n)] for r in range(world)]

    def comm_barrier(self):
        self._check(self._lib.tmk_comm_barrier(self._h))

    def comm_destroy(self):
        self._check(self._lib.tmk_comm_destroy(self._h))


class IpReassembler:
    """IpReassembler kept across receive polls (src/mac/acoustic_interface.rs:316-378), for `n_links` links at once:
    push(records, payload) of each decode call -> the packets that complete now."""

    def __init__(self, engine, n_links=1):
        self._eng = engine
        h = C.c_void_p()
        engine._check(engine._lib.tmk_reasm_create(engine._h, int(n_links), C.byref(h)))
        self._r = h

    def push(self, recs, payload, pkt_cap=None, byte_cap=None):
        r = np.ascontiguousarray(recs, dtype=REC_DTYPE)
        pl = np.ascontiguousarray(payload, dtype=np.uint8)
        pkt_cap = int(pkt_cap) if pkt_cap is not None else max(16, len(r) + 16)
        byte_cap = int(byte_cap) if byte_cap is not None else max(1 << 18, 4 * pl.size + 64 * len(r))
        pk = np.zeros(pkt_cap, _abi.PKT_DTYPE)
        by = np.zeros(byte_cap, np.uint8)
        n, nb = C.c_uint64(0), C.c_uint64(0)
        self._eng._check(self._eng._lib.tmk_reasm_push(self._r, _ptr(r), len(r), _ptr(pl), pl.size, _ptr(pk), pkt_cap, C.byref(n),
                                                       _ptr(by), byte_cap, C.byref(nb)))
        return pk[:n.value].copy(), by[:nb.value].copy()

    def pending(self, link=0):
        nf, nb = C.c_uint32(0), C.c_uint64(0)
        self._eng._check(self._eng._lib.tmk_reasm_pending(self._r, int(link), C.byref(nf), C.byref(nb)))
        return int(nf.value), int(nb.value)

    def reset(self, link=0xFFFFFFFF):
        self._eng._check(self._eng._lib.tmk_reasm_reset(self._r, int(link)))

    def close(self):
        if self._r:
            self._eng._lib.tmk_reasm_destroy(self._r)
            self._r = None


def comm_unique_id():
    """ncclGetUniqueId through the library (rank 0 calls it and hands the 128 bytes to the other ranks)"""
    buf = C.create_string_buffer(128)
    rc = _abi.load_library().tmk_comm_unique_id(buf)
    if rc != 0:
        raise TmkError(rc, "tmk_comm_unique_id failed (libnccl not loadable?)")
    return buf.raw


def _blob(files):
    """list of bytes -> (uint8 array with one pad byte, uint64 offsets[n + 1])"""
    off = np.zeros(len(files) + 1, np.uint64)
    if files:
        off[1:] = np.cumsum([len(f) for f in files], dtype=np.uint64)
    return np.frombuffer(b"".join(bytes(f) for f in files) + b"\0", dtype=np.uint8).copy(), off


def wav_parse(data):
    """tmk_wav_parse -> dict (format, channels, sample_rate, bits_per_sample, block_align, data_off, n_frames)"""
    b = np.frombuffer(bytes(data) + b"\0", dtype=np.uint8)
    info = _abi.WavInfo()
    rc = _abi.load_library().tmk_wav_parse(_ptr(b), len(data), C.byref(info))
    if rc != 0:
        raise ValueError("not a supported WAVE file")
    return {k: int(getattr(info, k)) for k, _ in _abi.WavInfo._fields_ if k != "_pad"}


def wav_bytes(pcm16, sample_rate=SAMPLE_RATE, channels=1):
    """the file hound writes for dump_to_wav (src/utils/dump.rs:23-46): 44-byte PCM header + int16 little endian"""
    import struct
    d = np.ascontiguousarray(pcm16, dtype="<i2").tobytes()
    return (b"RIFF" + struct.pack("<I", 36 + len(d)) + b"WAVEfmt " +
            struct.pack("<IHHIIHH", 16, 1, channels, sample_rate, sample_rate * channels * 2, channels * 2, 16) +
            b"data" + struct.pack("<I", len(d)) + d)


def audio_data_from_json(text):
    """AudioData as dump_to_json writes it (src/utils/dump.rs:4-21): {sample_rate, audio_data, duration, channels}.
    serde_json reads the numbers as f64 and narrows to f32; so does this. Container parsing is host work."""
    import json
    d = json.loads(text)
    return {"sample_rate": int(d["sample_rate"]), "channels": int(d["channels"]), "duration": float(d["duration"]),
            "audio_data": np.asarray(d["audio_data"], dtype=np.float64).astype(np.float32)}


def make_config(kind, local_addr=1, spl=SAMPLES_PER_LEVEL, preamble_bytes=PREAMBLE_PATTERN_BYTES, threshold=0.9,
                max_frame_bytes=2 * MAX_FRAME_DATA_SIZE, chunk=0):
    return PhyConfig(int(kind), spl, preamble_bytes, local_addr, threshold, max_frame_bytes, chunk)


def make_streams(lengths, cfg, align=4):
    """Contiguous ragged layout: stream i starts at the next multiple of `align` samples."""
    st = np.zeros(len(lengths), STREAM_DTYPE)
    off = 0
    for i, n in enumerate(lengths):
        st[i]["offset"] = off
        st[i]["n_samples"] = n
        c = cfg[i] if isinstance(cfg, (list, tuple)) else cfg
        st[i]["cfg"] = (c.line_code, c.spl, c.preamble_bytes, c.local_addr, c.threshold, c.max_frame_bytes, c.chunk)
        off += (n + align - 1) // align * align
    return st, off


def records_to_frames(recs, payload):
    """Only FRAME_OK records are Frames PhyDecoder::process_samples returns (decoder.rs:341-344)."""
    out = []
    for r in recs:
        if r["status"] == FRAME_OK:
            o = int(r["payload_off"])
            out.append(Frame(FrameType(int(r["type"])), int(r["seq"]), int(r["src"]), int(r["dst"]),
                             bytes(payload[o:o + int(r["len"])])))
    return out


class PhyEncoder:
    """PhyEncoder::new(samples_per_level, preamble_bytes, line_coding_kind)  encoder.rs:18-39"""

    def __init__(self, engine, samples_per_level, preamble_bytes, line_coding_kind):
        self.engine = engine
        self.cfg = make_config(line_coding_kind, spl=samples_per_level, preamble_bytes=preamble_bytes)

    def preamble_len(self):  # encoder.rs:94-96
        return int(self.engine._lib.tmk_preamble_len(C.byref(self.cfg)))

    def encode_frame(self, frame):  # encoder.rs:43-63
        return self.engine.modulate(self.cfg, [frame], 0)

    def encode_frames(self, frames, inter_frame_gap_samples):  # encoder.rs:70-92
        return self.engine.modulate(self.cfg, list(frames), inter_frame_gap_samples)


class PhyDecoder:
    """PhyDecoder::new(samples_per_level, preamble_bytes, line_coding_kind, local_addr)  decoder.rs:36-65

    STREAMING, like the reference (decoder.rs:68-112): every process_samples call appends the slice to the
    decoder's buffer, runs search / decode_frame until more data is needed, and returns the frames that completed
    in this call; a frame still being gathered is carried to the next call. It is a bank of one link
    (tmk_links_*); feed many decoders at once through PhyDecoderBank. reset() is decoder.rs:114-119."""

    def __init__(self, engine, samples_per_level, preamble_bytes, line_coding_kind, local_addr):
        self.engine = engine
        self.cfg = make_config(line_coding_kind, local_addr=local_addr, spl=samples_per_level,
                               preamble_bytes=preamble_bytes)
        self._bank = PhyDecoderBank(engine, [self.cfg])
        self.last_records = np.zeros(0, REC_DTYPE)  # verdict records of the last call (positions from the first sample fed)
        self.last_payload = np.zeros(0, np.uint8)

    def process_samples(self, samples):
        s = np.ascontiguousarray(samples, dtype=np.float32)
        self.last_records, self.last_payload = self._bank.process_samples([s])
        return records_to_frames(self.last_records, self.last_payload)

    def reset(self):
        self._bank.reset()
        self.last_records = np.zeros(0, REC_DTYPE)
        self.last_payload = np.zeros(0, np.uint8)

    def state(self):
        return self._bank.state(0)

    def close(self):
        self._bank.close()


def decode_capture(engine, cfg, samples):
    """One-shot decode of a whole capture with a fresh decoder (what src/main.rs:540 does with one process_samples
    call); cfg.chunk reproduces the verdicts of feeding it in chunk-sized calls. -> (frames, records, payload);
    a frame the capture cuts off is a TMK_FRAME_TRUNCATED record."""
    s = np.ascontiguousarray(samples, dtype=np.float32)
    st, _ = make_streams([s.size], cfg)
    recs, pl = engine.decode_batch(s, st)
    return records_to_frames(recs, pl), recs, pl


class PhyDecoderBank:
    """n persistent PhyDecoders (decoder.rs:16-33) fed with process_samples calls, all decoded in one
    batch per call (tmk_links_*): the live cadence of src/mac/csma.rs:446-465 for many links at once."""

    def __init__(self, engine, cfgs):
        self.engine = engine
        self._lib = engine._lib
        self.cfgs = (PhyConfig * len(cfgs))(*cfgs)
        self.n = len(cfgs)
        h = C.c_void_p()
        engine._check(self._lib.tmk_links_create(engine._h, C.byref(self.cfgs) if self.n else None, self.n, C.byref(h)))
        self._h = h

    def close(self):
        if getattr(self, "_h", None):
            self._lib.tmk_links_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def process_samples(self, chunks, rec_cap=None, payload_cap=None):
        """chunks: one float32 array per link (may be empty). -> (records ndarray[REC_DTYPE], payload ndarray[u8]);
        records[i]["stream"] is the link index, positions are absolute per link."""
        assert len(chunks) == self.n
        lens = np.array([len(c) for c in chunks], np.uint64)
        offs = np.zeros(self.n, np.uint64)
        if self.n:
            offs[1:] = np.cumsum(lens)[:-1]
        total = int(lens.sum())
        s = np.zeros(max(total, 1), np.float32)
        for c, o in zip(chunks, offs):
            s[int(o):int(o) + len(c)] = c
        rec_cap = int(rec_cap) if rec_cap is not None else max(1024, total // 96 + 4 * self.n)
        payload_cap = int(payload_cap) if payload_cap is not None else max(65536, total // 12 + 1100 * self.n)
        recs = np.zeros(rec_cap, REC_DTYPE)
        pl = np.zeros(payload_cap, np.uint8)
        nr, npl = C.c_uint64(0), C.c_uint64(0)
        rc = self._lib.tmk_links_process_samples(self._h, _ptr(s), total, _ptr(offs), _ptr(lens), _ptr(recs), rec_cap,
                                                 C.byref(nr), _ptr(pl), payload_cap, C.byref(npl))
        if rc == _abi.E_CAPACITY and (nr.value > rec_cap or npl.value > payload_cap):
            return self.process_samples(chunks, max(rec_cap, nr.value), max(payload_cap, npl.value))
        self.engine._check(rc)
        return recs[:nr.value].copy(), pl[:npl.value].copy()

    def process_frames(self, chunks):
        """-> list (per link) of the Frames this call returned, like PhyDecoder::process_samples"""
        recs, pl = self.process_samples(chunks)
        return [records_to_frames(recs[recs["stream"] == i], pl) for i in range(self.n)]

    def reset(self, link=None):  # decoder.rs:114-119
        self.engine._check(self._lib.tmk_links_reset(self._h, 0xffffffff if link is None else int(link)))

    def state(self, link):
        b, o, d = C.c_uint64(0), C.c_uint64(0), C.c_int(0)
        self.engine._check(self._lib.tmk_links_state(self._h, int(link), C.byref(b), C.byref(o), C.byref(d)))
        return {"buffered": b.value, "buffer_offset": o.value, "decoding": bool(d.value)}


class IpFragmenter:
    """Mirror of IpFragmenter (src/net/fragmentation.rs:60-204); every call runs tmk_fragment_batch."""

    def __init__(self, engine, mtu):
        self._e, self.mtu, self._next = engine, int(mtu), 0

    def fragment_packets(self, packets):
        recs, by, self._next = self._e.fragment_batch(list(packets), self.mtu, self._next)
        out = [[] for _ in packets]
        for r in recs:
            if r["status"] != _abi.PKT_OK:
                raise ValueError(f"packet {int(r['link'])}: fragment_packet error {int(r['status'])}")
            out[int(r["link"])].append(bytes(by[int(r["data_off"]):int(r["data_off"]) + int(r["len"])]))
        return out

    def fragment_packet(self, packet):  # fragmentation.rs:85-203
        return self.fragment_packets([packet])[0]


def packets_of(pkts, data):
    """[(link, bytes)] of the TMK_PKT_OK records of Engine.reassemble_batch."""
    return [(int(p["link"]), bytes(data[int(p["data_off"]):int(p["data_off"]) + int(p["len"])])) for p in pkts
            if p["status"] == _abi.PKT_OK]
