"""ctypes view of include/tmkphy.h (struct layouts, constants, symbol table).

The shared library is the product; this file is only the binding a Python caller needs.
There is no fallback: if libtmkphy.so is missing, importing the library raises.
"""
import ctypes as C
import os

import numpy as np

LINE_MANCHESTER = 0
LINE_4B5B = 1
FRAME_TYPE_DATA = 1
FRAME_TYPE_ACK = 2

OK, E_INVALID, E_CUDA, E_CAPACITY, E_NOMEM = 0, -1, -2, -3, -4
NOALLOC_MAX_STREAMS = 4096  # TMK_NOALLOC_MAX_STREAMS
(FRAME_OK, FRAME_HDR_PARSE, FRAME_BAD_LEN, FRAME_SHORT_DECODE, FRAME_NOT_FOR_US, FRAME_CRC_FAIL,
 FRAME_TRUNCATED) = range(7)
STATUS_NAMES = ["OK", "HDR_PARSE", "BAD_LEN", "SHORT_DECODE", "NOT_FOR_US", "CRC_FAIL", "TRUNCATED"]
STAT_NAMES = ["samples", "detections", "frames_ok", "crc_fail", "hdr_fail", "not_for_us", "exact_replays",
              "candidates"]
STAT_COUNT = 8


class PhyConfig(C.Structure):
    _fields_ = [("line_code", C.c_int32), ("spl", C.c_uint32), ("preamble_bytes", C.c_uint32),
                ("local_addr", C.c_int32), ("threshold", C.c_float), ("max_frame_bytes", C.c_uint32),
                ("chunk", C.c_uint64)]


class Stream(C.Structure):
    _fields_ = [("offset", C.c_uint64), ("n_samples", C.c_uint64), ("cfg", PhyConfig)]


class FrameRec(C.Structure):
    _fields_ = [("stream", C.c_uint32), ("status", C.c_uint32), ("preamble_start", C.c_uint64),
                ("frame_start", C.c_uint64), ("payload_off", C.c_uint64), ("corr", C.c_float),
                ("consumed", C.c_uint32), ("len", C.c_uint16), ("crc", C.c_uint8), ("type", C.c_uint8),
                ("seq", C.c_uint8), ("src", C.c_uint8), ("dst", C.c_uint8), ("hdr_valid", C.c_uint8)]


class FrameDesc(C.Structure):
    _fields_ = [("payload_off", C.c_uint64), ("len", C.c_uint32), ("type", C.c_uint8), ("seq", C.c_uint8),
                ("src", C.c_uint8), ("dst", C.c_uint8)]


class SynthParams(C.Structure):
    _fields_ = [("snr_db", C.c_float), ("amplitude", C.c_float), ("multipath", C.c_uint32), ("drift_ppm", C.c_float),
                ("gap_min", C.c_uint32), ("gap_max", C.c_uint32), ("amp_min", C.c_float), ("_pad", C.c_uint32)]


class WavInfo(C.Structure):
    _fields_ = [("format", C.c_uint32), ("channels", C.c_uint32), ("sample_rate", C.c_uint32),
                ("bits_per_sample", C.c_uint32), ("block_align", C.c_uint32), ("_pad", C.c_uint32),
                ("data_off", C.c_uint64), ("n_frames", C.c_uint64)]


assert C.sizeof(PhyConfig) == 32 and C.sizeof(Stream) == 48 and C.sizeof(FrameRec) == 48
assert C.sizeof(FrameDesc) == 16 and C.sizeof(SynthParams) == 32 and C.sizeof(WavInfo) == 40

# numpy mirrors (zero-copy views of record arrays)
CFG_DTYPE = np.dtype([("line_code", "<i4"), ("spl", "<u4"), ("preamble_bytes", "<u4"), ("local_addr", "<i4"),
                      ("threshold", "<f4"), ("max_frame_bytes", "<u4"), ("chunk", "<u8")])
STREAM_DTYPE = np.dtype([("offset", "<u8"), ("n_samples", "<u8"), ("cfg", CFG_DTYPE)])
REC_DTYPE = np.dtype([("stream", "<u4"), ("status", "<u4"), ("preamble_start", "<u8"), ("frame_start", "<u8"),
                      ("payload_off", "<u8"), ("corr", "<f4"), ("consumed", "<u4"), ("len", "<u2"),
                      ("crc", "u1"), ("type", "u1"), ("seq", "u1"), ("src", "u1"), ("dst", "u1"),
                      ("hdr_valid", "u1")])
DESC_DTYPE = np.dtype([("payload_off", "<u8"), ("len", "<u4"), ("type", "u1"), ("seq", "u1"), ("src", "u1"),
                       ("dst", "u1")])
ASK_REC_DTYPE = np.dtype([("stream", "<u4"), ("accepted", "<u4"), ("start_index", "<u8"), ("sync_power", "<f4"),
                          ("id", "u1"), ("bits", "u1", (14,)), ("_pad", "u1", (5,))])
PKT_DTYPE = np.dtype([("link", "<u4"), ("status", "<u4"), ("frame", "<u4"), ("len", "<u4"), ("data_off", "<u8"),
                      ("identification", "<u2"), ("n_fragments", "<u2"), ("_pad", "<u4")])
assert PKT_DTYPE.itemsize == 32
PKT_OK, PKT_E_SHORT, PKT_E_IHL, PKT_E_TOO_MANY, PKT_E_MTU = range(5)
MAX_FRAGMENTS = 512
assert STREAM_DTYPE.itemsize == 48 and REC_DTYPE.itemsize == 48 and DESC_DTYPE.itemsize == 16
assert ASK_REC_DTYPE.itemsize == 40
MCM_REC_DTYPE = np.dtype([("stream", "<u4"), ("status", "<u4"), ("start_index", "<u8"), ("payload_off", "<u8"),
                          ("sync_power", "<f4"), ("seq", "u1"), ("len", "u1"), ("crc", "u1"), ("mode", "u1")])
assert MCM_REC_DTYPE.itemsize == 32
MCM_FSK, MCM_OFDM = 0, 1
MCM_MAX_PAYLOAD = {MCM_FSK: 25, MCM_OFDM: 100}

_vp, _u64, _u32, _i32 = C.c_void_p, C.c_uint64, C.c_uint32, C.c_int32
_pu64 = C.POINTER(C.c_uint64)
_pcfg = C.POINTER(PhyConfig)

# every symbol include/tmkphy.h declares: name -> (restype, argtypes)
SYMBOLS = {
    "tmk_abi_version": (C.c_int, []),
    "tmk_engine_create": (C.c_int, [C.c_int, C.POINTER(_vp)]),
    "tmk_engine_destroy": (None, [_vp]),
    "tmk_last_error": (C.c_char_p, [_vp]),
    "tmk_default_config": (None, [_i32, _i32, _pcfg]),
    "tmk_samples_for_bits": (_u64, [_pcfg, _u64]),
    "tmk_preamble_len": (_u64, [_pcfg]),
    "tmk_modulated_len": (_u64, [_pcfg, _vp, _u32, _u32]),
    "tmk_modulate": (C.c_int, [_vp, _pcfg, _vp, _u32, _vp, _u64, _u32, _vp, _u64, _pu64]),
    "tmk_decode_batch": (C.c_int, [_vp, _vp, _u64, _vp, _u32, _vp, _u64, _pu64, _vp, _u64, _pu64]),
    "tmk_correlate": (C.c_int, [_vp, _vp, _u64, _vp, _u32, _vp]),
    "tmk_scan_probe": (C.c_int, [_vp, _vp, _u64, _vp, _u32, _vp, _vp, _u64, _pu64, _vp]),
    "tmk_scan_block": (_u32, []),
    "tmk_scan_dot_probe": (C.c_int, [_vp, _vp, _u64, _vp, _u32, _vp]),
    "tmk_replay_probe": (C.c_int, [_vp, _vp, _u64, _pcfg, _u64, _vp, _u32, _vp, _pu64]),
    "tmk_channel_busy": (C.c_int, [_vp, _vp, _u64, _vp, _u32, _vp]),
    "tmk_channel_busy_noalloc": (C.c_int, [_vp, _vp, _u64, _vp, _u32, _vp]),
    "tmk_stats": (C.c_int, [_vp, _pu64]),
    "tmk_stats_reset": (C.c_int, [_vp]),
    "tmk_debug_counters": (C.c_int, [_vp, _pu64]),
    "tmk_device_alloc": (C.c_int, [_vp, _u64, C.POINTER(_vp)]),
    "tmk_device_free": (C.c_int, [_vp, _vp]),
    "tmk_host_alloc_pinned": (C.c_int, [_vp, _u64, C.POINTER(_vp)]),
    "tmk_host_free_pinned": (C.c_int, [_vp, _vp]),
    "tmk_memcpy_h2d": (C.c_int, [_vp, _vp, _vp, _u64]),
    "tmk_memcpy_d2h": (C.c_int, [_vp, _vp, _vp, _u64]),
    "tmk_synchronize": (C.c_int, [_vp]),
    "tmk_engine_stream": (_vp, [_vp]),
    "tmk_last_kernel_ms": (C.c_int, [_vp, C.POINTER(C.c_float)]),
    "tmk_last_stage_ms": (C.c_int, [_vp, C.POINTER(C.c_float)]),
    "tmk_timer_start": (C.c_int, [_vp]),
    "tmk_timer_stop": (C.c_int, [_vp, C.POINTER(C.c_float)]),
    "tmk_device_pci_bus_id": (C.c_int, [_vp, C.c_char_p, C.c_int]),
    "tmk_launch_count": (_u64, [_vp]),
    "tmk_links_create": (C.c_int, [_vp, _vp, _u32, C.POINTER(_vp)]),
    "tmk_links_destroy": (None, [_vp]),
    "tmk_links_process_samples": (C.c_int, [_vp, _vp, _u64, _vp, _vp, _vp, _u64, _pu64, _vp, _u64, _pu64]),
    "tmk_links_reset": (C.c_int, [_vp, _u32]),
    "tmk_links_state": (C.c_int, [_vp, _u32, _pu64, _pu64, C.POINTER(C.c_int)]),
    "tmk_ask_tables": (C.c_int, [_vp, _vp, _vp, _u64]),
    "tmk_ask_modulated_len": (_u64, [_u32, _vp, _vp]),
    "tmk_ask_modulate": (C.c_int, [_vp, _vp, _u32, _vp, _vp, _vp, _u64, _pu64]),
    "tmk_ask_decode_batch": (C.c_int, [_vp, _vp, _u64, _vp, _u32, _vp, _u64, _pu64]),
    "tmk_ask_last_stage_ms": (C.c_int, [_vp, C.POINTER(C.c_float)]),
    "tmk_mcm_modulated_len": (_u64, [_u32, _vp, _vp]),
    "tmk_mcm_modulate": (C.c_int, [_vp, _i32, _vp, _vp, _vp, _u32, _vp, _vp, _vp, _u64, _pu64]),
    "tmk_mcm_decode_batch": (C.c_int, [_vp, _i32, _vp, _u64, _vp, _u32, _vp, _u64, _pu64, _vp, _u64, _pu64]),
    "tmk_mcm_synth_batch": (C.c_int, [_vp, _i32, _u64, _u32, _u64, C.c_float, _i32, _vp, _pu64]),
    "tmk_mcm_fft_probe": (C.c_int, [_vp, _u32, _vp, _u32, _vp]),
    "tmk_mcm_last_stage_ms": (C.c_int, [_vp, C.POINTER(C.c_float)]),
    "tmk_pcm16_encode": (C.c_int, [_vp, _vp, _u64, _vp]),
    "tmk_pcm16_to_f32": (C.c_int, [_vp, _vp, _u64, _vp]),
    "tmk_decode_batch_pcm16": (C.c_int, [_vp, _vp, _u64, _vp, _u32, _vp, _u64, _pu64, _vp, _u64, _pu64]),
    "tmk_reassemble_batch": (C.c_int, [_vp, _vp, _u64, _vp, _u64, _u32, _vp, _u64, _pu64, _vp, _u64, _pu64]),
    "tmk_fragment_batch": (C.c_int, [_vp, _vp, _vp, _u32, _u32, C.c_uint16, _vp, _u64, _pu64, _vp, _u64, _pu64,
                                     C.POINTER(C.c_uint16)]),
    "tmk_reasm_create": (C.c_int, [_vp, _u32, C.POINTER(_vp)]),
    "tmk_reasm_destroy": (None, [_vp]),
    "tmk_reasm_push": (C.c_int, [_vp, _vp, _u64, _vp, _u64, _vp, _u64, _pu64, _vp, _u64, _pu64]),
    "tmk_reasm_pending": (C.c_int, [_vp, _u32, C.POINTER(_u32), _pu64]),
    "tmk_reasm_reset": (C.c_int, [_vp, _u32]),
    "tmk_synth_batch": (C.c_int, [_vp, _pcfg, _u64, _u32, _u64, C.c_float, C.c_float, _vp, _vp, _u64, _pu64]),
    "tmk_synth_batch_ex": (C.c_int, [_vp, _pcfg, _u64, _u32, _u64, C.POINTER(SynthParams), _vp, _vp, _u64, _pu64]),
    "tmk_ask_synth_batch": (C.c_int, [_vp, _u64, _u32, _u64, C.c_float, _vp, _pu64]),
    "tmk_wav_parse": (C.c_int, [_vp, _u64, C.POINTER(WavInfo)]),
    "tmk_wav_to_f32": (C.c_int, [_vp, _vp, _u64, _vp, _u32, _vp, _u64, _vp, _vp, _pu64]),
    "tmk_decode_batch_wav": (C.c_int, [_vp, _vp, _u64, _vp, _u32, _vp, _vp, _vp, _u64, _pu64, _vp, _u64, _pu64]),
    "tmk_comm_unique_id": (C.c_int, [_vp]),
    "tmk_comm_init": (C.c_int, [_vp, _vp, C.c_int, C.c_int]),
    "tmk_comm_size": (C.c_int, [_vp]),
    "tmk_stats_allreduce": (C.c_int, [_vp, _pu64]),
    "tmk_comm_allreduce_sum_u64": (C.c_int, [_vp, _pu64, _u32]),
    "tmk_comm_allreduce_max_f64": (C.c_int, [_vp, C.POINTER(C.c_double), _u32]),
    "tmk_comm_allgather_u64": (C.c_int, [_vp, _pu64, _u32, _pu64]),
    "tmk_comm_barrier": (C.c_int, [_vp]),
    "tmk_comm_destroy": (C.c_int, [_vp]),
}

PKG_DIR = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.environ.get("TMKPHY_LIB") or os.path.join(PKG_DIR, "libtmkphy.so")

_lib = None


def load_library(path=None):
    """dlopen libtmkphy.so and bind every declared symbol. Raises if anything is missing."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or LIB_PATH
    if not os.path.exists(p):
        raise ImportError(f"libtmkphy.so not found at {p}: run `python __graft_entry__.py build` "
                          "(there is no CPU fallback)")
    lib = C.CDLL(p)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    if lib.tmk_abi_version() != 1:
        raise ImportError("libtmkphy ABI version mismatch")
    if path is None:
        _lib = lib
    return lib
