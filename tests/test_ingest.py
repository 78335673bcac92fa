"""Capture container parsing that needs no GPU: the RIFF chunk walk of tmk_wav_parse against files as
hound writes them for dump_to_wav (src/utils/dump.rs:23-46), the JSON AudioData reader (dump.rs:4-21), and the
mechanical check of the ABI struct layouts: include/tmkphy.h (compiled by gcc) == ctypes mirror == the #[repr(C)]
listings of INTEGRATION.md."""
import ctypes as C
import json
import os
import re
import struct
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def riff(fmt_chunk, data, extra=b""):
    body = b"WAVE" + b"fmt " + struct.pack("<I", len(fmt_chunk)) + fmt_chunk + extra + b"data" + struct.pack("<I", len(data)) + data
    return b"RIFF" + struct.pack("<I", len(body)) + body


def pcm_fmt(ch, rate, bits, tag=1):
    return struct.pack("<HHIIHH", tag, ch, rate, rate * ch * bits // 8, ch * bits // 8, bits)


def test_wav_parse_hound_header():
    import tmkphy
    pcm = np.arange(-5, 6, dtype=np.int16)
    f = tmkphy.wav_bytes(pcm, 48000, 1)
    assert len(f) == 44 + 22  # the canonical 44-byte header dump_to_wav's files carry
    info = tmkphy.wav_parse(f)
    assert info == {"format": 1, "channels": 1, "sample_rate": 48000, "bits_per_sample": 16, "block_align": 2,
                    "data_off": 44, "n_frames": 11}


def test_wav_parse_variants_and_errors(orc):
    import tmkphy
    d = bytes(range(48))
    # stereo 16 bit, a LIST chunk of odd size (word padding) in front of the data
    f = riff(pcm_fmt(2, 44100, 16), d, extra=b"LIST" + struct.pack("<I", 5) + b"abcde\0")
    info = tmkphy.wav_parse(f)
    assert (info["channels"], info["block_align"], info["n_frames"], info["data_off"]) == (2, 4, 12, 12 + 24 + 14 + 8)
    # WAVE_FORMAT_EXTENSIBLE (hound writes it for > 2 channels or > 16 bits): 40-byte fmt chunk, sub-format PCM
    ext = pcm_fmt(1, 48000, 24, tag=0xFFFE) + struct.pack("<HHI", 22, 24, 4) + struct.pack("<H", 1) + b"\0" * 14
    info = tmkphy.wav_parse(riff(ext, d))
    assert (info["format"], info["bits_per_sample"], info["n_frames"]) == (1, 24, 16)
    # float32
    info = tmkphy.wav_parse(riff(pcm_fmt(1, 48000, 32, tag=3), d))
    assert (info["format"], info["n_frames"]) == (3, 12)
    # a writer that was never finalised: the data chunk claims more than the file holds
    f = riff(pcm_fmt(1, 48000, 16), d)
    cut = f[:-10]
    assert tmkphy.wav_parse(cut)["n_frames"] == 19
    for bad in (b"", b"RIFF", f[:20], b"RIFX" + f[4:], f[:8] + b"AVI " + f[12:], riff(pcm_fmt(1, 48000, 12), d),
                riff(pcm_fmt(1, 48000, 16, tag=2), d), b"RIFF" + struct.pack("<I", 4) + b"WAVE",
                riff(pcm_fmt(0, 48000, 16), d)):
        with pytest.raises(ValueError):
            tmkphy.wav_parse(bad)
    # the oracle's numpy reader agrees on what the samples are
    x = orc.wav_to_f32(tmkphy.wav_bytes(np.array([0, 1, -1, 32767, -32768], np.int16)))
    assert np.array_equal(x, np.array([0, 1 / 32768, -1 / 32768, 32767 / 32768, -1.0], np.float32))


def test_audio_data_json():
    import tmkphy
    vals = [0.0, 1.0, -1.0, 0.1, 0.30000001192092896, 1e-7, 3.4028234663852886e+38]
    text = json.dumps({"sample_rate": 48000, "audio_data": vals, "duration": 1.5, "channels": 1}, indent=2)
    a = tmkphy.audio_data_from_json(text)
    assert a["sample_rate"] == 48000 and a["channels"] == 1 and a["audio_data"].dtype == np.float32
    assert np.array_equal(a["audio_data"], np.array(vals, np.float64).astype(np.float32))


# ---------------------------------------------------------------- ABI struct layouts, checked mechanically
STRUCTS = ["tmk_phy_config", "tmk_stream", "tmk_frame_rec", "tmk_frame_desc", "tmk_ask_rec", "tmk_packet_rec",
           "tmk_synth_params", "tmk_wav_info", "tmk_mcm_rec"]


def c_layouts(tmp_path):
    """sizeof / offsetof of every field of every ABI struct, from gcc compiling include/tmkphy.h"""
    hdr = open(os.path.join(ROOT, "include", "tmkphy.h")).read()
    fields = {}
    for name in STRUCTS:
        m = re.search(r"typedef struct %s \{(.*?)\} %s;" % (name, name), hdr, re.S)
        assert m, name
        body = re.sub(r"/\*.*?\*/", "", m.group(1), flags=re.S)
        fl = []
        for decl in body.split(";"):
            decl = decl.strip()
            if not decl:
                continue
            for part in decl.split(",")[0:1] + [p for p in decl.split(",")[1:]]:
                nm = re.sub(r"\[.*?\]", "", part.strip().split()[-1])
                fl.append(nm)
        fields[name] = fl
    src = ['#include <stddef.h>', '#include <stdio.h>', '#include "tmkphy.h"', "int main(void) {"]
    for name, fl in fields.items():
        src.append(f'printf("{name} %zu\\n", sizeof({name}));')
        for f in fl:
            src.append(f'printf("{name}.{f} %zu\\n", offsetof({name}, {f}));')
    src.append("return 0; }")
    c = tmp_path / "layout.c"
    c.write_text("\n".join(src))
    exe = tmp_path / "layout"
    subprocess.check_call(["gcc", "-std=c11", "-I", os.path.join(ROOT, "include"), str(c), "-o", str(exe)])
    out = subprocess.check_output([str(exe)], text=True)
    return {k: int(v) for k, v in (ln.split() for ln in out.splitlines())}, fields


RUST_TYPES = {"u8": (1, 1), "i8": (1, 1), "u16": (2, 2), "i16": (2, 2), "u32": (4, 4), "i32": (4, 4), "f32": (4, 4),
              "u64": (8, 8), "i64": (8, 8), "f64": (8, 8)}


def rust_layouts(known):
    """#[repr(C)] layout of every `pub struct` in INTEGRATION.md, computed from its field list"""
    text = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    out = {}
    for m in re.finditer(r"pub struct (tmk_\w+) \{(.*?)\n?\}", text, re.S):
        name, body = m.group(1), re.sub(r"//[^\n]*", "", m.group(2))
        if "_private" in body:
            continue
        off, align_max, lay = 0, 1, {}
        for fm in re.finditer(r"pub (?:r#)?(\w+): ([^,]+)", body):
            fname, ty = fm.group(1), fm.group(2).strip()
            am = re.match(r"\[(\w+); (\d+)\]", ty)
            if am:
                sz, al = RUST_TYPES[am.group(1)]
                sz *= int(am.group(2))
            elif ty in RUST_TYPES:
                sz, al = RUST_TYPES[ty]
            else:
                sz, al = known[ty]
            off = (off + al - 1) // al * al
            lay[fname] = off
            off += sz
            align_max = max(align_max, al)
        size = (off + align_max - 1) // align_max * align_max
        known[name] = (size, align_max)
        out[name] = (size, lay)
    return out


def test_abi_struct_layouts(tmp_path):
    from tmkphy import _abi
    cl, fields = c_layouts(tmp_path)
    # header == ctypes / numpy mirrors
    mirrors = {"tmk_phy_config": _abi.PhyConfig, "tmk_stream": _abi.Stream, "tmk_frame_rec": _abi.FrameRec,
               "tmk_frame_desc": _abi.FrameDesc, "tmk_synth_params": _abi.SynthParams, "tmk_wav_info": _abi.WavInfo}
    for name, cls in mirrors.items():
        assert C.sizeof(cls) == cl[name], name
        for f in fields[name]:
            assert getattr(cls, f).offset == cl[f"{name}.{f}"], (name, f)
    for name, dt in (("tmk_frame_rec", _abi.REC_DTYPE), ("tmk_stream", _abi.STREAM_DTYPE), ("tmk_frame_desc", _abi.DESC_DTYPE),
                     ("tmk_ask_rec", _abi.ASK_REC_DTYPE), ("tmk_packet_rec", _abi.PKT_DTYPE),
                     ("tmk_mcm_rec", _abi.MCM_REC_DTYPE)):
        assert dt.itemsize == cl[name], name
        for f in fields[name]:
            assert dt.fields[f][1] == cl[f"{name}.{f}"], (name, f)
    # header == the #[repr(C)] listings of INTEGRATION.md (the Rust shim cannot be compiled here: no rustc)
    rl = rust_layouts({})
    for name in STRUCTS:
        assert name in rl, f"INTEGRATION.md lacks a #[repr(C)] listing of {name}"
        size, lay = rl[name]
        assert size == cl[name], (name, size, cl[name])
        assert sorted(lay) == sorted(fields[name]), (name, sorted(lay), sorted(fields[name]))
        for f, o in lay.items():
            assert o == cl[f"{name}.{f}"], (name, f, o)
