"""The C-ABI library loads on a CPU-only box and exports exactly what include/tmkphy.h declares."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "tmkphy.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(tmk_[a-z0-9_]+)\s*\(", src)))


def test_header_and_bindings_agree():
    from tmkphy import _abi
    assert _declared() == sorted(_abi.SYMBOLS)


def test_library_exports_every_symbol():
    import __graft_entry__ as ge
    ge.build()
    from tmkphy import _abi
    lib = _abi.load_library()
    for name in _declared():
        assert hasattr(lib, name), name
    assert lib.tmk_abi_version() == 1


def test_struct_layout_matches_header():
    from tmkphy import _abi
    assert C.sizeof(_abi.PhyConfig) == 32 and C.sizeof(_abi.Stream) == 48
    assert C.sizeof(_abi.FrameRec) == 48 and C.sizeof(_abi.FrameDesc) == 16
    assert _abi.REC_DTYPE.fields["corr"][1] == _abi.FrameRec.corr.offset
    assert _abi.REC_DTYPE.fields["hdr_valid"][1] == _abi.FrameRec.hdr_valid.offset


def test_pure_host_helpers():
    """samples_for_bits / preamble_len / modulated_len need no GPU (line_coding.rs:125-127,295-300)."""
    import tmkphy
    from tmkphy import _abi
    lib = _abi.load_library()
    for kind, p, h in ((0, 96, 336), (1, 60, 210)):
        cfg = tmkphy.make_config(kind)
        assert lib.tmk_preamble_len(C.byref(cfg)) == p
        assert lib.tmk_samples_for_bits(C.byref(cfg), 56) == h
    cfg = _abi.PhyConfig()
    lib.tmk_default_config(1, 7, C.byref(cfg))
    assert (cfg.line_code, cfg.spl, cfg.preamble_bytes, cfg.local_addr, cfg.max_frame_bytes, cfg.chunk) == (1, 3, 2, 7, 256, 0)
    assert abs(cfg.threshold - 0.9) < 1e-7


def test_no_cpu_fallback():
    """Without a GPU the engine must refuse to exist (the product path never computes on the CPU)."""
    import tmkphy
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = os.path.exists("/dev/nvidia0")
    if has_gpu:
        pytest.skip("GPU present")
    with pytest.raises(tmkphy.TmkError):
        tmkphy.Engine(0)
