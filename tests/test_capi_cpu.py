"""CPU: the C-ABI library loads and exports every symbol include/grb200.h declares; without a GPU every
compute entry point fails loudly (no CPU fallback); the host-side block contract mirrors the reference."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from conftest import ROOT, has_gpu


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "grb200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(grb200_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported_and_bound():
    from gnuradio_b200 import _lib

    syms = declared_symbols()
    assert len(syms) >= 45
    lib = C.CDLL(_lib.LIB_PATH)
    for s in syms:
        assert hasattr(lib, s), "libgrb200.so does not export " + s
    assert sorted(_lib.SIGNATURES) == syms, "ctypes binding and header disagree"
    out = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], capture_output=True, text=True).stdout
    exported = set(re.findall(r" T (grb200_[a-z0-9_]+)", out))
    assert set(syms) <= exported


def test_library_is_sm100a_only():
    from gnuradio_b200 import _lib

    out = subprocess.run(["cuobjdump", "-lelf", _lib.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs


def test_cxx_block_wrappers_link():
    drv = os.path.join(ROOT, "gnuradio_b200", "cxx", "libgrb200_blocks.so")
    lib = C.CDLL(drv)
    for s in ("grbx_fir_run", "grbx_fft_filter_run", "grbx_fft_vcc_run", "grbx_magsq_run", "grbx_pfb_run", "grbx_what",
              "grbx_fir_fff_run", "grbx_fft_vfc_run", "grbx_pfb_synth_run", "grbx_filterbank_run"):
        assert hasattr(lib, s)


@pytest.mark.skipif(has_gpu(), reason="checks the no-GPU behaviour")
def test_fails_loudly_without_gpu():
    import gnuradio_b200 as grb

    assert grb.device_count() == 0
    with pytest.raises(RuntimeError):
        grb.filter.fir_filter_ccf(1, [1.0, 2.0])
    with pytest.raises(RuntimeError):
        grb.fft.fft_vcc(64, True, [], False)
    with pytest.raises(RuntimeError):
        grb.filter.pfb_channelizer_ccf(4, np.ones(8, np.float32), 1.0)
    with pytest.raises(RuntimeError):
        grb.filter.fft_filter_ccc(1, [1 + 0j])
    with pytest.raises(RuntimeError):
        grb.blocks.complex_to_mag_squared().process(np.ones(4, np.complex64))
    # argument errors are still the reference's exception types, checked before any CUDA call
    with pytest.raises(ValueError):
        grb.filter.fir_filter_ccf(0, [1.0])
    with pytest.raises(ValueError):
        grb.filter.pfb_channelizer_ccf(36, np.ones(8, np.float32), 10.1334)
    with pytest.raises(IndexError):
        grb.fft.fft_vcc(0, True, [], False)
    # C++ wrappers: same exception classes (-1 invalid_argument, -2 out_of_range, -3 runtime_error)
    lib = C.CDLL(os.path.join(ROOT, "gnuradio_b200", "cxx", "libgrb200_blocks.so"))
    lib.grbx_fir_run.restype = C.c_longlong
    lib.grbx_what.restype = C.c_char_p
    t = np.ones(4, np.float32)
    x = np.zeros(64, np.complex64)
    y = np.zeros(64, np.complex64)
    rc = lib.grbx_fir_run(0, 1, t.ctypes.data_as(C.c_void_p), 4, x.ctypes.data_as(C.c_void_p),
                          y.ctypes.data_as(C.c_void_p), C.c_longlong(10), 4096, 0)
    assert rc == -3 and b"CUDA" in lib.grbx_what() or b"cuda" in lib.grbx_what()
    rc = lib.grbx_fir_run(0, 0, t.ctypes.data_as(C.c_void_p), 4, x.ctypes.data_as(C.c_void_p),
                          y.ctypes.data_as(C.c_void_p), C.c_longlong(10), 4096, 0)
    assert rc == -1


def test_product_never_touches_oracle():
    """The product path must not import, link or execute anything under oracle/."""
    pkg = os.path.join(ROOT, "gnuradio_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cc", ".hpp", ".h")) or f == "Makefile":
                text = open(os.path.join(dp, f), errors="replace").read()
                for pat in (r"^\s*(from|import)\s+oracle", r"pyoracle", r"libgrb_oracle", r"libgrref", r"grbo_", r"grref_",
                            r"oracle/_ref", r"#include\s+[<\"].*oracle"):
                    assert not re.search(pat, text, flags=re.M), (os.path.join(dp, f), pat)
    out = subprocess.run(["ldd", os.path.join(pkg, "libgrb200.so")], capture_output=True, text=True).stdout
    assert "grb_oracle" not in out and "grref" not in out


def test_runtime_contract_mirror():
    from gnuradio_b200 import runtime as rt

    d = rt.sync_decimator("d", rt.io_signature.make(1, 1, 8), rt.io_signature.make(1, 1, 8), 4)
    d.set_history(64)
    assert d.relative_rate() == 0.25 and d.fixed_rate()
    # sync_decimator.cc:28-43
    assert d.fixed_rate_noutput_to_ninput(100) == 100 * 4 + 63
    assert d.fixed_rate_ninput_to_noutput(100 * 4 + 63) == 100
    assert d.forecast(10, 1) == [10 * 4 + 63]
    with pytest.raises(ValueError):
        d.set_output_multiple(0)
    b = rt.block("b", rt.io_signature.make(2, 2, 8), rt.io_signature.make(1, 2, 8))
    assert not b.fixed_rate() and b.input_signature().max_streams() == 2
    assert rt.WORK_DONE == -1 and rt.WORK_CALLED_PRODUCE == -2
