"""GPU: the C++ block classes of include/grb200/blocks.hpp (make/set_taps/work, the host side a GNU Radio
maintainer compiles) driven in scheduler-sized chunks through gnuradio_b200/cxx/libgrb200_blocks.so."""
import ctypes as C
import os

import numpy as np
import pytest

from common import assert_parity, crand
from conftest import ROOT

pytestmark = pytest.mark.gpu
LL = C.c_longlong


@pytest.fixture(scope="module")
def drv():
    lib = C.CDLL(os.path.join(ROOT, "gnuradio_b200", "cxx", "libgrb200_blocks.so"))
    for f in ("grbx_fir_run", "grbx_freq_xlating_run", "grbx_pfb_dec_run", "grbx_interp_fir_run", "grbx_rational_resampler_run", "grbx_fft_filter_run", "grbx_fft_vcc_run", "grbx_magsq_run", "grbx_pfb_run", "grbx_fir_fff_run", "grbx_fft_vfc_run", "grbx_pfb_synth_run", "grbx_filterbank_run", "grbx_fir_lag_run"):
        getattr(lib, f).restype = LL
    lib.grbx_what.restype = C.c_char_p
    return lib


def p(a):
    return a.ctypes.data_as(C.c_void_p)


def test_cxx_fir(drv, orc):
    rng = np.random.default_rng(1)
    x = crand(rng, 30000)
    for ct, nt, D in ((0, 64, 1), (0, 1024, 8), (1, 33, 2)):
        taps = crand(rng, nt) if ct else rng.uniform(-1, 1, nt).astype(np.float32)
        xin = np.concatenate([np.zeros(nt - 1, np.complex64), x])
        nout = len(x) // D
        out = np.zeros(nout, np.complex64)
        rc = drv.grbx_fir_run(ct, D, p(taps), nt, p(xin), p(out), LL(nout), 4096, 0 if ct else 1)
        assert rc == nout, drv.grbx_what()
        assert_parity(out, orc.fir(D, taps, xin, nout), "c++ fir ct=%d %d/%d" % (ct, nt, D))
    rc = drv.grbx_fir_run(0, 0, p(taps), 4, p(xin), p(out), LL(10), 4096, 0)
    assert rc == -1  # std::invalid_argument


def test_cxx_freq_xlating(drv, orc):
    rng = np.random.default_rng(6)
    for ct, nt, D in ((0, 64, 1), (1, 33, 4)):
        nout = 4000
        x = crand(rng, nout * D)
        taps = crand(rng, nt) if ct else rng.uniform(-1, 1, nt).astype(np.float32)
        xin = np.concatenate([np.zeros(nt - 1, np.complex64), x])
        out = np.zeros(nout, np.complex64)
        rc = drv.grbx_freq_xlating_run(ct, D, p(taps), nt, C.c_double(-3e3), C.c_double(48e3), p(xin), p(out), LL(nout), 1024)
        assert rc == nout, drv.grbx_what()
        ref, _ = orc.freq_xlating_fir(D, taps, -3e3, 48e3, xin, nout, chunk_out=1024)
        from test_gpu_freq_xlating import assert_vs_reference  # tolerance explained there

        assert_vs_reference(out, ref, "c++ freq_xlating ct=%d" % ct)
    rc = drv.grbx_freq_xlating_run(0, 0, p(taps), 4, C.c_double(0), C.c_double(1), p(xin), p(out), LL(10), 1024)
    assert rc == -1  # std::invalid_argument


def test_cxx_fft_filter(drv, orc):
    rng = np.random.default_rng(2)
    x, taps = crand(rng, 50000), crand(rng, 200)
    ns = C.c_int(0)
    nout = (len(x) // 313) * 313
    out = np.zeros(nout, np.complex64)
    rc = drv.grbx_fft_filter_run(1, p(taps), 200, p(x), p(out), LL(nout), 4096, C.byref(ns))
    assert ns.value == 313 and rc == nout, drv.grbx_what()
    assert_parity(out, orc.fft_filter_ccc(1, taps, x), "c++ fft_filter")


def test_cxx_fft_vcc_and_magsq(drv, orc):
    rng = np.random.default_rng(3)
    N, nvec = 4096, 50
    x = crand(rng, N * nvec)
    w = orc.window(5, N)
    out = np.zeros(N * nvec, np.complex64)
    rc = drv.grbx_fft_vcc_run(N, 1, p(w), N, 1, p(x), p(out), LL(nvec), 1)  # one vector per work(), as in SURVEY 3.4
    assert rc == nvec, drv.grbx_what()
    assert_parity(out, orc.fft_vcc(N, 1, w, 1, x), "c++ fft_vcc")
    assert drv.grbx_fft_vcc_run(N, 1, p(w), N - 1, 1, p(x), p(out), LL(nvec), 1) == -3  # std::runtime_error
    assert drv.grbx_fft_vcc_run(0, 1, p(w), 0, 1, p(x), p(out), LL(nvec), 1) == -2      # std::out_of_range
    m = np.zeros(N * nvec, np.float32)
    assert drv.grbx_magsq_run(N, p(out), p(m), LL(nvec), 8) == nvec
    assert_parity(m, orc.mag_squared(out), "c++ mag^2", 1e-6, 1e-5)


def test_cxx_pfb(drv, orc):
    rng = np.random.default_rng(4)
    for M, nt, osr, nouts in ((64, 1024, 1.0, 64), (8, 50, 2.0, 5)):
        taps = rng.uniform(-1, 1, nt).astype(np.float32)
        T = orc.pfb_taps_per_filter(M, nt)
        cons = 4000
        nout = int(cons * osr)
        ins = np.concatenate([np.zeros((M, T), np.complex64), crand(rng, M * cons).reshape(M, cons)], axis=1)
        outs = np.zeros((nouts, nout), np.complex64)
        cmap = rng.permutation(M).astype(np.int32)
        inp = (C.c_void_p * M)(*[ins[j].ctypes.data for j in range(M)])
        outp = (C.c_void_p * nouts)(*[outs[j].ctypes.data for j in range(nouts)])
        ctot = LL(0)
        rc = drv.grbx_pfb_run(M, p(taps), nt, C.c_float(osr), p(cmap), M, inp, outp, nouts, LL(nout), 1000, C.byref(ctot))
        assert rc == nout and ctot.value == cons, drv.grbx_what()
        ref, _ = orc.pfb_channelizer(M, taps, osr, ins, nout, channel_map=cmap, nouts=nouts)
        assert_parity(outs, ref, "c++ pfb M=%d" % M)
    assert drv.grbx_pfb_run(36, p(taps), nt, C.c_float(10.1334), None, 0, inp, outp, 1, LL(0), 10, None) == -1


def test_cxx_pfb_decimator(drv, orc):
    rng = np.random.default_rng(7)
    M, nt, n = 6, 100, 17 * 16 * 5
    x = crand(rng, M * n).reshape(M, n)
    taps = rng.uniform(-1, 1, nt).astype(np.float32)
    ins = (C.c_void_p * M)(*[x[s].ctypes.data for s in range(M)])
    ref = orc.pfb_decimator(M, taps, 4, x)
    for rot, filt in ((0, 0), (1, 1)):  # history 17 / history 1 with the FFT filters' output multiple 64 - 17 + 1 = 48
        out = np.zeros(n, np.complex64)
        rc = drv.grbx_pfb_dec_run(M, p(taps), nt, 4, rot, filt, ins, p(out), LL(n), 480)
        assert rc == (n if not filt else n // 48 * 48), drv.grbx_what()
        assert_parity(out[:rc], ref[:rc], "c++ pfb_decimator rot=%d filt=%d" % (rot, filt), 2e-6, 2e-5)
    rc = drv.grbx_pfb_dec_run(M, p(taps), nt, 6, 0, 0, ins, p(out), LL(n), 480)
    assert rc == -1  # channel out of range -> std::invalid_argument


def test_cxx_interp_fir(drv, orc):
    rng = np.random.default_rng(8)
    x = crand(rng, 5000)
    for ct, L, nt in ((0, 4, 50), (1, 3, 33)):
        taps = crand(rng, nt) if ct else rng.uniform(-1, 1, nt).astype(np.float32)
        out = np.zeros(L * len(x), np.complex64)
        rc = drv.grbx_interp_fir_run(ct, L, p(taps), nt, p(x), p(out), LL(len(x)), 1024)
        assert rc == L * len(x), drv.grbx_what()
        assert_parity(out, orc.interp_fir(L, taps, x), "c++ interp_fir ct=%d" % ct, 2e-6, 2e-5)
    rc = drv.grbx_interp_fir_run(0, 0, p(taps), 4, p(x), p(out), LL(10), 1024)
    assert rc == -2  # std::out_of_range


def test_cxx_rational_resampler(drv, orc):
    rng = np.random.default_rng(9)
    x = crand(rng, 6000)
    for ct, L, D, nt in ((0, 4, 7, 50), (1, 3, 2, 33), (0, 1, 1, 2)):
        taps = crand(rng, nt) if ct else rng.uniform(-1, 1, nt).astype(np.float32)
        exp, _, _ = orc.rational_resampler(L, D, taps, x)
        out = np.zeros(len(exp) + 8, np.complex64)
        rc = drv.grbx_rational_resampler_run(ct, L, D, p(taps), nt, p(x), p(out), LL(len(x)), 1000)
        assert len(exp) - 2 * (L // D + 2) <= rc <= len(exp), drv.grbx_what()
        assert_parity(out[:rc], exp[:rc], "c++ rational_resampler ct=%d" % ct, 2e-6, 2e-5)
    rc = drv.grbx_rational_resampler_run(0, 0, 1, p(taps), 2, p(x), p(out), LL(10), 1000)
    assert rc == -2  # std::out_of_range
    rc = drv.grbx_rational_resampler_run(0, 1, 0, p(taps), 2, p(x), p(out), LL(10), 1000)
    assert rc == -2


def test_cxx_fir_fff(drv, orc):
    rng = np.random.default_rng(21)
    for nt, D in ((256, 1), (33, 4)):
        nout = 20000
        x = rng.uniform(-1, 1, nout * D).astype(np.float32)
        taps = rng.uniform(-1, 1, nt).astype(np.float32)
        xin = np.concatenate([np.zeros(nt - 1, np.float32), x])
        out = np.zeros(nout, np.float32)
        rc = drv.grbx_fir_fff_run(D, p(taps), nt, p(xin), p(out), LL(nout), 4096)
        assert rc == nout, drv.grbx_what()
        assert_parity(out, orc.fir_fff(D, taps, xin, nout), "c++ fir_filter_fff %d/%d" % (nt, D))


def test_cxx_fft_vfc(drv, orc):
    rng = np.random.default_rng(22)
    N, nvec = 1024, 130
    x = rng.uniform(-1, 1, N * nvec).astype(np.float32)
    w = rng.uniform(0.1, 1, N).astype(np.float32)
    out = np.zeros(N * nvec, np.complex64)
    rc = drv.grbx_fft_vfc_run(N, 1, p(w), N, p(x), p(out), LL(nvec), 7)
    assert rc == nvec, drv.grbx_what()
    assert_parity(out, orc.fft_vfc(N, True, w, x), "c++ fft_vfc")
    assert drv.grbx_fft_vfc_run(N, 1, p(w), N - 1, p(x), p(out), LL(1), 7) == -3  # std::runtime_error: window length


def test_cxx_pfb_synthesizer(drv, orc):
    rng = np.random.default_rng(23)
    for M, T, twox, nin in ((64, 16, 0, 64), (8, 5, 1, 11), (5, 7, 0, 3)):
        n = 3000
        taps = rng.uniform(-1, 1, M * T - 1).astype(np.float32)
        ins = crand(rng, nin * n).reshape(nin, n)
        cmap = rng.permutation((2 if twox else 1) * M).astype(np.int32)
        out = np.zeros(n * M, np.complex64)
        inp = (C.c_void_p * nin)(*[ins[i].ctypes.data for i in range(nin)])
        rc = drv.grbx_pfb_synth_run(M, p(taps), len(taps), twox, p(cmap), len(cmap), inp, nin, p(out), LL(n * M), 4096 // M * M)
        assert rc == n * M, drv.grbx_what()
        assert_parity(out, orc.pfb_synthesizer(M, taps, bool(twox), ins, channel_map=cmap), "c++ pfb_synthesizer %d" % M)
    assert drv.grbx_pfb_synth_run(5, p(taps), 10, 1, None, 0, inp, 1, p(out), LL(5), 5) == -1  # odd channels with 2x


def test_cxx_filterbank(drv, orc):
    rng = np.random.default_rng(24)
    nf, nt, n = 32, 9, 6000
    t1 = rng.uniform(-1, 1, (nf, nt)).astype(np.float32)
    t2 = rng.uniform(-1, 1, (nf, nt)).astype(np.float32)
    t1[4] = 0
    x = crand(rng, nf * (n + nt))
    out = np.zeros(nf * n, np.complex64)
    rc = drv.grbx_filterbank_run(nf, nt, p(t1), None, LL(0), p(x), p(out), LL(n), 512)
    assert rc == n, drv.grbx_what()
    assert_parity(out.reshape(n, nf), orc.filterbank(t1, x, n), "c++ filterbank")
    # set_taps mid-stream: work() returns 0 once, then the new taps apply from that item on
    rc = drv.grbx_filterbank_run(nf, nt, p(t1), p(t2), LL(2048), p(x), p(out), LL(n), 512)
    assert rc == n, drv.grbx_what()
    got = out.reshape(n, nf)
    assert_parity(got[:2048], orc.filterbank(t1, x, 2048), "c++ filterbank before set_taps")
    assert_parity(got[2048:], orc.filterbank(t2, x[2048 * nf:], n - 2048), "c++ filterbank after set_taps")


def test_cxx_fir_lag_by_one(drv, orc):
    """fir_filter_ccf with set_lag(): general_work consumes the call it was offered and returns the previous call's outputs;
    everything but the last call (dropped at stop()) comes out, in order, equal to the blocking block's."""
    rng = np.random.default_rng(31)
    nt, D, nout, lag = 64, 2, 200000, 16384
    taps = rng.uniform(-1, 1, nt).astype(np.float32)
    x = crand(rng, nout * D)
    xin = np.concatenate([np.zeros(nt - 1, np.complex64), x])
    out = np.zeros(nout, np.complex64)
    rc = drv.grbx_fir_lag_run(D, p(taps), nt, p(xin), p(out), LL(nout), lag, 4 * lag)
    assert rc > 0, drv.grbx_what()
    assert nout - 4 * lag <= rc < nout  # at most the calls still in flight are missing
    assert_parity(out[:rc], orc.fir(D, taps, xin, rc), "c++ fir lag-by-one")
