"""The window checker used by the full-size GPU parity tests and by bench.py's "parity" key (oracle/spotcheck.py) is itself
checked here on the CPU: outputs computed by the oracle over a whole (small) stream must pass every window, and a single
corrupted output item or an off-by-one in the stream must fail."""
import numpy as np
import pytest

from common import blackman_harris, crand, hamming, lowpass_taps
from oracle import pyoracle as po
from oracle import spotcheck as sc


def _fetch(a):
    return lambda i, j: np.ascontiguousarray(a[i:j])


@pytest.mark.parametrize("decim,nt,cplx", [(1, 64, False), (8, 1024, False), (1, 33, True), (3, 100, True)])
def test_spotcheck_fir(orc, decim, nt, cplx):
    rng = np.random.default_rng(nt)
    nout = 30000
    x = crand(rng, nout * decim + nt - 1)
    taps = crand(rng, nt) if cplx else rng.uniform(-1, 1, nt).astype(np.float32)
    y = orc.fir(decim, taps, x, nout)
    res = sc.fir(_fetch(x), _fetch(y), decim, taps, nout, width=300, count=16, seed=1, orc=orc)
    assert res["windows"] >= 10 and sc.ok(res), res
    assert res["nrms_max"] < 1e-6  # same oracle; its summation order depends on the window's alignment (fir_filter.cc:67-74)
    bad = y.copy()
    bad[sc.window_starts(nout, 300, 16, 1)[3] + 7] += 1e-2 * np.abs(y).mean()
    assert not sc.ok(sc.fir(_fetch(x), _fetch(bad), decim, taps, nout, width=300, count=16, seed=1, orc=orc))
    shifted = np.concatenate([y[1:], y[:1]])
    assert not sc.ok(sc.fir(_fetch(x), _fetch(shifted), decim, taps, nout, width=300, count=16, seed=1, orc=orc))


def test_spotcheck_fft_filter(orc):
    rng = np.random.default_rng(5)
    nt = 300
    ns = orc.fft_filter_nsamples(nt)
    nout = 40 * ns
    taps = crand(rng, nt)
    x = crand(rng, nout + nt - 1)  # ntaps-1 history items in front, like grb200_fftfilt_work_dev
    y = orc.fir(1, taps, x, nout)  # the device form is stateless: the plain convolution
    res = sc.fft_filter(_fetch(x), _fetch(y), taps, nout, count=12, seed=2, orc=orc)
    assert res["windows"] >= 8 and sc.ok(res), res
    assert res["nrms_max"] > 0.0  # overlap-add FFT arithmetic vs direct sum: close, not identical
    bad = y.copy()
    bad[sc.window_starts(nout, 2 * ns - (nt - 1), 12, 2)[2] + 5] += 1e-2 * np.abs(y).mean()
    assert not sc.ok(sc.fft_filter(_fetch(x), _fetch(bad), taps, nout, count=12, seed=2, orc=orc))


@pytest.mark.parametrize("M,T", [(64, 16), (16, 8)])
def test_spotcheck_pfb_channelizer(orc, M, T):
    rng = np.random.default_rng(M)
    nout = 3000
    taps = lowpass_taps(M * T, 0.5 / M, blackman_harris(M * T)) * np.float32(M)
    rows = crand(rng, (nout + T) * M).reshape(-1, M)
    ins = np.ascontiguousarray(rows.T)
    outs, _ = orc.pfb_channelizer(M, taps, 1.0, ins, nout)
    f_in = lambda a, b: rows[a:b]
    f_out = lambda a, b: np.ascontiguousarray(outs[:, a:b])
    res = sc.pfb_channelizer(f_in, f_out, M, taps, nout, width=32, count=10, seed=3, orc=orc)
    assert res["windows"] >= 8 and sc.ok(res), res
    bad = outs.copy()
    bad[M // 3, sc.window_starts(nout, 32, 10, 3)[1] + 2] *= 1.01
    assert not sc.ok(sc.pfb_channelizer(f_in, lambda a, b: np.ascontiguousarray(bad[:, a:b]), M, taps, nout, width=32,
                                        count=10, seed=3, orc=orc))


@pytest.mark.parametrize("magsq", [False, True])
def test_spotcheck_fft_vcc(orc, magsq):
    rng = np.random.default_rng(9)
    N, nvec = 256, 50
    win = blackman_harris(N)
    x = crand(rng, N * nvec)
    y = orc.fft_vcc(N, True, win, True, x)
    if magsq:
        y = orc.mag_squared(y)
    xv, yv = x.reshape(nvec, N), y.reshape(nvec, N)
    res = sc.fft_vcc(lambda a, b: xv[a:b], lambda a, b: yv[a:b], N, True, win, True, nvec, magsq=magsq, count=10, seed=4, orc=orc)
    assert res["windows"] >= 8 and sc.ok(res), res


def test_spotcheck_pfb_decimator(orc):
    rng = np.random.default_rng(13)
    M, T, nout = 8, 16, 5000
    taps = lowpass_taps(M * T, 0.5 / M, hamming(M * T))
    streams = crand(rng, M * nout).reshape(M, nout)
    y = orc.pfb_decimator(M, taps, 3, streams)  # fresh streams: T-1 zeros of history
    hist = np.concatenate([np.zeros((M, T - 1), np.complex64), streams], axis=1)
    res = sc.pfb_decimator(lambda s, a, b: hist[s, a:b], _fetch(y), M, taps, 3, nout, width=100, count=8, seed=5, orc=orc)
    assert res["windows"] >= 6 and sc.ok(res), res


@pytest.mark.parametrize("L,D,ctr0", [(3, 2, 0), (4, 1, 0), (5, 7, 3)])
def test_spotcheck_rational_resampler(orc, L, D, ctr0):
    """The window checker against one whole-stream run of the oracle (a window restarts the phase counter where the
    stream's closed form puts it); a corrupted output must fail."""
    rng = np.random.default_rng(L * 10 + D)
    T, nout = 9, 6000
    taps = rng.uniform(-1, 1, L * T - 1).astype(np.float32)
    nin = (ctr0 + (nout - 1) * D) // L + 1
    x = crand(rng, nin + T - 1)
    y, _, _ = orc.rational_resampler(L, D, taps, x, noutput=nout, ctr=ctr0)
    assert len(y) == nout
    res = sc.rational_resampler(_fetch(x), _fetch(y), L, D, taps, nout, ctr0=ctr0, width=90, count=8, seed=2, orc=orc)
    assert res["windows"] >= 6 and sc.ok(res), res
    y[0] += 1e-2
    assert not sc.ok(sc.rational_resampler(_fetch(x), _fetch(y), L, D, taps, nout, ctr0=ctr0, width=90, count=8, seed=2, orc=orc))


def test_window_starts_cover_ends_and_boundaries():
    s = sc.window_starts(1 << 30, 1024, 64, 0)
    assert s[0] == 0 and s[-1] == (1 << 30) - 1024 and len(s) >= 66
    assert any(abs(v + 512 - (1 << 32) // 8) < 1024 for v in s)  # a window across the 4 GiB byte offset
    assert sc.window_starts(10, 100, 64, 0) == [0]
