"""Shared helpers for the parity tests: fixtures, tolerance (north_star), QA helpers restated from the reference."""
import json
import math
import os
import random

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
NRMS_TOL = 1e-5    # north_star: normalised RMS error
MAXABS_TOL = 1e-4  # north_star: max abs error / signal RMS


def fixtures():
    return np.load(os.path.join(HERE, "golden", "ref_fixtures.npz"))


def qa_vectors():
    d = json.load(open(os.path.join(HERE, "golden", "qa_vectors.json")))
    out = {}
    for k, v in d.items():
        out[k] = np.array(v["re"]) + 1j * np.array(v["im"]) if isinstance(v, dict) else np.array(v)
    return out


def assert_parity(got, ref, what="", nrms_tol=NRMS_TOL, max_tol=MAXABS_TOL):
    got, ref = np.asarray(got), np.asarray(ref)
    assert got.shape == ref.shape, (what, got.shape, ref.shape)
    if ref.size == 0:
        return
    rms = float(np.sqrt(np.mean(np.abs(ref.astype(np.complex128)) ** 2))) or 1.0
    d = np.abs(got.astype(np.complex128) - ref.astype(np.complex128))
    nrms, mx = float(np.sqrt(np.mean(d ** 2))) / rms, float(d.max()) / rms
    assert nrms <= nrms_tol and mx <= max_tol, "%s: nrms %.3g (tol %g), max/rms %.3g (tol %g)" % (
        what, nrms, nrms_tol, mx, max_tol)


def assert_tuples_almost_equal2(ref, x, abs_eps=1e-12, rel_eps=1e-6):
    """gr_unittest.assertComplexTuplesAlmostEqual2 (gnuradio-runtime/python/gnuradio/gr_unittest.py:46-61)."""
    assert len(ref) == len(x)
    for r, v in zip(ref, x):
        if abs(r - v) < abs_eps:
            continue
        assert abs(r - v) / max(abs(r), 1e-300) <= rel_eps, (r, v)


def crand(rng, n):
    return (rng.uniform(-1, 1, n) + 1j * rng.uniform(-1, 1, n)).astype(np.complex64)


# --- qa_fft_filter.py:20-23: the reference QA draws its data from CPython's Mersenne Twister, seed 0 ------
def make_random_complex_tuple(L):
    return [complex(2 * random.random() - 1, 2 * random.random() - 1) for _ in range(L)]


# --- qa_pfb_channelizer.py:16-20 ---------------------------------------------------------------------
def sig_source_c(samp_rate, freq, amp, N):
    t = [float(x) / samp_rate for x in range(N)]
    return [math.cos(2. * math.pi * freq * x) + 1j * math.sin(2. * math.pi * freq * x) for x in t]


def lowpass_taps(ntaps, cutoff, window):
    """Windowed-sinc low-pass with unity DC gain (the formula of gr-filter/lib/firdes.cc:77-119) at any length."""
    M = ntaps - 1
    n = np.arange(ntaps) - M / 2.0
    h = np.sinc(2 * cutoff * n) * 2 * cutoff * np.asarray(window, np.float64)
    return (h / h.sum()).astype(np.float32)


def coswindow(n, c0, c1, c2, c3=0.0):
    """gr-fft/lib/window.cc:94-104 in float64."""
    M = n - 1
    k = np.arange(n)
    return c0 - c1 * np.cos(2 * np.pi * k / M) + c2 * np.cos(4 * np.pi * k / M) - c3 * np.cos(6 * np.pi * k / M)


def blackman_harris(n):
    return coswindow(n, 0.35875, 0.48829, 0.14128, 0.01168).astype(np.float32)  # window.cc:170-185, atten 92


def hamming(n):
    return coswindow(n, 0.54, 0.46, 0.0).astype(np.float32)
