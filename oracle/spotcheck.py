"""oracle/spotcheck.py -- TEST INFRASTRUCTURE ONLY (the checker, never the thing measured or shipped).

Parity of a BASELINE-sized device-resident run against the CPU oracle WITHOUT running the oracle over 2^30 samples
(last row): the first and the last segment of the output plus >= 64 seeded random interior windows
are pulled back from the device, each with the input context it depends on (ntaps-1 items / T rows / one vector), and compared
with (i) the C oracle (oracle/_ref = the reference's own sources when present, else the C port) and (ii) the float64 closed
form.  Used by tests/test_gpu_fullsize.py and, OUTSIDE the timed region, by bench.py's "parity" key.

fetch_in(a, b) / fetch_out(a, b) return numpy arrays of items [a, b) of the device buffers (the caller slices its torch
tensors and copies only those bytes).  Every check returns
    {"windows": n, "items": total compared, "nrms_max": worst window NRMS vs the C oracle, "max_over_rms": worst max-abs / RMS,
     "nrms_f64_max": ..., "max_over_rms_f64": ..., "oracle": "reference" | "port"}
north_star tolerance: NRMS <= 1e-5, max-abs <= 1e-4 x signal RMS.
"""
import numpy as np

from . import pyoracle as po

NRMS_TOL = 1e-5
MAXABS_TOL = 1e-4


def window_starts(total, width, count=64, seed=0, align=1):
    """First, last and `count` seeded random interior window starts in [0, total - width], multiples of `align`."""
    width = min(width, total)
    last = ((total - width) // align) * align
    rng = np.random.default_rng(seed)
    s = {0, last}
    if last > 0:
        s.update(int(v) // align * align for v in rng.integers(0, last + 1, count))
        # one window straddling every power-of-two byte boundary that index arithmetic could trip over (2^31, 2^32 bytes ...)
        for p in (28, 29, 30, 31, 32, 33):
            q = ((1 << p) // 8 - width // 2) // align * align
            if 0 < q < last:
                s.add(q)
    return sorted(s)


class _Acc:
    def __init__(self, kind):
        self.d = {"windows": 0, "items": 0, "nrms_max": 0.0, "max_over_rms": 0.0, "nrms_f64_max": 0.0, "max_over_rms_f64": 0.0,
                  "oracle": kind}

    def add(self, got, ref, f64):
        assert np.isfinite(np.asarray(got).view(np.float32)).all(), "non-finite output"
        a, b = po.nrms(got, ref)
        self.d["nrms_max"] = max(self.d["nrms_max"], a)
        self.d["max_over_rms"] = max(self.d["max_over_rms"], b)
        if f64 is not None:
            a, b = po.nrms(np.asarray(got, np.complex128), f64)
            self.d["nrms_f64_max"] = max(self.d["nrms_f64_max"], a)
            self.d["max_over_rms_f64"] = max(self.d["max_over_rms_f64"], b)
        self.d["windows"] += 1
        self.d["items"] += int(np.asarray(got).size)

    def result(self):
        return self.d


def ok(res, nrms_tol=NRMS_TOL, max_tol=MAXABS_TOL):
    return (res["nrms_max"] <= nrms_tol and res["max_over_rms"] <= max_tol and res["nrms_f64_max"] <= nrms_tol and
            res["max_over_rms_f64"] <= max_tol)


# ---------------------------------------------------------------------------------------------------------------------
def fir(fetch_in, fetch_out, decim, taps, nout, width=512, count=64, seed=0, orc=None, rotator=None):
    """fir_filter_ccf/ccc work(): in[] carries ntaps-1 history items in front; out[i] = sum_k taps[k] in[i D + ntaps-1-k]
    (gr-filter/lib/fir_filter_blk_impl.cc:67-90).  rotator = (theta0, delta): out[i] *= e^{j(theta0 + i delta)} (freq_xlating)."""
    orc = orc or po.best()
    acc = _Acc(orc.kind)
    nt = len(taps)
    t64 = np.asarray(taps, np.complex128 if np.iscomplexobj(taps) else np.float64)
    for o0 in window_starts(nout, width, count, seed):
        w = min(width, nout - o0)
        xin = fetch_in(o0 * decim, (o0 + w) * decim + nt - 1)
        got = fetch_out(o0, o0 + w)
        ref = orc.fir(decim, taps, xin, w)
        f64 = np.convolve(xin.astype(np.complex128), t64)[nt - 1: nt - 1 + w * decim: decim][:w]
        if rotator is not None:
            ph = np.exp(1j * (rotator[0] + rotator[1] * (o0 + np.arange(w))))
            ref = (ref.astype(np.complex128) * ph).astype(np.complex64)
            f64 = f64 * ph
        acc.add(got, ref, f64)
    return acc.result()


def fft_filter(fetch_in, fetch_out, taps, nout, count=64, seed=0, orc=None):
    """fft_filter_ccc device form (stateless, ntaps-1 history items in front, decimation 1): every window is re-computed by the
    reference's OWN overlap-add kernel (gr-filter/lib/fft_filter.cc:241-285) started ntaps-1 items early -- its first ntaps-1
    outputs are the start-up transient and are dropped -- and by the float64 convolution."""
    orc = orc or po.best()
    acc = _Acc(orc.kind)
    nt = len(taps)
    ns = orc.fft_filter_nsamples(nt)
    blocks = max(2, -(-(nt - 1 + 64) // ns))  # enough reference blocks to get past the transient
    span = blocks * ns                        # input items handed to the reference kernel
    w = span - (nt - 1)                       # outputs compared per window
    t64 = np.asarray(taps, np.complex128)
    for o0 in window_starts(nout, w, count, seed):
        ww = min(w, nout - o0)
        xin = fetch_in(o0, o0 + span) if o0 + w <= nout else np.concatenate(
            [fetch_in(o0, o0 + ww + nt - 1), np.zeros(span - (ww + nt - 1), np.complex64)])
        got = fetch_out(o0, o0 + ww)
        ref = orc.fft_filter_ccc(1, taps, xin)[nt - 1: nt - 1 + ww]
        f64 = np.convolve(xin[: ww + nt - 1].astype(np.complex128), t64)[nt - 1: nt - 1 + ww]
        acc.add(got, ref, f64)
    return acc.result()


def pfb_channelizer(fetch_in_rows, fetch_out, nfilts, taps, nout, width=64, count=64, seed=0, orc=None):
    """pfb_channelizer_ccf at oversample_rate 1, identity map, interleaved input (stream_to_streams absorbed): output time o of
    every channel depends on interleaved rows [o, o + T] (T history rows in front).  fetch_in_rows(a, b) -> [b-a, M] rows;
    fetch_out(a, b) -> [M, b-a].  Reference: gr-filter/lib/pfb_channelizer_ccf_impl.cc:128-193."""
    orc = orc or po.best()
    acc = _Acc(orc.kind)
    M = nfilts
    T = orc.pfb_taps_per_filter(M, len(taps))
    h = np.zeros(M * T, np.float64)
    h[: len(taps)] = np.asarray(taps, np.float64)
    n = np.arange(M * T)
    for o0 in window_starts(nout, width, count, seed):
        w = min(width, nout - o0)
        rows = fetch_in_rows(o0, o0 + w + T)
        ins = np.ascontiguousarray(rows.T)  # [M, w + T]: stream j = column j of the interleaved rows
        got = fetch_out(o0, o0 + w)
        ref, _ = orc.pfb_channelizer(M, taps, 1.0, ins, w)
        # float64 closed form: out_c[o] = sum_n h[n] e^{+j 2 pi c n / M} x[(o + T + 1) M - 1 - n], x = the rows flattened
        x = rows.reshape(-1).astype(np.complex128)
        idx = (np.arange(w)[:, None] + T + 1) * M - 1 - n[None, :]  # [w, M T]
        xw = x[idx] * h[None, :]
        # sum over u for each branch i (n = i + u M), then the M-point DFT with e^{+j 2 pi i c / M}
        v = xw.reshape(w, T, M).sum(axis=1)  # [w, M]: v_i[o]
        f64 = (np.fft.ifft(v, axis=1) * M).T  # [M, w]
        acc.add(got, ref, f64)
    return acc.result()


def fft_vcc(fetch_in, fetch_out, fft_size, forward, window, shift, nvec, magsq=False, count=64, seed=0, orc=None):
    """fft_vcc (+ complex_to_mag_squared): vectors are independent; first, last and `count` random vectors.
    fetch_in(a, b) / fetch_out(a, b) are in VECTORS.  Reference: gr-fft/lib/fft_vcc_fftw.cc:63-122."""
    orc = orc or po.best()
    acc = _Acc(orc.kind)
    for v0 in window_starts(nvec, 1, count, seed):
        x = fetch_in(v0, v0 + 1).reshape(-1)
        got = fetch_out(v0, v0 + 1).reshape(-1)
        ref = orc.fft_vcc(fft_size, forward, window, shift, x)
        f64 = po.f64_fft_vcc(fft_size, forward, window, shift, x)
        if magsq:
            ref = orc.mag_squared(ref)
            f64 = np.abs(f64) ** 2
            a, b = po.nrms(got, ref)
            c, d = po.nrms(got.astype(np.float64), f64)
            assert np.isfinite(got).all()
            r = acc.d
            r["nrms_max"], r["max_over_rms"] = max(r["nrms_max"], a), max(r["max_over_rms"], b)
            r["nrms_f64_max"], r["max_over_rms_f64"] = max(r["nrms_f64_max"], c), max(r["max_over_rms_f64"], d)
            r["windows"] += 1
            r["items"] += got.size
        else:
            acc.add(got, ref, f64)
    return acc.result()


def rational_resampler(fetch_in, fetch_out, interpolation, decimation, taps, nout, ctr0=0, width=384, count=64, seed=0, orc=None):
    """rational_resampler_base general_work() (gr-filter/lib/rational_resampler_base_impl.cc:151-173) in closed form:
    out[i] = firs[(ctr0 + i D) % L].filter(&in[(ctr0 + i D) / L]); interp_fir_filter is the case D = 1, ctr0 = 0
    (interp_fir_filter_impl.cc:139-147).  in[] starts at the first window item (no history in front)."""
    orc = orc or po.best()
    acc = _Acc(orc.kind)
    L, D = interpolation, decimation
    T = -(-len(taps) // L)
    for o0 in window_starts(nout, width, count, seed):
        w = min(width, nout - o0)
        s0, s1 = ctr0 + o0 * D, ctr0 + (o0 + w - 1) * D
        q0, q1 = s0 // L, s1 // L
        xin = fetch_in(q0, q1 + T)
        got = fetch_out(o0, o0 + w)
        ref, _, _ = orc.rational_resampler(L, D, taps, xin, noutput=w, ctr=s0 % L)
        f64 = po.f64_rational_resampler(L, D, taps, xin, w, ctr=s0 % L)
        acc.add(got, ref[:w], f64)
    return acc.result()


def pfb_decimator(fetch_in, fetch_out, decim, taps, channel, nout, width=256, count=32, seed=0, orc=None):
    """pfb_decimator_ccf: fetch_in(s, a, b) -> items [a, b) of input stream s (T-1 history items in front)."""
    orc = orc or po.best()
    acc = _Acc(orc.kind)
    M = decim
    T = -(-len(taps) // M)
    for o0 in window_starts(nout, width, count, seed):
        w = min(width, nout - o0)
        streams = np.stack([fetch_in(s, o0, o0 + w + T - 1) for s in range(M)])  # history first
        got = fetch_out(o0, o0 + w)
        # the oracle call prepends T-1 zeros itself: hand it the window as a fresh stream and drop the first T-1 outputs
        ref = orc.pfb_decimator(M, taps, channel, streams)[T - 1:]
        f64 = po.f64_pfb_decimator(M, taps, channel, streams)[T - 1:]
        acc.add(got, ref, f64)
    return acc.result()


def fir_fff(fetch_in, fetch_out, decim, taps, nout, width=1024, count=64, seed=0, orc=None):
    """fir_filter_fff: float stream with ntaps-1 history floats in front."""
    orc = orc or po.best()
    acc = _Acc(orc.kind)
    nt = len(taps)
    t64 = np.asarray(taps, np.float64)
    for o0 in window_starts(nout, width, count, seed):
        w = min(width, nout - o0)
        xin = fetch_in(o0 * decim, (o0 + w) * decim + nt - 1)
        got = fetch_out(o0, o0 + w)
        ref = orc.fir_fff(decim, taps, xin, w)
        f64 = np.convolve(xin.astype(np.float64), t64)[nt - 1: nt - 1 + w * decim: decim][:w]
        assert np.isfinite(got).all()
        a, b = po.nrms(got, ref)
        c, d = po.nrms(got.astype(np.float64), f64)
        r = acc.d
        r["nrms_max"], r["max_over_rms"] = max(r["nrms_max"], a), max(r["max_over_rms"], b)
        r["nrms_f64_max"], r["max_over_rms_f64"] = max(r["nrms_f64_max"], c), max(r["max_over_rms_f64"], d)
        r["windows"] += 1
        r["items"] += got.size
    return acc.result()


def pfb_synthesizer(fetch_in, fetch_out, numchans, taps, ntimes, width=128, count=64, seed=0, orc=None):
    """pfb_synthesizer_ccf (1x), stateless device form: fetch_in(a, b) -> [nin][b-a] items of every input stream (T-1 history
    items in front); fetch_out(a, b) -> output items [a*M, b*M).  A window is re-synthesised by the oracle from a zero state
    T-1 items early; its first (T-1)*M outputs (the start-up transient) are dropped."""
    orc = orc or po.best()
    acc = _Acc(orc.kind)
    M = numchans
    T = -(-len(taps) // M)
    for t0 in window_starts(ntimes, width, count, seed):
        w = min(width, ntimes - t0)
        ins = fetch_in(t0, t0 + w + T - 1)
        got = fetch_out(t0, t0 + w)
        ref = orc.pfb_synthesizer(M, taps, False, ins)[(T - 1) * M:]
        f64 = po.f64_pfb_synthesizer(M, taps, False, ins)[(T - 1) * M:]
        acc.add(got, ref, f64)
    return acc.result()


def filterbank(fetch_in, fetch_out, taps, nout, width=128, count=64, seed=0, orc=None):
    """filterbank_vcvcf: fetch_in(a, b) -> vectors [a, b) of the input (ntaps history vectors in front), fetch_out likewise."""
    orc = orc or po.best()
    acc = _Acc(orc.kind)
    t = np.asarray(taps, np.float32)
    nt = t.shape[1]
    for j0 in window_starts(nout, width, count, seed):
        w = min(width, nout - j0)
        xin = fetch_in(j0, j0 + w + nt)
        got = fetch_out(j0, j0 + w)
        acc.add(got, orc.filterbank(t, xin, w), po.f64_filterbank(t, xin, w))
    return acc.result()
