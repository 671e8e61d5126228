"""oracle/pyoracle.py -- TEST INFRASTRUCTURE ONLY.

Python face of the CPU oracle for the gr-filter / gr-fft hot path.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference`` legs may import
this module; the product package ``gnuradio_b200`` never does.

Three layers, from most to least authoritative:

* ``Ref``   -- oracle/_ref/libgrref.so: the reference's OWN kernel sources (fir_filter.cc,
  fft_filter.cc, polyphase_filterbank.cc, firdes.cc, window.cc) compiled unmodified from
  /root/reference against oracle/shim by oracle/Makefile (block shells restated in oracle/ref_capi.cc).
* ``Port``  -- oracle/libgrb_oracle.so: the plain-C restatement oracle/port.c.
* ``f64_*`` -- float64 numpy closed forms.

Both C layers expose the same calls through ``_CLib`` so tests can run either against the other.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SO = os.path.join(HERE, "_ref", "libgrref.so")
PORT_SO = os.path.join(HERE, "libgrb_oracle.so")


def build(ref=True, port=True):
    """Compile the checker libraries (building the checker is not using it)."""
    if port:
        subprocess.check_call(["make", "-s", "-C", HERE, "port"])
    if ref and os.path.isdir("/root/reference/gr-filter"):
        subprocess.check_call(["make", "-s", "-C", HERE, "ref"])


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _c64(a):
    return np.ascontiguousarray(a, dtype=np.complex64)


def _headroom(a):
    """Copy of 1-D/2-D complex64 ``a`` whose every row is preceded by 64 zero bytes of the same allocation.
    kernel::fir_filter::filter() (gr-filter/lib/fir_filter.cc:133-141) rounds the input pointer DOWN to the
    VOLK alignment and multiplies what lies in front by zero taps; inside a gr::buffer that memory is always
    finite, in a bare numpy allocation it may hold NaN bit patterns (0*NaN = NaN)."""
    a = _c64(a)
    if a.ndim == 1:
        buf = np.zeros(len(a) + 8, np.complex64)
        buf[8:] = a
        return buf[8:]
    buf = np.zeros((a.shape[0], a.shape[1] + 8), np.complex64)
    buf[:, 8:] = a
    return buf[:, 8:]


class _CLib:
    """Uniform wrapper over libgrref.so (prefix grref_) and libgrb_oracle.so (prefix grbo_)."""

    def __init__(self, path, prefix, kind):
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        self.lib = C.CDLL(path)
        self.prefix = prefix
        self.kind = kind  # "reference" | "port"

    def _f(self, name):
        return getattr(self.lib, self.prefix + name)

    # -- fir_filter_ccf / ccc work(): xin carries ntaps-1 history items in front ------------------
    def fir(self, decim, taps, xin, nout):
        taps = np.ascontiguousarray(taps)
        xin = _headroom(xin)
        ntaps = len(taps)
        assert len(xin) >= nout * decim + ntaps - 1
        out = np.empty(nout, np.complex64)
        if np.iscomplexobj(taps):
            t = _c64(taps)
            rc = self._f("fir_ccc")(decim, _p(t), ntaps, _p(xin), _p(out), C.c_long(nout))
        else:
            t = np.ascontiguousarray(taps, np.float32)
            rc = self._f("fir_ccf")(decim, _p(t), ntaps, _p(xin), _p(out), C.c_long(nout))
        assert rc == 0
        return out

    def fir_stream(self, decim, taps, x):
        """Whole finite stream as a flowgraph would see it: ntaps-1 zeros preloaded."""
        ntaps = len(taps)
        xin = np.concatenate([np.zeros(ntaps - 1, np.complex64), _c64(x)])
        return self.fir(decim, taps, xin, len(x) // decim)

    # -- freq_xlating_fir_filter_ccf / ccc -----------------------------------------------------------
    def freq_xlating_fir(self, decim, taps, center_freq, sampling_freq, xin, nout, chunk_out=0, phase=None):
        """work() over xin (ntaps-1 history items in front), `chunk_out` outputs per call (0 = one call).
        Returns (out, phase) with phase the rotator state after the run (complex64)."""
        taps = np.ascontiguousarray(taps)
        xin = _headroom(xin)
        ntaps = len(taps)
        assert len(xin) >= nout * decim + ntaps - 1
        cplx = bool(np.iscomplexobj(taps))
        t = _c64(taps) if cplx else np.ascontiguousarray(taps, np.float32)
        out = np.empty(nout, np.complex64)
        ph = np.array([1.0, 0.0] if phase is None else [np.real(phase), np.imag(phase)], np.float32)
        rc = self._f("freq_xlating_fir")(int(cplx), int(decim), _p(t), ntaps, C.c_double(center_freq),
                                          C.c_double(sampling_freq), _p(xin), _p(out), C.c_long(nout),
                                          C.c_long(chunk_out), _p(ph))
        assert rc == 0
        return out, np.complex64(complex(ph[0], ph[1]))

    # -- fft_filter_ccc ---------------------------------------------------------------------------
    def fft_filter_nsamples(self, ntaps):
        return int(self._f("fft_filter_nsamples")(int(ntaps)))

    def fft_filter_ccc(self, decim, taps, x, chunk_out=0):
        t = _c64(taps)
        x = _c64(x)
        ns = self.fft_filter_nsamples(len(t))
        nitems = (len(x) // decim // ns) * ns
        out = np.empty(nitems, np.complex64)
        if nitems == 0:
            return out
        rc = self._f("fft_filter_ccc")(decim, _p(t), len(t), _p(x), _p(out), C.c_long(nitems),
                                       C.c_long(chunk_out))
        assert rc == ns, rc
        return out

    def fft_filter_ccf(self, decim, taps, x, chunk_out=0):
        """kernel::fft_filter_ccf (gr-filter/lib/fft_filter.cc:291-419) driven like fft_filter_ccf_impl::work."""
        t = np.ascontiguousarray(taps, np.float32)
        x = _c64(x)
        ns = self.fft_filter_nsamples(len(t))
        nitems = (len(x) // decim // ns) * ns
        out = np.empty(nitems, np.complex64)
        if nitems == 0:
            return out
        rc = self._f("fft_filter_ccf")(decim, _p(t), len(t), _p(x), _p(out), C.c_long(nitems),
                                       C.c_long(chunk_out))
        assert rc == ns, rc
        return out

    # -- pfb_channelizer_ccf: one general_work() call ---------------------------------------------
    def pfb_taps_per_filter(self, nfilts, ntaps):
        return int(self._f("pfb_taps_per_filter")(C.c_uint(nfilts), int(ntaps)))

    def pfb_channelizer(self, nfilts, taps, osr, in_streams, noutput, channel_map=None, nouts=None):
        """in_streams: [nfilts, consumed + T] complex64 (T history items first).
        Returns ([nouts, noutput] complex64, consumed) or raises ValueError."""
        t = np.ascontiguousarray(taps, np.float32)
        ins = _headroom(in_streams)
        assert ins.shape[0] == nfilts
        nouts = nfilts if nouts is None else nouts
        outs = np.zeros((nouts, noutput), np.complex64)
        inp = (C.c_void_p * nfilts)(*[ins[j].ctypes.data for j in range(nfilts)])
        outp = (C.c_void_p * nouts)(*[outs[j].ctypes.data for j in range(nouts)])
        if channel_map is None:
            cm, ncm = None, 0
        else:
            cma = np.ascontiguousarray(channel_map, np.int32)
            cm, ncm = _p(cma), len(cma)
        rc = self._f("pfb_channelizer_ccf")(C.c_uint(nfilts), _p(t), len(t), C.c_float(osr), cm, ncm,
                                            inp, outp, nouts, int(noutput))
        if rc < 0:
            raise ValueError("pfb_channelizer_ccf oracle error %d" % rc)
        return outs, rc

    # -- interp_fir_filter_ccf / ccc ----------------------------------------------------------------
    def interp_fir(self, interpolation, taps, x):
        """Fresh stream x (zero history): returns len(x)*interpolation output items."""
        taps = np.ascontiguousarray(taps)
        cplx = bool(np.iscomplexobj(taps))
        t = _c64(taps) if cplx else np.ascontiguousarray(taps, np.float32)
        T = -(-len(t) // interpolation)
        xin = _headroom(np.concatenate([np.zeros(T - 1, np.complex64), _c64(x)]))
        out = np.empty(len(x) * interpolation, np.complex64)
        rc = self._f("interp_fir")(int(cplx), C.c_uint(interpolation), _p(t), len(t), _p(xin), _p(out), C.c_long(len(x)))
        assert rc == T, rc
        return out

    # -- rational_resampler_base_ccf / ccc ----------------------------------------------------------
    def rational_resampler(self, interpolation, decimation, taps, x, noutput=None, ctr=0, nin=None):
        """One general_work() call over x (no history in front; the reference class shadows history()).  `nin` is
        ninput_items[0] (default: everything whose window fits in x); returns (out[:produced], consumed, ctr_after)."""
        taps = np.ascontiguousarray(taps)
        cplx = bool(np.iscomplexobj(taps))
        t = _c64(taps) if cplx else np.ascontiguousarray(taps, np.float32)
        T = -(-len(t) // interpolation)
        x = _c64(x)
        if nin is None:
            nin = max(0, len(x) - (T - 1))
        assert nin + T - 1 <= len(x) or nin == 0
        if noutput is None:
            noutput = (nin * interpolation) // decimation + 2
        xin = _headroom(x)
        out = np.empty(max(noutput, 1), np.complex64)
        c, used = C.c_uint(ctr), C.c_long(0)
        fn = self._f("rational_resampler")
        fn.restype = C.c_long
        n = fn(int(cplx), C.c_uint(interpolation), C.c_uint(decimation), _p(t), len(t), _p(xin), C.c_long(nin), _p(out),
               C.c_long(noutput), C.byref(c), C.byref(used))
        if n < 0:
            raise ValueError("rational_resampler_base oracle error %d" % n)
        return out[:n].copy(), used.value, c.value

    # -- pfb_decimator_ccf ------------------------------------------------------------------------
    def pfb_decimator(self, decim, taps, channel, streams, mode=0):
        """streams: [decim][n] fresh input streams (zero history).  mode bit 0 = use_fft_rotator, bit 1 =
        use_fft_filters (reference layer only; the port restates work_fir_exp, which the others equal up to rounding).
        FFT-filter modes emit whole multiples of the block's output multiple only."""
        taps = np.ascontiguousarray(taps, np.float32)
        streams = np.ascontiguousarray(streams, np.complex64)
        M, n = streams.shape
        assert M == decim
        T = -(-len(taps) // M)
        if self.kind == "reference" and (mode & 2):
            ns = self._f("pfb_decimator_nsamples")(C.c_uint(decim), _p(taps), len(taps))
            n = (n // ns) * ns
            xs = [_headroom(np.ascontiguousarray(streams[s, :n])) for s in range(M)]
        else:
            xs = [_headroom(np.concatenate([np.zeros(T - 1, np.complex64), streams[s]])) for s in range(M)]
        out = np.empty(n, np.complex64)
        inp = (C.c_void_p * M)(*[x.ctypes.data for x in xs])
        if self.kind == "reference":
            rc = self._f("pfb_decimator_ccf")(C.c_uint(decim), _p(taps), len(taps), C.c_uint(channel), int(mode), inp,
                                              _p(out), int(n), None)
        else:
            rc = self._f("pfb_decimator_ccf")(C.c_uint(decim), _p(taps), len(taps), C.c_uint(channel), inp, _p(out),
                                              C.c_long(n))
        assert rc == n, rc
        return out

    # -- fft_vcc -----------------------------------------------------------------------------------
    def fft_vcc(self, fft_size, forward, window, shift, x):
        x = _c64(x)
        nvec = len(x) // fft_size
        out = np.empty(nvec * fft_size, np.complex64)
        if window is None or len(window) == 0:
            w, nw = None, 0
        else:
            wa = np.ascontiguousarray(window, np.float32)
            w, nw = _p(wa), len(wa)
        rc = self._f("fft_vcc")(int(fft_size), int(bool(forward)), w, nw, int(bool(shift)), _p(x),
                                _p(out), C.c_long(nvec))
        if rc != 0:
            raise RuntimeError("fft_vcc: window not the same length as fft_size")
        return out

    def mag_squared(self, x):
        x = _c64(x)
        out = np.empty(len(x), np.float32)
        self._f("mag_squared")(_p(x), _p(out), C.c_long(len(x)))
        return out

    # -- fir_filter_fff ------------------------------------------------------------------------------
    def fir_fff(self, decim, taps, xin, nout):
        """fir_filter_fff work(): xin carries ntaps-1 history floats in front."""
        t = np.ascontiguousarray(taps, np.float32)
        buf = np.zeros(len(xin) + 16, np.float32)  # headroom in front: filter() rounds its pointer down
        buf[16:] = np.asarray(xin, np.float32)
        x = buf[16:]
        assert len(x) >= nout * decim + len(t) - 1
        out = np.empty(nout, np.float32)
        rc = self._f("fir_fff")(int(decim), _p(t), len(t), _p(x), _p(out), C.c_long(nout))
        assert rc == 0
        return out

    # -- fft_vfc -------------------------------------------------------------------------------------
    def fft_vfc(self, fft_size, forward, window, x):
        x = np.ascontiguousarray(x, np.float32)
        nvec = len(x) // fft_size
        out = np.empty(nvec * fft_size, np.complex64)
        if window is None or len(window) == 0:
            w, nw = None, 0
        else:
            wa = np.ascontiguousarray(window, np.float32)
            w, nw = _p(wa), len(wa)
        rc = self._f("fft_vfc")(int(fft_size), int(bool(forward)), w, nw, _p(x), _p(out), C.c_long(nvec))
        if rc != 0:
            raise RuntimeError("fft_vfc: window not the same length as fft_size")
        return out

    # -- logpwrfft_c / logpwrfft_f ---------------------------------------------------------------------
    def logpwrfft(self, complex_input, fft_size, ref_scale, decim, avg_alpha, average, window, x, count=None, iir=None):
        """The whole hier block over the stream x.  count / iir: keep_one_in_n counter and filter memory to continue from
        (None = a fresh block).  Returns (frames [nout, fft_size] float32, count, iir)."""
        x = np.ascontiguousarray(x, np.complex64 if complex_input else np.float32)
        w = np.ascontiguousarray(window, np.float32)
        assert len(w) == fft_size
        nvec = len(x) // fft_size
        out = np.empty((max(nvec, 1), fft_size), np.float32)
        cnt = C.c_int(int(decim) if count is None else int(count))
        st = np.zeros(fft_size, np.float32) if iir is None else np.array(iir, np.float32)
        fn = self._f("logpwrfft")
        fn.restype = C.c_long
        n = fn(int(bool(complex_input)), int(fft_size), C.c_double(ref_scale), int(decim), C.c_double(avg_alpha),
               int(bool(average)), _p(w), _p(x), C.c_long(len(x)), _p(out), C.c_long(nvec), C.byref(cnt), _p(st))
        if n < 0:
            raise ValueError("logpwrfft oracle error %d" % n)
        return out[:n].copy(), cnt.value, st

    # -- pfb_synthesizer_ccf -------------------------------------------------------------------------
    def pfb_synthesizer(self, numchans, taps, twox, in_streams, channel_map=None, chunk=0):
        """A fresh block fed in_streams [nin][n]: returns n*numchans output items (ValueError on ctor/map errors)."""
        t = np.ascontiguousarray(taps, np.float32)
        ins = np.ascontiguousarray(in_streams, np.complex64)
        nin, n = ins.shape
        out = np.zeros(n * numchans, np.complex64)
        inp = (C.c_void_p * nin)(*[ins[j].ctypes.data for j in range(nin)])
        if channel_map is None:
            cm, ncm = None, 0
        else:
            cma = np.ascontiguousarray(channel_map, np.int32)
            cm, ncm = _p(cma), len(cma)
        rc = self._f("pfb_synthesizer_ccf")(C.c_uint(numchans), _p(t), len(t), int(bool(twox)), cm, ncm, inp, nin, _p(out),
                                            int(n * numchans), int(chunk))
        if rc < 0:
            raise ValueError("pfb_synthesizer_ccf oracle error %d" % rc)
        return out

    # -- filterbank_vcvcf ----------------------------------------------------------------------------
    def filterbank(self, taps, xin, nout):
        """taps [nfilts][ntaps]; xin: (nout + ntaps) vectors of nfilts items, history (ntaps vectors) first."""
        t = np.ascontiguousarray(taps, np.float32)
        nfilts, ntaps = t.shape
        pad = np.zeros(len(np.asarray(xin).reshape(-1)) + 8 * nfilts, np.complex64)
        pad[8 * nfilts:] = np.asarray(xin, np.complex64).reshape(-1)
        x = pad[8 * nfilts:]
        assert len(x) >= (nout + ntaps) * nfilts
        out = np.empty(nout * nfilts, np.complex64)
        rc = self._f("filterbank_vcvcf")(int(nfilts), int(ntaps), _p(t), _p(x), _p(out), int(nout))
        if rc < 0:
            raise ValueError("filterbank_vcvcf oracle error %d" % rc)
        return out.reshape(nout, nfilts)

    # -- window / firdes ---------------------------------------------------------------------------
    def window(self, wtype, ntaps, beta=6.76):
        out = np.empty(ntaps, np.float32)
        if self.kind == "reference":
            rc = self._f("window_build")(int(wtype), int(ntaps), C.c_double(beta), _p(out))
        else:
            rc = self._f("window_build")(int(wtype), int(ntaps), _p(out))
        if rc < 0:
            raise ValueError("window type %d not available in %s oracle" % (wtype, self.kind))
        return out

    def low_pass_2(self, gain, fs, cutoff, tw, atten_db, wtype=0, beta=6.76):
        buf = np.empty(1 << 16, np.float32)
        if self.kind == "reference":
            n = self._f("firdes_low_pass_2")(C.c_double(gain), C.c_double(fs), C.c_double(cutoff),
                                             C.c_double(tw), C.c_double(atten_db), int(wtype),
                                             C.c_double(beta), _p(buf), len(buf))
        else:
            n = self._f("firdes_low_pass_2")(C.c_double(gain), C.c_double(fs), C.c_double(cutoff),
                                             C.c_double(tw), C.c_double(atten_db), int(wtype),
                                             _p(buf), len(buf))
        if n < 0:
            raise ValueError("firdes error %d" % n)
        return buf[:n].copy()


_cache = {}


def Ref():
    if "ref" not in _cache:
        _cache["ref"] = _CLib(REF_SO, "grref_", "reference")
    return _cache["ref"]


def Port():
    if "port" not in _cache:
        if not os.path.exists(PORT_SO):
            build(ref=False)
        _cache["port"] = _CLib(PORT_SO, "grbo_", "port")
    return _cache["port"]


def have_ref():
    return os.path.exists(REF_SO)


def best():
    """The most authoritative C oracle available on this machine."""
    return Ref() if have_ref() else Port()


# ------------------------------------------------------------------------------------------------
# float64 closed forms
# ------------------------------------------------------------------------------------------------
def f64_fir(decim, taps, x):
    """out[i] = sum_k taps[k] x[i*D-k], x[<0]=0 (gr-filter/python/filter/qa_fir_filter.py:14-22)."""
    x = np.asarray(x, np.complex128)
    t = np.asarray(taps, np.complex128 if np.iscomplexobj(taps) else np.float64)
    n = (len(x) // decim)
    return np.convolve(x, t)[: n * decim: decim][:n]


_libm = None


def xlat_increment_angle(fwT0, decim):
    """Angle of the float phasor exp(gr_complex(0, -fwT0*decim)) the rotator steps by (:92, rotator.h:31).  One ulp of
    its cosf/sinf moves the angle by ~6e-8 rad per item -- 2e-4 rad over 4000 items -- so the C library's own
    cosf/sinf are called here rather than numpy's float32 kernels, which may round differently."""
    global _libm
    if _libm is None:
        _libm = C.CDLL("libm.so.6")
        for f in (_libm.cosf, _libm.sinf):
            f.restype, f.argtypes = C.c_float, [C.c_float]
    a = np.float32(-np.float32(fwT0) * np.float32(decim))
    return float(np.arctan2(float(_libm.sinf(C.c_float(a))), float(_libm.cosf(C.c_float(a)))))


def f64_freq_xlating_fir(decim, taps, center_freq, sampling_freq, x, n0=0):
    """Closed form in float64 over the float32 quantities the block derives (freq_xlating_fir_filter_impl.cc:85-92):
    y[n] = e^{j delta (n0+n)} sum_k taps[k] e^{j k fwT0} x[nD-k], with float fwT0 and delta the angle of the float
    rotator increment exp(-j fwT0 D) the block really steps by (xlat_increment_angle).  Same quantity as the expected_result of
    gr-filter/python/filter/qa_freq_xlating_fir_filter.py:96-127 (mix down, filter, decimate)."""
    fwT0 = np.float32(2 * np.pi * center_freq / sampling_freq)
    arg = np.arange(len(taps), dtype=np.float32) * fwT0  # float product as in :87; it defines the composite taps
    t = np.asarray(taps, np.complex128) * np.exp(1j * arg.astype(np.float64))
    y = f64_fir(decim, t, x)
    delta = xlat_increment_angle(fwT0, decim)
    return y * np.exp(1j * delta * (n0 + np.arange(len(y))))


def f64_fft_filter(decim, taps, x, nsamples):
    x = np.asarray(x, np.complex128)
    t = np.asarray(taps, np.complex128)
    nitems = (len(x) // decim // nsamples) * nsamples
    return np.convolve(x[: nitems * decim], t)[: nitems * decim: decim][:nitems]


def f64_channelizer(nfilts, taps, x, nout, osr=1.0):
    """Closed form of pfb_channelizer_ccf for an interleaved input stream x (zeros before t=0).

    out_c[o] = rot(c,o) * (x conv g_c)[P(o)],  g_c[n] = h[n] e^{+j2pi c n/M},
    P(o) = (o+1)*R - 1, R = M/osr, rot = e^{-j2pi c (o+2) R / M}   (==1 when osr == 1).
    Derived from gr-filter/lib/pfb_channelizer_ccf_impl.cc:73-77,157-189."""
    M = int(nfilts)
    R = int(round(M / osr))
    h = np.asarray(taps, np.float64)
    x = np.asarray(x, np.complex128)
    n = np.arange(len(h))
    out = np.zeros((M, nout), np.complex128)
    o = np.arange(nout)
    P = (o + 1) * R - 1
    assert P.max() < len(x) + 0 or nout == 0
    for c in range(M):
        g = h * np.exp(2j * np.pi * c * n / M)
        y = np.convolve(x, g)[: len(x)]
        out[c] = y[P] * np.exp(-2j * np.pi * c * ((o + 2) * R % M) / M)
    return out


def f64_interp_fir(interpolation, taps, x):
    """y = taps (*) zero-stuffed x: y[m] = sum_k taps[k] u[m-k], u[i*L] = x[i] (what qa_interp_fir_filter.py:24-45 spells
    out digit by digit)."""
    L = interpolation
    x = np.asarray(x, np.complex128)
    u = np.zeros(len(x) * L, np.complex128)
    u[::L] = x
    t = np.asarray(taps, np.complex128 if np.iscomplexobj(taps) else np.float64)
    return np.convolve(u, t)[: len(u)]


def f64_rational_resampler(interpolation, decimation, taps, x, nout, ctr=0):
    """out[i] = y[(T-1)*L + ctr + i*D], y = taps (*) zero-stuffed x (u[j*L] = x[j]): the polyphase walk of
    rational_resampler_base_impl.cc:156-172 with no history in front of the stream."""
    L, D = interpolation, decimation
    x = np.asarray(x, np.complex128)
    u = np.zeros(len(x) * L, np.complex128)
    u[::L] = x
    t = np.asarray(taps, np.complex128 if np.iscomplexobj(taps) else np.float64)
    T = -(-len(t) // L)
    y = np.convolve(u, t)
    idx = (T - 1) * L + ctr + np.arange(nout) * D
    return y[idx]


def f64_pfb_decimator(decim, taps, channel, streams):
    """out[i] = sum_k h[k] e^{j 2 pi k chan / M} x[iM + M-1 - k] with x the stream_to_streams-interleaved input
    (stream s holds x[s], x[s+M], ...): the closed form of every pfb_decimator_ccf work() variant."""
    M = decim
    streams = np.asarray(streams, np.complex128)
    n = streams.shape[1]
    x = streams.T.reshape(-1)  # x[i*M + s] = streams[s][i]
    k = np.arange(len(taps))
    # rot[j] = gr_expj(float(j * chan * 2 pi / M)) (pfb_decimator_ccf_impl.cc:49-51): the float rounding of the phase
    # (up to 4e-5 rad at j*chan ~ 10^4) belongs to the block's definition, like a tap value
    ph = ((k % M) * channel * 2 * np.pi / M).astype(np.float32).astype(np.float64)
    g = np.asarray(taps, np.float64) * np.exp(1j * ph)
    y = np.convolve(x, g)
    return y[M - 1: M - 1 + n * M: M][:n]


def f64_fft_vcc(fft_size, forward, window, shift, x):
    N = fft_size
    v = np.asarray(x, np.complex128)[: (len(x) // N) * N].reshape(-1, N)
    if window is not None and len(window):
        v = v * np.asarray(window, np.float64)[None, :]
    if forward:
        y = np.fft.fft(v, axis=1)
        if shift:
            y = np.fft.fftshift(y, axes=1)
    else:
        if shift:
            v = np.fft.ifftshift(v, axes=1)
        y = np.fft.ifft(v, axis=1) * N
    return y.reshape(-1)


def f64_fft_vfc(fft_size, forward, window, x):
    return f64_fft_vcc(fft_size, forward, window, False, np.asarray(x, np.float64).astype(np.complex128))


def f64_logpwrfft(complex_input, fft_size, ref_scale, decim, avg_alpha, average, window, x, count=None, iir=None):
    """logpwrfft in float64 (gr-fft/python/fft/logpwrfft.py:29-64): every decim-th vector (the LAST of each group,
    keep_one_in_n_impl.cc:66-74) -> window -> FFT -> |.|^2 -> y = a x + (1-a) y -> 10 log10(y) + k."""
    N = fft_size
    w = np.asarray(window, np.float64)
    v = np.asarray(x, np.complex128 if complex_input else np.float64)
    v = v[: (len(v) // N) * N].reshape(-1, N)
    cnt = decim if count is None else count
    keep = []
    for i in range(len(v)):
        cnt -= 1
        if cnt <= 0:
            keep.append(i)
            cnt = decim
    p = np.abs(np.fft.fft(v[keep] * w[None, :], axis=1)) ** 2
    a = avg_alpha if average else 1.0
    y = np.zeros(N) if iir is None else np.asarray(iir, np.float64).copy()
    out = np.empty_like(p)
    for j in range(len(p)):
        y = a * p[j] + (1 - a) * y
        out[j] = y
    k = -20 * np.log10(N) - 10 * np.log10(np.sum(w * w) / N) - 20 * np.log10(ref_scale / 2)
    return 10 * np.log10(np.maximum(out, np.finfo(np.float32).tiny)) + k


def f64_pfb_synthesizer(numchans, taps, twox, in_streams, channel_map=None):
    """Closed form of pfb_synthesizer_ccf_impl::work (:221-284) from a zero state:
    V_q[m] = sum_i X_i[m] e^{+j 2 pi map[i] q / (tw M)};  out[n M + i] = sum_j h[i + j M] V_{q(n,i)}[n - j],
    q = i (1x);  q = (((n+1) mod 2) M + i) (2x: the filter pair of branch i reads the IFFT half that alternates with n)."""
    M = numchans
    tw = 2 if twox else 1
    X = np.asarray(in_streams, np.complex128)
    nin, n = X.shape
    cmap = np.arange(tw * M) if channel_map is None else np.asarray(channel_map)
    NF = tw * M
    F = np.zeros((n, NF), np.complex128)
    for i in range(nin):
        F[:, cmap[i]] = X[i]  # a later stream overwrites an earlier one mapped to the same bin (:236-239)
    V = np.fft.ifft(F, axis=1) * NF  # [time][q]
    T = -(-len(taps) // M)
    h = np.zeros(M * T)
    h[: len(taps)] = np.asarray(taps, np.float64)
    out = np.zeros((n, M), np.complex128)
    for j in range(T):
        hj = h[j * M: (j + 1) * M]
        if tw == 1:
            out[j:] += hj[None, :] * V[: n - j, :M]
        else:
            for par in (0, 1):
                rows = np.arange(j, n)
                rows = rows[(rows % 2) == par]
                half = ((par + 1) % 2) * M
                out[rows] += hj[None, :] * V[rows - j, half: half + M]
    return out.reshape(-1)


def f64_filterbank(taps, xin, nout):
    """out[j][i] = sum_k taps[i][k] xin[j + ntaps-1-k][i] (filterbank_vcvcf_impl.cc:73-82)."""
    t = np.asarray(taps, np.float64)
    nfilts, ntaps = t.shape
    x = np.asarray(xin, np.complex128).reshape(-1, nfilts)
    out = np.zeros((nout, nfilts), np.complex128)
    for k in range(ntaps):
        out += t[None, :, k] * x[ntaps - 1 - k: ntaps - 1 - k + nout]
    return out


def nrms(a, ref):
    """Normalised RMS error and max-abs error relative to the reference RMS (north_star tolerance)."""
    a = np.asarray(a)
    ref = np.asarray(ref)
    rms = np.sqrt(np.mean(np.abs(ref) ** 2)) if ref.size else 1.0
    rms = rms if rms > 0 else 1.0
    d = np.abs(a - ref)
    return float(np.sqrt(np.mean(d ** 2)) / rms) if ref.size else 0.0, float(d.max() / rms) if ref.size else 0.0
