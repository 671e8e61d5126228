"""gr-fft hot-path blocks re-hosted on libgrb200.

  fft_vcc      gr-fft/include/gnuradio/fft/fft_vcc.h:47-70, gr-fft/lib/fft_vcc_fftw.cc:33-122
               (make(fft_size, forward, window, shift=False, nthreads=1), set_window -> bool, work per vector)
  fft_vfc      gr-fft/include/gnuradio/fft/fft_vfc.h, gr-fft/lib/fft_vfc_fftw.cc:30-92
               (make(fft_size, forward, window, nthreads=1): float vectors in, complex vectors out)
  logpwrfft_c / logpwrfft_f   gr-fft/python/fft/logpwrfft.py:29-177 (the hier block as one fused device chain)
  window.blackmanharris       gr-fft/lib/window.cc:94-104,170-185 (host side, float arithmetic like the reference)
"""
import ctypes as C

import numpy as np

from . import _lib
from .runtime import as_items, io_signature, sync_block

_c64 = np.dtype(np.complex64)
_f32 = np.dtype(np.float32)


class window:
    """The slice of gr::fft::window the path's callers use (one-shot host side tap/window design stays on the CPU)."""

    @staticmethod
    def coswindow(ntaps, c0, c1, c2, c3=0.0):
        # window.cc:83-104: float arithmetic, cosf of a float argument
        M = np.float32(ntaps - 1)
        n = np.arange(ntaps, dtype=np.float32)
        pi = np.float32(np.pi)
        w = (np.float32(c0) - np.float32(c1) * np.cos((np.float32(2.0) * pi * n) / M, dtype=np.float32) +
             np.float32(c2) * np.cos((np.float32(4.0) * pi * n) / M, dtype=np.float32))
        if c3:
            w = w - np.float32(c3) * np.cos((np.float32(6.0) * pi * n) / M, dtype=np.float32)
        return w.astype(np.float32)

    @staticmethod
    def blackman_harris(ntaps, atten=92):
        # window.cc:170-185
        co = {61: (0.42323, 0.49755, 0.07922, 0.0), 67: (0.44959, 0.49364, 0.05677, 0.0),
              74: (0.40271, 0.49703, 0.09392, 0.00183), 92: (0.35875, 0.48829, 0.14128, 0.01168)}
        if atten not in co:
            raise IndexError("window::blackman_harris: unknown attenuation value (must be 61, 67, 74, or 92)")
        return window.coswindow(ntaps, *co[atten])

    blackmanharris = blackman_harris

    @staticmethod
    def hamming(ntaps):
        return window.coswindow(ntaps, 0.54, 0.46, 0.0)  # window.cc:125-131

    @staticmethod
    def rectangular(ntaps):
        return np.ones(ntaps, np.float32)


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class fft_vcc(sync_block):
    def __init__(self, fft_size, forward, window, shift=False, nthreads=1, device=0):
        L = _lib.lib()
        fft_size = int(fft_size)
        sync_block.__init__(self, "fft_vcc_fftw", io_signature.make(1, 1, max(fft_size, 0) * 8),
                            io_signature.make(1, 1, max(fft_size, 0) * 8))
        self._h = C.c_void_p()
        self.d_fft_size, self.d_forward, self.d_shift = fft_size, bool(forward), bool(shift)
        self._nthreads = int(nthreads)
        w = np.ascontiguousarray(window if window is not None else [], np.float32)
        # a wrong-length window makes the constructor throw std::runtime_error (fft_vcc_fftw.cc:46-47)
        _lib.check(L.grb200_fft_vcc_create(fft_size, int(self.d_forward), _ptr(w), len(w), int(self.d_shift),
                                           int(device), C.byref(self._h)))
        self.d_window = w.copy()

    @classmethod
    def make(cls, fft_size, forward, window, shift=False, nthreads=1, device=0):
        return cls(fft_size, forward, window, shift, nthreads, device)

    def set_nthreads(self, n):
        self._nthreads = int(n)

    def nthreads(self):
        return self._nthreads

    def set_window(self, window):
        """True when accepted (empty or exactly fft_size long), False otherwise (fft_vcc_fftw.cc:54-61)."""
        w = np.ascontiguousarray(window if window is not None else [], np.float32)
        if len(w) != 0 and len(w) != self.d_fft_size:
            return False
        _lib.check(_lib.lib().grb200_fft_vcc_set_window(self._h, _ptr(w), len(w)))
        self.d_window = w.copy()
        return True

    def work(self, noutput_items, input_items, output_items):
        n, N = int(noutput_items), self.d_fft_size
        xin = as_items(input_items[0], _c64, n * N, "input_items[0]")
        out = as_items(output_items[0], _c64, n * N, "output_items[0]")
        _lib.check(_lib.lib().grb200_fft_vcc_work(self._h, _ptr(xin), _ptr(out), n))
        return n

    def work_magsq(self, noutput_items, input_items, output_items):
        """fft_vcc -> complex_to_mag_squared(vlen=fft_size) fused: float32 out."""
        n, N = int(noutput_items), self.d_fft_size
        xin = as_items(input_items[0], _c64, n * N, "input_items[0]")
        out = as_items(output_items[0], _f32, n * N, "output_items[0]")
        _lib.check(_lib.lib().grb200_fft_vcc_work_magsq(self._h, _ptr(xin), _ptr(out), n))
        return n

    def work_dev(self, in_dev, out_dev, nvectors, stream=0, magsq=False):
        fn = _lib.lib().grb200_fft_vcc_work_magsq_dev if magsq else _lib.lib().grb200_fft_vcc_work_dev
        _lib.check(fn(self._h, C.c_void_p(in_dev), C.c_void_p(out_dev), int(nvectors), C.c_void_p(stream)))
        return int(nvectors)

    def process(self, x, magsq=False):
        x = np.ascontiguousarray(x, np.complex64).reshape(-1)
        n = len(x) // self.d_fft_size
        out = np.empty(n * self.d_fft_size, np.float32 if magsq else np.complex64)
        (self.work_magsq if magsq else self.work)(n, [x], [out])
        return out

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                _lib.lib().grb200_fft_vcc_destroy(h)
            except Exception:
                pass


class fft_vfc(fft_vcc):
    """fft_vfc: float32 vectors in, the full fft_size-point complex transform out (fft_vfc_fftw.cc:57-92)."""

    def __init__(self, fft_size, forward, window, nthreads=1, device=0):
        L = _lib.lib()
        fft_size = int(fft_size)
        sync_block.__init__(self, "fft_vfc_fftw", io_signature.make(1, 1, max(fft_size, 0) * 4),
                            io_signature.make(1, 1, max(fft_size, 0) * 8))
        self._h = C.c_void_p()
        self.d_fft_size, self.d_forward, self.d_shift = fft_size, bool(forward), False
        self._nthreads = int(nthreads)
        w = np.ascontiguousarray(window if window is not None else [], np.float32)
        _lib.check(L.grb200_fft_vfc_create(fft_size, int(self.d_forward), _ptr(w), len(w), int(device), C.byref(self._h)))
        self.d_window = w.copy()

    @classmethod
    def make(cls, fft_size, forward, window, nthreads=1, device=0):
        return cls(fft_size, forward, window, nthreads, device)

    def work(self, noutput_items, input_items, output_items):
        n, N = int(noutput_items), self.d_fft_size
        xin = as_items(input_items[0], _f32, n * N, "input_items[0]")
        out = as_items(output_items[0], _c64, n * N, "output_items[0]")
        _lib.check(_lib.lib().grb200_fft_vcc_work(self._h, _ptr(xin), _ptr(out), n))
        return n

    def work_magsq(self, noutput_items, input_items, output_items):
        n, N = int(noutput_items), self.d_fft_size
        xin = as_items(input_items[0], _f32, n * N, "input_items[0]")
        out = as_items(output_items[0], _f32, n * N, "output_items[0]")
        _lib.check(_lib.lib().grb200_fft_vcc_work_magsq(self._h, _ptr(xin), _ptr(out), n))
        return n

    def process(self, x, magsq=False):
        x = np.ascontiguousarray(x, np.float32).reshape(-1)
        n = len(x) // self.d_fft_size
        out = np.empty(n * self.d_fft_size, np.float32 if magsq else np.complex64)
        (self.work_magsq if magsq else self.work)(n, [x], [out])
        return out


class _logpwrfft_base:
    """logpwrfft_c / logpwrfft_f: stream in, frames of fft_size float32 (dB) out (gr-fft/python/fft/logpwrfft.py:29-177).
    general_work() follows the gr::block contract (consume / produce counts); the setters are the reference's."""
    _name = "logpwrfft"
    _complex = True

    def __init__(self, sample_rate, fft_size, ref_scale, frame_rate, avg_alpha, average, win=None, device=0):
        L = _lib.lib()
        self._item = _c64 if self._complex else _f32
        self._fft_size = int(fft_size)
        if win is None:
            win = window.blackmanharris  # logpwrfft.py:54
        w = np.ascontiguousarray(win(self._fft_size), np.float32)
        self._h = C.c_void_p()
        _lib.check(L.grb200_logpwrfft_create(int(self._complex), float(sample_rate), self._fft_size, float(ref_scale),
                                             float(frame_rate), float(avg_alpha), int(bool(average)), _ptr(w), len(w),
                                             int(device), C.byref(self._h)))
        self._consumed = 0

    def name(self):
        return self._name

    def set_decimation(self, decim):
        _lib.check(_lib.lib().grb200_logpwrfft_set_decimation(self._h, float(decim)))

    def set_vec_rate(self, vec_rate):
        _lib.check(_lib.lib().grb200_logpwrfft_set_vec_rate(self._h, float(vec_rate)))

    def set_sample_rate(self, sample_rate):
        _lib.check(_lib.lib().grb200_logpwrfft_set_sample_rate(self._h, float(sample_rate)))

    def set_average(self, average):
        _lib.check(_lib.lib().grb200_logpwrfft_set_average(self._h, int(bool(average))))

    def set_avg_alpha(self, avg_alpha):
        _lib.check(_lib.lib().grb200_logpwrfft_set_avg_alpha(self._h, float(avg_alpha)))

    def sample_rate(self):
        return _lib.lib().grb200_logpwrfft_sample_rate(self._h)

    def decimation(self):
        return _lib.lib().grb200_logpwrfft_decimation(self._h)

    def frame_rate(self):
        return _lib.lib().grb200_logpwrfft_frame_rate(self._h)

    def average(self):
        return bool(_lib.lib().grb200_logpwrfft_average(self._h))

    def avg_alpha(self):
        return _lib.lib().grb200_logpwrfft_avg_alpha(self._h)

    def general_work(self, noutput_items, ninput_items, input_items, output_items):
        nin = int(ninput_items[0])
        xin = as_items(input_items[0], self._item, nin, "input_items[0]")
        out = as_items(output_items[0], _f32, int(noutput_items) * self._fft_size, "output_items[0]")
        used, made = C.c_longlong(0), C.c_longlong(0)
        _lib.check(_lib.lib().grb200_logpwrfft_work(self._h, _ptr(xin), nin, _ptr(out), int(noutput_items), C.byref(used),
                                                   C.byref(made)))
        self._consumed = used.value
        return made.value

    def work_dev(self, in_dev, nitems, out_dev, max_out, stream=0):
        used, made = C.c_longlong(0), C.c_longlong(0)
        _lib.check(_lib.lib().grb200_logpwrfft_work_dev(self._h, C.c_void_p(in_dev), int(nitems), C.c_void_p(out_dev),
                                                       int(max_out), C.byref(used), C.byref(made), C.c_void_p(stream)))
        return used.value, made.value

    def work_dev_static(self, in_dev, nitems, out_dev, max_out, stream=0):
        """work_dev over the library's static device scratch: allocates nothing (grb200.h: grb200_logpwrfft_work_dev_static;
        at most 8 * (2^18 // fft_size) kept frames per averaging call, calls that overlap on the GPU race)."""
        used, made = C.c_longlong(0), C.c_longlong(0)
        _lib.check(_lib.lib().grb200_logpwrfft_work_dev_static(self._h, C.c_void_p(in_dev), int(nitems), C.c_void_p(out_dev),
                                                              int(max_out), C.byref(used), C.byref(made), C.c_void_p(stream)))
        return used.value, made.value

    def process(self, x):
        """All whole vectors of x through one general_work(): returns frames [nframes, fft_size]."""
        x = np.ascontiguousarray(x, self._item).reshape(-1)
        nvec = len(x) // self._fft_size
        out = np.empty(max(nvec, 1) * self._fft_size, np.float32)
        n = self.general_work(nvec, [len(x)], [x], [out])
        return out[: n * self._fft_size].reshape(n, self._fft_size).copy()

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                _lib.lib().grb200_logpwrfft_destroy(h)
            except Exception:
                pass


class logpwrfft_c(_logpwrfft_base):
    _name = "logpwrfft_c"
    _complex = True


class logpwrfft_f(_logpwrfft_base):
    _name = "logpwrfft_f"
    _complex = False
