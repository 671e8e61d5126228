"""gr-fft hot-path block re-hosted on libgrb200.

  fft_vcc   gr-fft/include/gnuradio/fft/fft_vcc.h:47-70, gr-fft/lib/fft_vcc_fftw.cc:33-122
            (make(fft_size, forward, window, shift=False, nthreads=1), set_window -> bool, work per vector)
"""
import ctypes as C

import numpy as np

from . import _lib
from .runtime import as_items, io_signature, sync_block

_c64 = np.dtype(np.complex64)
_f32 = np.dtype(np.float32)


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
