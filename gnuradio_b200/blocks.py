"""gr-blocks pieces adjacent to the hot path, re-hosted on libgrb200.

  complex_to_mag_squared   gr-blocks/lib/complex_to_mag_squared_impl.cc:22-48 (make(vlen=1), work)
"""
import ctypes as C

import numpy as np

from . import _lib
from .runtime import as_items, io_signature, sync_block


class complex_to_mag_squared(sync_block):
    def __init__(self, vlen=1, device=0):
        _lib.lib()
        self.d_vlen = int(vlen)
        self._device = int(device)
        sync_block.__init__(self, "complex_to_mag_squared", io_signature.make(1, 1, 8 * self.d_vlen),
                            io_signature.make(1, 1, 4 * self.d_vlen))

    @classmethod
    def make(cls, vlen=1, device=0):
        return cls(vlen, device)

    def work(self, noutput_items, input_items, output_items):
        n = int(noutput_items) * self.d_vlen
        xin = as_items(input_items[0], np.dtype(np.complex64), n, "input_items[0]")
        out = as_items(output_items[0], np.dtype(np.float32), n, "output_items[0]")
        _lib.check(_lib.lib().grb200_magsq_work(self._device, xin.ctypes.data_as(C.c_void_p),
                                                out.ctypes.data_as(C.c_void_p), n))
        return int(noutput_items)

    def work_dev(self, in_dev, out_dev, nitems, stream=0):
        _lib.check(_lib.lib().grb200_magsq_work_dev(self._device, C.c_void_p(in_dev), C.c_void_p(out_dev),
                                                    int(nitems) * self.d_vlen, C.c_void_p(stream)))
        return int(nitems)

    def process(self, x):
        x = np.ascontiguousarray(x, np.complex64).reshape(-1)
        out = np.empty(len(x), np.float32)
        self.work(len(x) // self.d_vlen, [x], [out])
        return out
