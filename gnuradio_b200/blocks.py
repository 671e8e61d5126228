"""gr-blocks pieces adjacent to the hot path, re-hosted on libgrb200.

  complex_to_mag_squared   gr-blocks/lib/complex_to_mag_squared_impl.cc:22-48 (make(vlen=1), work)
  stream_to_vector         gr-blocks/lib/stream_to_vector_impl.cc:25-58   (a memcpy: on the device, the same pointer)
  vector_to_stream         gr-blocks/lib/vector_to_stream_impl.cc         (the same, read backwards)
  vector_to_streams        gr-blocks/lib/vector_to_streams_impl.cc:25-55  (out_j[n] = in[n*nstreams + j])
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


class stream_to_vector(sync_block):
    """stream_to_vector(itemsize, nitems_per_block): a sync_decimator whose work() is memcpy(out, in, noutput*block_size)
    (stream_to_vector_impl.cc:46-58).  The bytes do not move: work_dev() returns the input address -- chain the next device
    block straight onto the upstream buffer."""

    def __init__(self, itemsize, nitems_per_block):
        self.d_itemsize, self.d_n = int(itemsize), int(nitems_per_block)
        sync_block.__init__(self, "stream_to_vector", io_signature.make(1, 1, self.d_itemsize),
                            io_signature.make(1, 1, self.d_itemsize * self.d_n))

    @classmethod
    def make(cls, itemsize, nitems_per_block):
        return cls(itemsize, nitems_per_block)

    def decimation(self):
        return self.d_n

    def work(self, noutput_items, input_items, output_items):
        nbytes = int(noutput_items) * self.d_itemsize * self.d_n
        src = input_items[0].reshape(-1).view(np.uint8)
        dst = output_items[0].reshape(-1).view(np.uint8)
        dst[:nbytes] = src[:nbytes]
        return int(noutput_items)

    def work_dev(self, in_dev, noutput_items=0, stream=0):
        return in_dev  # zero copy: a stream of items is the stream of vectors over it


class vector_to_stream(stream_to_vector):
    """vector_to_stream(itemsize, nitems_per_block): the same bytes read as items again (a sync_interpolator memcpy)."""

    def __init__(self, itemsize, nitems_per_block):
        stream_to_vector.__init__(self, itemsize, nitems_per_block)
        self._name = "vector_to_stream"

    def work(self, noutput_items, input_items, output_items):
        nbytes = int(noutput_items) * self.d_itemsize
        output_items[0].reshape(-1).view(np.uint8)[:nbytes] = input_items[0].reshape(-1).view(np.uint8)[:nbytes]
        return int(noutput_items)


class vector_to_streams(sync_block):
    """vector_to_streams(itemsize, nstreams): one stream of nstreams-item vectors -> nstreams item streams (a transpose,
    vector_to_streams_impl.cc:34-55).  itemsize 4 (float) or 8 (gr_complex) on the device."""

    def __init__(self, itemsize, nstreams, device=0):
        _lib.lib()
        self.d_itemsize, self.d_nstreams, self._device = int(itemsize), int(nstreams), int(device)
        sync_block.__init__(self, "vector_to_streams", io_signature.make(1, 1, self.d_itemsize * self.d_nstreams),
                            io_signature.make(self.d_nstreams, self.d_nstreams, self.d_itemsize))

    @classmethod
    def make(cls, itemsize, nstreams, device=0):
        return cls(itemsize, nstreams, device)

    def work(self, noutput_items, input_items, output_items):
        n = int(noutput_items)
        dt = np.dtype(np.complex64) if self.d_itemsize == 8 else np.dtype(np.float32)
        xin = as_items(input_items[0], dt, n * self.d_nstreams, "input_items[0]")
        outs = [as_items(o, dt, n, "output_items[%d]" % j) for j, o in enumerate(output_items)]
        if len(outs) != self.d_nstreams:
            raise ValueError("vector_to_streams: %d output streams, %d needed" % (len(outs), self.d_nstreams))
        outp = (C.c_void_p * self.d_nstreams)(*[o.ctypes.data for o in outs])
        _lib.check(_lib.lib().grb200_vector_to_streams_work(self._device, self.d_itemsize, self.d_nstreams,
                                                            xin.ctypes.data_as(C.c_void_p), outp, n))
        return n

    def work_dev(self, in_dev, out_dev, out_stride, noutput_items, stream=0):
        """Stream j lands at out_dev + j*out_stride items."""
        _lib.check(_lib.lib().grb200_vector_to_streams_work_dev(self._device, self.d_itemsize, self.d_nstreams,
                                                                C.c_void_p(in_dev), C.c_void_p(out_dev), int(out_stride),
                                                                int(noutput_items), C.c_void_p(stream)))
        return int(noutput_items)

    def process(self, x):
        dt = np.complex64 if self.d_itemsize == 8 else np.float32
        x = np.ascontiguousarray(x, dt).reshape(-1)
        n = len(x) // self.d_nstreams
        outs = np.empty((self.d_nstreams, n), dt)
        self.work(n, [x], [outs[j] for j in range(self.d_nstreams)])
        return outs
