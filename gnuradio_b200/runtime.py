"""The slice of the gr::block contract the hot-path blocks rely on (host side, no compute).

Mirrors gnuradio-runtime/include/gnuradio/block.h (history, output_multiple, relative_rate, consume_each,
WORK_DONE), sync_block / sync_decimator (gnuradio-runtime/lib/sync_block.cc:44-53,
lib/sync_decimator.cc:28-54) and io_signature, so the re-hosted blocks expose the same attributes the
unmodified scheduler queries.  It is NOT a scheduler: tests/sched_mock.py plays block_executor.
"""
import threading

import numpy as np

WORK_CALLED_PRODUCE = -2
WORK_DONE = -1


class io_signature:
    def __init__(self, min_streams, max_streams, sizeof_stream_item):
        self._min, self._max, self._size = int(min_streams), int(max_streams), int(sizeof_stream_item)

    @classmethod
    def make(cls, min_streams, max_streams, sizeof_stream_item):
        return cls(min_streams, max_streams, sizeof_stream_item)

    def min_streams(self):
        return self._min

    def max_streams(self):
        return self._max

    def sizeof_stream_item(self, _index=0):
        return self._size


class block:
    """gr::block: general_work() + forecast(); fixed_rate false."""

    def __init__(self, name, in_sig, out_sig):
        self._name = name
        self._in_sig, self._out_sig = in_sig, out_sig
        self._history = 1
        self._output_multiple = 1
        self._relative_rate = 1.0
        self._consumed = 0
        self.d_setlock = threading.RLock()

    def name(self):
        return self._name

    def input_signature(self):
        return self._in_sig

    def output_signature(self):
        return self._out_sig

    def history(self):
        return self._history

    def set_history(self, h):
        self._history = int(h)

    def output_multiple(self):
        return self._output_multiple

    def set_output_multiple(self, m):
        if m < 1:
            raise ValueError("block::set_output_multiple")
        self._output_multiple = int(m)

    def relative_rate(self):
        return self._relative_rate

    def set_relative_rate(self, r):
        if r <= 0.0:
            raise ValueError("block::set_relative_rate: relative rate must be > 0.0")
        self._relative_rate = float(r)

    def fixed_rate(self):
        return False

    def consume_each(self, n):
        self._consumed = int(n)

    def forecast(self, noutput_items, ninputs):
        # default gr::block::forecast: 1:1 plus history
        return [noutput_items + self.history() - 1] * ninputs


class sync_block(block):
    def fixed_rate(self):
        return True

    def decimation(self):
        return 1

    def forecast(self, noutput_items, ninputs):
        return [self.fixed_rate_noutput_to_ninput(noutput_items)] * ninputs

    def fixed_rate_noutput_to_ninput(self, noutput_items):
        return noutput_items * self.decimation() + self.history() - 1

    def fixed_rate_ninput_to_noutput(self, ninput_items):
        return max(0, ninput_items - self.history() + 1) // self.decimation()

    def general_work(self, noutput_items, ninput_items, input_items, output_items):
        r = self.work(noutput_items, input_items, output_items)
        if r > 0:
            self.consume_each(r * self.decimation())
        else:
            self.consume_each(0)
        return r


class sync_decimator(sync_block):
    def __init__(self, name, in_sig, out_sig, decimation):
        super().__init__(name, in_sig, out_sig)
        self._decimation = int(decimation)
        self.set_relative_rate(1.0 / decimation)

    def decimation(self):
        return self._decimation


class sync_interpolator(sync_block):
    """gr::sync_interpolator: noutput is a multiple of the interpolation, ninput = noutput / interpolation."""

    def __init__(self, name, in_sig, out_sig, interpolation):
        super().__init__(name, in_sig, out_sig)
        self._interpolation = int(interpolation)
        self.set_relative_rate(float(interpolation))
        self.set_output_multiple(self._interpolation)

    def interpolation(self):
        return self._interpolation

    def forecast(self, noutput_items, ninputs):
        return [self.fixed_rate_noutput_to_ninput(noutput_items)] * ninputs

    def fixed_rate_noutput_to_ninput(self, noutput_items):
        return noutput_items // self._interpolation + self.history() - 1

    def fixed_rate_ninput_to_noutput(self, ninput_items):
        return max(0, ninput_items - self.history() + 1) * self._interpolation

    def general_work(self, noutput_items, ninput_items, input_items, output_items):
        r = self.work(noutput_items, input_items, output_items)
        self.consume_each(r // self._interpolation if r > 0 else 0)
        return r


def as_items(a, dtype, min_items, what):
    """Borrowed host buffer check: contiguous, right dtype, at least min_items items."""
    if not isinstance(a, np.ndarray) or a.dtype != dtype or not a.flags["C_CONTIGUOUS"]:
        raise TypeError("%s must be a C-contiguous numpy array of %s" % (what, np.dtype(dtype).name))
    if a.size < min_items:
        raise ValueError("%s holds %d items, %d needed" % (what, a.size, min_items))
    return a
