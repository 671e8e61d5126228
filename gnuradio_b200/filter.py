"""gr-filter hot-path blocks re-hosted on libgrb200 (same names, arguments and error behaviour as the reference).

  fir_filter_ccf / ccc / fff        gr-filter/include/gnuradio/filter/fir_filter_blk.h:46-68,
                                    gr-filter/lib/fir_filter_blk_impl.cc:22-90
  freq_xlating_fir_filter_ccf/ccc   gr-filter/include/gnuradio/filter/freq_xlating_fir_filter.h:56-90,
                                    gr-filter/lib/freq_xlating_fir_filter_impl.cc:26-180
  fft_filter_ccc                    gr-filter/include/gnuradio/filter/fft_filter_ccc.h:43-70,
                                    gr-filter/lib/fft_filter_ccc_impl.cc:22-78
  pfb_channelizer_ccf               gr-filter/include/gnuradio/filter/pfb_channelizer_ccf.h:100-192,
                                    gr-filter/lib/pfb_channelizer_ccf_impl.cc:34-193
  pfb.channelizer_ccf               gr-filter/python/filter/pfb.py:23-50 (hier block; stream_to_streams absorbed)
  interp_fir_filter_ccf/ccc         gr-filter/include/gnuradio/filter/interp_fir_filter.h:43-72,
                                    gr-filter/lib/interp_fir_filter_impl.cc:24-152
  rational_resampler_base_ccf/ccc   gr-filter/include/gnuradio/filter/rational_resampler_base.h,
                                    gr-filter/lib/rational_resampler_base_impl.cc:24-176
  pfb_decimator_ccf                 gr-filter/include/gnuradio/filter/pfb_decimator_ccf.h:94-138,
                                    gr-filter/lib/pfb_decimator_ccf_impl.cc:24-257

work()/general_work() take borrowed HOST numpy buffers like the scheduler's; *_dev variants take raw
device pointers (ints, e.g. torch.Tensor.data_ptr()) for chained GPU blocks.  All arithmetic runs in the
CUDA library; this module holds no compute.
"""
import ctypes as C

import numpy as np

from . import _lib
from .runtime import as_items, block, io_signature, sync_block, sync_decimator, sync_interpolator

_c64 = np.dtype(np.complex64)


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class _fir_filter_blk(sync_decimator):
    _complex_taps = False
    _item = _c64  # stream item type (fir_filter_fff: float32)
    _name = "fir_filter"

    def __init__(self, decimation, taps, device=0):
        L = _lib.lib()
        if int(decimation) < 1:
            raise ValueError("fir_filter: decimation must be >= 1")
        isz = self._item.itemsize
        sync_decimator.__init__(self, self._name, io_signature.make(1, 1, isz), io_signature.make(1, 1, isz), decimation)
        self._device = int(device)
        self._h = C.c_void_p()
        t = self._conv_taps(taps)
        create = (L.grb200_fir_fff_create if self._item is not _c64 else
                  L.grb200_fir_ccc_create if self._complex_taps else L.grb200_fir_ccf_create)
        _lib.check(create(int(decimation), _ptr(t), len(t), self._device, C.byref(self._h)))
        self._taps = t.copy()
        self.d_updated = False
        self.set_history(len(t))  # fir_filter_blk_impl.cc:41

    @classmethod
    def make(cls, decimation, taps, device=0):
        return cls(decimation, taps, device)

    def _conv_taps(self, taps):
        t = np.ascontiguousarray(taps, np.complex64 if self._complex_taps else np.float32)
        if t.ndim != 1 or t.size < 1:
            raise ValueError("fir_filter: at least one tap is required")
        return t

    def set_taps(self, taps):
        t = self._conv_taps(taps)
        with self.d_setlock:  # fir_filter_blk_impl.cc:53-59
            _lib.check(_lib.lib().grb200_fir_set_taps(self._h, _ptr(t), len(t)))
            self._taps = t.copy()
            self.d_updated = True

    def taps(self):
        return self._taps.tolist()

    def work(self, noutput_items, input_items, output_items):
        with self.d_setlock:
            if self.d_updated:  # fir_filter_blk_impl.cc:77-81
                self.set_history(len(self._taps))
                self.d_updated = False
                return 0
            n = int(noutput_items)
            need = n * self.decimation() + len(self._taps) - 1
            xin = as_items(input_items[0], self._item, need if n else 0, "input_items[0]")
            out = as_items(output_items[0], self._item, n, "output_items[0]")
            _lib.check(_lib.lib().grb200_fir_work(self._h, _ptr(xin), _ptr(out), n))
            return n

    def work_dev(self, in_dev, out_dev, noutput_items, stream=0):
        """Device-resident mode: in_dev/out_dev are device addresses; asynchronous on `stream`."""
        _lib.check(_lib.lib().grb200_fir_work_dev(self._h, C.c_void_p(in_dev), C.c_void_p(out_dev),
                                                  int(noutput_items), C.c_void_p(stream)))
        return int(noutput_items)

    def set_halo_source(self, peer_tail):
        """Time-sharded stream: work_dev() fetches the ntaps-1 history items in front of its input from `peer_tail` (the left
        neighbour's newest items, a device address valid in this process -- sharding.PeerHalo) itself.  None clears."""
        _lib.check(_lib.lib().grb200_fir_set_halo_source(self._h, C.c_void_p(peer_tail or 0)))

    def general_work_lag(self, noutput_items, input_items, output_items):
        """Lag-by-one host mode (include/grb200.h "Host staging"): takes the input of `noutput_items` outputs now, writes the
        PREVIOUS call's outputs into output_items[0] and returns how many -- 0 on the first call.  consume = this call's
        noutput_items * decimation (self._consumed), like a gr::block that consumes and produces independently."""
        with self.d_setlock:
            if self.d_updated:
                self.set_history(len(self._taps))
                self.d_updated = False
                self._consumed = 0
                return 0
            n = int(noutput_items)
            need = n * self.decimation() + len(self._taps) - 1
            xin = as_items(input_items[0], self._item, need if n else 0, "input_items[0]")
            out = output_items[0]
            got = C.c_longlong(0)
            _lib.check(_lib.lib().grb200_fir_work_lag(self._h, _ptr(xin), n, _ptr(out), out.size, C.byref(got)))
            self._consumed = n * self.decimation()
            return got.value

    def flush(self, output_items):
        """End of stream in lag mode: the outputs of the call still in flight."""
        out = output_items[0]
        got = C.c_longlong(0)
        _lib.check(_lib.lib().grb200_fir_flush(self._h, _ptr(out), out.size, C.byref(got)))
        return got.value

    def process(self, xin, noutput_items):
        """Convenience: one work() call on `xin` (history items first); returns the output array."""
        xin = np.ascontiguousarray(xin, self._item)
        out = np.empty(int(noutput_items), self._item)
        if self.d_updated:
            self.work(0, [xin], [out])
        r = self.work(noutput_items, [xin], [out])
        return out[:r]

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                _lib.lib().grb200_fir_destroy(h)
            except Exception:
                pass


class fir_filter_ccf(_fir_filter_blk):
    _complex_taps = False
    _name = "fir_filter_ccf"


class fir_filter_ccc(_fir_filter_blk):
    _complex_taps = True
    _name = "fir_filter_ccc"


class fir_filter_fff(_fir_filter_blk):
    """fir_filter_blk<float,float,float> (gr-filter/include/gnuradio/filter/fir_filter_blk.h:66): float stream, float taps."""
    _complex_taps = False
    _item = np.dtype(np.float32)
    _name = "fir_filter_fff"


class _freq_xlating_fir_filter(sync_decimator):
    """freq_xlating_fir_filter: band-pass FIR (prototype taps moved up to center_freq), decimate, derotate.

    set_taps()/set_center_freq() only mark the block (freq_xlating_fir_filter_impl.cc:96-113); the next work()
    rebuilds the composite taps and returns 0 (:155-160).  The rotator phase runs on across work() calls."""
    _complex_taps = False
    _name = "freq_xlating_fir_filter"

    def __init__(self, decimation, taps, center_freq, sampling_freq, device=0):
        L = _lib.lib()
        if int(decimation) < 1:
            raise ValueError("freq_xlating_fir_filter: decimation must be >= 1")
        sync_decimator.__init__(self, self._name, io_signature.make(1, 1, 8), io_signature.make(1, 1, 8), decimation)
        self._device = int(device)
        self._h = C.c_void_p()
        t = self._conv_taps(taps)
        create = L.grb200_freq_xlating_fir_ccc_create if self._complex_taps else L.grb200_freq_xlating_fir_ccf_create
        _lib.check(create(int(decimation), _ptr(t), len(t), float(center_freq), float(sampling_freq), self._device,
                          C.byref(self._h)))
        self._proto_taps = t.copy()
        self._center_freq = float(center_freq)
        self._sampling_freq = float(sampling_freq)
        self.d_updated = False
        self.set_history(len(t))  # freq_xlating_fir_filter_impl.cc:53

    @classmethod
    def make(cls, decimation, taps, center_freq, sampling_freq, device=0):
        return cls(decimation, taps, center_freq, sampling_freq, device)

    def _conv_taps(self, taps):
        t = np.ascontiguousarray(taps, np.complex64 if self._complex_taps else np.float32)
        if t.ndim != 1 or t.size < 1:
            raise ValueError("freq_xlating_fir_filter: at least one tap is required")
        return t

    def set_center_freq(self, center_freq):
        self._center_freq = float(center_freq)
        self.d_updated = True

    def center_freq(self):
        return self._center_freq

    def set_taps(self, taps):
        self._proto_taps = self._conv_taps(taps).copy()
        self.d_updated = True

    def taps(self):
        return self._proto_taps.tolist()

    def handle_set_center_freq(self, msg):
        """The "freq" message port (freq_xlating_fir_filter_impl.cc:121-141): a dict with key "freq" or a
        (key, value) pair; anything else is ignored."""
        if isinstance(msg, dict):
            v = msg.get("freq")
        elif isinstance(msg, tuple) and len(msg) == 2:
            v = msg[1]
        else:
            return
        if isinstance(v, (int, float)) and not isinstance(v, bool):
            self.set_center_freq(float(v))

    def advance(self, noutput_items):
        """Move the rotator on by noutput_items outputs without data: a shard that starts at output item n0 of the
        stream (sharding.py) calls advance(n0) once so that its phase continues the global stream's."""
        _lib.check(_lib.lib().grb200_freq_xlating_fir_advance(self._h, int(noutput_items)))

    def composite_taps(self):
        t = np.empty(len(self._proto_taps), np.complex64)
        _lib.check(_lib.lib().grb200_freq_xlating_fir_composite_taps(self._h, _ptr(t), len(t)))
        return t

    def _rebuild(self):
        L = _lib.lib()
        # the library rebuilds on each call; push the frequency first so one rebuild sees both new values
        if len(self._proto_taps) == L.grb200_freq_xlating_fir_ntaps(self._h):
            _lib.check(L.grb200_freq_xlating_fir_set_center_freq(self._h, self._center_freq))
            _lib.check(L.grb200_freq_xlating_fir_set_taps(self._h, _ptr(self._proto_taps), len(self._proto_taps)))
        else:
            _lib.check(L.grb200_freq_xlating_fir_set_taps(self._h, _ptr(self._proto_taps), len(self._proto_taps)))
            _lib.check(L.grb200_freq_xlating_fir_set_center_freq(self._h, self._center_freq))

    def work(self, noutput_items, input_items, output_items):
        if self.d_updated:  # freq_xlating_fir_filter_impl.cc:155-160
            self.set_history(len(self._proto_taps))
            self._rebuild()
            self.d_updated = False
            return 0
        n = int(noutput_items)
        need = n * self.decimation() + len(self._proto_taps) - 1
        xin = as_items(input_items[0], _c64, need if n else 0, "input_items[0]")
        out = as_items(output_items[0], _c64, n, "output_items[0]")
        _lib.check(_lib.lib().grb200_freq_xlating_fir_work(self._h, _ptr(xin), _ptr(out), n))
        return n

    def work_dev(self, in_dev, out_dev, noutput_items, stream=0):
        """Device-resident mode: in_dev/out_dev are device addresses; asynchronous on `stream`."""
        if self.d_updated:
            self._rebuild()
            self.d_updated = False
        _lib.check(_lib.lib().grb200_freq_xlating_fir_work_dev(self._h, C.c_void_p(in_dev), C.c_void_p(out_dev),
                                                               int(noutput_items), C.c_void_p(stream)))
        return int(noutput_items)

    def process(self, xin, noutput_items):
        """Convenience: one work() call on `xin` (history items first); returns the output array."""
        xin = np.ascontiguousarray(xin, np.complex64)
        out = np.empty(int(noutput_items), np.complex64)
        if self.d_updated:
            self.work(0, [xin], [out])
        r = self.work(noutput_items, [xin], [out])
        return out[:r]

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                _lib.lib().grb200_freq_xlating_fir_destroy(h)
            except Exception:
                pass


class freq_xlating_fir_filter_ccf(_freq_xlating_fir_filter):
    _complex_taps = False
    _name = "freq_xlating_fir_filter_ccf"


class freq_xlating_fir_filter_ccc(_freq_xlating_fir_filter):
    _complex_taps = True
    _name = "freq_xlating_fir_filter_ccc"


class _interp_fir_filter(sync_interpolator):
    """interp_fir_filter: `interpolation` polyphase FIRs over one input stream, outputs interleaved.  set_taps() is
    picked up by the next work(), which installs the taps, re-declares the history and returns 0
    (interp_fir_filter_impl.cc:72-90, 136-139)."""
    _complex_taps = False
    _name = "interp_fir_filter"

    def __init__(self, interpolation, taps, device=0):
        L = _lib.lib()
        if int(interpolation) <= 0:
            raise IndexError("interp_fir_filter_impl: interpolation must be > 0")  # std::out_of_range, :46-48
        t = self._conv_taps(taps)
        sync_interpolator.__init__(self, self._name, io_signature.make(1, 1, 8), io_signature.make(1, 1, 8),
                                   interpolation)
        self._h = C.c_void_p()
        create = L.grb200_interp_fir_ccc_create if self._complex_taps else L.grb200_interp_fir_ccf_create
        _lib.check(create(int(interpolation), _ptr(t), len(t), int(device), C.byref(self._h)))
        self._new_taps = self._padded(t)
        self.d_updated = False
        self.set_history(L.grb200_interp_fir_taps_per_filter(self._h))  # install_taps, :115

    @classmethod
    def make(cls, interpolation, taps, device=0):
        return cls(interpolation, taps, device)

    def _conv_taps(self, taps):
        t = np.ascontiguousarray(taps, np.complex64 if self._complex_taps else np.float32)
        if t.ndim != 1 or t.size < 1:
            raise RuntimeError("interp_fir_filter_impl: no filter taps provided.")  # std::runtime_error, :50-52
        return t

    def _padded(self, t):  # :78-84: rounded up to a multiple of the interpolation with zeros
        n = (-len(t)) % self.interpolation()
        return np.concatenate([t, np.zeros(n, t.dtype)])

    def set_taps(self, taps):
        self._new_taps = self._padded(self._conv_taps(taps))
        self.d_updated = True

    def taps(self):
        return self._new_taps.tolist()  # d_new_taps, padding included (:120-124)

    def work(self, noutput_items, input_items, output_items):
        L = _lib.lib()
        if self.d_updated:  # :136-139
            _lib.check(L.grb200_interp_fir_set_taps(self._h, _ptr(self._new_taps), len(self._new_taps)))
            self.set_history(L.grb200_interp_fir_taps_per_filter(self._h))
            self.d_updated = False
            return 0
        n = int(noutput_items)
        ni = n // self.interpolation()
        n = ni * self.interpolation()
        xin = as_items(input_items[0], _c64, ni + self.history() - 1 if n else 0, "input_items[0]")
        out = as_items(output_items[0], _c64, n, "output_items[0]")
        _lib.check(L.grb200_interp_fir_work(self._h, _ptr(xin), _ptr(out), n))
        return n

    def work_dev(self, in_dev, out_dev, noutput_items, stream=0):
        """Device-resident mode: in_dev/out_dev are device addresses; asynchronous on `stream`."""
        _lib.check(_lib.lib().grb200_interp_fir_work_dev(self._h, C.c_void_p(in_dev), C.c_void_p(out_dev),
                                                         int(noutput_items), C.c_void_p(stream)))
        return int(noutput_items)

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                _lib.lib().grb200_interp_fir_destroy(h)
            except Exception:
                pass


class interp_fir_filter_ccf(_interp_fir_filter):
    _complex_taps = False
    _name = "interp_fir_filter_ccf"


class interp_fir_filter_ccc(_interp_fir_filter):
    _complex_taps = True
    _name = "interp_fir_filter_ccc"


class _rational_resampler_base(block):
    """rational_resampler_base: `interpolation` polyphase FIRs walked by a phase counter that advances `decimation` per
    output (rational_resampler_base_impl.cc:141-173).  The reference class shadows history() (…_impl.h:48-49): the
    scheduler sees history 1 and puts no zeros in front of the stream; history() here returns the taps per phase like the
    shadowing accessor does.  set_taps() is picked up by the next general_work(), which installs the taps and returns 0
    (:151-154)."""
    _complex_taps = False
    _name = "rational_resampler_base"

    def __init__(self, interpolation, decimation, taps, device=0):
        L = _lib.lib()
        if int(interpolation) <= 0:  # std::out_of_range, :46-48
            raise IndexError("rational_resampler_base_impl<IN_T,OUT_T,TAP_T>: interpolation must be > 0")
        if int(decimation) <= 0:  # std::out_of_range, :49-51
            raise IndexError("rational_resampler_base_impl<IN_T,OUT_T,TAP_T>: decimation must be > 0")
        block.__init__(self, self._name, io_signature.make(1, 1, 8), io_signature.make(1, 1, 8))
        self._interpolation, self._decimation = int(interpolation), int(decimation)
        self.set_relative_rate(self._interpolation / self._decimation)  # :53
        self.set_output_multiple(1)
        t = self._conv_taps(taps)
        self._h = C.c_void_p()
        create = L.grb200_rational_resampler_ccc_create if self._complex_taps else L.grb200_rational_resampler_ccf_create
        _lib.check(create(self._interpolation, self._decimation, _ptr(t), len(t), int(device), C.byref(self._h)))
        self._new_taps = self._padded(t)
        self.d_updated = False
        self._nt = L.grb200_rational_resampler_taps_per_filter(self._h)  # install_taps, :115

    @classmethod
    def make(cls, interpolation, decimation, taps, device=0):
        return cls(interpolation, decimation, taps, device)

    def _conv_taps(self, taps):
        t = np.ascontiguousarray(taps, np.complex64 if self._complex_taps else np.float32)
        if t.ndim != 1 or t.size < 1:
            raise RuntimeError("rational_resampler_base_impl: no filter taps provided.")
        return t

    def _padded(self, t):  # :83-90: rounded up to a multiple of the interpolation with zeros
        n = (-len(t)) % self._interpolation
        return np.concatenate([t, np.zeros(n, t.dtype)])

    def interpolation(self):
        return self._interpolation

    def decimation(self):
        return self._decimation

    def history(self):  # the shadowing accessor of rational_resampler_base_impl.h:48
        return self._nt

    def set_history(self, h):
        self._nt = int(h)

    def set_taps(self, taps):
        self._new_taps = self._padded(self._conv_taps(taps))
        self.d_updated = True

    def taps(self):
        return self._new_taps.tolist()  # d_new_taps, padding included (:120-124)

    def forecast(self, noutput_items, ninputs):  # :128-138
        nreqd = max(1, int((noutput_items + 1) * self._decimation / self._interpolation) + self.history() - 1)
        return [nreqd] * ninputs

    def general_work(self, noutput_items, ninput_items, input_items, output_items):
        L = _lib.lib()
        if self.d_updated:  # :151-154
            _lib.check(L.grb200_rational_resampler_set_taps(self._h, _ptr(self._new_taps), len(self._new_taps)))
            self.set_history(L.grb200_rational_resampler_taps_per_filter(self._h))
            self.d_updated = False
            return 0
        xin = as_items(input_items[0], _c64, 0, "input_items[0]")
        # the reference's loop runs while count < ninput_items[0] and filter() reads history() items from there -- past
        # ninput_items[0] into the circular buffer; a borrowed array ends, so the walk stops where the array does
        nin = max(0, min(int(ninput_items[0]), xin.size - (self.history() - 1)))
        out = as_items(output_items[0], _c64, int(noutput_items), "output_items[0]")
        consumed, produced = C.c_longlong(0), C.c_longlong(0)
        _lib.check(L.grb200_rational_resampler_work(self._h, _ptr(xin), nin, _ptr(out), int(noutput_items),
                                                    C.byref(consumed), C.byref(produced)))
        self.consume_each(consumed.value)
        return produced.value

    def work_dev(self, in_dev, ninput_items, out_dev, noutput_items, stream=0):
        """Device-resident mode: in_dev (ninput_items + history()-1 readable items) / out_dev are device addresses;
        asynchronous on `stream`.  Returns (consumed, produced)."""
        consumed, produced = C.c_longlong(0), C.c_longlong(0)
        _lib.check(_lib.lib().grb200_rational_resampler_work_dev(
            self._h, C.c_void_p(in_dev), int(ninput_items), C.c_void_p(out_dev), int(noutput_items), C.byref(consumed),
            C.byref(produced), C.c_void_p(stream)))
        return consumed.value, produced.value

    def ctr(self):
        return int(_lib.lib().grb200_rational_resampler_ctr(self._h))

    def set_ctr(self, ctr):
        """Phase counter hand-over for a time-sharded stream (d_ctr, :156,171)."""
        _lib.check(_lib.lib().grb200_rational_resampler_set_ctr(self._h, int(ctr)))

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                _lib.lib().grb200_rational_resampler_destroy(h)
            except Exception:
                pass


class rational_resampler_base_ccf(_rational_resampler_base):
    _complex_taps = False
    _name = "rational_resampler_base_ccf"


class rational_resampler_base_ccc(_rational_resampler_base):
    _complex_taps = True
    _name = "rational_resampler_base_ccc"


class fft_filter_ccc(sync_decimator):
    _name = "fft_filter_ccc"
    _create = "grb200_fftfilt_ccc_create"
    _set_taps = "grb200_fftfilt_set_taps"
    _tap_dtype = np.complex64

    def __init__(self, decimation, taps, nthreads=1, device=0):
        L = _lib.lib()
        if int(decimation) < 1:
            raise ValueError("%s: decimation must be >= 1" % self._name)
        sync_decimator.__init__(self, self._name, io_signature.make(1, 1, 8), io_signature.make(1, 1, 8),
                                decimation)
        self._h = C.c_void_p()
        t = self._conv_taps(taps)
        _lib.check(getattr(L, self._create)(int(decimation), _ptr(t), len(t), int(device), C.byref(self._h)))
        self._taps = t.copy()
        self._new_taps = t.copy()
        self._nthreads = int(nthreads)
        self.d_updated = False
        self.set_history(1)  # fft_filter_ccc_impl.cc:42
        self.d_nsamples = L.grb200_fftfilt_nsamples(self._h)
        self.set_output_multiple(self.d_nsamples)  # :46

    @classmethod
    def make(cls, decimation, taps, nthreads=1, device=0):
        return cls(decimation, taps, nthreads, device)

    @classmethod
    def _conv_taps(cls, taps):
        t = np.ascontiguousarray(taps, cls._tap_dtype)
        if t.ndim != 1 or t.size < 1:
            raise ValueError("%s: at least one tap is required" % cls._name)
        return t

    def set_taps(self, taps):  # fft_filter_ccc_impl.cc:49-53: deferred to the next work()
        self._new_taps = self._conv_taps(taps).copy()
        self.d_updated = True

    def taps(self):
        return self._new_taps.tolist()

    def set_nthreads(self, n):  # FFTW threads in the reference; the device transform has no such knob
        self._nthreads = int(n)

    def nthreads(self):
        return self._nthreads

    def nsamples(self):
        return self.d_nsamples

    def work(self, noutput_items, input_items, output_items):
        L = _lib.lib()
        if self.d_updated:  # fft_filter_ccc_impl.cc:68-73
            t = self._new_taps
            _lib.check(getattr(L, self._set_taps)(self._h, _ptr(t), len(t)))
            self._taps = t.copy()
            self.d_nsamples = L.grb200_fftfilt_nsamples(self._h)
            self.d_updated = False
            self.set_output_multiple(self.d_nsamples)
            return 0
        n = int(noutput_items)
        xin = as_items(input_items[0], _c64, n * self.decimation(), "input_items[0]")
        out = as_items(output_items[0], _c64, n, "output_items[0]")
        _lib.check(L.grb200_fftfilt_work(self._h, _ptr(xin), _ptr(out), n))
        return n

    def work_dev(self, in_dev, out_dev, noutput_items, stream=0):
        """Stateless device form: in_dev[0] is the oldest of ntaps-1 history items."""
        _lib.check(_lib.lib().grb200_fftfilt_work_dev(self._h, C.c_void_p(in_dev), C.c_void_p(out_dev),
                                                      int(noutput_items), C.c_void_p(stream)))
        return int(noutput_items)

    def process(self, x, noutput_items):
        x = np.ascontiguousarray(x, np.complex64)
        out = np.empty(int(noutput_items), np.complex64)
        if self.d_updated:
            self.work(0, [x], [out])
        r = self.work(noutput_items, [x], [out])
        return out[:r]

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                _lib.lib().grb200_fftfilt_destroy(h)
            except Exception:
                pass


class fft_filter_ccf(fft_filter_ccc):
    """gr::filter::fft_filter_ccf (gr-filter/lib/fft_filter_ccf_impl.cc:32-78 over kernel::fft_filter_ccf,
    fft_filter.cc:291-419): real taps on a complex stream; same geometry, state and work() contract as ccc."""
    _name = "fft_filter_ccf"
    _create = "grb200_fftfilt_ccf_create"
    _set_taps = "grb200_fftfilt_set_taps_real"
    _tap_dtype = np.float32


class pfb_channelizer_ccf(block):
    def __init__(self, numchans, taps, oversample_rate=1.0, device=0):
        L = _lib.lib()
        numchans = int(numchans)
        block.__init__(self, "pfb_channelizer_ccf", io_signature.make(numchans, numchans, 8),
                       io_signature.make(1, numchans, 8))
        self._h = C.c_void_p()
        self._M = numchans
        t = self._conv_taps(taps)
        _lib.check(L.grb200_pfb_chan_ccf_create(numchans, _ptr(t), len(t), float(oversample_rate), int(device),
                                                C.byref(self._h)))
        self.d_oversample_rate = float(oversample_rate)
        self.set_relative_rate(self.d_oversample_rate)  # pfb_channelizer_ccf_impl.cc:55
        self.set_output_multiple(L.grb200_pfb_chan_output_multiple(self._h))  # :81-84
        self.d_taps_per_filter = L.grb200_pfb_chan_taps_per_filter(self._h)
        self.set_history(self.d_taps_per_filter + 1)  # :101
        self.d_updated = True  # the constructor goes through set_taps (:87, :102)

    @classmethod
    def make(cls, numchans, taps, oversample_rate=1.0, device=0):
        return cls(numchans, taps, oversample_rate, device)

    @staticmethod
    def _conv_taps(taps):
        t = np.ascontiguousarray(taps, np.float32)
        if t.ndim != 1 or t.size < 1:
            raise ValueError("pfb_channelizer_ccf: at least one tap is required")
        return t

    def set_taps(self, taps):
        t = self._conv_taps(taps)
        with self.d_setlock:  # :96-103
            L = _lib.lib()
            _lib.check(L.grb200_pfb_chan_set_taps(self._h, _ptr(t), len(t)))
            self.d_taps_per_filter = L.grb200_pfb_chan_taps_per_filter(self._h)
            self.set_history(self.d_taps_per_filter + 1)
            self.d_updated = True

    def taps(self):
        out = np.empty((self._M, self.d_taps_per_filter), np.float32)
        _lib.check(_lib.lib().grb200_pfb_chan_branch_taps(self._h, _ptr(out)))
        return out.tolist()

    def print_taps(self):  # polyphase_filterbank::print_taps
        for i, row in enumerate(self.taps()):
            print("filter[%d]: [%s]" % (i, " ".join("%.4e" % v for v in row)))

    def set_channel_map(self, chan_map):
        m = np.ascontiguousarray(chan_map, np.int32)
        with self.d_setlock:
            _lib.check(_lib.lib().grb200_pfb_chan_set_channel_map(self._h, _ptr(m), len(m)))

    def channel_map(self):
        m = np.empty(self._M, np.int32)
        n = _lib.lib().grb200_pfb_chan_get_channel_map(self._h, _ptr(m), self._M)
        return m[:n].tolist()

    def taps_per_filter(self):
        return self.d_taps_per_filter

    def _toconsume(self, noutput_items):
        return int(np.rint(np.float32(noutput_items) / np.float32(self.d_oversample_rate)))  # :158

    def forecast(self, noutput_items, ninputs):
        return [self._toconsume(noutput_items) + self.history() - 1] * ninputs

    def general_work(self, noutput_items, ninput_items, input_items, output_items):
        with self.d_setlock:
            if self.d_updated:  # :140-143
                self.d_updated = False
                self.consume_each(0)
                return 0
            n = int(noutput_items)
            M, T = self._M, self.d_taps_per_filter
            if len(input_items) != M:
                raise ValueError("pfb_channelizer_ccf: %d input streams, %d needed" % (len(input_items), M))
            nouts = len(output_items)
            need = self._toconsume(n) + T
            ins = [as_items(a, _c64, need if n else 0, "input_items[%d]" % j) for j, a in enumerate(input_items)]
            outs = [as_items(a, _c64, n, "output_items[%d]" % j) for j, a in enumerate(output_items)]
            inp = (C.c_void_p * M)(*[a.ctypes.data for a in ins])
            outp = (C.c_void_p * nouts)(*[a.ctypes.data for a in outs])
            consumed = C.c_longlong(0)
            _lib.check(_lib.lib().grb200_pfb_chan_work(self._h, inp, outp, nouts, n, C.byref(consumed)))
            self.consume_each(consumed.value)
            return n

    def work_interleaved(self, in_interleaved, output_items, noutput_items):
        """Host form with stream_to_streams absorbed (see pfb.channelizer_ccf)."""
        n = int(noutput_items)
        M, T = self._M, self.d_taps_per_filter
        need = (self._toconsume(n) + T) * M
        xin = as_items(in_interleaved, _c64, need if n else 0, "input")
        nouts = len(output_items)
        outs = [as_items(a, _c64, n, "output_items[%d]" % j) for j, a in enumerate(output_items)]
        outp = (C.c_void_p * nouts)(*[a.ctypes.data for a in outs])
        consumed = C.c_longlong(0)
        with self.d_setlock:
            self.d_updated = False
            _lib.check(_lib.lib().grb200_pfb_chan_work_interleaved(self._h, _ptr(xin), outp, nouts, n,
                                                                  C.byref(consumed)))
        self.consume_each(consumed.value)
        return n

    def set_halo_source(self, peer_tail):
        """Time-sharded stream: work_interleaved_dev() reads its taps_per_filter history rows from `peer_tail` (the left
        neighbour's newest rows, in ITS memory over NVLink -- sharding.PeerHalo) inside the kernel.  None clears."""
        _lib.check(_lib.lib().grb200_pfb_chan_set_halo_source(self._h, C.c_void_p(peer_tail or 0)))

    def work_lag(self, input_items, output_items, noutput_items, interleaved=False):
        """Lag-by-one host mode: queues this call (M streams, or ONE interleaved stream when `interleaved`), writes the
        previous call's outputs to output_items and returns how many items per stream (0 on the first call)."""
        n = int(noutput_items)
        M, T = self._M, self.d_taps_per_filter
        nouts = len(output_items)
        outp = (C.c_void_p * nouts)(*[a.ctypes.data for a in output_items])
        cap = min(a.size for a in output_items)
        consumed, got = C.c_longlong(0), C.c_longlong(0)
        with self.d_setlock:
            self.d_updated = False
            if interleaved:
                xin = as_items(input_items, _c64, (self._toconsume(n) + T) * M if n else 0, "input")
                _lib.check(_lib.lib().grb200_pfb_chan_work_interleaved_lag(self._h, _ptr(xin), n, outp, nouts, cap,
                                                                          C.byref(consumed), C.byref(got)))
            else:
                ins = [as_items(a, _c64, self._toconsume(n) + T if n else 0, "input_items[%d]" % j)
                       for j, a in enumerate(input_items)]
                inp = (C.c_void_p * M)(*[a.ctypes.data for a in ins])
                _lib.check(_lib.lib().grb200_pfb_chan_work_lag(self._h, inp, n, outp, nouts, cap, C.byref(consumed),
                                                              C.byref(got)))
        self.consume_each(consumed.value)
        return got.value

    def flush(self, output_items):
        nouts = len(output_items)
        outp = (C.c_void_p * nouts)(*[a.ctypes.data for a in output_items])
        got = C.c_longlong(0)
        _lib.check(_lib.lib().grb200_pfb_chan_flush(self._h, outp, nouts, min(a.size for a in output_items), C.byref(got)))
        return got.value

    def work_interleaved_dev(self, in_dev, out_dev, out_stride, nout_streams, noutput_items, stream=0):
        _lib.check(_lib.lib().grb200_pfb_chan_work_interleaved_dev(
            self._h, C.c_void_p(in_dev), C.c_void_p(out_dev), int(out_stride), int(nout_streams),
            int(noutput_items), C.c_void_p(stream)))
        return int(noutput_items)

    def process(self, in_streams, noutput_items, nouts=None):
        """Convenience: in_streams [M, consumed+T] (history first) -> [nouts, noutput] array."""
        ins = np.ascontiguousarray(in_streams, np.complex64)
        nouts = self._M if nouts is None else int(nouts)
        outs = np.zeros((nouts, int(noutput_items)), np.complex64)
        il, ol = [ins[j] for j in range(self._M)], [outs[j] for j in range(nouts)]
        if self.d_updated:
            self.general_work(0, [], il, ol)
        r = self.general_work(noutput_items, [], il, ol)
        return outs[:, :r]

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                _lib.lib().grb200_pfb_chan_destroy(h)
            except Exception:
                pass


class pfb_decimator_ccf(sync_block):
    """Polyphase filterbank decimator: `decim` input streams (a stream_to_streams in front), one output = channel
    `channel` of the bank, at 1/decim of the total input rate.

    use_fft_rotator / use_fft_filters pick among the reference's four work() variants (pfb_decimator_ccf_impl.cc:
    120-131), which differ only in rounding; here they only decide what the scheduler is told: with use_fft_filters
    the block declares history 1 and an output multiple (:57-59) and the overlap is carried in the handle, otherwise
    history = taps_per_filter (:61)."""

    def __init__(self, decim, taps, channel, use_fft_rotator=True, use_fft_filters=True, device=0):
        L = _lib.lib()
        decim = int(decim)
        if decim < 1:
            raise ValueError("pfb_decimator_ccf: decimation must be >= 1")
        sync_block.__init__(self, "pfb_decimator_ccf", io_signature.make(decim, decim, 8), io_signature.make(1, 1, 8))
        self._h = C.c_void_p()
        self._M = decim
        t = pfb_channelizer_ccf._conv_taps(taps)
        _lib.check(L.grb200_pfb_dec_ccf_create(decim, _ptr(t), len(t), int(channel), int(device), C.byref(self._h)))
        self._proto = t.copy()
        self.d_chan = int(channel)
        self.d_use_fft_rotator = bool(use_fft_rotator)
        self.d_use_fft_filters = bool(use_fft_filters)
        self.d_taps_per_filter = L.grb200_pfb_dec_taps_per_filter(self._h)
        self.d_updated = False
        if self.d_use_fft_filters:
            # fft_filter_ccf geometry for taps_per_filter taps (fft_filter.cc:336-341): filtersize - ntaps + 1
            T = self.d_taps_per_filter
            fftsize = int(2 * 2 ** np.ceil(np.log(float(T)) / np.log(2.0)))
            self.set_history(1)
            self.set_output_multiple(fftsize - T + 1)
        else:
            self.set_history(self.d_taps_per_filter)

    @classmethod
    def make(cls, decim, taps, channel, use_fft_rotator=True, use_fft_filters=True, device=0):
        return cls(decim, taps, channel, use_fft_rotator, use_fft_filters, device)

    def set_taps(self, taps):
        t = pfb_channelizer_ccf._conv_taps(taps)
        with self.d_setlock:  # pfb_decimator_ccf_impl.cc:88-95
            L = _lib.lib()
            _lib.check(L.grb200_pfb_dec_set_taps(self._h, _ptr(t), len(t)))
            self._proto = t.copy()
            self.d_taps_per_filter = L.grb200_pfb_dec_taps_per_filter(self._h)
            self.set_history(self.d_taps_per_filter)
            self.d_updated = True

    def taps(self):
        M, T = self._M, self.d_taps_per_filter
        padded = np.zeros(M * T, np.float32)
        padded[: len(self._proto)] = self._proto
        return padded.reshape(T, M).T.tolist()  # d_taps[i][j] = padded[i + j*M], polyphase_filterbank.cc:69-73

    def print_taps(self):
        for i, row in enumerate(self.taps()):
            print("filter[%d]: [%s]" % (i, " ".join("%.4e" % v for v in row)))

    def set_channel(self, chan):
        with self.d_setlock:  # :106-110
            _lib.check(_lib.lib().grb200_pfb_dec_set_channel(self._h, int(chan)))
            self.d_chan = int(chan)

    def work(self, noutput_items, input_items, output_items):
        with self.d_setlock:
            if self.d_updated:  # :114-117
                self.d_updated = False
                return 0
            n = int(noutput_items)
            M = self._M
            if len(input_items) != M:
                raise ValueError("pfb_decimator_ccf: %d input streams, %d needed" % (len(input_items), M))
            need = n + self.history() - 1
            ins = [as_items(a, _c64, need if n else 0, "input_items[%d]" % j) for j, a in enumerate(input_items)]
            out = as_items(output_items[0], _c64, n, "output_items[0]")
            inp = (C.c_void_p * M)(*[a.ctypes.data for a in ins])
            carry = 1 if self.history() == 1 and self.d_taps_per_filter > 1 else 0
            _lib.check(_lib.lib().grb200_pfb_dec_work(self._h, inp, _ptr(out), n, carry))
            return n

    def work_dev(self, x_dev, stride, out_dev, noutput_items, stream=0):
        """Device-resident, stateless: stream s at x_dev + 8*s*stride bytes, taps_per_filter-1 history items first."""
        _lib.check(_lib.lib().grb200_pfb_dec_work_dev(self._h, C.c_void_p(x_dev), int(stride), C.c_void_p(out_dev),
                                                      int(noutput_items), C.c_void_p(stream)))
        return int(noutput_items)

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                _lib.lib().grb200_pfb_dec_destroy(h)
            except Exception:
                pass


class pfb_synthesizer_ccf(sync_interpolator):
    """pfb_synthesizer_ccf (gr-filter/include/gnuradio/filter/pfb_synthesizer_ccf.h, lib/pfb_synthesizer_ccf_impl.cc:23-284):
    up to twox*numchans input streams -> one stream at numchans times the rate.  work() takes the NEW items only
    (history 1); the filters' delay lines live in the library handle."""

    def __init__(self, numchans, taps, twox=False, device=0):
        L = _lib.lib()
        numchans = int(numchans)
        if numchans < 1:
            raise ValueError("pfb_synthesizer_ccf: bad channel count")
        sync_interpolator.__init__(self, "pfb_synthesizer_ccf", io_signature.make(1, numchans, 8), io_signature.make(1, 1, 8),
                                   numchans)
        self._h = C.c_void_p()
        self._M = numchans
        self.d_twox = 2 if twox else 1
        t = pfb_channelizer_ccf._conv_taps(taps)
        _lib.check(L.grb200_pfb_synth_ccf_create(numchans, _ptr(t), len(t), int(bool(twox)), int(device), C.byref(self._h)))
        self.d_taps_per_filter = L.grb200_pfb_synth_taps_per_filter(self._h)
        self.set_history(1)  # :92
        self.set_output_multiple(numchans)  # :68
        self.d_updated = True  # the constructor goes through set_taps (:63, :93)

    @classmethod
    def make(cls, numchans, taps, twox=False, device=0):
        return cls(numchans, taps, twox, device)

    def set_taps(self, taps):
        t = pfb_channelizer_ccf._conv_taps(taps)
        with self.d_setlock:  # :79-94
            L = _lib.lib()
            _lib.check(L.grb200_pfb_synth_set_taps(self._h, _ptr(t), len(t)))
            self.d_taps_per_filter = L.grb200_pfb_synth_taps_per_filter(self._h)
            self.set_history(1)
            self.d_updated = True

    def taps(self):
        out = np.zeros((self.d_twox * self._M, self.d_taps_per_filter), np.float32)
        _lib.check(_lib.lib().grb200_pfb_synth_branch_taps(self._h, _ptr(out)))
        return out.tolist()

    def print_taps(self):
        for i, row in enumerate(self.taps()):
            print("filter[%d]: [%s]" % (i, " ".join("%.4e" % v for v in row)))

    def set_channel_map(self, chan_map):
        m = np.ascontiguousarray(chan_map, np.int32)
        with self.d_setlock:
            _lib.check(_lib.lib().grb200_pfb_synth_set_channel_map(self._h, _ptr(m), len(m)))

    def channel_map(self):
        m = np.empty(4 * self._M + 16, np.int32)
        n = _lib.lib().grb200_pfb_synth_get_channel_map(self._h, _ptr(m), len(m))
        if n < 0:
            _lib.check(n)
        return m[:n].tolist()

    def taps_per_filter(self):
        return self.d_taps_per_filter

    def work(self, noutput_items, input_items, output_items):
        with self.d_setlock:
            if self.d_updated:  # :229-232
                self.d_updated = False
                return 0
            n = int(noutput_items)
            nin = len(input_items)
            ins = [as_items(a, _c64, n // self._M, "input_items[%d]" % j) for j, a in enumerate(input_items)]
            out = as_items(output_items[0], _c64, n, "output_items[0]")
            inp = (C.c_void_p * nin)(*[a.ctypes.data for a in ins])
            _lib.check(_lib.lib().grb200_pfb_synth_work(self._h, inp, nin, _ptr(out), n))
            return n

    def work_dev(self, x_dev, stride, nin, out_dev, noutput_items, parity=0, stream=0):
        """Device-resident, stateless: stream i at x_dev + 8*i*stride bytes, taps_per_filter-1 history items first."""
        _lib.check(_lib.lib().grb200_pfb_synth_work_dev(self._h, C.c_void_p(x_dev), int(stride), int(nin), C.c_void_p(out_dev),
                                                        int(noutput_items), int(parity), C.c_void_p(stream)))
        return int(noutput_items)

    def process(self, in_streams):
        """Convenience: all of in_streams ([nin][n]) through work(); returns n*numchans items."""
        ins = [np.ascontiguousarray(a, np.complex64) for a in in_streams]
        n = min(len(a) for a in ins) * self._M
        out = np.empty(n, np.complex64)
        if self.d_updated:
            self.work(0, ins, [out])
        r = self.work(n, ins, [out])
        return out[:r]

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                _lib.lib().grb200_pfb_synth_destroy(h)
            except Exception:
                pass


class filterbank_vcvcf(block):
    """filterbank_vcvcf (gr-filter/include/gnuradio/filter/filterbank_vcvcf.h, lib/filterbank_vcvcf_impl.cc:23-96 over
    lib/filterbank.cc): items are vectors of len(taps) complex; one FIR per vector element; history ntaps+1."""

    def __init__(self, taps, device=0):
        L = _lib.lib()
        t = self._conv_taps(taps)
        nf, nt = t.shape
        block.__init__(self, "filterbank_vcvcf", io_signature.make(1, 1, 8 * nf), io_signature.make(1, 1, 8 * nf))
        self._h = C.c_void_p()
        _lib.check(L.grb200_filterbank_vcvcf_create(nf, nt, _ptr(t), int(device), C.byref(self._h)))
        self.d_nfilts, self.d_ntaps = nf, nt
        self._taps = t.copy()
        self.d_updated = False
        self.set_history(nt + 1)  # :34

    @classmethod
    def make(cls, taps, device=0):
        return cls(taps, device)

    @staticmethod
    def _conv_taps(taps):
        rows = [np.asarray(r, np.float32).reshape(-1) for r in taps]
        if len(rows) == 0:
            raise ValueError("The taps vector may not be empty.")  # filterbank.cc:30-32 (std::invalid_argument)
        if any(len(r) != len(rows[0]) for r in rows):
            raise RuntimeError("All sets of taps must be of the same length.")  # :60-64 (std::runtime_error)
        if len(rows[0]) == 0:
            raise ValueError("filterbank_vcvcf: at least one tap per filter is required")
        return np.ascontiguousarray(np.stack(rows), np.float32)

    def set_taps(self, taps):
        rows = list(taps)
        if len(rows) != self.d_nfilts:
            raise RuntimeError("The number of filters is incorrect.")  # filterbank.cc:54-56
        t = self._conv_taps(rows)
        with self.d_setlock:  # filterbank_vcvcf_impl.cc:39-45
            _lib.check(_lib.lib().grb200_filterbank_vcvcf_set_taps(self._h, t.shape[0], t.shape[1], _ptr(t)))
            self.d_ntaps = t.shape[1]
            self._taps = t.copy()
            self.set_history(self.d_ntaps + 1)
            self.d_updated = True

    def taps(self):
        return self._taps.tolist()

    def print_taps(self):
        for i, row in enumerate(self.taps()):
            print("filter[%d]: [%s]" % (i, " ".join("%.4e" % v for v in row)))

    def forecast(self, noutput_items, ninputs):
        return [int(noutput_items) + self.history() - 1] * ninputs

    def general_work(self, noutput_items, ninput_items, input_items, output_items):
        with self.d_setlock:
            if self.d_updated:  # :67-70
                self.d_updated = False
                return 0
            n, nf = int(noutput_items), self.d_nfilts
            xin = as_items(input_items[0], _c64, (n + self.d_ntaps) * nf if n else 0, "input_items[0]")
            out = as_items(output_items[0], _c64, n * nf, "output_items[0]")
            _lib.check(_lib.lib().grb200_filterbank_vcvcf_work(self._h, _ptr(xin), _ptr(out), n))
            self.consume_each(n)  # :94
            return n

    def work_dev(self, in_dev, out_dev, noutput_items, stream=0):
        """Device-resident: in_dev holds (noutput + ntaps) vectors, history first."""
        _lib.check(_lib.lib().grb200_filterbank_vcvcf_work_dev(self._h, C.c_void_p(in_dev), C.c_void_p(out_dev),
                                                               int(noutput_items), C.c_void_p(stream)))
        return int(noutput_items)

    def process(self, xin, noutput_items):
        """One general_work() on xin ((noutput + ntaps) vectors, history first)."""
        xin = np.ascontiguousarray(xin, np.complex64).reshape(-1)
        out = np.empty(int(noutput_items) * self.d_nfilts, np.complex64)
        if self.d_updated:
            self.general_work(0, [0], [xin], [out])
        r = self.general_work(noutput_items, [len(xin) // self.d_nfilts], [xin], [out])
        return out[: r * self.d_nfilts].reshape(r, self.d_nfilts)

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                _lib.lib().grb200_filterbank_vcvcf_destroy(h)
            except Exception:
                pass


class _pfb_namespace:
    """gnuradio.filter.pfb: the Python hier blocks of gr-filter/python/filter/pfb.py."""

    class channelizer_ccf:
        """pfb.channelizer_ccf(numchans, taps, oversample_rate): ONE interleaved input stream -> numchans
        outputs (pfb.py:30-50).  stream_to_streams + pfb_channelizer_ccf collapse into one device pass."""

        def __init__(self, numchans, taps=None, oversample_rate=1, atten=100, device=0):
            if taps is None or len(taps) == 0:
                # pfb.py:65-72 would design taps with optfir; out of scope here
                raise ValueError("pfb.channelizer_ccf: taps must be supplied")
            self._numchans = int(numchans)
            self._oversample_rate = oversample_rate
            self.pfb = pfb_channelizer_ccf(numchans, taps, float(oversample_rate), device)
            self.pfb.d_updated = False  # the hier block is driven whole; no first-call skip
            self._hist = None

        def set_channel_map(self, newmap):
            self.pfb.set_channel_map(newmap)

        def set_taps(self, taps):
            self.pfb.set_taps(taps)
            self.pfb.d_updated = False
            self._hist = None

        def taps(self):
            return self.pfb.taps()

        def run(self, x, nouts=None):
            """Whole finite interleaved stream, zeros before t=0 (what tb.run() delivers)."""
            M, T = self._numchans, self.pfb.taps_per_filter()
            x = np.ascontiguousarray(x, np.complex64)
            rows = len(x) // M
            om = self.pfb.output_multiple()
            R = int(round(M / float(self._oversample_rate)))
            noutput = (rows * M // R) // om * om
            xin = np.concatenate([np.zeros(T * M, np.complex64), x[: rows * M]])
            nouts = M if nouts is None else int(nouts)
            outs = np.zeros((nouts, noutput), np.complex64)
            self.pfb.work_interleaved(xin, [outs[j] for j in range(nouts)], noutput)
            return outs

    class channelizer_hier_ccf:
        """pfb.channelizer_hier_ccf(n_chans, n_filterbanks, taps, outchans) (gr-filter/python/filter/pfb.py:406-480):
        stream_to_vector -> vector_map split -> n_filterbanks x filterbank_vcvcf (branch taps reversed(taps[i::n_chans]))
        -> vector_map combine -> fft_vcc(n_chans, forward, window of ones) -> selector -> vector_to_streams.

        On the CPU the split into n_filterbanks blocks buys threads; it does not change a single output value, so here ONE
        filterbank_vcvcf handle carries all n_chans branch filters, stream_to_vector is the pointer reinterpretation it is
        on a device, and the three stages are chained device-resident (no host round trip between them)."""

        def __init__(self, n_chans, n_filterbanks=1, taps=None, outchans=None, atten=100, bw=1.0, tb=0.2, ripple=0.1,
                     device=0):
            from . import blocks as _blocks
            from . import fft as _fft

            if taps is None or len(taps) == 0:
                # pfb.py:434-435 designs taps with optfir; tap design is out of scope here
                raise ValueError("pfb.channelizer_hier_ccf: taps must be supplied")
            n_chans = int(n_chans)
            self.n_chans = n_chans
            self.n_filterbanks = min(int(n_filterbanks), n_chans)  # :423-424
            self.outchans = list(range(n_chans)) if outchans is None else [int(c) for c in outchans]
            if any(c < 0 or c >= n_chans for c in self.outchans):
                raise ValueError("pfb.channelizer_hier_ccf: outchans out of range")
            taps = [float(v) for v in taps]
            extra = int(np.ceil(1.0 * len(taps) / n_chans) * n_chans - len(taps))  # :437-438
            taps = taps + [0.0] * extra
            chantaps = [list(reversed(taps[i: len(taps): n_chans])) for i in range(n_chans)]  # :440
            self._device = int(device)
            self.fb = filterbank_vcvcf(chantaps, device)
            self.fft = _fft.fft_vcc(n_chans, True, [1.0] * n_chans, False, 1, device)  # :466
            self.v2ss = _blocks.vector_to_streams(8, n_chans, device)
            self.ntaps = self.fb.d_ntaps

        def run(self, x):
            """Whole finite stream x, zeros before t=0 (what tb.run() delivers): returns [len(outchans)][n] complex64."""
            M, nt = self.n_chans, self.ntaps
            x = np.ascontiguousarray(x, np.complex64)
            n = len(x) // M
            xin = np.concatenate([np.zeros(nt * M, np.complex64), x[: n * M]])  # history ntaps+1 -> ntaps zero vectors
            L = _lib.lib()
            dev = self._device
            bufs = []

            def dalloc(nbytes):
                p = C.c_void_p()
                _lib.check(L.grb200_dev_alloc(dev, C.byref(p), max(int(nbytes), 16)))
                bufs.append(p)
                return p

            try:
                d_in, d_fb, d_fft, d_out = dalloc(xin.nbytes), dalloc(n * M * 8), dalloc(n * M * 8), dalloc(n * M * 8)
                _lib.check(L.grb200_memcpy_h2d(dev, d_in, _ptr(xin), xin.nbytes))
                # one stream (0 = the legacy default stream) orders the three launches; no host round trip in between
                self.fb.work_dev(d_in.value, d_fb.value, n, 0)
                self.fft.work_dev(d_fb.value, d_fft.value, n, 0)  # stream_to_vector / the vector stream: same bytes
                self.v2ss.work_dev(d_fft.value, d_out.value, n, n, 0)
                outs = np.empty((M, n), np.complex64)
                _lib.check(L.grb200_dev_sync(dev))
                _lib.check(L.grb200_memcpy_d2h(dev, _ptr(outs), d_out, outs.nbytes))
            finally:
                for p in bufs:
                    L.grb200_dev_free(dev, p)
            return outs[self.outchans]


pfb = _pfb_namespace()
