"""gnuradio_b200 -- B200-native (sm_100a) drop-in for GNU Radio's filtering hot path.

Only what the path needs: csrc/ (CUDA kernels + the C-ABI library libgrb200.so), the ctypes binding,
and the host-side mirror of the reference block interfaces (filter, fft, blocks).  No CPU fallback.
"""
from . import _lib
from ._lib import load_library


def device_count():
    return int(_lib.lib().grb200_device_count())


def launch_count():
    """Kernels launched by libgrb200 in this process so far."""
    return int(_lib.lib().grb200_launch_count())


def host_register(array):
    """Page-lock a numpy array the caller owns (what an integration does once for a gr::buffer) so that the staging path
    copies from / to it directly instead of through the pinned bounce.  Undo with host_unregister() before freeing it."""
    import ctypes as C

    _lib.check(_lib.lib().grb200_host_register(C.c_void_p(array.ctypes.data), array.nbytes))


def host_unregister(array):
    import ctypes as C

    _lib.check(_lib.lib().grb200_host_unregister(C.c_void_p(array.ctypes.data)))


def version():
    return _lib.lib().grb200_version().decode()


from . import blocks, fft, filter  # noqa: E402,F401
