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


def version():
    return _lib.lib().grb200_version().decode()


from . import blocks, fft, filter  # noqa: E402,F401
