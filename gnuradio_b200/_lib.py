"""ctypes binding of the C-ABI library libgrb200.so (include/grb200.h).

There is no CPU fallback: if the library is missing or no sm_100 device is usable every call raises.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libgrb200.so")

OK, EINVAL, ERANGE, ECUDA, ENOMEM, EWINDOW, ERUNTIME = 0, -1, -2, -3, -4, -5, -6

_lib = None

vp, i32, i64, u64, f32, f64 = C.c_void_p, C.c_int, C.c_longlong, C.c_ulonglong, C.c_float, C.c_double
pvp = C.POINTER(C.c_void_p)

# name -> (restype, argtypes); every symbol include/grb200.h declares
SIGNATURES = {
    "grb200_last_error": (C.c_char_p, []),
    "grb200_version": (C.c_char_p, []),
    "grb200_device_count": (i32, []),
    "grb200_launch_count": (u64, []),
    "grb200_host_alloc": (i32, [pvp, u64]),
    "grb200_host_free": (i32, [vp]),
    "grb200_host_register": (i32, [vp, u64]),
    "grb200_host_unregister": (i32, [vp]),
    "grb200_fir_work_lag": (i32, [vp, vp, i64, vp, i64, C.POINTER(i64)]),
    "grb200_fir_flush": (i32, [vp, vp, i64, C.POINTER(i64)]),
    "grb200_pfb_chan_work_lag": (i32, [vp, pvp, i64, pvp, i32, i64, C.POINTER(i64), C.POINTER(i64)]),
    "grb200_pfb_chan_work_interleaved_lag": (i32, [vp, vp, i64, pvp, i32, i64, C.POINTER(i64), C.POINTER(i64)]),
    "grb200_pfb_chan_flush": (i32, [vp, pvp, i32, i64, C.POINTER(i64)]),
    "grb200_dev_alloc": (i32, [i32, pvp, u64]),
    "grb200_dev_free": (i32, [i32, vp]),
    "grb200_memcpy_h2d": (i32, [i32, vp, vp, u64]),
    "grb200_memcpy_d2h": (i32, [i32, vp, vp, u64]),
    "grb200_dev_sync": (i32, [i32]),
    "grb200_ipc_export": (i32, [i32, vp, vp]),
    "grb200_ipc_open": (i32, [i32, vp, pvp]),
    "grb200_ipc_close": (i32, [i32, vp]),
    "grb200_enable_peer": (i32, [i32, i32]),
    "grb200_fir_set_halo_source": (i32, [vp, vp]),
    "grb200_pfb_chan_set_halo_source": (i32, [vp, vp]),
    "grb200_fir_ccf_create": (i32, [i32, vp, i32, i32, pvp]),
    "grb200_fir_ccc_create": (i32, [i32, vp, i32, i32, pvp]),
    "grb200_fir_fff_create": (i32, [i32, vp, i32, i32, pvp]),
    "grb200_fir_set_taps": (i32, [vp, vp, i32]),
    "grb200_fir_ntaps": (i32, [vp]),
    "grb200_fir_decimation": (i32, [vp]),
    "grb200_fir_work": (i32, [vp, vp, vp, i64]),
    "grb200_fir_work_dev": (i32, [vp, vp, vp, i64, vp]),
    "grb200_fir_destroy": (None, [vp]),
    "grb200_pfb_synth_ccf_create": (i32, [C.c_uint, vp, i32, i32, i32, pvp]),
    "grb200_pfb_synth_set_taps": (i32, [vp, vp, i32]),
    "grb200_pfb_synth_set_channel_map": (i32, [vp, vp, i32]),
    "grb200_pfb_synth_get_channel_map": (i32, [vp, vp, i32]),
    "grb200_pfb_synth_taps_per_filter": (i32, [vp]),
    "grb200_pfb_synth_numchans": (i32, [vp]),
    "grb200_pfb_synth_twox": (i32, [vp]),
    "grb200_pfb_synth_branch_taps": (i32, [vp, vp]),
    "grb200_pfb_synth_work": (i32, [vp, pvp, i32, vp, i64]),
    "grb200_pfb_synth_work_dev": (i32, [vp, vp, i64, i32, vp, i64, i32, vp]),
    "grb200_pfb_synth_destroy": (None, [vp]),
    "grb200_filterbank_vcvcf_create": (i32, [i32, i32, vp, i32, pvp]),
    "grb200_filterbank_vcvcf_set_taps": (i32, [vp, i32, i32, vp]),
    "grb200_filterbank_vcvcf_nfilts": (i32, [vp]),
    "grb200_filterbank_vcvcf_ntaps": (i32, [vp]),
    "grb200_filterbank_vcvcf_work": (i32, [vp, vp, vp, i64]),
    "grb200_filterbank_vcvcf_work_dev": (i32, [vp, vp, vp, i64, vp]),
    "grb200_filterbank_vcvcf_destroy": (None, [vp]),
    "grb200_interp_fir_ccf_create": (i32, [C.c_uint, vp, i32, i32, pvp]),
    "grb200_interp_fir_ccc_create": (i32, [C.c_uint, vp, i32, i32, pvp]),
    "grb200_interp_fir_set_taps": (i32, [vp, vp, i32]),
    "grb200_interp_fir_taps_per_filter": (i32, [vp]),
    "grb200_interp_fir_interpolation": (i32, [vp]),
    "grb200_interp_fir_work": (i32, [vp, vp, vp, i64]),
    "grb200_interp_fir_work_dev": (i32, [vp, vp, vp, i64, vp]),
    "grb200_interp_fir_destroy": (None, [vp]),
    "grb200_rational_resampler_ccf_create": (i32, [C.c_uint, C.c_uint, vp, i32, i32, pvp]),
    "grb200_rational_resampler_ccc_create": (i32, [C.c_uint, C.c_uint, vp, i32, i32, pvp]),
    "grb200_rational_resampler_set_taps": (i32, [vp, vp, i32]),
    "grb200_rational_resampler_taps_per_filter": (i32, [vp]),
    "grb200_rational_resampler_interpolation": (C.c_uint, [vp]),
    "grb200_rational_resampler_decimation": (C.c_uint, [vp]),
    "grb200_rational_resampler_ctr": (C.c_uint, [vp]),
    "grb200_rational_resampler_set_ctr": (i32, [vp, C.c_uint]),
    "grb200_rational_resampler_work": (i32, [vp, vp, i64, vp, i64, vp, vp]),
    "grb200_rational_resampler_work_dev": (i32, [vp, vp, i64, vp, i64, vp, vp, vp]),
    "grb200_rational_resampler_destroy": (None, [vp]),
    "grb200_pfb_dec_ccf_create": (i32, [C.c_uint, vp, i32, C.c_uint, i32, pvp]),
    "grb200_pfb_dec_set_taps": (i32, [vp, vp, i32]),
    "grb200_pfb_dec_set_channel": (i32, [vp, C.c_uint]),
    "grb200_pfb_dec_taps_per_filter": (i32, [vp]),
    "grb200_pfb_dec_decimation": (i32, [vp]),
    "grb200_pfb_dec_work": (i32, [vp, vp, vp, i64, i32]),
    "grb200_pfb_dec_work_dev": (i32, [vp, vp, i64, vp, i64, vp]),
    "grb200_pfb_dec_destroy": (None, [vp]),
    "grb200_freq_xlating_fir_ccf_create": (i32, [i32, vp, i32, f64, f64, i32, pvp]),
    "grb200_freq_xlating_fir_ccc_create": (i32, [i32, vp, i32, f64, f64, i32, pvp]),
    "grb200_freq_xlating_fir_set_taps": (i32, [vp, vp, i32]),
    "grb200_freq_xlating_fir_set_center_freq": (i32, [vp, f64]),
    "grb200_freq_xlating_fir_advance": (i32, [vp, i64]),
    "grb200_freq_xlating_fir_center_freq": (f64, [vp]),
    "grb200_freq_xlating_fir_ntaps": (i32, [vp]),
    "grb200_freq_xlating_fir_decimation": (i32, [vp]),
    "grb200_freq_xlating_fir_composite_taps": (i32, [vp, vp, i32]),
    "grb200_freq_xlating_fir_work": (i32, [vp, vp, vp, i64]),
    "grb200_freq_xlating_fir_work_dev": (i32, [vp, vp, vp, i64, vp]),
    "grb200_freq_xlating_fir_destroy": (None, [vp]),
    "grb200_fftfilt_ccc_create": (i32, [i32, vp, i32, i32, pvp]),
    "grb200_fftfilt_set_taps": (i32, [vp, vp, i32]),
    "grb200_fftfilt_ccf_create": (i32, [i32, vp, i32, i32, pvp]),
    "grb200_fftfilt_set_taps_real": (i32, [vp, vp, i32]),
    "grb200_fftfilt_nsamples": (i32, [vp]),
    "grb200_fftfilt_ntaps": (i32, [vp]),
    "grb200_fftfilt_work": (i32, [vp, vp, vp, i64]),
    "grb200_fftfilt_work_dev": (i32, [vp, vp, vp, i64, vp]),
    "grb200_fftfilt_destroy": (None, [vp]),
    "grb200_pfb_chan_ccf_create": (i32, [C.c_uint, vp, i32, f32, i32, pvp]),
    "grb200_pfb_chan_set_taps": (i32, [vp, vp, i32]),
    "grb200_pfb_chan_set_channel_map": (i32, [vp, vp, i32]),
    "grb200_pfb_chan_get_channel_map": (i32, [vp, vp, i32]),
    "grb200_pfb_chan_taps_per_filter": (i32, [vp]),
    "grb200_pfb_chan_output_multiple": (i32, [vp]),
    "grb200_pfb_chan_branch_taps": (i32, [vp, vp]),
    "grb200_pfb_chan_work": (i32, [vp, pvp, pvp, i32, i64, C.POINTER(i64)]),
    "grb200_pfb_chan_work_interleaved": (i32, [vp, vp, pvp, i32, i64, C.POINTER(i64)]),
    "grb200_pfb_chan_work_interleaved_dev": (i32, [vp, vp, vp, i64, i32, i64, vp]),
    "grb200_pfb_chan_destroy": (None, [vp]),
    "grb200_fft_vcc_create": (i32, [i32, i32, vp, i32, i32, i32, pvp]),
    "grb200_fft_vcc_set_window": (i32, [vp, vp, i32]),
    "grb200_fft_vcc_fft_size": (i32, [vp]),
    "grb200_fft_vcc_work": (i32, [vp, vp, vp, i64]),
    "grb200_fft_vcc_work_dev": (i32, [vp, vp, vp, i64, vp]),
    "grb200_fft_vcc_work_magsq": (i32, [vp, vp, vp, i64]),
    "grb200_fft_vcc_work_magsq_dev": (i32, [vp, vp, vp, i64, vp]),
    "grb200_fft_vcc_destroy": (None, [vp]),
    "grb200_fft_vfc_create": (i32, [i32, i32, vp, i32, i32, pvp]),
    "grb200_logpwrfft_create": (i32, [i32, f64, i32, f64, f64, f64, i32, vp, i32, i32, pvp]),
    "grb200_logpwrfft_set_decimation": (i32, [vp, f64]),
    "grb200_logpwrfft_set_vec_rate": (i32, [vp, f64]),
    "grb200_logpwrfft_set_sample_rate": (i32, [vp, f64]),
    "grb200_logpwrfft_set_average": (i32, [vp, i32]),
    "grb200_logpwrfft_set_avg_alpha": (i32, [vp, f64]),
    "grb200_logpwrfft_decimation": (i32, [vp]),
    "grb200_logpwrfft_sample_rate": (f64, [vp]),
    "grb200_logpwrfft_frame_rate": (f64, [vp]),
    "grb200_logpwrfft_average": (i32, [vp]),
    "grb200_logpwrfft_avg_alpha": (f64, [vp]),
    "grb200_logpwrfft_fft_size": (i32, [vp]),
    "grb200_logpwrfft_work": (i32, [vp, vp, i64, vp, i64, C.POINTER(i64), C.POINTER(i64)]),
    "grb200_logpwrfft_work_dev": (i32, [vp, vp, i64, vp, i64, C.POINTER(i64), C.POINTER(i64), vp]),
    "grb200_logpwrfft_work_dev_static": (i32, [vp, vp, i64, vp, i64, C.POINTER(i64), C.POINTER(i64), vp]),
    "grb200_logpwrfft_destroy": (None, [vp]),
    "grb200_vector_to_streams_work": (i32, [i32, i32, i32, vp, pvp, i64]),
    "grb200_vector_to_streams_work_dev": (i32, [i32, i32, i32, vp, vp, i64, i64, vp]),
    "grb200_streams_to_vector_work_dev": (i32, [i32, i32, i32, vp, i64, vp, i64, vp]),
    "grb200_magsq_work": (i32, [i32, vp, vp, i64]),
    "grb200_magsq_work_dev": (i32, [i32, vp, vp, i64, vp]),
}


def load_library():
    """Load libgrb200.so and bind every declared symbol.  Raises if the CUDA extension is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            "gnuradio_b200: %s is missing -- run `python __graft_entry__.py` (nvcc, sm_100a) first; "
            "there is no CPU fallback" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the library does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def lib():
    return load_library()


def last_error():
    return lib().grb200_last_error().decode("utf-8", "replace")


def check(rc):
    """Status -> the exception the reference block would raise."""
    if rc == OK:
        return
    msg = last_error()
    if rc == EINVAL:
        raise ValueError(msg)  # std::invalid_argument
    if rc == ERANGE:
        raise IndexError(msg)  # std::out_of_range
    if rc == ENOMEM:
        raise MemoryError(msg)
    raise RuntimeError(msg)  # std::runtime_error (CUDA failure, window mismatch)
