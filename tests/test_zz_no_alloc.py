"""grb200_logpwrfft_work_dev_static -- the entry point that allocates nothing (static device scratch inside the library) --
against grb200_logpwrfft_work_dev on the same input, through the C-ABI, two calls in a row (the decimator counter and the
averaging filter's memory carry from the first call into the second; the shared scratch array is reused as the first call
left it).  Same kernels and run lengths, so the tolerance of the other GPU tests (NRMS <= 1e-5, max-abs <= 1e-4 x RMS) is
met with room to spare; the outputs are in fact expected to be equal bit for bit, which is asserted too.
Named test_zz_* so that pytest collects it last."""
import ctypes as C

import numpy as np
import pytest

from common import assert_parity, blackman_harris

pytestmark = pytest.mark.gpu

N = 1024
FRAMES = (2000, 1500)  # > 16 each: the three-kernel path that uses the scratch; <= 8 * (2^18 // N) = 2048: fits the array


@pytest.fixture(scope="module")
def grb():
    import gnuradio_b200 as g

    assert g.device_count() > 0, "no CUDA device: libgrb200 has no CPU fallback"
    return g


def _make(grb, cin=True, decim=1, alpha=0.05):
    w = blackman_harris(N)
    cls = grb.fft.logpwrfft_c if cin else grb.fft.logpwrfft_f
    return cls(float(N * decim), N, 2.0, 1.0, alpha, True, lambda n: w)


@pytest.mark.parametrize("cin,decim", [(True, 1), (False, 2)])
def test_static_scratch_entry_matches_work_dev_on_two_calls_in_a_row(grb, cin, decim):
    import torch

    rng = np.random.default_rng(77 + decim)
    old, new = _make(grb, cin, decim), _make(grb, cin, decim)
    assert old.decimation() == decim and new.decimation() == decim
    torch.cuda.synchronize()
    for call, frames in enumerate(FRAMES):
        nitems = frames * decim * N + (call * 37)  # the second call drags a partial vector along (not consumed)
        if cin:
            x = (rng.standard_normal(nitems) + 1j * rng.standard_normal(nitems)).astype(np.complex64)
            xd = torch.view_as_real(torch.from_numpy(x)).cuda()
        else:
            x = rng.standard_normal(nitems).astype(np.float32)
            xd = torch.from_numpy(x).cuda()
        a = torch.full((frames, N), float("nan"), dtype=torch.float32, device="cuda")
        b = torch.full((frames, N), float("nan"), dtype=torch.float32, device="cuda")
        torch.cuda.synchronize()
        ua, ma = old.work_dev(xd.data_ptr(), nitems, a.data_ptr(), frames)
        torch.cuda.synchronize()
        ub, mb = new.work_dev_static(xd.data_ptr(), nitems, b.data_ptr(), frames)
        torch.cuda.synchronize()
        assert (ua, ma) == (ub, mb) == (frames * decim * N, frames)
        ha, hb = a.cpu().numpy(), b.cpu().numpy()
        assert np.isfinite(ha).all() and np.isfinite(hb).all()
        # dB values: compare the powers, as the other logpwrfft tests do
        assert_parity(10.0 ** (hb.astype(np.float64) / 10), 10.0 ** (ha.astype(np.float64) / 10), "static call %d" % call)
        np.testing.assert_array_equal(hb, ha)


def test_static_scratch_entry_refuses_what_does_not_fit_and_leaves_the_handle_alone(grb):
    import torch

    L = grb._lib.lib()
    blk, ref = _make(grb), _make(grb)
    too_many = 8 * ((1 << 18) // N) + 8  # one run more than the array holds
    x = torch.randn(too_many * N, 2, device="cuda")
    out = torch.zeros(too_many, N, device="cuda")
    used, made = C.c_longlong(-1), C.c_longlong(-1)
    rc = L.grb200_logpwrfft_work_dev_static(blk._h, C.c_void_p(x.data_ptr()), too_many * N, C.c_void_p(out.data_ptr()),
                                            too_many, C.byref(used), C.byref(made), None)
    assert rc == grb._lib.ERANGE
    assert "static" in grb._lib.last_error() and "scratch" in grb._lib.last_error()
    torch.cuda.synchronize()
    assert float(out.abs().max()) == 0.0  # nothing was launched
    # the refused call changed nothing: the handle now behaves like a fresh one
    n = 64
    a = torch.empty(n, N, device="cuda")
    b = torch.empty(n, N, device="cuda")
    assert blk.work_dev_static(x.data_ptr(), n * N, a.data_ptr(), n) == ref.work_dev(x.data_ptr(), n * N, b.data_ptr(), n)
    torch.cuda.synchronize()
    np.testing.assert_array_equal(a.cpu().numpy(), b.cpu().numpy())
