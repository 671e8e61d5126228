"""GPU parity (through the C-ABI): interp_fir_filter_ccf/ccc against the oracle, the committed reference fixtures, the
reference QA's golden vector (gr-filter/python/filter/qa_interp_fir_filter.py) and float64.
Tolerance: north_star (NRMS <= 1e-5, max-abs <= 1e-4 x RMS); 2e-6 / 2e-5 where stated; the QA vector exactly."""
import numpy as np
import pytest

import sched_mock
from common import assert_parity, crand, fixtures
from oracle import pyoracle as po
from test_oracle import INTERP_QA_EXPECTED, INTERP_QA_SRC, INTERP_QA_TAPS

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def grb():
    import gnuradio_b200 as g

    assert g.device_count() > 0, "no CUDA device: libgrb200 has no CPU fallback"
    return g


def _cls(grb, taps):
    return grb.filter.interp_fir_filter_ccc if np.iscomplexobj(taps) else grb.filter.interp_fir_filter_ccf


def test_interp_fir_qa_golden_vector(grb):
    src = np.array(INTERP_QA_SRC, np.float32)
    exp = np.array(INTERP_QA_EXPECTED, np.float32)
    for taps in (np.array(INTERP_QA_TAPS, np.float32), np.array(INTERP_QA_TAPS, np.complex64)):
        blk = _cls(grb, taps)(3, taps)
        assert blk.history() == 2 and blk.taps() == list(taps) + [0]  # 5 taps padded to 6 (:78-84)
        got = sched_mock.run_sync(blk, (src + 1j * 2 * src).astype(np.complex64))
        assert len(got) == 3 * len(src)
        assert np.array_equal(got[: len(exp)].real, exp) and np.array_equal(got[: len(exp)].imag, 2 * exp)


def test_interp_fir_reference_fixtures(grb):
    fx = fixtures()
    for key, L, tk in (("interp_ccf_l4", 4, "interp_taps_f"), ("interp_ccc_l3", 3, "interp_taps_c")):
        got = sched_mock.run_sync(_cls(grb, fx[tk])(L, fx[tk]), fx["interp_x"], max_noutput=L * 333)
        assert_parity(got, fx[key], key, 2e-6, 2e-5)


@pytest.mark.parametrize("L,ntaps,cplx", [(1, 1, False), (1, 64, True), (2, 7, False), (3, 33, True), (4, 64, False),
                                          (4, 128, False), (8, 64, True), (5, 55, False), (16, 400, False), (7, 300, True)])
def test_interp_fir_sweep(grb, orc, L, ntaps, cplx):
    rng = np.random.default_rng(L * 1000 + ntaps)
    x = crand(rng, 20011)
    taps = crand(rng, ntaps) if cplx else rng.uniform(-1, 1, ntaps).astype(np.float32)
    blk = _cls(grb, taps)(L, taps)
    assert blk.history() == -(-ntaps // L) and blk.interpolation() == L and blk.output_multiple() == L
    got = sched_mock.run_sync(blk, x, max_noutput=L * 4096)
    assert len(got) == L * len(x)
    assert_parity(got, orc.interp_fir(L, taps, x), "oracle %d/%d" % (L, ntaps), 2e-6, 2e-5)
    assert_parity(got, po.f64_interp_fir(L, taps, x), "f64 %d/%d" % (L, ntaps), 2e-6, 2e-5)


@pytest.mark.parametrize("L,ntaps,cplx,shift", [(2, 64, False, 0), (3, 33, True, 1), (4, 64, False, 1), (4, 31, False, 0),
                                                (8, 64, True, 0), (5, 55, False, 1), (16, 400, False, 0), (7, 100, True, 0),
                                                (32, 64, False, 1), (4, 61, False, 0), (6, 96, True, 1)])
def test_interp_fir_streaming_kernel(grb, orc, L, ntaps, cplx, shift):
    """One device-resident call long enough for interp_stream_kernel (>= 8192 input positions): every ring depth, phase counts
    that do not divide the warp, complex taps, an input pointer that is only 8-byte aligned, ragged last blocks and runs."""
    import torch

    rng = np.random.default_rng(L * 977 + ntaps)
    ni = 30011 + 2 * L
    T = -(-ntaps // L)
    x = crand(rng, ni)
    taps = crand(rng, ntaps) if cplx else rng.uniform(-1, 1, ntaps).astype(np.float32)
    xin = np.concatenate([np.zeros(shift + T - 1, np.complex64), x])
    xd = torch.from_numpy(xin.view(np.float32).copy()).cuda()
    yd = torch.full((2 * ni * L + 2,), float("nan"), dtype=torch.float32, device="cuda")
    blk = _cls(grb, taps)(L, taps)
    blk.work_dev(xd.data_ptr() + 8 * shift, yd.data_ptr(), ni * L, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    y = yd.cpu().numpy()
    assert np.isnan(y[-2:]).all()  # nothing written past the end
    got = y[:-2].view(np.complex64)
    assert_parity(got, orc.interp_fir(L, taps, x), "oracle %d/%d" % (L, ntaps), 2e-6, 2e-5)
    assert_parity(got, po.f64_interp_fir(L, taps, x), "f64 %d/%d" % (L, ntaps), 2e-6, 2e-5)


def test_interp_fir_set_taps_midstream_and_errors(grb, orc):
    rng = np.random.default_rng(6)
    x = crand(rng, 4000)
    t1, t2 = rng.uniform(-1, 1, 12).astype(np.float32), rng.uniform(-1, 1, 41).astype(np.float32)
    blk = grb.filter.interp_fir_filter_ccf(4, t1)
    xin = np.concatenate([np.zeros(2, np.complex64), x])
    out1 = np.empty(8000, np.complex64)
    assert blk.work(8000, [xin], [out1]) == 8000
    blk.set_taps(t2)  # picked up by the next work(): installs, re-declares the history, returns 0 (:136-139)
    assert len(blk.taps()) == 44
    assert blk.work(400, [xin], [np.empty(400, np.complex64)]) == 0
    assert blk.history() == 11
    out2 = np.empty(8000, np.complex64)
    assert blk.work(8000, [x[2000 - 10:]], [out2]) == 8000
    assert_parity(out1, orc.interp_fir(4, t1, x)[:8000], "before", 2e-6, 2e-5)
    assert_parity(out2, orc.interp_fir(4, t2, x)[8000:16000], "after", 2e-6, 2e-5)
    with pytest.raises(IndexError):  # std::out_of_range
        grb.filter.interp_fir_filter_ccf(0, t1)
    with pytest.raises(RuntimeError):  # std::runtime_error
        grb.filter.interp_fir_filter_ccc(2, [])
    with pytest.raises(ValueError):  # too few input items for the request
        blk.work(400, [np.zeros(50, np.complex64)], [np.empty(400, np.complex64)])
    assert blk.work(0, [xin], [np.empty(0, np.complex64)]) == 0


def test_interp_fir_large_device_resident(grb):
    import torch

    rng = np.random.default_rng(7)
    L, T, ni = 4, 16, 1 << 20
    taps = rng.uniform(-1, 1, L * T).astype(np.float32)
    x = crand(rng, ni)
    xin = np.concatenate([np.zeros(T - 1, np.complex64), x])
    xd = torch.from_numpy(xin.view(np.float32)).cuda()
    yd = torch.empty(2 * ni * L, dtype=torch.float32, device="cuda")
    blk = grb.filter.interp_fir_filter_ccf(L, taps)
    blk.work_dev(xd.data_ptr(), yd.data_ptr(), ni * L, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert_parity(yd.cpu().numpy().view(np.complex64), po.f64_interp_fir(L, taps, x), "2^22 outputs", 2e-6, 2e-5)
