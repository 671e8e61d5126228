"""GPU parity (through the C-ABI): rational_resampler_base_ccf/ccc against the oracle, the committed reference fixtures,
the reference QA's golden vectors (gr-filter/python/filter/qa_rational_resampler.py:71-109) and float64.
Tolerance: 2e-6 NRMS / 2e-5 max-abs over RMS (north_star allows 1e-5 / 1e-4); the QA vectors exactly."""
import numpy as np
import pytest

from common import assert_parity, crand, fixtures
from oracle import pyoracle as po
from test_oracle import INTERP_QA_SRC, INTERP_QA_TAPS, RESAMP_QA0_EXPECTED, RESAMP_QA0_SRC, RESAMP_QA0_TAPS, RESAMP_QA1_EXPECTED

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def grb():
    import gnuradio_b200 as g

    assert g.device_count() > 0, "no CUDA device: libgrb200 has no CPU fallback"
    return g


def _cls(grb, taps):
    f = grb.filter
    return f.rational_resampler_base_ccc if np.iscomplexobj(taps) else f.rational_resampler_base_ccf


def run_stream(blk, x, max_noutput=4096):
    """vector_source -> block -> vector_sink the way block_executor drives a general block whose scheduler-visible
    history is 1 (the class shadows history()): noutput is stepped down until forecast() fits what is available."""
    x = np.ascontiguousarray(x, np.complex64)
    outs, consumed, zero_returns = [], 0, 0
    while True:
        avail = len(x) - consumed
        nout = max_noutput
        while nout > 0 and blk.forecast(nout, 1)[0] > avail:
            nout -= 1
        if nout <= 0:
            break
        ob = np.empty(nout, np.complex64)
        r = blk.general_work(nout, [avail], [x[consumed:]], [ob])
        if r == 0:  # once after set_taps(); twice in a row: the last window does not fit what is left
            zero_returns += 1
            if zero_returns >= 2:
                break
            continue
        zero_returns = 0
        outs.append(ob[:r])
        consumed += blk._consumed
    return np.concatenate(outs) if outs else np.empty(0, np.complex64)


def test_rational_resampler_qa_golden_vectors(grb):
    for dt in (np.float32, np.complex64):
        src = np.array(RESAMP_QA0_SRC, np.float32)
        blk = _cls(grb, np.zeros(1, dt))(1, 1, np.array(RESAMP_QA0_TAPS, dt))
        assert blk.history() == 2 and blk.relative_rate() == 1.0
        got = run_stream(blk, (src + 2j * src).astype(np.complex64))  # 6 outputs from 8 inputs, like the QA
        exp = np.array(RESAMP_QA0_EXPECTED, np.float32)
        assert np.array_equal(got.real, exp) and np.array_equal(got.imag, 2 * exp)
        # test_001_interp: the last three of the 24 windows read one item past the 8 the scheduler counted
        taps = np.array(INTERP_QA_TAPS, dt)
        blk = _cls(grb, taps)(3, 1, taps)
        assert blk.taps() == list(taps) + [0] and blk.history() == 2
        src = np.array(INTERP_QA_SRC + (0,), np.float32)
        ob = np.empty(24, np.complex64)
        assert blk.general_work(24, [8], [(src + 2j * src).astype(np.complex64)], [ob]) == 24 and blk._consumed == 8
        exp = np.array(RESAMP_QA1_EXPECTED, np.float32)
        assert np.array_equal(ob.real, exp) and np.array_equal(ob.imag, 2 * exp)


def test_rational_resampler_reference_fixtures(grb):
    fx = fixtures()
    for key, L, D, tk in (("resamp_ccf_4_7", 4, 7, "resamp_taps_f"), ("resamp_ccc_3_2", 3, 2, "resamp_taps_c")):
        blk = _cls(grb, fx[tk])(L, D, fx[tk])
        x = fx["resamp_x"]
        ob = np.empty(len(fx[key]) + 5, np.complex64)
        nin = len(x) - (blk.history() - 1)
        assert blk.general_work(len(ob), [nin], [x], [ob]) == len(fx[key])  # one call, input-limited
        assert_parity(ob[: len(fx[key])], fx[key], key, 2e-6, 2e-5)
        got = run_stream(_cls(grb, fx[tk])(L, D, fx[tk]), x, max_noutput=333)  # many calls: d_ctr carried
        assert 0 < len(got) <= len(fx[key])
        assert_parity(got, fx[key][: len(got)], key + " streamed", 2e-6, 2e-5)


@pytest.mark.parametrize("L,D,ntaps,cplx", [(1, 1, 1, False), (1, 1, 64, True), (1, 8, 64, False), (3, 1, 33, True),
                                            (2, 3, 7, False), (3, 2, 48, False), (4, 7, 128, False), (7, 4, 300, True),
                                            (160, 147, 1600, False), (147, 160, 1470, True), (5, 64, 55, False),
                                            (16, 1, 400, False)])
def test_rational_resampler_sweep(grb, orc, L, D, ntaps, cplx):
    rng = np.random.default_rng(L * 100000 + D * 1000 + ntaps)
    x = crand(rng, 20011)
    taps = crand(rng, ntaps) if cplx else rng.uniform(-1, 1, ntaps).astype(np.float32)
    blk = _cls(grb, taps)(L, D, taps)
    assert blk.history() == -(-ntaps // L) and blk.interpolation() == L and blk.decimation() == D
    got = run_stream(blk, x, max_noutput=4096)
    exp, _, _ = orc.rational_resampler(L, D, taps, x)
    assert len(exp) - 2 * (L // D + 2) <= len(got) <= len(exp)  # the forecast's +1 holds back the last few outputs
    assert_parity(got, exp[: len(got)], "oracle %d/%d/%d" % (L, D, ntaps), 2e-6, 2e-5)
    assert_parity(got, po.f64_rational_resampler(L, D, taps, x, len(got)), "f64 %d/%d/%d" % (L, D, ntaps), 2e-6, 2e-5)


def test_rational_resampler_counts_match_oracle_loop(grb, orc):
    # produced / consumed / d_ctr of every call agree with the literal loop (:160-172), including ragged requests
    rng = np.random.default_rng(5)
    x = crand(rng, 5000)
    taps = rng.uniform(-1, 1, 37).astype(np.float32)
    for L, D in ((5, 3), (3, 5), (1, 4), (4, 1), (7, 23)):
        blk = grb.filter.rational_resampler_base_ccf(L, D, taps)
        T = blk.history()
        pos, ctr = 0, 0
        for nout, nin in ((1, 1), (3, 2), (17, 100), (1000, 10), (10, 1000), (256, 256), (0, 5), (5, 0)):
            exp, used, ctr = orc.rational_resampler(L, D, taps, x[pos:], noutput=nout, ctr=ctr, nin=nin)
            ob = np.empty(nout, np.complex64)
            r = blk.general_work(nout, [nin], [x[pos:]], [ob])
            assert (r, blk._consumed, blk.ctr()) == (len(exp), used, ctr), (L, D, nout, nin)
            assert_parity(ob[:r], exp, "%d/%d call" % (L, D), 2e-6, 2e-5)
            pos += used
        assert T == -(-37 // L)


def test_rational_resampler_wide_index_path(grb, orc):
    # one launch whose s_i = ctr + i*D passes 2^32 (the kernel's 64-bit divide): 40000 one-tap phases, D = 39999
    rng = np.random.default_rng(12)
    L, D, nout = 40000, 39999, 120000
    taps = rng.uniform(-1, 1, L).astype(np.float32)
    x = crand(rng, nout + 8)
    blk = grb.filter.rational_resampler_base_ccf(L, D, taps)
    blk.set_ctr(17)
    ob = np.empty(nout, np.complex64)
    assert blk.general_work(nout, [len(x)], [x], [ob]) == nout and (17 + nout * D) >= 1 << 32
    exp, used, ctr = orc.rational_resampler(L, D, taps, x, noutput=nout, ctr=17)
    assert (blk._consumed, blk.ctr()) == (used, ctr)
    assert_parity(ob, exp, "wide index", 2e-6, 2e-5)


def test_rational_resampler_set_taps_midstream_and_errors(grb, orc):
    rng = np.random.default_rng(6)
    x = crand(rng, 4000)
    t1, t2 = rng.uniform(-1, 1, 12).astype(np.float32), rng.uniform(-1, 1, 41).astype(np.float32)
    blk = grb.filter.rational_resampler_base_ccf(4, 3, t1)
    ob = np.empty(1000, np.complex64)
    assert blk.general_work(1000, [2000], [x], [ob]) == 1000
    pos, ctr = blk._consumed, blk.ctr()
    assert (pos, ctr) == (750, 0)
    assert_parity(ob, orc.rational_resampler(4, 3, t1, x, noutput=1000)[0], "before", 2e-6, 2e-5)
    blk.set_taps(t2)  # picked up by the next general_work(): installs, re-declares the history, returns 0 (:151-154)
    assert len(blk.taps()) == 44
    assert blk.general_work(400, [100], [x[pos:]], [np.empty(400, np.complex64)]) == 0
    assert blk.history() == 11 and blk.ctr() == 0
    assert blk.general_work(1000, [2000], [x[pos:]], [ob]) == 1000
    assert_parity(ob, orc.rational_resampler(4, 3, t2, x[pos:], noutput=1000)[0], "after", 2e-6, 2e-5)
    with pytest.raises(IndexError):  # std::out_of_range
        grb.filter.rational_resampler_base_ccf(0, 1, t1)
    with pytest.raises(IndexError):
        grb.filter.rational_resampler_base_ccc(1, 0, t1)
    with pytest.raises(RuntimeError):
        grb.filter.rational_resampler_base_ccc(2, 3, [])
    with pytest.raises(Exception):
        blk.set_ctr(4)
    assert blk.general_work(0, [100], [x], [np.empty(0, np.complex64)]) == 0


@pytest.mark.parametrize("L,D,ntaps,cplx,ctr0,shift", [(3, 2, 48, False, 0, 0), (2, 3, 64, False, 1, 1), (4, 3, 61, True, 2, 0),
                                                       (5, 4, 100, False, 3, 1), (4, 1, 64, False, 0, 0), (32, 3, 96, False, 7, 0),
                                                       (7, 2, 100, True, 5, 1), (3, 4, 96, False, 1, 0), (2, 1, 64, False, 1, 1),
                                                       (3, 2, 20, False, 2, 0), (2, 3, 32, False, 0, 0), (3, 4, 40, True, 1, 1),
                                                       (2, 4, 30, False, 1, 0),  # (the last three: 16-group rows)
                                                       # interpolation 1 = a decimating FIR: the FIR handle's kernels (direct,
                                                       # tensor cores, overlap-save)
                                                       (1, 4, 64, False, 0, 0), (1, 1, 48, False, 0, 1), (1, 8, 700, False, 0, 0),
                                                       (1, 3, 200, True, 0, 1), (1, 2, 9, True, 0, 0)])
def test_rational_resampler_streaming_kernel(grb, orc, L, D, ntaps, cplx, ctr0, shift):
    """One device-resident call long enough for resamp_stream_kernel (small interpolation and decimation): every ring depth
    and decimation step, interpolations that do not divide the warp, complex taps, a starting phase counter, an input
    pointer that is only 8-byte aligned, an output count that is not a multiple of the interpolation."""
    import torch

    rng = np.random.default_rng(L * 131 + D * 17 + ntaps)
    nout = 200001
    T = -(-ntaps // L)
    taps = crand(rng, ntaps) if cplx else rng.uniform(-1, 1, ntaps).astype(np.float32)
    ni = (ctr0 + (nout - 1) * D) // L + 1
    x = crand(rng, ni + T - 1)
    xd = torch.from_numpy(np.concatenate([np.zeros(shift, np.complex64), x]).view(np.float32).copy()).cuda()
    yd = torch.full((2 * nout + 2,), float("nan"), dtype=torch.float32, device="cuda")
    cls = grb.filter.rational_resampler_base_ccc if cplx else grb.filter.rational_resampler_base_ccf
    blk = cls(L, D, taps)
    blk.set_ctr(ctr0)
    used, made = blk.work_dev(xd.data_ptr() + 8 * shift, ni, yd.data_ptr(), nout, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert made == nout
    y = yd.cpu().numpy()
    assert np.isnan(y[-2:]).all()  # nothing written past the end
    got = y[:-2].view(np.complex64)
    exp, used_ref, ctr_ref = orc.rational_resampler(L, D, taps, x, noutput=nout, ctr=ctr0)
    assert (used, blk.ctr()) == (used_ref, ctr_ref)
    assert_parity(got, exp, "oracle %d/%d/%d" % (L, D, ntaps), 2e-6, 2e-5)


def test_rational_resampler_large_device_resident_and_host_chunks(grb):
    import torch

    rng = np.random.default_rng(7)
    L, D, T, ni = 3, 2, 16, 1 << 21
    taps = rng.uniform(-1, 1, L * T).astype(np.float32)
    x = crand(rng, ni + T - 1)
    nout = ni * L // D
    xd = torch.from_numpy(x.view(np.float32)).cuda()
    yd = torch.empty(2 * nout, dtype=torch.float32, device="cuda")
    blk = grb.filter.rational_resampler_base_ccf(L, D, taps)
    used, made = blk.work_dev(xd.data_ptr(), ni, yd.data_ptr(), nout, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert (used, made) == (ni, nout) and blk.ctr() == 0
    ref = po.f64_rational_resampler(L, D, taps, x, nout)
    dev = yd.cpu().numpy().view(np.complex64)
    assert_parity(dev, ref, "2^21 inputs, device", 2e-6, 2e-5)
    # host-staged path over several chunks (heavy decimation makes the input span the chunk limit) is the same bits
    blk2 = grb.filter.rational_resampler_base_ccf(L, D, taps)
    ob = np.empty(nout, np.complex64)
    assert blk2.general_work(nout, [ni], [x], [ob]) == nout and blk2._consumed == ni
    assert np.array_equal(ob.view(np.uint32), dev.view(np.uint32))
    D2 = 37
    n2 = ni * L // D2
    blk3 = grb.filter.rational_resampler_base_ccf(L, D2, taps)
    ob = np.empty(n2, np.complex64)
    assert blk3.general_work(n2, [ni], [x], [ob]) == n2
    assert_parity(ob, po.f64_rational_resampler(L, D2, taps, x, n2), "3/37 host", 2e-6, 2e-5)
