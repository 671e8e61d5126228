"""GPU parity (through the C-ABI): freq_xlating_fir_filter_ccf/ccc against the oracle, the committed reference
fixtures, the reference QA cases (gr-filter/python/filter/qa_freq_xlating_fir_filter.py) and float64.

Tolerance.  north_star: NRMS <= 1e-5, max-abs <= 1e-4 x RMS.  The device rotator evaluates the closed form
e^{j(theta + n delta)} (angle in float64, delta = the angle of the float increment phasor the block builds), so against
FLOAT64 it is held to 2e-6 / 2e-5 at any length.  The reference's rotator is a float phasor recurrence
(rotator.h:40-41): its |phase| wanders by a few 1e-6 between the renormalisations every 512 items, and its angle
drifts off the increment's by up to one float ulp per item, systematically (measured: -1.15e-8 rad/item at
fc/fs = 0.1, i.e. 3.5e-5 rad after 3000 items; tests/test_oracle.py pins this against float64).  Against the
reference layer the comparison is therefore split (assert_vs_reference): magnitudes to north_star's bound, the phase
difference to the drift bound 2e-6 + n * 2^-24 rad, and the first 512 items -- before drift matters -- in full."""
import numpy as np
import pytest

import sched_mock
from common import assert_parity, crand, fixtures
from oracle import pyoracle as po

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def grb():
    import gnuradio_b200 as g

    assert g.device_count() > 0, "no CUDA device: libgrb200 has no CPU fallback"
    return g


def assert_vs_reference(got, ref, what):
    got, ref = np.asarray(got), np.asarray(ref)
    assert got.shape == ref.shape, (what, got.shape, ref.shape)
    assert_parity(np.abs(got), np.abs(ref), what + " |.|", 1e-5, 1e-4)
    assert_parity(got[:512], ref[:512], what + " first 512", 1e-5, 1e-4)
    big = np.abs(ref) > 0.1 * np.sqrt(np.mean(np.abs(ref) ** 2))
    dphi = np.abs(np.angle(got[big].astype(np.complex128) * np.conj(ref[big].astype(np.complex128))))
    n = np.flatnonzero(big) + 1.0
    assert np.all(dphi <= 2e-6 + n * 2.0 ** -24), "%s: phase off by %.3g rad at item %d" % (
        what, dphi.max(), int(n[np.argmax(dphi)]))


def _cls(grb, taps):
    return grb.filter.freq_xlating_fir_filter_ccc if np.iscomplexobj(taps) else grb.filter.freq_xlating_fir_filter_ccf


def test_freq_xlating_reference_fixtures(grb):
    fx = fixtures()
    fc, fs, chunk = fx["xlat_meta"]
    for D in (1, 4):
        for tag in ("f", "c"):
            t = fx["xlat_taps_" + tag]
            blk = _cls(grb, t)(D, t, fc, fs)
            got = sched_mock.run_sync(blk, fx["xlat_x"], max_noutput=int(chunk) // D)
            assert_vs_reference(got, fx["xlat_cc%s_d%d" % (tag, D)], "fixture cc%s d%d" % (tag, D))
            assert_parity(got, po.f64_freq_xlating_fir(D, t, fc, fs, fx["xlat_x"]), "f64 cc%s d%d" % (tag, D),
                          2e-6, 2e-5)


def test_freq_xlating_qa_cases(grb, orc):
    # qa_freq_xlating_fir_filter.py:96-160 (ccf_001/002, ccc_001/002): expected = fir_filter(mix(lo, src), taps, decim),
    # assertComplexTuplesAlmostEqual(..., 5)
    fs, fc, bw = 1.0, 0.3, 0.1
    lp = orc.low_pass_2(1, fs, bw, bw / 4, 53)
    src = np.exp(-2j * np.pi * fc / fs * (np.arange(100) / 100.0)).astype(np.complex64)
    lo = np.exp(-2j * np.pi * fc / fs * np.arange(100))
    for taps in (lp, (lp * np.exp(2j * np.pi * bw / 8 * np.arange(len(lp)))).astype(np.complex64)):
        for D in (1, 4):
            got = sched_mock.run_sync(_cls(grb, taps)(D, taps, fc, fs), src)
            assert len(got) == 100 // D
            np.testing.assert_allclose(got, po.f64_fir(D, taps, lo * src.astype(np.complex128)), atol=0.5e-5, rtol=0)


@pytest.mark.parametrize("ntaps,decim,cplx", [(1, 1, False), (7, 1, True), (64, 1, False), (64, 1, True), (63, 2, False),
                                              (200, 5, True), (1024, 8, False), (3001, 25, False)])
def test_freq_xlating_sweep(grb, orc, ntaps, decim, cplx):
    rng = np.random.default_rng(ntaps * 31 + decim)
    nout = 4000
    x = crand(rng, nout * decim + decim - 1)  # ragged tail: the last decim-1 items yield no output
    taps = crand(rng, ntaps) if cplx else rng.uniform(-1, 1, ntaps).astype(np.float32)
    fc, fs = 12345.678, 250e3
    got = sched_mock.run_sync(_cls(grb, taps)(decim, taps, fc, fs), x, max_noutput=1500)
    xin = np.concatenate([np.zeros(ntaps - 1, np.complex64), x])
    ref, _ = orc.freq_xlating_fir(decim, taps, fc, fs, xin, nout, chunk_out=1500)
    assert_vs_reference(got, ref, "reference %d/%d" % (ntaps, decim))
    assert_parity(got, po.f64_freq_xlating_fir(decim, taps, fc, fs, x), "f64 %d/%d" % (ntaps, decim), 2e-6, 2e-5)


@pytest.mark.parametrize("cplx", [False, True])
def test_freq_xlating_long_run_device_resident(grb, cplx):
    """2^22 outputs through the tcgen05 complex-tap FIR + rotator, device-resident, in three uneven calls: the phase
    carries across calls and the result stays within tolerance of float64 at a length where a float phasor
    recurrence has long since drifted."""
    import torch

    rng = np.random.default_rng(11)
    n, nt = 1 << 22, 64
    data = crand(rng, n)
    x = np.concatenate([np.zeros(nt - 1, np.complex64), data])  # a fresh stream: zero history
    taps = crand(rng, nt) if cplx else rng.uniform(-1, 1, nt).astype(np.float32)
    fc, fs = -0.1234, 1.0
    blk = _cls(grb, taps)(1, taps, fc, fs)
    xd = torch.from_numpy(x.view(np.float32)).cuda()
    yd = torch.empty(2 * n, dtype=torch.float32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    cuts = [0, 64 * 63 * 100 + 17, 3000001, n]
    for a, b in zip(cuts[:-1], cuts[1:]):
        blk.work_dev(xd.data_ptr() + 8 * a, yd.data_ptr() + 8 * a, b - a, st)
    torch.cuda.synchronize()
    got = yd.cpu().numpy().view(np.complex64)
    assert_parity(got, po.f64_freq_xlating_fir(1, taps, fc, fs, data), "2^22 device-resident vs f64")


@pytest.mark.parametrize("ntaps,decim", [(128, 8), (40, 4), (100, 3), (33, 1), (2000, 16), (200, 10), (401, 20), (250, 50)])
def test_freq_xlating_on_the_overlap_save_route(grb, monkeypatch, ntaps, decim):
    """Composite taps past the FIR kernels' crossover run on the overlap-save kernels (every store form of fftfilt_ols3_kernel:
    full-rate, decimation dividing 256, any other decimation, the ragged last block; 2000 taps: the same), rotate_kernel
    behind them; the same calls on the direct kernels (GRB200_FIR_FFT=0) and float64 agree, across two uneven device-resident
    calls."""
    import torch

    rng = np.random.default_rng(ntaps + decim)
    nout = ((1 << 21) + 12345) // decim
    data = crand(rng, nout * decim)
    x = np.concatenate([np.zeros(ntaps - 1, np.complex64), data])
    taps = (rng.uniform(-1, 1, ntaps) * (4.0 / ntaps)).astype(np.float32)
    fc, fs = 0.0731, 1.0
    xd = torch.from_numpy(x.view(np.float32)).cuda()
    st = torch.cuda.current_stream().cuda_stream
    res = []
    for route in (None, "0"):
        if route is None:
            monkeypatch.delenv("GRB200_FIR_FFT", raising=False)
        else:
            monkeypatch.setenv("GRB200_FIR_FFT", route)
        blk = grb.filter.freq_xlating_fir_filter_ccf(decim, taps, fc, fs)
        yd = torch.full((2 * (nout + 1),), float("nan"), dtype=torch.float32, device="cuda")
        cut = min(70001, nout // 3)
        blk.work_dev(xd.data_ptr(), yd.data_ptr(), cut, st)
        blk.work_dev(xd.data_ptr() + 8 * cut * decim, yd.data_ptr() + 8 * cut, nout - cut, st)
        torch.cuda.synchronize()
        y = yd.cpu().numpy()
        assert np.isnan(y[-2:]).all() and np.isfinite(y[:-2]).all()
        res.append(y[:-2].view(np.complex64))
    ref = po.f64_freq_xlating_fir(decim, taps, fc, fs, data)
    assert_parity(res[0], ref, "overlap-save route vs f64", 2e-6, 2e-5)
    assert_parity(res[1], ref, "direct kernels vs f64", 2e-6, 2e-5)
    assert_parity(res[0], res[1], "one against the other", 2e-6, 2e-5)


def test_freq_xlating_host_chunking_is_invisible(grb):
    """The host-staged path cuts a large work() into pipeline chunks; the rotator angle of each chunk comes from the
    closed form, so one big call equals many scheduler-sized ones bit for bit on the CUDA-core FIR."""
    rng = np.random.default_rng(12)
    nt, D, nout = 33, 3, 300000
    taps = crand(rng, nt)
    x = crand(rng, nout * D)
    a = sched_mock.run_sync(grb.filter.freq_xlating_fir_filter_ccc(D, taps, 1000.0, 48000.0), x, max_noutput=1 << 20)
    b = sched_mock.run_sync(grb.filter.freq_xlating_fir_filter_ccc(D, taps, 1000.0, 48000.0), x, max_noutput=4096)
    assert len(a) == len(b) == nout
    # the chunk's angle is reduced mod 2 pi in float64 before float sincos: equal up to that reduction's last bit
    np.testing.assert_allclose(a, b, atol=2e-6 * np.abs(a).max(), rtol=0)
    assert_parity(a, po.f64_freq_xlating_fir(D, taps, 1000.0, 48000.0, x), "f64", 2e-6, 2e-5)


def test_freq_xlating_set_center_freq_and_taps_midstream(grb, orc):
    # freq_xlating_fir_filter_impl.cc:96-113,155-160: the work() after a change rebuilds, re-declares history and
    # returns 0; the rotator keeps its phase and only its increment changes.
    rng = np.random.default_rng(13)
    x = crand(rng, 6000)
    t1, t2 = rng.uniform(-1, 1, 16).astype(np.float32), rng.uniform(-1, 1, 40).astype(np.float32)
    fs = 1e6
    blk = grb.filter.freq_xlating_fir_filter_ccf(1, t1, 1e5, fs)
    assert blk.center_freq() == 1e5 and blk.history() == 16
    xin = np.concatenate([np.zeros(15, np.complex64), x])
    out1 = np.empty(3000, np.complex64)
    assert blk.work(3000, [xin], [out1]) == 3000
    blk.set_center_freq(-2.5e5)
    blk.set_taps(t2)
    assert blk.taps() == t2.tolist() and blk.center_freq() == -2.5e5
    assert blk.work(100, [xin], [np.empty(100, np.complex64)]) == 0
    assert blk.history() == 40
    out2 = np.empty(3000, np.complex64)
    assert blk.work(3000, [x[3000 - 39:]], [out2]) == 3000
    ref1, ph = orc.freq_xlating_fir(1, t1, 1e5, fs, xin, 3000)
    ref2, _ = orc.freq_xlating_fir(1, t2, -2.5e5, fs, x[3000 - 39:], 3000, phase=ph)
    assert_vs_reference(out1, ref1, "before")
    # after the retune the rotator carries on from the phase it had reached: the reference's has drifted by then
    assert_parity(np.abs(out2), np.abs(ref2), "after |.|", 1e-5, 1e-4)
    dphi = np.angle(out2.astype(np.complex128) * np.conj(ref2.astype(np.complex128)))[np.abs(ref2) > 0.5]
    assert np.abs(dphi).max() <= 2e-6 + 6000 * 2.0 ** -24
    # composite taps as the reference builds them (float arithmetic, :85-88)
    fwT0 = np.float32(2 * np.pi * -2.5e5 / fs)
    arg = (np.arange(40, dtype=np.uint32).astype(np.float32) * fwT0).astype(np.float32)
    exp = t2 * (np.cos(arg, dtype=np.float32) + 1j * np.sin(arg, dtype=np.float32)).astype(np.complex64)
    np.testing.assert_allclose(blk.composite_taps(), exp, atol=3e-7, rtol=0)


def test_freq_xlating_message_port_and_errors(grb):
    blk = grb.filter.freq_xlating_fir_filter_ccf(2, [1.0, 0.5], 10.0, 100.0)
    blk.handle_set_center_freq({"freq": 20.0})  # dict form, :124-130
    assert blk.center_freq() == 20.0 and blk.d_updated
    blk.handle_set_center_freq(("freq", 30))  # pair form, :131-137
    assert blk.center_freq() == 30.0
    blk.handle_set_center_freq("nonsense")
    blk.handle_set_center_freq({"other": 1.0})
    assert blk.center_freq() == 30.0
    with pytest.raises(ValueError):
        grb.filter.freq_xlating_fir_filter_ccf(0, [1.0], 0.0, 1.0)
    with pytest.raises(ValueError):
        grb.filter.freq_xlating_fir_filter_ccc(1, [], 0.0, 1.0)
    blk2 = grb.filter.freq_xlating_fir_filter_ccf(1, [1.0, 2.0], 0.0, 1.0)
    assert blk2.work(0, [np.zeros(1, np.complex64)], [np.empty(0, np.complex64)]) == 0
    with pytest.raises(ValueError):  # too few input items for the request
        blk2.work(10, [np.zeros(5, np.complex64)], [np.empty(10, np.complex64)])


def test_freq_xlating_zero_frequency_is_plain_fir(grb):
    rng = np.random.default_rng(14)
    taps = rng.uniform(-1, 1, 64).astype(np.float32)
    x = crand(rng, 100000)
    a = sched_mock.run_sync(grb.filter.freq_xlating_fir_filter_ccf(1, taps, 0.0, 1.0), x, max_noutput=1 << 20)
    assert_parity(a, po.f64_fir(1, taps, x), "fc=0", 2e-6, 2e-5)


def test_freq_xlating_time_sharded_segments(grb):
    """Time-segment sharding: each shard filters its own segment (ntaps-1 halo items in front) with
    its own block, rotator advanced to the segment's first output; the concatenation equals the one-block result."""
    from gnuradio_b200 import sharding

    rng = np.random.default_rng(15)
    nt, D, nout, world = 48, 4, 30000, 3
    taps = crand(rng, nt)
    x = crand(rng, nout * D)
    xin = np.concatenate([np.zeros(nt - 1, np.complex64), x])
    whole = np.empty(nout, np.complex64)
    one = grb.filter.freq_xlating_fir_filter_ccc(D, taps, 2.5e3, 48e3)
    assert one.work(nout, [xin], [whole]) == nout
    parts = []
    for r in range(world):
        seg = sharding.segment(r, world, nout, decimation=D, history=nt)
        o0, o1 = seg.out_begin, seg.out_end
        assert seg.in_begin == o0 * D
        blk = grb.filter.freq_xlating_fir_filter_ccc(D, taps, 2.5e3, 48e3)
        blk.advance(o0)
        out = np.empty(o1 - o0, np.complex64)
        assert blk.work(o1 - o0, [xin[o0 * D:]], [out]) == o1 - o0
        parts.append(out)
    got = np.concatenate(parts)
    np.testing.assert_allclose(got, whole, atol=2e-6 * np.abs(whole).max(), rtol=0)
