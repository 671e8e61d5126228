"""GPU parity (through the C-ABI): fir_filter_ccf/ccc and fft_filter_ccc against the oracle, the committed
reference fixtures and the reference QA cases.  Tolerance: north_star (NRMS <= 1e-5, max-abs <= 1e-4 x RMS)."""
import random

import numpy as np
import pytest

import sched_mock
from common import assert_parity, assert_tuples_almost_equal2, crand, fixtures, make_random_complex_tuple
from oracle import pyoracle as po

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def grb():
    import gnuradio_b200 as g

    assert g.device_count() > 0, "no CUDA device: libgrb200 has no CPU fallback"
    return g


def test_fir_reference_fixtures(grb):
    fx = fixtures()
    for D in (1, 4):
        for kind, cls in (("ccf", grb.filter.fir_filter_ccf), ("ccc", grb.filter.fir_filter_ccc)):
            taps = fx["fir_taps_f"] if kind == "ccf" else fx["fir_taps_c"]
            got = sched_mock.run_sync(cls(D, taps), fx["fir_x"], max_noutput=700)
            assert_parity(got, fx["fir_%s_d%d" % (kind, D)], "%s d%d" % (kind, D))


def test_fir_qa_cases(grb):
    # qa_fir_filter.py:60-129
    src = np.array(40 * [1 + 1j, 2 + 2j, 3 + 3j, 4 + 4j], np.complex64)
    for D in (1, 4):
        taps = np.array(20 * [0.5, 0.5], np.float32)
        got = sched_mock.run_sync(grb.filter.fir_filter_ccf(D, taps), src)
        np.testing.assert_allclose(got, po.f64_fir(D, taps, src), atol=0.5e-5 * 80)
        tc = np.array(20 * [0.5 + 1j, 0.5 + 1j], np.complex64)
        got = sched_mock.run_sync(grb.filter.fir_filter_ccc(D, tc), src)
        np.testing.assert_allclose(got, po.f64_fir(D, tc, src), atol=0.5e-5 * 200)


@pytest.mark.parametrize("ntaps", [1, 2, 7, 63, 64, 65, 200, 1024, 3001])
@pytest.mark.parametrize("decim", [1, 2, 3, 8, 25])
def test_fir_ccf_sweep(grb, orc, ntaps, decim):
    rng = np.random.default_rng(ntaps * 100 + decim)
    x = crand(rng, 40000 + 13)
    taps = rng.uniform(-1, 1, ntaps).astype(np.float32)
    got = sched_mock.run_sync(grb.filter.fir_filter_ccf(decim, taps), x, max_noutput=100000)
    ref = orc.fir_stream(decim, taps, x)
    assert_parity(got, ref, "ccf %d/%d" % (ntaps, decim))
    assert_parity(got, po.f64_fir(decim, taps, x), "ccf f64 %d/%d" % (ntaps, decim))


@pytest.mark.parametrize("ntaps,decim", [(1, 1), (5, 1), (64, 1), (64, 4), (257, 3), (1024, 8), (2500, 1)])
def test_fir_ccc_sweep(grb, orc, ntaps, decim):
    rng = np.random.default_rng(ntaps * 7 + decim)
    x = crand(rng, 30000 + 5)
    taps = crand(rng, ntaps)
    got = sched_mock.run_sync(grb.filter.fir_filter_ccc(decim, taps), x, max_noutput=100000)
    assert_parity(got, orc.fir_stream(decim, taps, x), "ccc %d/%d" % (ntaps, decim))
    assert_parity(got, po.f64_fir(decim, taps, x), "ccc f64")


@pytest.mark.parametrize("ntaps,decim,cplx", [(1024, 8, False), (1000, 1, False), (777, 3, False), (600, 2, True), (2048, 5, True)])
def test_fir_long_taps_many_tiles(grb, orc, ntaps, decim, cplx):
    """Long filters take the split-tap kernel (taps spread over the warps of a CTA, partial sums reduced in
    shared memory); enough outputs that every persistent CTA walks several double-buffered tiles."""
    rng = np.random.default_rng(ntaps + decim)
    nout = 296 * 512 * 3 + 777
    x = crand(rng, nout * decim)
    taps = crand(rng, ntaps) if cplx else rng.uniform(-1, 1, ntaps).astype(np.float32)
    blk = (grb.filter.fir_filter_ccc if cplx else grb.filter.fir_filter_ccf)(decim, taps)
    got = sched_mock.run_sync(blk, x, max_noutput=1 << 24)
    assert len(got) == nout
    # float64 check on three windows (a full float64 convolution of this size would take minutes)
    xin = np.concatenate([np.zeros(ntaps - 1, np.complex64), x])
    for o0 in (0, nout // 2 - 100, nout - 3000):
        seg = xin[o0 * decim: (o0 + 3000) * decim + ntaps - 1]
        assert_parity(got[o0:o0 + 3000], orc.fir(decim, taps, seg, 3000), "oracle window %d" % o0)
        ref64 = po.f64_fir(decim, taps, seg[ntaps - 1:] if o0 == 0 else seg)
        if o0 == 0:
            assert_parity(got[:3000], ref64[:3000], "f64 window 0")


def test_fir_scheduler_sized_chunks_equal_one_call(grb, monkeypatch):
    """Chunking must not change the result.  On the CUDA-core kernels that holds bit for bit; the tcgen05 path
    (large, 16-byte aligned calls) sums each output in an order that depends on its position inside a 64-item row,
    so there the invariant is the parity tolerance (the reference's VOLK kernels are alignment dependent too:
    gr-filter/lib/fir_filter.cc:67-74,133-141)."""
    rng = np.random.default_rng(3)
    x, taps = crand(rng, 50000), rng.uniform(-1, 1, 64).astype(np.float32)
    monkeypatch.setenv("GRB200_FIR_TC", "0")
    a = sched_mock.run_sync(grb.filter.fir_filter_ccf(1, taps), x, max_noutput=4096)
    b = sched_mock.run_sync(grb.filter.fir_filter_ccf(1, taps), x, max_noutput=1 << 20)
    c = sched_mock.run_sync(grb.filter.fir_filter_ccf(1, taps), x, max_noutput=1)[:64]
    np.testing.assert_array_equal(a, b)
    np.testing.assert_array_equal(a[:64], c)
    monkeypatch.delenv("GRB200_FIR_TC")
    d = sched_mock.run_sync(grb.filter.fir_filter_ccf(1, taps), x, max_noutput=1 << 20)
    assert_parity(d, a, "tensor-core path vs chunked CUDA-core path")
    assert_parity(d, po.f64_fir(1, taps, x), "tensor-core path vs float64")


@pytest.mark.parametrize("cplx", [False, True])
@pytest.mark.parametrize("ntaps,nout", [(64, 64 * 63), (64, 64 * 63 + 1), (64, 64 * 63 * 148 * 2 + 4095), (63, 100001),
                                        (37, 250000), (2, 8191), (1, 70000), (64, 1 << 21)])
def test_fir_tensor_core_path(grb, orc, monkeypatch, ntaps, nout, cplx):
    """fir_tc_kernel (tcgen05 block-Toeplitz FIR, 3xTF32), real and complex taps: ragged last tile, short filters
    (zero-padded Toeplitz), one tile, several tiles per CTA; against the C oracle and float64, and against the
    CUDA-core kernel."""
    monkeypatch.setenv("GRB200_FIR_TC", "1")
    rng = np.random.default_rng(1000 * ntaps + nout % 997 + cplx)
    x = crand(rng, nout)
    taps = crand(rng, ntaps) if cplx else rng.uniform(-1, 1, ntaps).astype(np.float32)
    cls = grb.filter.fir_filter_ccc if cplx else grb.filter.fir_filter_ccf
    xin = np.concatenate([np.zeros(ntaps - 1, np.complex64), x])
    got = cls(1, taps).process(xin, nout)
    assert len(got) == nout and np.isfinite(got.view(np.float32)).all()
    assert_parity(got, po.f64_fir(1, taps, x), "tc f64 %d/%d" % (ntaps, nout))
    n0 = min(nout, 200000)
    assert_parity(got[:n0], orc.fir(1, taps, xin[:n0 + ntaps - 1], n0), "tc oracle")
    monkeypatch.setenv("GRB200_FIR_TC", "0")
    ref = cls(1, taps).process(xin, nout)
    assert_parity(got, ref, "tc vs cuda-core")


@pytest.mark.parametrize("ntaps,nout", [(1024, 384 * 148 * 2 + 5), (1024, 4032), (1023, 9999), (969, 50000), (968, 50000),
                                        (513, 70001), (256, 40000), (1000, 384 * 3 + 1 + 4032)])
def test_fir_tensor_core_decimate_by_8(grb, orc, monkeypatch, ntaps, nout):
    """fir_tc_kernel<M_DEC8> (BASELINE configs[3] geometry: decimation 8, up to 1024 real taps): 16 tap blocks stacked in the
    MMA lanes, the partial 17th block (taps 968..) on the CUDA cores; ragged last tile, tap counts either side of 968."""
    monkeypatch.setenv("GRB200_FIR_TC", "1")
    rng = np.random.default_rng(ntaps + nout % 1009)
    D = 8
    x = crand(rng, nout * D)
    taps = rng.uniform(-1, 1, ntaps).astype(np.float32)
    xin = np.concatenate([np.zeros(ntaps - 1, np.complex64), x])
    got = grb.filter.fir_filter_ccf(D, taps).process(xin, nout)
    assert len(got) == nout and np.isfinite(got.view(np.float32)).all()
    for o0 in sorted({0, max(0, nout // 2 - 500), max(0, nout - 2000)}):
        n0 = min(2000, nout - o0)
        seg = xin[o0 * D: (o0 + n0) * D + ntaps - 1]
        assert_parity(got[o0:o0 + n0], orc.fir(D, taps, seg, n0), "dec8 oracle window %d" % o0)
        if o0 == 0:
            assert_parity(got[:n0], po.f64_fir(D, taps, x[: n0 * D]), "dec8 f64 window 0")
    monkeypatch.setenv("GRB200_FIR_TC", "0")
    ref = grb.filter.fir_filter_ccf(D, taps).process(xin, nout)
    assert_parity(got, ref, "dec8 tc vs cuda-core")


@pytest.mark.parametrize("mode", ["ccf", "ccc", "dec8"])
def test_fir_tensor_core_run_to_run_identical(grb, monkeypatch, mode):
    """The warp-specialised pipeline (TMA ring, converter / MMA / epilogue hand-offs over mbarriers, TMEM accumulator ring)
    must be race free: the same call repeated gives bit-identical output, at a size where every CTA walks many tiles."""
    monkeypatch.setenv("GRB200_FIR_TC", "1")
    rng = np.random.default_rng(5)
    D, ntaps = (8, 1024) if mode == "dec8" else (1, 64)
    nout = (1 << 22) // D + 777
    taps = crand(rng, ntaps) if mode == "ccc" else rng.uniform(-1, 1, ntaps).astype(np.float32)
    blk = (grb.filter.fir_filter_ccc if mode == "ccc" else grb.filter.fir_filter_ccf)(D, taps)
    xin = crand(rng, nout * D + ntaps - 1)
    first = blk.process(xin, nout)
    for _ in range(8):
        np.testing.assert_array_equal(blk.process(xin, nout), first)


def test_fir_tensor_core_linearity_and_impulse(grb, monkeypatch):
    """Size-independent properties at a BASELINE-sized call (2^24): impulse -> taps, linearity, shift invariance."""
    monkeypatch.setenv("GRB200_FIR_TC", "1")
    rng = np.random.default_rng(77)
    n, ntaps = 1 << 24, 64
    taps = rng.uniform(-1, 1, ntaps).astype(np.float32)
    blk = grb.filter.fir_filter_ccf(1, taps)
    imp = np.zeros(n + ntaps - 1, np.complex64)
    pos = [ntaps - 1, ntaps - 1 + 4031, ntaps - 1 + 64 * 63 * 148 + 5, n - 100]
    for q in pos:
        imp[q] = 1.0 + 2.0j
    y = blk.process(imp, n)
    for q in pos:
        o = q - (ntaps - 1)
        seg = y[o:o + ntaps]
        assert_parity(seg, (taps * np.complex64(1 + 2j))[:len(seg)], "impulse at %d" % q)
    y[np.concatenate([np.arange(q - (ntaps - 1), min(q + 1, n)) for q in pos])] = 0
    assert not y.any()
    a, b = crand(rng, n + ntaps - 1), crand(rng, n + ntaps - 1)
    ya, yb, yab = blk.process(a, n), blk.process(b, n), blk.process((a + np.complex64(0.5) * b).astype(np.complex64), n)
    assert_parity(yab, ya + np.complex64(0.5) * yb, "linearity")
    ys = blk.process(a[64 * 7:], n - 64 * 7)
    assert_parity(ys, ya[64 * 7:], "shift by 7 rows")


def test_fir_empty_and_errors(grb):
    blk = grb.filter.fir_filter_ccf(1, [1.0, 2.0])
    out = np.empty(0, np.complex64)
    assert blk.work(0, [np.zeros(1, np.complex64)], [out]) == 0
    with pytest.raises(ValueError):
        grb.filter.fir_filter_ccf(0, [1.0])
    with pytest.raises(ValueError):
        grb.filter.fir_filter_ccf(1, [])
    with pytest.raises(ValueError):  # too few input items for the request
        blk.work(10, [np.zeros(5, np.complex64)], [np.empty(10, np.complex64)])


def test_fir_set_taps_midstream(grb, orc):
    # fir_filter_blk_impl.cc:53-59,77-81: next work() returns 0 once and re-declares history
    rng = np.random.default_rng(4)
    x = crand(rng, 6000)
    t1, t2 = rng.uniform(-1, 1, 16).astype(np.float32), rng.uniform(-1, 1, 40).astype(np.float32)
    blk = grb.filter.fir_filter_ccf(1, t1)
    xin = np.concatenate([np.zeros(15, np.complex64), x])
    out1 = np.empty(3000, np.complex64)
    assert blk.work(3000, [xin], [out1]) == 3000
    blk.set_taps(t2)
    assert blk.taps() == t2.tolist()
    assert blk.work(100, [xin], [np.empty(100, np.complex64)]) == 0
    assert blk.history() == 40
    xin2 = x[3000 - 39:]  # the scheduler now presents 39 history items
    out2 = np.empty(3000, np.complex64)
    assert blk.work(3000, [xin2], [out2]) == 3000
    assert_parity(out1, orc.fir_stream(1, t1, x)[:3000], "before")
    assert_parity(out2, orc.fir_stream(1, t2, x)[3000:], "after")


def test_fir_large_device_resident(grb, orc):
    import torch

    rng = np.random.default_rng(5)
    n = 1 << 22
    x = crand(rng, n + 63)
    taps = rng.uniform(-1, 1, 64).astype(np.float32)
    blk = grb.filter.fir_filter_ccf(1, taps)
    xd = torch.from_numpy(x.view(np.float32)).cuda()
    yd = torch.empty(2 * n, dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    blk.work_dev(xd.data_ptr(), yd.data_ptr(), n, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    got = yd.cpu().numpy().view(np.complex64)
    assert_parity(got, orc.fir(1, taps, x, n), "2^22 device-resident")


# ---------------------------------------------------------------------------------------------------
def test_fft_filter_fixtures(grb):
    fx = fixtures()
    for D in (1, 3):
        blk = grb.filter.fft_filter_ccc(D, fx["fftfilt_taps"])
        assert blk.nsamples() == 725 and blk.output_multiple() == 725
        got = sched_mock.run_sync(blk, fx["fftfilt_x"], max_noutput=725 * 2)
        assert_parity(got, fx["fftfilt_d%d" % D], "fixture d%d" % D)


def test_fft_filter_qa_random_trials(grb, orc):
    # qa_fft_filter.py:144-184: seeded random data, fft_filter_ccc against fir_filter_ccc, rel_eps 4e-4
    random.seed(0)
    for i in range(25):
        src = np.array(make_random_complex_tuple(4096), np.complex64)
        ntaps = int(random.uniform(2, 1000))
        taps = np.array(make_random_complex_tuple(ntaps), np.complex64)
        exp_gpu = sched_mock.run_sync(grb.filter.fir_filter_ccc(1, taps), src)
        got = sched_mock.run_sync(grb.filter.fft_filter_ccc(1, taps), src)
        assert len(got) == (4096 // orc.fft_filter_nsamples(ntaps)) * orc.fft_filter_nsamples(ntaps)
        assert_parity(got, exp_gpu[: len(got)], "trial %d vs gpu fir" % i)
        assert_parity(got, orc.fft_filter_ccc(1, taps, src), "trial %d vs oracle" % i)
    random.seed(0)
    for i in range(25):
        dec = i + 1
        src = np.array(make_random_complex_tuple(4096), np.complex64)
        taps = np.array(make_random_complex_tuple(int(random.uniform(2, 100))), np.complex64)
        got = sched_mock.run_sync(grb.filter.fft_filter_ccc(dec, taps, 2), src)
        ref = orc.fft_filter_ccc(dec, taps, src)
        assert_parity(got, ref, "dec %d" % dec)


def test_fft_filter_ccf(grb, orc):
    """fft_filter_ccf (qa_fft_filter.py:214-283): real taps through grb200_fftfilt_ccf_create /
    set_taps_real, against kernel::fft_filter_ccf (oracle), fir_filter_ccf and float64; decimation and set_taps too."""
    src = np.arange(8).astype(np.complex64)
    for g in (1.0, 2.0):
        blk = grb.filter.fft_filter_ccf(1, [g], 2)
        assert blk.nthreads() == 2 and blk.nsamples() == 2
        assert_tuples_almost_equal2((g * src).tolist(), sched_mock.run_sync(blk, src).tolist(), abs_eps=1e-5, rel_eps=1e-5)
    random.seed(0)
    for i in range(6):
        src = np.array(make_random_complex_tuple(4096), np.complex64)
        ntaps = int(random.uniform(2, 1000))
        taps = np.array([2 * random.random() - 1 for _ in range(ntaps)], np.float32)
        dec = 1 + i % 3
        blk = grb.filter.fft_filter_ccf(dec, taps)
        got = sched_mock.run_sync(blk, src)
        ref = orc.fft_filter_ccf(dec, taps, src)
        assert len(got) == len(ref) > 0
        assert_parity(got, ref, "ccf trial %d vs oracle" % i)
        assert_parity(got, orc.fir_stream(dec, taps, src)[: len(got)], "ccf trial %d vs fir" % i)
        assert_parity(got, po.f64_fft_filter(dec, taps, src, blk.nsamples()), "ccf trial %d f64" % i)
    blk = grb.filter.fft_filter_ccf(1, np.ones(5, np.float32))
    x = crand(np.random.default_rng(9), 12 * 40)
    a = sched_mock.run_sync(blk, x)
    t2 = np.linspace(-1, 1, 37).astype(np.float32)
    blk.set_taps(t2)  # deferred to the next work(): output multiple changes from 12 to 92
    assert blk.taps() == t2.tolist()
    b = sched_mock.run_sync(blk, x)
    assert blk.nsamples() == 128 - 37 + 1
    assert_parity(a, orc.fft_filter_ccf(1, np.ones(5, np.float32), x), "before set_taps")
    assert_parity(b, orc.fft_filter_ccf(1, t2, x)[: len(b)], "after set_taps")


@pytest.mark.parametrize("ntaps", [1, 2, 31, 1024, 4096, 6000])
def test_fft_filter_sizes(grb, orc, ntaps):
    rng = np.random.default_rng(ntaps)
    x, taps = crand(rng, 60000), crand(rng, ntaps) * np.float32(4.0 / max(ntaps, 4))
    blk = grb.filter.fft_filter_ccc(1, taps)
    assert blk.nsamples() == orc.fft_filter_nsamples(ntaps)
    got = sched_mock.run_sync(blk, x, max_noutput=1 << 20)
    ref = orc.fft_filter_ccc(1, taps, x)
    assert_parity(got, ref, "ntaps %d" % ntaps)
    assert_parity(got, po.f64_fft_filter(1, taps, x, blk.nsamples()), "f64 ntaps %d" % ntaps)


@pytest.mark.parametrize("ntaps,decim", [(31, 1), (1500, 1), (4096, 1), (4096, 3)])
def test_fft_filter_many_blocks_per_cta(grb, orc, ntaps, decim):
    """2^22 samples: more overlap-save blocks than persistent CTAs, so the streaming kernel's TMA prefetch
    chain (odd and even block bases, clipped last copy) is exercised end to end."""
    import scipy.signal

    rng = np.random.default_rng(77 + ntaps)
    x, taps = crand(rng, (1 << 22) + 1234), crand(rng, ntaps) * np.float32(4.0 / max(ntaps, 4))
    blk = grb.filter.fft_filter_ccc(decim, taps)
    ns = blk.nsamples()
    got = sched_mock.run_sync(blk, x, max_noutput=1 << 23)
    nitems = (len(x) // decim // ns) * ns
    assert len(got) == nitems
    ref64 = scipy.signal.fftconvolve(x.astype(np.complex128), taps.astype(np.complex128))[: nitems * decim: decim]
    assert_parity(got, ref64, "f64 ntaps %d decim %d" % (ntaps, decim))
    n_c = (200000 // decim // ns) * ns  # the C oracle on a prefix
    assert_parity(got[:n_c], orc.fft_filter_ccc(decim, taps, x[: n_c * decim]), "oracle prefix")


@pytest.mark.parametrize("ntaps,decim", [(4096, 1), (4095, 1), (3001, 2), (2050, 1), (1026, 5), (2049, 1), (4097, 1), (6000, 3),
                                         (8192, 1), (10240, 1)])
def test_fft_filter_partitioned_kernel(grb, orc, monkeypatch, ntaps, decim):
    """K = ceil(ntaps/2048) tap partitions on the 4096-point kernel, the K-1 previous spectra parked in tensor memory (the
    default plan for 2050..10240 taps; forced here for shorter filters too).  2^21 samples = many blocks per CTA run plus a ragged tail;
    odd and even tap counts give odd and even slice bases for the TMA ring."""
    import scipy.signal

    monkeypatch.setenv("GRB200_FFTFILT_UPOLS", "1")
    rng = np.random.default_rng(99 + ntaps)
    x, taps = crand(rng, (1 << 21) + 777), crand(rng, ntaps) * np.float32(4.0 / ntaps)
    blk = grb.filter.fft_filter_ccc(decim, taps)
    ns = blk.nsamples()
    assert ns == orc.fft_filter_nsamples(ntaps)
    got = sched_mock.run_sync(blk, x, max_noutput=1 << 23)
    nitems = (len(x) // decim // ns) * ns
    assert len(got) == nitems
    ref64 = scipy.signal.fftconvolve(x.astype(np.complex128), taps.astype(np.complex128))[: nitems * decim: decim]
    assert_parity(got, ref64, "f64 ntaps %d decim %d" % (ntaps, decim))
    n_c = max((100000 // decim // ns), 1) * ns  # the C oracle on a prefix
    assert_parity(got[:n_c], orc.fft_filter_ccc(decim, taps, x[: n_c * decim]), "oracle prefix")
    # scheduler-sized calls (one reference block each): the overlap state carries across work() calls
    small = sched_mock.run_sync(grb.filter.fft_filter_ccc(decim, taps), x[: 40 * ns * decim], max_noutput=ns)
    assert_parity(small, got[: len(small)], "one block per call")
    monkeypatch.setenv("GRB200_FFTFILT_UPOLS", "0")
    plain = sched_mock.run_sync(grb.filter.fft_filter_ccc(decim, taps), x, max_noutput=1 << 23)
    assert_parity(got, plain, "vs plain overlap-save", 2e-6, 2e-5)


def test_fft_filter_partitioned_device_resident_unaligned(grb):
    """work_dev with an input pointer that is only 8-byte aligned: the partitioned kernel starts one item into the
    aligned stream; same result as the aligned call."""
    import torch

    rng = np.random.default_rng(101)
    nt, n = 4096, 1 << 20
    taps = crand(rng, nt) * np.float32(4.0 / nt)
    x = crand(rng, n + nt - 1 + 1)
    xd = torch.from_numpy(x.view(np.float32)).cuda()
    y0 = torch.zeros(2 * n, dtype=torch.float32, device="cuda")
    y1 = torch.zeros(2 * n, dtype=torch.float32, device="cuda")
    blk = grb.filter.fft_filter_ccc(1, taps)
    st = torch.cuda.current_stream().cuda_stream
    blk.work_dev(xd.data_ptr() + 8, y1.data_ptr(), n, st)   # stream starts at item 1: 8-byte aligned
    torch.cuda.synchronize()
    xs = torch.from_numpy(x[1:].copy().view(np.float32)).cuda()  # the same stream, 16-byte aligned
    blk.work_dev(xs.data_ptr(), y0.data_ptr(), n, st)
    torch.cuda.synchronize()
    a, b = y0.cpu().numpy().view(np.complex64), y1.cpu().numpy().view(np.complex64)
    assert np.array_equal(a, b)
    import scipy.signal

    ref = scipy.signal.fftconvolve(x[1:].astype(np.complex128), taps.astype(np.complex128))[nt - 1: nt - 1 + n]
    assert_parity(a, ref, "f64")


@pytest.mark.parametrize("ntaps,decim,shift", [(31, 1, 0), (128, 1, 1), (1025, 1, 0), (2048, 1, 1), (2047, 1, 0), (600, 2, 0),
                                               (333, 3, 1), (100, 4, 1), (512, 8, 0), (77, 16, 0), (1, 1, 1), (2, 2, 0),
                                               (300, 32, 1), (1000, 64, 0), (40, 256, 0), (2048, 16, 1), (129, 128, 0),
                                               (100, 10, 0), (333, 20, 1), (64, 24, 0), (500, 6, 1), (50, 100, 0), (700, 40, 1),
                                               (90, 512, 0)])
def test_fft_filter_three_group_kernel(grb, ntaps, decim, shift):
    """fftfilt_ols3_kernel (plain overlap-save, up to 2048 taps): device resident, many blocks per group and a ragged last one,
    16- and 8-byte aligned input, decimations 2 / 4 / 8 (spectrum folded, L/D-point inverse), 16 .. 256 (folded by 8, every
    D/8-th result kept), other even decimations (folded by their largest factor 2 / 4 / 8, the rest dropped per result) and odd
    ones (full inverse, decimated on store); nothing is written past the last output."""
    import scipy.signal
    import torch

    rng = np.random.default_rng(5 * ntaps + decim)
    nout = ((1 << 21) + 777) // decim
    x = crand(rng, nout * decim + ntaps - 1 + shift)
    taps = crand(rng, ntaps) * np.float32(4.0 / max(ntaps, 4))
    xd = torch.from_numpy(x.view(np.float32)).cuda()
    yd = torch.full((2 * (nout + 2),), float("nan"), dtype=torch.float32, device="cuda")
    blk = grb.filter.fft_filter_ccc(decim, taps)
    blk.work_dev(xd.data_ptr() + 8 * shift, yd.data_ptr(), nout, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    y = yd.cpu().numpy()
    assert np.isnan(y[-4:]).all() and np.isfinite(y[:-4]).all()
    ref = scipy.signal.fftconvolve(x[shift:].astype(np.complex128), taps.astype(np.complex128))[ntaps - 1: ntaps - 1 + nout * decim: decim]
    assert_parity(y[:-4].view(np.complex64), ref, "f64 %d/%d" % (ntaps, decim))


@pytest.mark.parametrize("kind,ntaps,decim", [("ccf", 100, 1), ("ccf", 128, 2), ("ccf", 256, 4), ("ccc", 24, 1), ("ccc", 64, 1),
                                              ("ccc", 48, 2), ("ccc", 100, 4), ("ccc", 200, 8)])
@pytest.mark.parametrize("route", ["0", None])
def test_fir_direct_kernels_and_the_overlap_save_route_agree(grb, orc, monkeypatch, kind, ntaps, decim, route):
    """Tap counts the handle sends to the overlap-save kernels on its own (past 16 complex / 48..64 real taps per decimation
    step): the direct FIR kernels (GRB200_FIR_FFT=0, read when the handle is made) and the routed call both meet the oracle."""
    if route is None:
        monkeypatch.delenv("GRB200_FIR_FFT", raising=False)
    else:
        monkeypatch.setenv("GRB200_FIR_FFT", route)
    rng = np.random.default_rng(ntaps * 3 + decim)
    x = crand(rng, 150000 + 7)
    taps = crand(rng, ntaps) if kind == "ccc" else rng.uniform(-1, 1, ntaps).astype(np.float32)
    cls = grb.filter.fir_filter_ccc if kind == "ccc" else grb.filter.fir_filter_ccf
    got = sched_mock.run_sync(cls(decim, taps), x, max_noutput=1 << 20)
    assert_parity(got, orc.fir_stream(decim, taps, x), "%s %d/%d oracle" % (kind, ntaps, decim))
    assert_parity(got, po.f64_fir(decim, taps, x), "%s %d/%d f64" % (kind, ntaps, decim))


def test_fft_filter_state_across_calls_and_set_taps(grb, orc):
    rng = np.random.default_rng(8)
    x, taps = crand(rng, 20000), crand(rng, 100)
    a = sched_mock.run_sync(grb.filter.fft_filter_ccc(1, taps), x, max_noutput=157)       # one block per call
    b = sched_mock.run_sync(grb.filter.fft_filter_ccc(1, taps), x, max_noutput=1 << 20)   # one call
    np.testing.assert_allclose(a, b, atol=1e-5)
    # set_taps is deferred to the next work(), which returns 0, and it zeroes the overlap state (fft_filter.cc:177-179)
    blk = grb.filter.fft_filter_ccc(1, taps)
    ns = blk.nsamples()
    out = np.empty(ns * 10, np.complex64)
    assert blk.work(ns * 10, [x], [out]) == ns * 10
    t2 = crand(rng, 300)
    blk.set_taps(t2)
    assert blk.work(ns, [x], [out]) == 0
    ns2 = blk.nsamples()
    assert ns2 == orc.fft_filter_nsamples(300) == blk.output_multiple()
    seg = x[ns * 10: ns * 10 + ns2 * 5]
    out2 = np.empty(ns2 * 5, np.complex64)
    assert blk.work(ns2 * 5, [seg], [out2]) == ns2 * 5
    assert_parity(out2, orc.fft_filter_ccc(1, t2, seg), "fresh state after set_taps")
    with pytest.raises(ValueError):
        blk.work(ns2 + 1, [x], [np.empty(ns2 + 1, np.complex64)])
