"""GPU parity (through the C-ABI): pfb_decimator_ccf against the oracle (all four work() variants of the reference),
the committed reference fixtures, the reference QA tones (gr-filter/python/filter/qa_pfb_decimator.py) and float64.
Tolerance: north_star (NRMS <= 1e-5, max-abs <= 1e-4 x RMS); tighter where stated."""
import numpy as np
import pytest

from common import assert_parity, crand, fixtures
from oracle import pyoracle as po
from test_oracle import pfb_decimator_qa_case, pfb_decimator_qa_expected

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def grb():
    import gnuradio_b200 as g

    assert g.device_count() > 0, "no CUDA device: libgrb200 has no CPU fallback"
    return g


def run_block(blk, streams, max_noutput=4096):
    """M vector_sources -> pfb_decimator_ccf -> vector_sink, the scheduler's way: history items in front, noutput a
    multiple of the output multiple, work() may return 0 once after set_taps."""
    streams = np.ascontiguousarray(streams, np.complex64)
    M, n = streams.shape
    hist = blk.history()
    buf = np.concatenate([np.zeros((M, hist - 1), np.complex64), streams], axis=1)
    om = blk.output_multiple()
    out, consumed = [], 0
    while True:
        avail = buf.shape[1] - consumed
        nout = min(avail - (hist - 1), max_noutput) // om * om
        if nout <= 0:
            break
        ob = np.empty(nout, np.complex64)
        r = blk.work(nout, [np.ascontiguousarray(buf[s, consumed:]) for s in range(M)], [ob])
        if r == 0:
            continue
        out.append(ob[:r])
        consumed += r
    return np.concatenate(out) if out else np.empty(0, np.complex64)


MODES = [(False, False), (True, False), (False, True), (True, True)]  # (use_fft_rotator, use_fft_filters)


def test_pfb_decimator_reference_fixtures(grb):
    fx = fixtures()
    taps, x = fx["lp2_qa_taps"], fx["pfbdec_x"]
    for ch in (0, 3):
        for rot, filt in MODES:
            mode = int(rot) | (int(filt) << 1)
            blk = grb.filter.pfb_decimator_ccf(5, taps, ch, rot, filt)
            got = run_block(blk, x, max_noutput=500)
            ref = fx["pfbdec_c%d_m%d" % (ch, mode)]
            assert len(got) == len(ref)
            assert_parity(got, ref, "fixture ch %d mode %d" % (ch, mode), 2e-6, 2e-5)
    got = run_block(grb.filter.pfb_decimator_ccf(8, fx["pfbdec_b_taps"], 5, False, False), fx["pfbdec_b_x"])
    assert_parity(got, fx["pfbdec_b_c5"], "ragged taps", 2e-6, 2e-5)


@pytest.mark.parametrize("channel", [0, 1, 2, 3, 4])
def test_pfb_decimator_qa_tones(grb, channel):
    # qa_pfb_decimator.py:78-141: last 50 outputs against the channel's tone, 4 places, all four variants
    taps = fixtures()["lp2_qa_taps"]
    streams, f, phase, fs = pfb_decimator_qa_case(channel)
    for rot, filt in MODES:
        got = run_block(grb.filter.pfb_decimator_ccf(5, taps, channel, rot, filt), streams)
        exp = pfb_decimator_qa_expected(len(taps), len(got), f, phase, fs)
        np.testing.assert_allclose(got[-50:], exp[-50:], atol=0.5e-4, rtol=0)


@pytest.mark.parametrize("M,ntaps,chan", [(1, 7, 0), (2, 9, 1), (5, 93, 2), (16, 256, 11), (64, 1024, 37), (64, 1000, 0),
                                          (7, 300, 6), (128, 4096, 100)])
def test_pfb_decimator_sweep(grb, orc, M, ntaps, chan):
    rng = np.random.default_rng(M * 1000 + ntaps)
    n = 3000
    x = crand(rng, M * n).reshape(M, n)
    taps = rng.uniform(-1, 1, ntaps).astype(np.float32) / np.float32(np.sqrt(ntaps))
    blk = grb.filter.pfb_decimator_ccf(M, taps, chan, False, False)
    assert blk.history() == -(-ntaps // M) and len(blk.taps()) == M and len(blk.taps()[0]) == blk.history()
    got = run_block(blk, x, max_noutput=1024)
    assert_parity(got, orc.pfb_decimator(M, taps, chan, x), "oracle %d/%d" % (M, ntaps), 2e-6, 2e-5)
    assert_parity(got, po.f64_pfb_decimator(M, taps, chan, x), "f64 %d/%d" % (M, ntaps), 2e-6, 2e-5)


def test_pfb_decimator_carried_overlap_equals_history_mode(grb):
    """use_fft_filters declares history 1 and an output multiple (pfb_decimator_ccf_impl.cc:57-59): the overlap then lives
    in the handle.  Scheduler-sized calls in that mode equal one call in history mode."""
    rng = np.random.default_rng(3)
    M, ntaps, n = 8, 200, 40 * 39
    x = crand(rng, M * n).reshape(M, n)
    taps = rng.uniform(-1, 1, ntaps).astype(np.float32)
    a = grb.filter.pfb_decimator_ccf(M, taps, 3, True, True)
    assert a.history() == 1 and a.output_multiple() == 64 - 25 + 1  # fft_filter geometry for 25 taps per branch
    got = run_block(a, x, max_noutput=3 * 40)
    ref = run_block(grb.filter.pfb_decimator_ccf(M, taps, 3, False, False), x, max_noutput=1 << 20)
    assert len(got) == n
    assert np.array_equal(got, ref)


def test_pfb_decimator_set_channel_set_taps_and_errors(grb, orc):
    rng = np.random.default_rng(4)
    M, n = 6, 2000
    x = crand(rng, M * n).reshape(M, n)
    t1, t2 = rng.uniform(-1, 1, 60).astype(np.float32), rng.uniform(-1, 1, 100).astype(np.float32)
    blk = grb.filter.pfb_decimator_ccf(M, t1, 1, False, False)
    blk.set_channel(4)  # pfb_decimator_ccf_impl.cc:106-110: immediate
    assert_parity(run_block(blk, x), orc.pfb_decimator(M, t1, 4, x), "after set_channel", 2e-6, 2e-5)
    blk.set_taps(t2)  # :88-95: history changes, the next work() returns 0 once
    assert blk.history() == 17
    ins = [np.zeros(100, np.complex64)] * M
    assert blk.work(10, ins, [np.empty(10, np.complex64)]) == 0
    assert_parity(run_block(blk, x), orc.pfb_decimator(M, t2, 4, x), "after set_taps", 2e-6, 2e-5)
    with pytest.raises(ValueError):
        grb.filter.pfb_decimator_ccf(0, t1, 0)
    with pytest.raises(ValueError):
        grb.filter.pfb_decimator_ccf(4, [], 0)
    with pytest.raises(ValueError):
        grb.filter.pfb_decimator_ccf(4, t1, 4)  # channel out of range
    with pytest.raises(ValueError):
        blk.work(10, ins[:3], [np.empty(10, np.complex64)])  # wrong number of input streams
    with pytest.raises(ValueError):
        blk.work(200, ins, [np.empty(200, np.complex64)])  # too few input items
    assert blk.work(0, ins, [np.empty(0, np.complex64)]) == 0


def test_pfb_decimator_large_device_resident(grb):
    """64 channels x 16 taps per branch, 2^16 outputs = 2^22 input items, device-resident, against float64."""
    import torch

    rng = np.random.default_rng(5)
    M, T, n = 64, 16, 1 << 16
    taps = rng.uniform(-1, 1, M * T).astype(np.float32) / np.float32(32)
    x = crand(rng, M * n).reshape(M, n)
    blk = grb.filter.pfb_decimator_ccf(M, taps, 21, False, False)
    ref = po.f64_pfb_decimator(M, taps, 21, x)
    # even stride -> the two-outputs-per-thread kernel (16-byte loads); odd stride -> the one-output kernel
    for stride, nn in ((n + T - 1 + 5, n), (n + T - 1 + 4, n), (n + T - 1 + 5, n - 1)):
        buf = np.zeros((M, stride), np.complex64)
        buf[:, T - 1: T - 1 + n] = x
        xd = torch.from_numpy(buf.view(np.float32)).cuda()
        yd = torch.zeros(2 * n, dtype=torch.float32, device="cuda")
        blk.work_dev(xd.data_ptr(), stride, yd.data_ptr(), nn, torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        got = yd.cpu().numpy().view(np.complex64)
        assert_parity(got[:nn], ref[:nn], "2^22 device-resident, stride %d" % stride, 2e-6, 2e-5)
        assert np.all(got[nn:] == 0)


@pytest.mark.parametrize("M,T,n", [(4, 8, 50001), (16, 32, 30000), (5, 64, 12345), (1, 16, 20000), (3, 16, 9216), (64, 16, 9217)])
def test_pfb_decimator_tiled_kernel(grb, orc, M, T, n):
    """Sizes that take the tiled kernel (TMA ring + register window, even T, >= 4 tiles): ragged last tile, a single
    stream, odd stream counts (the two-slot ring ends on either slot), and the exact 4-tile threshold."""
    rng = np.random.default_rng(M * 100 + T)
    taps = rng.uniform(-1, 1, M * T).astype(np.float32) / np.float32(np.sqrt(M * T))
    x = crand(rng, M * n).reshape(M, n)
    chan = M // 2
    got = run_block(grb.filter.pfb_decimator_ccf(M, taps, chan, False, False), x, max_noutput=1 << 20)
    assert_parity(got, po.f64_pfb_decimator(M, taps, chan, x), "f64 %dx%d" % (M, T), 2e-6, 2e-5)
    m = min(n, 20000)
    assert_parity(got[:m], orc.pfb_decimator(M, taps, chan, x[:, :m]), "oracle %dx%d" % (M, T), 2e-6, 2e-5)
