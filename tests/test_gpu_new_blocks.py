"""GPU parity (through the C-ABI) for the section 8(f) blocks added in round 2: fir_filter_fff, fft_vfc, logpwrfft_c/f,
pfb_synthesizer_ccf, filterbank_vcvcf, pfb.channelizer_hier_ccf, stream_to_vector / vector_to_streams.
Against the C oracle (the reference's own sources when oracle/_ref is present), the committed reference fixtures, the reference
QA cases and float64.  Tolerance: north_star (NRMS <= 1e-5, max-abs <= 1e-4 x RMS)."""
import math
import random

import numpy as np
import pytest

import sched_mock
from common import assert_parity, blackman_harris, crand, fixtures, hamming, lowpass_taps, sig_source_c
from oracle import pyoracle as po

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def grb():
    import gnuradio_b200 as g

    assert g.device_count() > 0, "no CUDA device: libgrb200 has no CPU fallback"
    return g


# ---------------------------------------------------------------------------------------------------------------------
# fir_filter_fff
# ---------------------------------------------------------------------------------------------------------------------
def test_fir_fff_fixtures_and_qa(grb):
    fx = fixtures()
    for D in (1, 3):
        got = sched_mock.run_sync(grb.filter.fir_filter_fff(D, fx["fff_taps"]), fx["fff_x"], max_noutput=700, out_dtype=np.float32)
        assert_parity(got, fx["fff_d%d" % D], "fff fixture d%d" % D)
    src = np.array(40 * [1, 2, 3, 4], np.float32)  # qa_fir_filter.py:32-58
    taps = np.array(20 * [0.5, 0.5], np.float32)
    for D in (1, 4):
        got = sched_mock.run_sync(grb.filter.fir_filter_fff(D, taps), src, out_dtype=np.float32)
        np.testing.assert_allclose(got, po.f64_fir(D, taps, src).real, atol=0.5e-5 * 80)


@pytest.mark.parametrize("ntaps", [1, 2, 15, 16, 17, 64, 255, 256, 257, 1000, 8192, 9000])
@pytest.mark.parametrize("decim", [1, 2, 7])
def test_fir_fff_sweep(grb, orc, ntaps, decim):
    rng = np.random.default_rng(ntaps * 10 + decim)
    nout = 20011 if ntaps < 5000 else 6007
    x = rng.uniform(-1, 1, nout * decim + ntaps - 1).astype(np.float32)
    taps = rng.uniform(-1, 1, ntaps).astype(np.float32)
    got = grb.filter.fir_filter_fff(decim, taps).process(x, nout)
    assert len(got) == nout and np.isfinite(got).all()
    assert_parity(got, orc.fir_fff(decim, taps, x, nout), "fff oracle %d/%d" % (ntaps, decim))
    f64 = np.convolve(x.astype(np.float64), taps.astype(np.float64))[ntaps - 1: ntaps - 1 + nout * decim: decim]
    assert_parity(got, f64, "fff f64 %d/%d" % (ntaps, decim))


@pytest.mark.parametrize("ntaps,decim,nout,shift", [(97, 1, 200001, 0), (256, 1, 131073, 0), (1000, 3, 70001, 0), (4097, 1, 66000, 0),
                                                    (8192, 2, 65537, 0), (300, 2, 65536, 0), (17, 1, 1 << 21, 0), (24, 4, 300001, 0),
                                                    (2049, 1, 400003, 0), (9, 2, 500002, 0), (64, 8, 100000, 0), (130, 1, 300000, 1),
                                                    (130, 5, 99999, 2), (40, 1, 65536, 3), (200, 16, 70000, 0), (33, 32, 65536, 0)])
def test_fir_fff_long_filters_on_the_overlap_save_kernel(grb, orc, ntaps, decim, nout, shift):
    """More than 16 taps (8 at decimation 2) and >= 65536 outputs: the two halves of the call go through the overlap-save kernels
    as re / im of one complex stream (odd output counts: the second half is one shorter) -- fftfilt_ols3_kernel<RP> up to 2049
    taps when the float stream is 16-byte aligned (two bulk copies per block, each up to three floats early), else
    fftfilt_ols_kernel<L, RP>."""
    import torch

    rng = np.random.default_rng(ntaps + decim)
    x = rng.uniform(-1, 1, nout * decim + ntaps - 1 + shift).astype(np.float32)
    taps = (rng.uniform(-1, 1, ntaps) * (4.0 / ntaps)).astype(np.float32)
    xd = torch.from_numpy(x).cuda()
    x = x[shift:]
    yd = torch.full((nout + 2,), float("nan"), dtype=torch.float32, device="cuda")
    blk = grb.filter.fir_filter_fff(decim, taps)
    blk.work_dev(xd.data_ptr() + 4 * shift, yd.data_ptr(), nout, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    y = yd.cpu().numpy()
    assert np.isnan(y[-2:]).all() and np.isfinite(y[:-2]).all()
    f64 = np.convolve(x.astype(np.float64), taps.astype(np.float64))[ntaps - 1: ntaps - 1 + nout * decim: decim]
    assert_parity(y[:-2], f64, "fff f64 %d/%d" % (ntaps, decim))
    assert_parity(y[:-2], orc.fir_fff(decim, taps, x, nout), "fff oracle %d/%d" % (ntaps, decim))


def test_fir_fff_tiles_ragged_ends_and_unaligned(grb, orc):
    """Several tiles per CTA, an odd output count (half a pair at the end), 4-byte aligned input and output pointers."""
    import torch

    rng = np.random.default_rng(5)
    nt = 256
    taps = rng.uniform(-1, 1, nt).astype(np.float32)
    blk = grb.filter.fir_filter_fff(1, taps)
    for nout in (4096 * 148 * 3 + 1, 4097, 1, 2, 4095):
        x = rng.uniform(-1, 1, nout + nt - 1).astype(np.float32)
        got = blk.process(x, nout)
        k = min(nout, 30000)
        if nout < 64:  # a handful of outputs: no RMS to speak of, compare against the size of the terms instead
            np.testing.assert_allclose(got, orc.fir_fff(1, taps, x, nout), atol=1e-5 * np.sqrt(nt))
            continue
        assert_parity(got[:k], orc.fir_fff(1, taps, x[: k + nt - 1], k), "head %d" % nout)
        assert_parity(got[-k:], orc.fir_fff(1, taps, x[nout - k:], k), "tail %d" % nout)
    n = 100000
    x = rng.uniform(-1, 1, n + nt + 3).astype(np.float32)
    d_x = torch.from_numpy(x).cuda()
    d_y = torch.zeros(n + 8, dtype=torch.float32, device="cuda")
    for off_in, off_out in ((1, 0), (0, 1), (3, 3)):
        d_y.zero_()
        blk.work_dev(d_x.data_ptr() + 4 * off_in, d_y.data_ptr() + 4 * off_out, n)
        torch.cuda.synchronize()
        got = d_y.cpu().numpy()
        assert_parity(got[off_out: off_out + n], orc.fir_fff(1, taps, x[off_in:], n), "unaligned %d/%d" % (off_in, off_out))
        assert not got[:off_out].any() and not got[off_out + n:].any()


def test_fir_fff_set_taps_and_errors(grb):
    blk = grb.filter.fir_filter_fff(2, [1.0, 2.0, 3.0])
    assert blk.history() == 3 and blk.taps() == [1.0, 2.0, 3.0]
    blk.set_taps([0.5] * 7)
    x = np.arange(100, dtype=np.float32)
    out = np.empty(10, np.float32)
    assert blk.work(10, [x], [out]) == 0 and blk.history() == 7  # fir_filter_blk_impl.cc:77-81
    got = sched_mock.run_sync(blk, x, out_dtype=np.float32)
    np.testing.assert_allclose(got, po.f64_fir(2, [0.5] * 7, x).real, rtol=1e-6)
    with pytest.raises(ValueError):
        grb.filter.fir_filter_fff(0, [1.0])
    with pytest.raises(TypeError):
        blk.work(1, [np.zeros(100, np.complex64)], [out])


# ---------------------------------------------------------------------------------------------------------------------
# fft_vfc
# ---------------------------------------------------------------------------------------------------------------------
def test_fft_vfc_fixtures(grb):
    fx = fixtures()
    assert_parity(grb.fft.fft_vfc(64, True, fx["win_bh_64"]).process(fx["vfc_x"]), fx["vfc64_f1_w"], "fwd window")
    assert_parity(grb.fft.fft_vfc(64, False, []).process(fx["vfc_x"]), fx["vfc64_f0"], "reverse")
    with pytest.raises(RuntimeError):
        grb.fft.fft_vfc(64, True, np.ones(63, np.float32))
    f = grb.fft.fft_vfc(64, True, [])
    assert f.set_window(np.ones(64, np.float32)) and not f.set_window(np.ones(5, np.float32))


@pytest.mark.parametrize("N,nvec", [(2, 5), (16, 100), (64, 1000), (512, 70), (1024, 300), (4096, 200), (8192, 70), (16384, 9),
                                    (100, 11), (1000, 3)])
@pytest.mark.parametrize("forward", [True, False])
def test_fft_vfc_sizes(grb, orc, N, nvec, forward):
    rng = np.random.default_rng(N + nvec)
    x = rng.uniform(-1, 1, N * nvec).astype(np.float32)
    w = rng.uniform(0.1, 1, N).astype(np.float32)
    got = grb.fft.fft_vfc(N, forward, w).process(x)
    assert_parity(got, po.f64_fft_vfc(N, forward, w, x), "vfc f64 %d" % N)
    k = min(nvec, 20) * N
    assert_parity(got[:k], orc.fft_vfc(N, forward, w, x[:k]), "vfc oracle %d" % N)
    p = grb.fft.fft_vfc(N, forward, w).process(x, magsq=True)
    assert_parity(p, np.abs(po.f64_fft_vfc(N, forward, w, x)) ** 2, "vfc magsq %d" % N)


# ---------------------------------------------------------------------------------------------------------------------
# logpwrfft
# ---------------------------------------------------------------------------------------------------------------------
def _ref_win(orc, N):
    return orc.window(5, N)  # WIN_BLACKMAN_hARRIS, what window.blackmanharris(N) returns (logpwrfft.py:54)


def test_logpwrfft_fixtures(grb):
    fx = fixtures()
    w = fx["win_bh_64"]
    win = lambda n: w
    blk = grb.fft.logpwrfft_c(64 * 3 * 10.0, 64, 2.0, 10.0, 0.25, True, win)
    assert blk.decimation() == 3 and abs(blk.frame_rate() - 10.0) < 1e-12
    assert_parity(blk.process(fx["logpwr_xc"]), fx["logpwr_c_d3_avg"], "c d3 avg", 1e-5, 2e-4)
    blk = grb.fft.logpwrfft_c(64.0, 64, 2.0, 1.0, 1.0, False, win)
    assert_parity(blk.process(fx["logpwr_xc"]), fx["logpwr_c_d1"], "c d1", 1e-5, 2e-4)
    blk = grb.fft.logpwrfft_f(128.0, 64, 1.0, 1.0, 0.1, True, win)
    assert blk.decimation() == 2
    assert_parity(blk.process(fx["logpwr_xf"]), fx["logpwr_f_d2_avg"], "f d2 avg", 1e-5, 2e-4)


@pytest.mark.parametrize("cin", [True, False])
@pytest.mark.parametrize("N,dec,avg,alpha,nvec", [(64, 1, False, 1.0, 300), (1024, 1, True, 0.2, 500), (4096, 4, True, 0.05, 300),
                                                  (4096, 1, True, 0.01, 600), (512, 7, True, 0.5, 100), (4096, 1, False, 0.3, 200),
                                                  (100, 2, True, 0.3, 50)])
def test_logpwrfft_against_oracle(grb, orc, cin, N, dec, avg, alpha, nvec):
    rng = np.random.default_rng(N + dec + nvec)
    w = _ref_win(orc, N)
    x = crand(rng, N * nvec) if cin else rng.uniform(-1, 1, N * nvec).astype(np.float32)
    tone = np.exp(2j * np.pi * 0.123 * np.arange(N * nvec)) * 5
    x = (x + tone).astype(np.complex64) if cin else (x + 5 * tone.real).astype(np.float32)
    cls = grb.fft.logpwrfft_c if cin else grb.fft.logpwrfft_f
    blk = cls(float(N * dec), N, 2.0, 1.0, alpha, avg, lambda n: w)
    assert blk.decimation() == dec
    got = blk.process(x)
    ref, cnt, st = orc.logpwrfft(cin, N, 2.0, dec, alpha, avg, w, x)
    assert got.shape == ref.shape
    # the block's output is a power in dB: north_star's tolerance applies to the power; on the dB scale every bin within
    # 60 dB of the strongest must agree to 2e-3 dB (a bin in a deep null carries the transform's rounding floor, fp32 like the
    # reference's -- there the two differ by what two fp32 FFTs differ)
    f64 = po.f64_logpwrfft(cin, N, 2.0, dec, alpha, avg, w, x)
    pw = lambda d: 10.0 ** (np.asarray(d, np.float64) / 10.0)
    assert_parity(pw(got), pw(f64), "power vs f64", 1e-5, 2e-4)
    assert_parity(pw(got), pw(ref), "power vs oracle", 1e-5, 2e-4)
    strong = f64 > f64.max() - 60.0
    assert np.abs(got - ref)[strong].max() < 2e-3, np.abs(got - ref)[strong].max()
    assert np.abs(got - f64)[strong].max() < 2e-3
    # the stream handed over in scheduler-sized pieces: counter and filter memory carry on
    blk2 = cls(float(N * dec), N, 2.0, 1.0, alpha, avg, lambda n: w)
    parts, pos = [], 0
    for k in (1, 3, 2, 17, 5, 1000000):
        seg = x[pos * N: (pos + k) * N]
        parts.append(blk2.process(seg))
        pos += len(seg) // N
        if pos >= nvec:
            break
    got2 = np.concatenate([p for p in parts if len(p)])
    assert got2.shape == got.shape
    assert_parity(pw(got2), pw(f64), "pieces: power vs f64", 1e-5, 2e-4)
    assert np.abs(got2 - got)[strong].max() < 2e-3


@pytest.mark.parametrize("cin,N,dec,nvec,first", [(True, 4096, 1, 1500, 1450), (True, 4096, 1, 1500, 37), (False, 4096, 2, 1300, 1250),
                                                  (True, 2048, 1, 2600, 2550), (True, 8192, 1, 700, 650), (False, 2048, 3, 2500, 60),
                                                  (False, 8192, 1, 700, 652)])
def test_logpwrfft_long_averaging_calls_filter_in_the_fft_kernel(grb, orc, cin, N, dec, nvec, first):
    """Calls of >= 4 frames per resident CTA at 2048 / 4096 / 8192 bins: fft_vcc_stream_kernel mode 3 filters each CTA's run of
    consecutive frames in its epilogue, logpwr_carry_kernel / logpwr_fixup_kernel finish.  Device resident, two calls in a row
    -- a long one and a short one that takes the three-kernel form, in either order; the second continues the first's filter
    memory and decimator phase --, against float64 and the oracle."""
    import torch

    rng = np.random.default_rng(N + dec + nvec)
    w = _ref_win(orc, N)
    n = N * nvec * dec
    x = crand(rng, n) if cin else rng.uniform(-1, 1, n).astype(np.float32)
    tone = np.exp(2j * np.pi * 0.0731 * np.arange(n)) * (3 + 2 * np.sin(np.arange(n) * 1e-5))
    x = (x + tone).astype(np.complex64) if cin else (x + tone.real).astype(np.float32)
    alpha = 0.07
    cls = grb.fft.logpwrfft_c if cin else grb.fft.logpwrfft_f
    blk = cls(float(N * dec), N, 2.0, 1.0, alpha, True, lambda k: w)
    xd = torch.from_numpy(x.view(np.float32) if cin else x).cuda()
    yd = torch.full((nvec + 1, N), float("nan"), dtype=torch.float32, device="cuda")
    isz = 8 if cin else 4
    cut = first * dec * N  # items in the first call
    u1, m1 = blk.work_dev(xd.data_ptr(), cut, yd.data_ptr(), nvec)
    u2, m2 = blk.work_dev(xd.data_ptr() + u1 * isz, n - u1, yd.data_ptr() + m1 * N * 4, nvec - m1)
    torch.cuda.synchronize()
    assert m1 + m2 == nvec and u1 + u2 == n
    got = yd.cpu().numpy()
    assert np.isnan(got[-1]).all() and np.isfinite(got[:-1]).all()
    got = got[:-1]
    f64 = po.f64_logpwrfft(cin, N, 2.0, dec, alpha, True, w, x)
    assert got.shape == f64.shape
    pw = lambda d: 10.0 ** (np.asarray(d, np.float64) / 10.0)
    assert_parity(pw(got), pw(f64), "power vs f64", 1e-5, 2e-4)
    strong = f64 > f64.max() - 60.0
    assert np.abs(got - f64)[strong].max() < 2e-3
    k = 200  # the C oracle on the first frames
    ref, cnt, st = orc.logpwrfft(cin, N, 2.0, dec, alpha, True, w, x[: k * dec * N])
    assert_parity(pw(got[:k]), pw(ref), "power vs oracle", 1e-5, 2e-4)


def test_logpwrfft_setters_and_limits(grb, orc):
    N = 256
    w = _ref_win(orc, N)
    blk = grb.fft.logpwrfft_c(1e6, N, 2.0, 30.0, 0.1, True, lambda n: w)
    assert blk.decimation() == max(1, int(round(1e6 / N / 30.0))) and blk.average() and blk.avg_alpha() == 0.1
    blk.set_decimation(5)
    assert blk.decimation() == 5
    blk.set_vec_rate(100.0)
    assert blk.decimation() == round(1e6 / N / 100.0)
    blk.set_sample_rate(2e6)
    assert blk.sample_rate() == 2e6 and blk.decimation() == round(2e6 / N / 100.0)
    blk.set_average(False)
    assert not blk.average()
    with pytest.raises(IndexError):
        blk.set_avg_alpha(1.5)  # single_pole_iir::set_taps throws std::out_of_range
    # max_out smaller than what the input holds: the block stops after max_out frames and reports what it consumed
    blk = grb.fft.logpwrfft_c(float(N * 2), N, 2.0, 1.0, 1.0, False, lambda n: w)
    x = crand(np.random.default_rng(0), N * 20)
    out = np.empty(3 * N, np.float32)
    n = blk.general_work(3, [len(x)], [x], [out])
    assert n == 3 and blk._consumed == 6 * N  # vectors 1, 3, 5 kept; the loop stops right after the third (keep_one_in_n_impl.cc:66)


# ---------------------------------------------------------------------------------------------------------------------
# pfb_synthesizer_ccf
# ---------------------------------------------------------------------------------------------------------------------
def test_pfb_synthesizer_fixtures_and_qa(grb):
    fx = fixtures()
    got = grb.filter.pfb_synthesizer_ccf(5, fx["synth_taps"]).process(fx["synth_in5"])
    assert_parity(got, fx["synth_5_1x"], "1x fixture")
    blk = grb.filter.pfb_synthesizer_ccf(5, fx["synth_taps"])
    blk.set_channel_map(fx["synth_map"])
    assert blk.channel_map() == list(fx["synth_map"])
    assert_parity(blk.process(fx["synth_in5"][:3]), fx["synth_5_1x_map3"], "3 inputs, map")
    got = grb.filter.pfb_synthesizer_ccf(8, fx["synth_taps8"], True).process(fx["synth_in8"])
    assert_parity(got, fx["synth_8_2x"], "2x fixture")
    # qa_pfb_synthesizer.py:31-84
    N, M, fs = 10000, 5, 1000
    ofs = M * fs
    ins = np.array([sig_source_c(fs, f, 1, N) for f in [0, 100, 200, -200, -100]], np.complex64)
    got = grb.filter.pfb_synthesizer_ccf(M, fx["synth_taps"]).process(ins)
    p = [2 * math.pi * k / M for k in range(M)]
    t = np.arange(len(got)) / ofs
    exp = sum(np.exp(1j * (2 * np.pi * f * t + q)) for f, q in zip([-2200, -1100, 0, 1100, 2200], [p[3], p[4], p[0], p[1], p[2]]))
    np.testing.assert_allclose(got[2000:3000], exp[2000:3000], atol=0.5e-4 * 10)


@pytest.mark.parametrize("M,T,twox,nin,n", [(64, 16, False, 64, 5000), (64, 16, False, 40, 3001), (16, 8, False, 16, 20000),
                                            (256, 16, False, 256, 700), (64, 32, False, 64, 1000), (64, 5, False, 64, 1000),
                                            (64, 1, False, 64, 300), (64, 16, True, 64, 2001), (64, 16, True, 128, 500),
                                            (5, 9, False, 5, 1000), (6, 4, True, 9, 999), (100, 3, False, 7, 200),
                                            (64, 40, False, 64, 400),
                                            # channel counts M = R1 x R2 with R1 != R2, and M = 8, on the fast path
                                            (8, 12, False, 8, 9000), (32, 16, False, 32, 4001), (128, 10, False, 128, 1500),
                                            (32, 7, False, 20, 3000), (128, 32, False, 77, 900)])
def test_pfb_synthesizer_sweep(grb, orc, M, T, twox, nin, n):
    """tensor-map TMA fast path (1x; M = 16 / 64 / 256; up to 32 taps per branch; fewer inputs than channels; random channel
    map) and the general path (2x, other channel counts, more than 32 taps per branch)."""
    rng = np.random.default_rng(M * 1000 + T + n)
    taps = rng.uniform(-1, 1, M * T - (M // 3)).astype(np.float32)
    ins = crand(rng, nin * n).reshape(nin, n)
    NF = (2 if twox else 1) * M
    cmap = rng.permutation(NF).astype(np.int32)
    blk = grb.filter.pfb_synthesizer_ccf(M, taps, twox)
    assert blk.taps_per_filter() == T and blk.output_multiple() == M
    np.testing.assert_array_equal(np.array(blk.taps(), np.float32), _synth_taps(M, taps, twox))
    blk.set_channel_map(cmap)
    got = blk.process(ins)
    assert len(got) == n * M and np.isfinite(got.view(np.float32)).all()
    assert_parity(got, orc.pfb_synthesizer(M, taps, twox, ins, channel_map=cmap), "synth oracle")
    assert_parity(got, po.f64_pfb_synthesizer(M, taps, twox, ins, channel_map=cmap), "synth f64")


def _synth_taps(M, taps, twox):
    T = -(-len(taps) // M)
    h = np.zeros(M * T, np.float32)
    h[: len(taps)] = taps
    d = h.reshape(T, M).T  # d[i][j] = h[i + j*M]
    if not twox:
        return d
    out = np.zeros((2 * M, T), np.float32)
    out[:M, 1::2] = d[:, 1::2]
    out[M:, 0::2] = d[:, 0::2]
    return out


@pytest.mark.parametrize("M,T,twox", [(64, 16, False), (8, 6, True)])
def test_pfb_synthesizer_scheduler_chunks_carry_the_delay_lines(grb, orc, M, T, twox):
    """work() sees only NEW items (history 1): the delay lines of the reference's filters are state in the handle.  Any
    chunking -- including chunks shorter than the history, and odd chunk lengths in 2x mode -- gives the one-call result."""
    rng = np.random.default_rng(M + T)
    taps = rng.uniform(-1, 1, M * T).astype(np.float32)
    n = 3000
    ins = crand(rng, M * n).reshape(M, n)
    whole = grb.filter.pfb_synthesizer_ccf(M, taps, twox).process(ins)
    blk = grb.filter.pfb_synthesizer_ccf(M, taps, twox)
    parts, pos = [], 0
    for k in [1, 2, 3, 7, T - 2, 100, 1, 511, 13, 10 ** 6]:
        seg = [np.ascontiguousarray(ins[i, pos: pos + k]) for i in range(M)]
        parts.append(blk.process(seg))
        pos += len(seg[0])
        if pos >= n:
            break
    got = np.concatenate(parts)
    assert_parity(got, whole, "chunked vs one call", 1e-6, 1e-5)
    assert_parity(got, orc.pfb_synthesizer(M, taps, twox, ins), "chunked vs oracle")
    # set_taps re-creates the filters: zero delay lines again (fir_filter_with_buffer.cc:79-86)
    blk.set_taps(taps)
    assert_parity(blk.process(ins), whole, "after set_taps", 1e-6, 1e-5)


def test_pfb_synthesizer_large_device_resident(grb, orc):
    """A call the size of the bench (2^22 times x 64 channels), device-resident, spot windows against the oracle."""
    import torch

    rng = np.random.default_rng(3)
    M, T, n = 64, 16, 1 << 20
    taps = (lowpass_taps(M * T, 0.5 / M, blackman_harris(M * T)) * np.float32(M)).astype(np.float32)
    stride = n + T - 1 + 1
    g = torch.Generator(device="cuda")
    g.manual_seed(9)
    x = torch.empty(2 * M * stride, dtype=torch.float32, device="cuda")
    x.uniform_(-1, 1, generator=g)
    y = torch.empty(2 * M * n, dtype=torch.float32, device="cuda")
    blk = grb.filter.pfb_synthesizer_ccf(M, taps)
    blk.work_dev(x.data_ptr(), stride, M, y.data_ptr(), n * M)
    torch.cuda.synchronize()
    xv = x.view(M, 2 * stride)
    for t0 in (0, 12345, n // 2 - 7, n - 300):
        w = 300
        seg = xv[:, 2 * t0: 2 * (t0 + w + T - 1)].cpu().numpy().view(np.complex64)  # T-1 history items in front
        ref = orc.pfb_synthesizer(M, taps, False, seg)[(T - 1) * M:]
        got = y[2 * t0 * M: 2 * (t0 + w) * M].cpu().numpy().view(np.complex64)
        assert_parity(got, ref, "window at %d" % t0)


def test_pfb_synthesizer_errors(grb):
    with pytest.raises(ValueError):
        grb.filter.pfb_synthesizer_ccf(5, np.ones(20, np.float32), True)  # odd channel count with 2x
    blk = grb.filter.pfb_synthesizer_ccf(4, np.ones(8, np.float32))
    with pytest.raises(ValueError):
        blk.set_channel_map([0, 1, 2, 4])
    out = np.empty(8, np.complex64)
    ins = [np.zeros(2, np.complex64)] * 4
    assert blk.work(8, ins, [out]) == 0  # d_updated after construction (:229-232)
    assert blk.work(8, ins, [out]) == 8


# ---------------------------------------------------------------------------------------------------------------------
# filterbank_vcvcf, pfb.channelizer_hier_ccf, stream_to_vector / vector_to_streams
# ---------------------------------------------------------------------------------------------------------------------
def test_filterbank_fixtures_and_qa(grb):
    fx = fixtures()
    got = grb.filter.filterbank_vcvcf(fx["fbank_taps"]).process(fx["fbank_x"], 200)
    assert_parity(got, fx["fbank_out"], "fixture")
    assert not got[:, 3].any() and not got[:, 7].any()
    # qa_filterbank.py:38-118 incl. the set_taps half
    myrand = random.Random(123).random
    nfilts, ntaps, nd = 10, 5, 100
    data_sets = [[(myrand() - 0.5) + (myrand() - 0.5) * 1j for _ in range(nd)] for _ in range(nfilts)]
    taps1, taps2 = [], []
    for i in range(nfilts):
        taps1.append([0] * ntaps if i in (3, 7) else [myrand() - 0.5 for _ in range(ntaps)])
        taps2.append([0] * ntaps if i in (1, 6, 9) else [myrand() - 0.5 for _ in range(ntaps)])
    vecs = np.array(data_sets, np.complex64).T
    xin = np.concatenate([vecs[-ntaps:], vecs]).reshape(-1)
    fb = grb.filter.filterbank_vcvcf(taps1)
    for taps in (taps1, taps2):
        if taps is taps2:
            fb.set_taps(taps2)
        exp = []
        for ds, ts in zip(data_sets, taps):
            A = ds[-len(ts):] + ds[:-1]
            exp.append([sum(A[i - j] * b for j, b in enumerate(ts)) for i in range(len(ts) - 1, len(A))])
        np.testing.assert_allclose(fb.process(xin, nd), np.array(exp).T, atol=0.5e-6 * 10)
    with pytest.raises(ValueError):
        grb.filter.filterbank_vcvcf([])
    with pytest.raises(RuntimeError):
        grb.filter.filterbank_vcvcf([[1.0, 2.0], [1.0]])
    with pytest.raises(RuntimeError):
        fb.set_taps(taps1[:5])


@pytest.mark.parametrize("nf,nt,n", [(64, 16, 30000), (10, 5, 1000), (1, 1, 77), (3, 8, 5000), (100, 17, 3000), (64, 32, 2000),
                                     (7, 33, 900), (256, 4, 4001),
                                     # the TMA-staged streaming kernel (nf in 16/64/256, <= 32 taps, >= 4096 vectors): all three
                                     # ring depths, one tap, ragged last block, a partial last CTA
                                     (16, 8, 10001), (64, 32, 9000), (256, 4, 5000), (64, 1, 4096), (16, 17, 33333),
                                     (256, 32, 6000), (64, 9, 4097), (256, 16, 20011), (8, 16, 9000), (32, 11, 7001), (128, 20, 5003)])
def test_filterbank_sweep(grb, orc, nf, nt, n):
    rng = np.random.default_rng(nf * 100 + nt)
    taps = rng.uniform(-1, 1, (nf, nt)).astype(np.float32)
    if nf > 2:
        taps[1] = 0
    x = crand(rng, nf * (n + nt))
    x[nf * 5 + 1] = np.nan  # an inactive filter never looks at its input (filterbank_vcvcf_impl.cc:83-88)
    got = grb.filter.filterbank_vcvcf(taps).process(x, n) if nf > 2 else grb.filter.filterbank_vcvcf(taps).process(np.nan_to_num(x), n)
    xr = np.nan_to_num(x)
    assert np.isfinite(got.view(np.float32)).all()
    assert_parity(got, orc.filterbank(taps, xr, n), "filterbank oracle")
    assert_parity(got, po.f64_filterbank(taps, xr, n), "filterbank f64")


def test_filterbank_through_scheduler_mock(grb, orc):
    rng = np.random.default_rng(8)
    nf, nt, n = 16, 9, 7000
    taps = rng.uniform(-1, 1, (nf, nt)).astype(np.float32)
    x = crand(rng, nf * n)
    blk = grb.filter.filterbank_vcvcf(taps)
    got = sched_mock.run_general(blk, [x], 1, max_noutput=512) if False else None
    # one vector stream: drive general_work directly with history = ntaps + 1 (ntaps zero vectors in front)
    xin = np.concatenate([np.zeros(nf * nt, np.complex64), x])
    outs, pos = [], 0
    while pos < n:
        k = min(513, n - pos)
        seg = xin[pos * nf: (pos + k + nt) * nf]
        o = np.empty(k * nf, np.complex64)
        r = blk.general_work(k, [k + nt], [seg], [o])
        assert r == k and blk._consumed == k
        outs.append(o)
        pos += k
    got = np.concatenate(outs).reshape(n, nf)
    assert_parity(got, orc.filterbank(taps, xin, n), "chunked")


def test_channelizer_hier_qa_tones_and_oracle(grb, orc):
    # qa_pfb_channelizer.py:52-54 (test_0001)
    freqs, fs, N = [110., -513., 203., -230, 121], 5000, 1000
    M = len(freqs)
    taps = fixtures()["lp2_qa_taps"]
    x = np.sum([sig_source_c(M * fs, f + i * fs, 1, N) for i, f in enumerate(freqs)], axis=0).astype(np.complex64)
    got = grb.filter.pfb.channelizer_hier_ccf(M, n_filterbanks=1, taps=taps).run(x)
    L = got.shape[1]
    assert got.shape == (M, N // M)
    delay = int(-(np.ceil(len(taps) / float(M)) - 1.0) / 2.0)
    t = np.arange(delay, L + delay) / fs
    for i, f in enumerate(freqs):
        exp, rec = np.exp(2j * np.pi * f * t)[-50:], got[i][-50:]
        np.testing.assert_allclose(rec / rec[0], exp / exp[0], atol=1e-3)
    # 64 channels, several filterbanks (a CPU threading knob: same values), a channel selection
    rng = np.random.default_rng(2)
    M, T, n = 64, 12, 4000
    taps = rng.uniform(-1, 1, M * T - 5).astype(np.float32)
    x = crand(rng, M * n)
    outch = [5, 0, 63, 17]
    got = grb.filter.pfb.channelizer_hier_ccf(M, n_filterbanks=4, taps=taps, outchans=outch).run(x)
    tp = np.concatenate([taps, np.zeros(5, np.float32)])
    chantaps = np.array([tp[i::M][::-1] for i in range(M)], np.float32)
    xin = np.concatenate([np.zeros(M * T, np.complex64), x])
    fb = orc.filterbank(chantaps, xin, n)
    ref = orc.fft_vcc(M, True, np.ones(M, np.float32), False, fb.reshape(-1)).reshape(n, M).T
    assert_parity(got, ref[outch], "hier 64 channels")


def test_stream_to_vector_is_the_same_bytes_and_vector_to_streams(grb):
    import torch

    rng = np.random.default_rng(4)
    N, nvec = 64, 1000
    x = crand(rng, N * nvec)
    s2v = grb.blocks.stream_to_vector(8, N)
    out = np.empty(N * nvec, np.complex64)
    assert s2v.work(nvec, [x], [out]) == nvec
    np.testing.assert_array_equal(out, x)
    d_x = torch.from_numpy(x.view(np.float32)).cuda()
    assert s2v.work_dev(d_x.data_ptr(), nvec) == d_x.data_ptr()  # zero copy on the device
    # chained: items -> (stream_to_vector: nothing) -> fft_vcc on the very same buffer -> vector_to_streams
    f = grb.fft.fft_vcc(N, True, [], False)
    d_y = torch.empty(2 * N * nvec, dtype=torch.float32, device="cuda")
    d_z = torch.empty(2 * N * nvec, dtype=torch.float32, device="cuda")
    f.work_dev(s2v.work_dev(d_x.data_ptr(), nvec), d_y.data_ptr(), nvec)
    v2s = grb.blocks.vector_to_streams(8, N)
    v2s.work_dev(d_y.data_ptr(), d_z.data_ptr(), nvec, nvec)
    torch.cuda.synchronize()
    z = d_z.cpu().numpy().view(np.complex64).reshape(N, nvec)
    ref = po.f64_fft_vcc(N, True, None, False, x).reshape(nvec, N).T
    assert_parity(z, ref, "s2v -> fft -> v2ss")
    # host form, float and complex items, ragged sizes
    for isz, dt, ns, n in ((8, np.complex64, 5, 1001), (4, np.float32, 33, 77), (8, np.complex64, 64, 40000), (4, np.float32, 1, 10)):
        v = (crand(rng, ns * n) if dt is np.complex64 else rng.uniform(-1, 1, ns * n).astype(np.float32))
        outs = grb.blocks.vector_to_streams(isz, ns).process(v)
        np.testing.assert_array_equal(outs, v.reshape(n, ns).T)
    with pytest.raises(ValueError):
        grb.blocks.vector_to_streams(2, 4).process(np.zeros(16, np.float32))
