"""GPU parity (through the C-ABI): pfb_channelizer_ccf, fft_vcc (+ fused / standalone complex_to_mag_squared)."""
import numpy as np
import pytest

import sched_mock
from common import assert_parity, assert_tuples_almost_equal2, crand, fixtures, qa_vectors, sig_source_c
from oracle import pyoracle as po

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def grb():
    import gnuradio_b200 as g

    assert g.device_count() > 0, "no CUDA device: libgrb200 has no CPU fallback"
    return g


def test_pfb_reference_fixtures(grb):
    fx = fixtures()
    for tag in "abc":
        M, osr, nout, cons, T, nouts = fx["pfb_%s_meta" % tag]
        M, nout, cons, T, nouts = int(M), int(nout), int(cons), int(T), int(nouts)
        blk = grb.filter.pfb_channelizer_ccf(M, fx["pfb_%s_taps" % tag], float(osr))
        blk.set_channel_map(fx["pfb_%s_map" % tag])
        assert blk.taps_per_filter() == T and blk.history() == T + 1
        assert blk.channel_map() == fx["pfb_%s_map" % tag].tolist()
        got = blk.process(fx["pfb_%s_in" % tag], nout, nouts)
        assert blk._consumed == cons
        assert_parity(got, fx["pfb_%s_out" % tag], "fixture " + tag)


def test_pfb_ctor_contract_and_errors(grb):
    taps = np.ones(50, np.float32)
    blk = grb.filter.pfb_channelizer_ccf(8, taps, 2.0)
    assert blk.relative_rate() == 2.0 and blk.output_multiple() == 2 and blk.taps_per_filter() == 7
    t = np.array(blk.taps())
    padded = np.concatenate([taps, np.zeros(6, np.float32)])
    np.testing.assert_array_equal(t, padded.reshape(7, 8).T)  # d_taps[i][j] = padded[i + j*M]
    # qa_pfb_channelizer.py:56-66: 36/25 is fine, 10.1334 is refused
    grb.filter.pfb_channelizer_ccf(36, taps, 36 / 25.)
    with pytest.raises(ValueError):
        grb.filter.pfb_channelizer_ccf(36, taps, 10.1334)
    with pytest.raises(ValueError):
        blk.set_channel_map([0, 8])
    blk.set_channel_map([])  # empty map: unchanged
    assert blk.channel_map() == list(range(8))
    # the constructor goes through set_taps: the first general_work returns 0 (pfb_channelizer_ccf_impl.cc:140-143)
    ins = [np.zeros(100, np.complex64)] * 8
    outs = [np.empty(4, np.complex64) for _ in range(8)]
    assert blk.general_work(4, [], ins, outs) == 0
    assert blk.general_work(4, [], ins, outs) == 4


@pytest.mark.parametrize("M,ntaps,osr", [(64, 1024, 1.0), (16, 200, 1.0), (256, 2048, 1.0), (64, 1024, 2.0),
                                          (5, 43, 1.0), (6, 31, 1.5), (64, 130, 1.0), (64, 2048, 1.0), (3, 2, 3.0),
                                          # the streaming kernel's other channel counts (M = R1 x R2 with R1 != R2, and M = 8)
                                          (8, 100, 1.0), (32, 500, 1.0), (128, 2048, 1.0), (32, 1024, 1.0), (128, 700, 1.0),
                                          (8, 256, 1.0),
                                          # pfb_fftgen_kernel: oversampled, more than 32 taps per branch, R1 != R2
                                          (64, 64 * 40, 1.0), (32, 512, 2.0), (128, 640, 4.0 / 3.0), (16, 100, 4.0), (8, 64, 8.0),
                                          (256, 256 * 3, 2.0), (32, 32 * 33, 1.0),
                                          # 33 .. 64 taps per branch at rate 1: two launches of the streaming kernel, the second
                                          # one (the older taps, 1 .. 32 of them: every ring depth) adding to the first
                                          (8, 8 * 64, 1.0), (256, 256 * 48, 1.0), (16, 16 * 33 - 5, 1.0), (128, 128 * 41, 1.0),
                                          (64, 64 * 56, 1.0),
                                          # channel counts that are not a power of two: the two-stage DFT of the generic kernel
                                          (48, 48 * 7, 1.0), (100, 1000, 1.0), (12, 100, 1.5), (20, 200, 2.0), (7, 50, 1.0), (96, 960, 4.0 / 3.0),
                                          # oversample_rate 2 on the streaming kernel (8 .. 128 channels, every ring depth)
                                          (8, 64, 2.0), (16, 100, 2.0), (128, 1024, 2.0), (32, 32 * 30, 2.0), (64, 64 * 5, 2.0)])
def test_pfb_against_oracle(grb, orc, M, ntaps, osr):
    rng = np.random.default_rng(M * 1000 + ntaps)
    taps = rng.uniform(-1, 1, ntaps).astype(np.float32)
    rows = 3000
    srcs = crand(rng, M * rows).reshape(M, rows)
    blk = grb.filter.pfb_channelizer_ccf(M, taps, osr)
    T = blk.taps_per_filter()
    outs = sched_mock.run_general(blk, list(srcs), M, max_noutput=1024)
    nout = len(outs[0])
    assert nout > 0 and nout % blk.output_multiple() == 0
    ins = np.concatenate([np.zeros((M, T), np.complex64), srcs], axis=1)
    ref, cons = orc.pfb_channelizer(M, taps, osr, ins, nout)
    assert_parity(np.array(outs), ref, "M=%d nt=%d osr=%g" % (M, ntaps, osr))
    hp = np.concatenate([taps, np.zeros(M * T - ntaps, np.float32)])
    if M * nout <= 64 * 3000:
        f64 = po.f64_channelizer(M, hp, srcs.T.reshape(-1), nout, osr)
        assert_parity(np.array(outs), f64, "f64")


@pytest.mark.parametrize("M,osr", [(32, 1.0), (128, 2.0), (8, 1.0)])
def test_pfb_channel_map_other_channel_counts(grb, orc, M, osr):
    """A permuted channel map with fewer outputs than channels on the FFT-based general kernel."""
    rng = np.random.default_rng(M + 5)
    ntaps, nout = M * 9 - 3, 1200
    taps = rng.uniform(-1, 1, ntaps).astype(np.float32)
    blk = grb.filter.pfb_channelizer_ccf(M, taps, osr)
    T = blk.taps_per_filter()
    rows = int(nout / osr) + 2
    ins = np.concatenate([np.zeros((M, T), np.complex64), crand(rng, M * rows).reshape(M, rows)], axis=1)
    cmap = rng.permutation(M).astype(np.int32)
    blk.set_channel_map(cmap)
    nouts = max(1, M // 3)
    got = blk.process(ins, nout, nouts)
    ref, _ = orc.pfb_channelizer(M, taps, osr, ins, nout, channel_map=cmap, nouts=nouts)
    assert_parity(got, ref, "mapped M=%d osr=%g" % (M, osr))


def test_pfb_channel_map_and_fewer_outputs_fast_path(grb, orc):
    rng = np.random.default_rng(11)
    M, ntaps, nout = 64, 1024, 2000
    taps = rng.uniform(-1, 1, ntaps).astype(np.float32)
    ins = np.concatenate([np.zeros((M, 16), np.complex64), crand(rng, M * nout).reshape(M, nout)], axis=1)
    blk = grb.filter.pfb_channelizer_ccf(M, taps, 1.0)
    cmap = rng.permutation(M).astype(np.int32)
    blk.set_channel_map(cmap)
    got = blk.process(ins, nout, 10)
    ref, _ = orc.pfb_channelizer(M, taps, 1.0, ins, nout, channel_map=cmap, nouts=10)
    assert_parity(got, ref, "mapped, 10 outputs")
    blk.set_channel_map(np.arange(M, dtype=np.int32))
    got = blk.process(ins, nout, 7)
    ref, _ = orc.pfb_channelizer(M, taps, 1.0, ins, nout, nouts=7)
    assert_parity(got, ref, "identity, 7 outputs")


def test_pfb_hier_qa_tones(grb):
    # qa_pfb_channelizer.py: pfb.channelizer_ccf on ONE interleaved stream, analytic tones, 3 places
    freqs, fs, N = [110., -513., 203., -230, 121], 5000, 1000
    M = len(freqs)
    ifs = M * fs
    taps = fixtures()["lp2_qa_taps"]
    x = np.sum([sig_source_c(ifs, f + i * fs, 1, N) for i, f in enumerate(freqs)], axis=0).astype(np.complex64)
    got = grb.filter.pfb.channelizer_ccf(M, taps=taps, oversample_rate=1).run(x)
    L = got.shape[1]
    assert L == N // M
    delay = int(-(np.ceil(len(taps) / float(M)) - 1.0) / 2.0)
    t = np.arange(delay, L + delay) / fs
    for i, f in enumerate(freqs):
        exp, rec = np.exp(2j * np.pi * f * t)[-50:], got[i][-50:]
        np.testing.assert_allclose(rec / rec[0], exp / exp[0], atol=1e-3)
    with pytest.raises(ValueError):
        grb.filter.pfb.channelizer_ccf(36, taps=taps, oversample_rate=10.1334)


def test_pfb_interleaved_equals_streams_large(grb, orc):
    rng = np.random.default_rng(12)
    M, T, nout = 64, 16, 1 << 14
    taps = rng.uniform(-1, 1, M * T).astype(np.float32)
    x = crand(rng, M * nout)
    hier = grb.filter.pfb.channelizer_ccf(M, taps=taps)
    a = hier.run(x)
    ins = np.concatenate([np.zeros((M, T), np.complex64), x.reshape(nout, M).T], axis=1)
    b = grb.filter.pfb_channelizer_ccf(M, taps, 1.0).process(ins, nout)
    np.testing.assert_array_equal(a, b)
    ref, _ = orc.pfb_channelizer(M, taps, 1.0, ins, nout)
    assert_parity(a, ref, "2^20 samples")


# ---------------------------------------------------------------------------------------------------
def test_fft_golden_vectors(grb):
    # gr-fft/python/fft/qa_fft.py golden vectors through the block
    q = qa_vectors()
    primes = q["primes"]
    src = np.array([complex(primes[2 * i], primes[2 * i + 1]) for i in range(32)], np.complex64)
    rev_in = (q["primes_transformed"] / 32).astype(np.complex64)
    ham = fixtures()["win_hamming_32"]
    f = grb.fft.fft_vcc
    assert_tuples_almost_equal2(q["primes_transformed"], f(32, True, [], False).process(src), 1e-9, 4e-4)
    assert_tuples_almost_equal2(src, f(32, False, [], False).process(rev_in), 1e-9, 4e-4)
    assert_tuples_almost_equal2(src, f(32, False, [], False, 2).process(rev_in), 1e-9, 4e-4)
    assert_tuples_almost_equal2(q["test_window"], f(32, True, ham, False).process(src), 1e-9, 4e-4)
    assert_tuples_almost_equal2(q["test_reverse_window_shift"], f(32, False, ham, True).process(rev_in), 1e-9, 4e-4)


def test_fft_vcc_reference_fixtures(grb):
    fx = fixtures()
    for fwd in (0, 1):
        for sh in (0, 1):
            got = grb.fft.fft_vcc(64, fwd, fx["win_bh_64"], sh).process(fx["vcc_x"])
            assert_parity(got, fx["vcc64_f%d_s%d_w" % (fwd, sh)], "w f%d s%d" % (fwd, sh))
            got = grb.fft.fft_vcc(64, fwd, [], sh).process(fx["vcc_x"])
            assert_parity(got, fx["vcc64_f%d_s%d" % (fwd, sh)], "f%d s%d" % (fwd, sh))
        got = grb.fft.fft_vcc(15, fwd, [], True).process(fx["vcc_x15"])
        assert_parity(got, fx["vcc15_f%d_s1" % fwd], "odd size shift f%d" % fwd)


@pytest.mark.parametrize("N", [1, 2, 4, 8, 16, 32, 64, 128, 256, 512, 1024, 2048, 4096, 8192, 16384, 12, 100, 1000,
                               129, 600, 1536, 2049, 5000, 8191])  # (sizes that are not powers of two: Bluestein from 128 on)
def test_fft_vcc_sizes(grb, orc, N):
    rng = np.random.default_rng(N)
    nvec = 37 if N <= 4096 else 5
    x = crand(rng, N * nvec)
    w = rng.uniform(0.1, 1, N).astype(np.float32)
    for fwd in (True, False):
        for sh in (False, True):
            got = grb.fft.fft_vcc(N, fwd, w, sh).process(x)
            assert_parity(got, orc.fft_vcc(N, fwd, w, sh, x), "N=%d f%d s%d" % (N, fwd, sh))
            assert_parity(got, po.f64_fft_vcc(N, fwd, w, sh, x), "f64 N=%d f%d s%d" % (N, fwd, sh))


@pytest.mark.parametrize("N,nvec", [(32768, 3), (65536, 5), (131072, 2), (262144, 3)])
def test_fft_vcc_big_sizes_four_step(grb, N, nvec):
    """Powers of two above the single-CTA transform: columns -> batched rows -> gather (fft_vcc, fused |.|^2, fft_vfc)."""
    rng = np.random.default_rng(N + nvec)
    x = crand(rng, N * nvec)
    w = rng.uniform(0.1, 1, N).astype(np.float32)
    for fwd in (True, False):
        for sh in (False, True):
            got = grb.fft.fft_vcc(N, fwd, w, sh).process(x)
            assert_parity(got, po.f64_fft_vcc(N, fwd, w, sh, x), "f64 N=%d f%d s%d" % (N, fwd, sh))
    P = grb.fft.fft_vcc(N, True, w, True).process(x, magsq=True)
    ref = po.f64_fft_vcc(N, True, w, True, x)
    assert_parity(P, (ref.real ** 2 + ref.imag ** 2), "fused mag^2 N=%d" % N)
    xr = rng.uniform(-1, 1, N * nvec).astype(np.float32)
    got = grb.fft.fft_vfc(N, True, w).process(xr)
    assert_parity(got, po.f64_fft_vcc(N, True, w, False, xr.astype(np.complex64)), "vfc N=%d" % N)


@pytest.mark.parametrize("N", [32, 64, 128, 256, 512, 1024, 2048, 4096, 8192])
def test_fft_vcc_streaming_kernel(grb, orc, N):
    """>= 64 vectors per work() call take the persistent TMA-prefetch kernel (fft_vcc_stream_kernel);
    odd vector counts exercise the clipped last bulk copy."""
    rng = np.random.default_rng(1000 + N)
    nvec = 701 if N <= 2048 else 333  # (odd: below 1024 several vectors share a bulk copy, the last one is clipped)
    x = crand(rng, N * nvec)
    w = rng.uniform(0.1, 1, N).astype(np.float32)
    for fwd in (True, False):
        for sh in (False, True):
            for win in (w, []):
                got = grb.fft.fft_vcc(N, fwd, win, sh).process(x)
                wl = win if len(win) else None
                assert_parity(got, orc.fft_vcc(N, fwd, wl, sh, x), "N=%d f%d s%d" % (N, fwd, sh))
                assert_parity(got, po.f64_fft_vcc(N, fwd, wl, sh, x), "f64 N=%d f%d s%d" % (N, fwd, sh))
    P = grb.fft.fft_vcc(N, True, w, True).process(x, magsq=True)
    assert_parity(P, orc.mag_squared(orc.fft_vcc(N, True, w, True, x)), "fused mag^2 N=%d" % N)


def test_fft_vcc_config5_window_and_magsq(grb, orc):
    # BASELINE config 5 geometry: 4096-point Blackman-Harris, forward, + complex_to_mag_squared
    rng = np.random.default_rng(21)
    N, nvec = 4096, 300
    x = crand(rng, N * nvec)
    w = fixtures()["win_bh_4096"]
    for sh in (False, True):
        blk = grb.fft.fft_vcc(N, True, w, sh)
        X = sched_mock.run_sync(blk, x, max_noutput=7, item_in=N, item_out=N)
        ref = orc.fft_vcc(N, True, w, sh, x)
        assert_parity(X, ref, "fft_vcc 4096 BH shift=%d" % sh)
        P = blk.process(x, magsq=True)
        Pref = orc.mag_squared(ref)
        assert_parity(P, Pref, "fused mag^2")
        P2 = grb.blocks.complex_to_mag_squared(N).process(X)
        assert_parity(P2, orc.mag_squared(X), "standalone mag^2", 1e-6, 1e-5)


def test_fft_vcc_window_contract(grb):
    blk = grb.fft.fft_vcc(64, True, [], False)
    assert blk.set_window(np.ones(63, np.float32)) is False  # fft_vcc_fftw.cc:54-61
    assert blk.set_window(np.ones(64, np.float32)) is True
    assert blk.set_window([]) is True
    with pytest.raises(RuntimeError):
        grb.fft.fft_vcc(64, True, np.ones(63, np.float32), False)  # :46-47
    with pytest.raises(IndexError):
        grb.fft.fft_vcc(0, True, [], False)  # fft.cc:161-163 std::out_of_range
    assert blk.work(0, [np.zeros(0, np.complex64)], [np.zeros(0, np.complex64)]) == 0


def test_mag_squared_qa(grb):
    q = qa_vectors()
    got = grb.blocks.complex_to_mag_squared().process(q["magsq_src"].astype(np.complex64))
    np.testing.assert_allclose(got, q["magsq_expected"], rtol=1e-6)
