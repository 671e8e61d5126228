"""CPU: pins the oracle for the section 8(f) blocks added in round 2 -- fir_filter_fff, fft_vfc, logpwrfft_c/f,
pfb_synthesizer_ccf, filterbank_vcvcf, channelizer_hier_ccf -- against fixtures generated from the reference's own sources
(tests/golden/make_golden.py), the reference QA cases, and float64 closed forms; and the plain-C port against the reference layer."""
import math
import random

import numpy as np
import pytest

from common import assert_parity, crand, fixtures, sig_source_c
from oracle import pyoracle as po


@pytest.fixture(params=["port", "reference"])
def lay(request):
    if request.param == "reference":
        if not po.have_ref():
            pytest.skip("oracle/_ref not built on this machine (needs /root/reference)")
        return po.Ref()
    return po.Port()


def test_fir_fff_fixtures_qa_and_f64(lay):
    fx = fixtures()
    for D in (1, 3):
        xin = np.concatenate([np.zeros(100, np.float32), fx["fff_x"]])
        got = lay.fir_fff(D, fx["fff_taps"], xin, len(fx["fff_x"]) // D)
        assert_parity(got, fx["fff_d%d" % D], "fixture d%d" % D, 1e-6, 1e-5)
        assert_parity(got, po.f64_fir(D, fx["fff_taps"], fx["fff_x"]).real, "f64 d%d" % D)
    # qa_fir_filter.py:32-58 (fir_filter_fff, decimation 1 and 4)
    src = np.array(40 * [1, 2, 3, 4], np.float32)
    taps = np.array(20 * [0.5, 0.5], np.float32)
    for D in (1, 4):
        xin = np.concatenate([np.zeros(39, np.float32), src])
        got = lay.fir_fff(D, taps, xin, len(src) // D)
        np.testing.assert_allclose(got, po.f64_fir(D, taps, src).real, atol=0.5e-5 * 80)


def test_fft_vfc_fixtures_and_f64(lay):
    fx = fixtures()
    assert_parity(lay.fft_vfc(64, True, fx["win_bh_64"], fx["vfc_x"]), fx["vfc64_f1_w"], "fwd window", 2e-6, 2e-5)
    assert_parity(lay.fft_vfc(64, False, None, fx["vfc_x"]), fx["vfc64_f0"], "reverse", 2e-6, 2e-5)
    assert_parity(lay.fft_vfc(64, True, fx["win_bh_64"], fx["vfc_x"]), po.f64_fft_vfc(64, True, fx["win_bh_64"], fx["vfc_x"]), "f64")
    with pytest.raises(RuntimeError):
        lay.fft_vfc(64, True, np.ones(63, np.float32), fx["vfc_x"])
    # a real input gives a Hermitian spectrum
    X = lay.fft_vfc(64, True, None, fx["vfc_x"][:64])
    np.testing.assert_allclose(X[1:], np.conj(X[1:][::-1]), atol=1e-4)


def test_logpwrfft_fixtures_f64_and_state(lay):
    fx = fixtures()
    w = fx["win_bh_64"]
    cases = (("logpwr_c_d3_avg", True, 2.0, 3, 0.25, True, "logpwr_xc"), ("logpwr_c_d1", True, 2.0, 1, 1.0, False, "logpwr_xc"),
             ("logpwr_f_d2_avg", False, 1.0, 2, 0.1, True, "logpwr_xf"))
    for key, cin, ref_scale, dec, alpha, avg, xk in cases:
        got, cnt, st = lay.logpwrfft(cin, 64, ref_scale, dec, alpha, avg, w, fx[xk])
        assert got.shape == fx[key].shape == (37 // dec, 64)
        assert_parity(got, fx[key], key, 2e-6, 1e-4)
        assert_parity(got, po.f64_logpwrfft(cin, 64, ref_scale, dec, alpha, avg, w, fx[xk]), key + " f64", 1e-5, 2e-4)
        # any split of the stream gives the same frames: the decimator's counter and the filter memory carry over
        a, c1, s1 = lay.logpwrfft(cin, 64, ref_scale, dec, alpha, avg, w, fx[xk][: 64 * 11])
        b, c2, s2 = lay.logpwrfft(cin, 64, ref_scale, dec, alpha, avg, w, fx[xk][64 * 11:], count=c1, iir=s1)
        np.testing.assert_array_equal(np.concatenate([a, b]), got)
        assert c2 == cnt
    # keep_one_in_n keeps the LAST vector of each group (keep_one_in_n_impl.cc:66-74): with decimation 3 frame 0 is vector 2
    one, _, _ = lay.logpwrfft(True, 64, 2.0, 3, 1.0, False, w, fx["logpwr_xc"][: 64 * 3])
    direct, _, _ = lay.logpwrfft(True, 64, 2.0, 1, 1.0, False, w, fx["logpwr_xc"][64 * 2: 64 * 3])
    np.testing.assert_array_equal(one, direct)


def test_pfb_synthesizer_fixtures_and_closed_form(lay):
    fx = fixtures()
    got = lay.pfb_synthesizer(5, fx["synth_taps"], False, fx["synth_in5"])
    assert_parity(got, fx["synth_5_1x"], "1x fixture", 2e-6, 2e-5)
    assert_parity(got, po.f64_pfb_synthesizer(5, fx["synth_taps"], False, fx["synth_in5"]), "1x f64")
    got = lay.pfb_synthesizer(5, fx["synth_taps"], False, fx["synth_in5"][:3], channel_map=fx["synth_map"])
    assert_parity(got, fx["synth_5_1x_map3"], "3 inputs, map", 2e-6, 2e-5)
    assert_parity(got, po.f64_pfb_synthesizer(5, fx["synth_taps"], False, fx["synth_in5"][:3], fx["synth_map"]), "map f64")
    got = lay.pfb_synthesizer(8, fx["synth_taps8"], True, fx["synth_in8"])
    assert_parity(got, fx["synth_8_2x"], "2x fixture", 2e-6, 2e-5)
    assert_parity(got, po.f64_pfb_synthesizer(8, fx["synth_taps8"], True, fx["synth_in8"]), "2x f64")
    with pytest.raises(ValueError):
        lay.pfb_synthesizer(5, fx["synth_taps"], True, fx["synth_in5"])  # odd channel count with 2x (:44-47)
    with pytest.raises(ValueError):
        lay.pfb_synthesizer(5, fx["synth_taps"], False, fx["synth_in5"], channel_map=[0, 1, 2, 3, 5])  # :207-212


def test_pfb_synthesizer_qa_tones(lay):
    """gr-filter/python/filter/qa_pfb_synthesizer.py:31-84: five baseband tones -> one wideband stream, 4 places."""
    N, M, fs = 10000, 5, 1000
    ofs = M * fs
    taps = fixtures()["synth_taps"]
    freqs = [0, 100, 200, -200, -100]
    ins = np.array([sig_source_c(fs, f, 1, N) for f in freqs], np.complex64)
    got = lay.pfb_synthesizer(M, taps, False, ins)
    p = [2 * math.pi * k / M for k in range(M)]
    t = np.arange(len(got)) / ofs
    ofreqs = [-2200, -1100, 0, 1100, 2200]
    ph = [p[3], p[4], p[0], p[1], p[2]]
    exp = sum(np.exp(1j * (2 * np.pi * f * t + q)) for f, q in zip(ofreqs, ph))
    np.testing.assert_allclose(got[2000:3000], exp[2000:3000], atol=0.5e-4 * 10)


def test_filterbank_fixtures_qa_and_f64(lay):
    fx = fixtures()
    got = lay.filterbank(fx["fbank_taps"], fx["fbank_x"], 200)
    assert_parity(got, fx["fbank_out"], "fixture", 2e-6, 2e-5)
    assert_parity(got, po.f64_filterbank(fx["fbank_taps"], fx["fbank_x"], 200), "f64")
    assert not got[:, 3].any() and not got[:, 7].any()  # all-zero filters write zeros (filterbank_vcvcf_impl.cc:83-88)
    # qa_filterbank.py:38-118: 10 filters x 5 taps on repeating data, against the QA's own python convolution
    myrand = random.Random(123).random
    nfilts, ntaps, nd = 10, 5, 100
    data_sets = [[(myrand() - 0.5) + (myrand() - 0.5) * 1j for _ in range(nd)] for _ in range(nfilts)]
    taps1 = []
    for i in range(nfilts):
        taps1.append([0] * ntaps if i in (3, 7) else [myrand() - 0.5 for _ in range(ntaps)])
        [myrand() for _ in range(0 if i in (1, 6, 9) else ntaps)]  # the QA draws taps2 in the same loop
    results = []
    for ds, ts in zip(data_sets, taps1):
        A = ds[-len(ts):] + ds[:-1]
        results.append([sum(A[i - j] * b for j, b in enumerate(ts)) for i in range(len(ts) - 1, len(A))])
    expected = np.array(results).T  # [time][filter]
    # a repeating vector source: the history in front of one period is the end of the previous one
    vecs = np.array(data_sets, np.complex64).T  # [time][filter]
    xin = np.concatenate([vecs[-ntaps:], vecs]).reshape(-1)
    got = lay.filterbank(np.array(taps1, np.float32), xin, nd)
    np.testing.assert_allclose(got, expected, atol=0.5e-6 * 10)


def test_channelizer_hier_is_filterbank_plus_fft(lay):
    """pfb.channelizer_hier_ccf (gr-filter/python/filter/pfb.py:406-480) composed from the oracle's filterbank_vcvcf and
    fft_vcc must pass the QA that the reference applies to it (qa_pfb_channelizer.py:52-54,87-153, 3 places)."""
    freqs, fs, N = [110., -513., 203., -230, 121], 5000, 1000
    M = len(freqs)
    ifs = M * fs
    taps = list(fixtures()["lp2_qa_taps"])
    x = np.sum([sig_source_c(ifs, f + i * fs, 1, N) for i, f in enumerate(freqs)], axis=0).astype(np.complex64)
    extra = int(math.ceil(1.0 * len(taps) / M) * M - len(taps))
    taps = taps + [0] * extra
    chantaps = np.array([list(reversed(taps[i::M])) for i in range(M)], np.float32)
    nt = chantaps.shape[1]
    n = N // M
    xin = np.concatenate([np.zeros(nt * M, np.complex64), x[: n * M]])
    fb = lay.filterbank(chantaps, xin, n)  # [n][M]
    spec = lay.fft_vcc(M, True, np.ones(M, np.float32), False, fb.reshape(-1)).reshape(n, M)
    got = spec.T
    delay = int(-(np.ceil(len(fixtures()["lp2_qa_taps"]) / float(M)) - 1.0) / 2.0)
    t = np.arange(delay, n + delay) / fs
    for i, f in enumerate(freqs):
        exp, rec = np.exp(2j * np.pi * f * t)[-50:], got[i][-50:]
        np.testing.assert_allclose(rec / rec[0], exp / exp[0], atol=1e-3)


def test_port_matches_reference_layer_new_blocks():
    if not po.have_ref():
        pytest.skip("oracle/_ref not built here")
    R, P = po.Ref(), po.Port()
    rng = np.random.default_rng(100)
    for D, nt in ((1, 256), (5, 33)):
        x = rng.uniform(-1, 1, 3000 * D + nt - 1).astype(np.float32)
        t = rng.uniform(-1, 1, nt).astype(np.float32)
        assert_parity(P.fir_fff(D, t, x, 3000), R.fir_fff(D, t, x, 3000), "fff", 1e-6, 1e-5)
    for M, T, twox in ((64, 16, False), (6, 5, True), (7, 3, False)):
        taps = rng.uniform(-1, 1, M * T - 1).astype(np.float32)
        ins = crand(rng, M * 200).reshape(M, 200)
        assert_parity(P.pfb_synthesizer(M, taps, twox, ins), R.pfb_synthesizer(M, taps, twox, ins), "synth", 2e-6, 2e-5)
        assert np.array_equal(R.pfb_synthesizer(M, taps, twox, ins, chunk=13 * M), R.pfb_synthesizer(M, taps, twox, ins))
    t = rng.uniform(-1, 1, (33, 17)).astype(np.float32)
    x = crand(rng, 33 * (300 + 17))
    assert_parity(P.filterbank(t, x, 300), R.filterbank(t, x, 300), "filterbank", 1e-6, 1e-5)
