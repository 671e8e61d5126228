"""CPU: pins the oracle.  Both C layers -- the reference's own compiled sources (oracle/_ref, build container
only) and the plain-C port (travels) -- against the committed fixtures generated from the reference sources,
the reference QA's golden vectors, and float64 closed forms."""
import random

import numpy as np
import pytest

from common import (assert_parity, assert_tuples_almost_equal2, crand, fixtures, make_random_complex_tuple,
                    qa_vectors, sig_source_c)
from oracle import pyoracle as po


def layers():
    out = [po.Port()]
    if po.have_ref():
        out.append(po.Ref())
    return out


@pytest.fixture(params=["port", "reference"])
def lay(request):
    if request.param == "reference":
        if not po.have_ref():
            pytest.skip("oracle/_ref not built on this machine (needs /root/reference)")
        return po.Ref()
    return po.Port()


def test_fir_against_reference_fixtures(lay):
    fx = fixtures()
    for D in (1, 4):
        assert_parity(lay.fir_stream(D, fx["fir_taps_f"], fx["fir_x"]), fx["fir_ccf_d%d" % D], "ccf", 1e-6, 1e-5)
        assert_parity(lay.fir_stream(D, fx["fir_taps_c"], fx["fir_x"]), fx["fir_ccc_d%d" % D], "ccc", 1e-6, 1e-5)
        assert_parity(fx["fir_ccf_d%d" % D], po.f64_fir(D, fx["fir_taps_f"], fx["fir_x"]), "ccf f64", 1e-6, 1e-5)


def test_fir_qa_cases(lay):
    # gr-filter/python/filter/qa_fir_filter.py:60-129: ccf_001/002, ccc_001/003 inputs; 5 places
    src = np.array(40 * [1 + 1j, 2 + 2j, 3 + 3j, 4 + 4j], np.complex64)
    for taps, decs in ((np.array(20 * [0.5, 0.5], np.float32), (1, 4)),
                       (np.array(20 * [0.5 + 1j, 0.5 + 1j], np.complex64), (1, 4))):
        for D in decs:
            got = lay.fir_stream(D, taps, src)
            exp = po.f64_fir(D, taps, src)
            assert len(got) == len(src) // D
            np.testing.assert_allclose(got, exp, atol=0.5e-5 * 80, rtol=0)  # values reach ~80: 5 places relative


def test_fft_filter_against_fixtures_and_f64(lay):
    fx = fixtures()
    assert lay.fft_filter_nsamples(300) == int(fx["fftfilt_nsamples_300"]) == 1024 - 300 + 1
    for nt, ns in ((1, 2), (2, 3), (64, 65), (1024, 1025), (4096, 4097), (5, 12)):
        assert lay.fft_filter_nsamples(nt) == ns  # gr-filter/lib/fft_filter.cc:208-209
    for D in (1, 3):
        got = lay.fft_filter_ccc(D, fx["fftfilt_taps"], fx["fftfilt_x"])
        assert_parity(got, fx["fftfilt_d%d" % D], "fixture", 2e-6, 2e-5)
        assert_parity(got, po.f64_fft_filter(D, fx["fftfilt_taps"], fx["fftfilt_x"], 725), "f64", 2e-6, 2e-5)


def test_fft_filter_qa_random_trials(lay):
    # qa_fft_filter.py:144-184 (test_ccc_004/005): seed 0, 4096 samples, fft filter vs fir filter, rel_eps 4e-4
    random.seed(0)
    for i in range(6):
        src = np.array(make_random_complex_tuple(4096), np.complex64)
        ntaps = int(random.uniform(2, 1000))
        taps = np.array(make_random_complex_tuple(ntaps), np.complex64)
        exp = lay.fir_stream(1, taps, src)
        got = lay.fft_filter_ccc(1, taps, src)
        assert_tuples_almost_equal2(exp[: len(got)], got, abs_eps=1e-9, rel_eps=4e-4 * 50)
        assert_parity(got, exp[: len(got)], "trial %d" % i, 2e-6, 4e-5)
    random.seed(0)
    for i in range(6):
        dec = i + 1
        src = np.array(make_random_complex_tuple(4096), np.complex64)
        taps = np.array(make_random_complex_tuple(int(random.uniform(2, 100))), np.complex64)
        exp = lay.fir_stream(dec, taps, src)
        got = lay.fft_filter_ccc(dec, taps, src)
        assert_parity(got, exp[: len(got)], "dec %d" % dec, 2e-6, 4e-5)


def test_fft_filter_ccf_qa(lay):
    """fft_filter_ccf: qa_fft_filter.py:214-283 -- identity / gain taps (test_ccf_001..003) and the
    seeded random trials against fir_filter_ccf (test_ccf_004, rel_eps 4e-4); both oracle layers, real taps."""
    src = np.arange(8).astype(np.complex64)
    assert_tuples_almost_equal2(src.tolist(), lay.fft_filter_ccf(1, [1.0], src).tolist(), abs_eps=1e-5, rel_eps=1e-5)
    assert_tuples_almost_equal2((2 * src).tolist(), lay.fft_filter_ccf(1, [2.0], src).tolist(), abs_eps=1e-5, rel_eps=1e-5)
    random.seed(0)
    for i in range(6):
        src = np.array(make_random_complex_tuple(4096), np.complex64)
        ntaps = int(random.uniform(2, 1000))
        taps = np.array([2 * random.random() - 1 for _ in range(ntaps)], np.float32)
        exp = lay.fir_stream(1, taps, src)
        got = lay.fft_filter_ccf(1, taps, src)
        assert len(got) == (4096 // lay.fft_filter_nsamples(ntaps)) * lay.fft_filter_nsamples(ntaps)
        assert_parity(got, exp[: len(got)], "ccf trial %d" % i, 2e-6, 4e-5)
        assert_parity(got, lay.fft_filter_ccc(1, taps.astype(np.complex64), src), "ccf == ccc with (t, 0) taps", 1e-7, 1e-6)
        assert_parity(got, po.f64_fft_filter(1, taps, src, lay.fft_filter_nsamples(ntaps)), "ccf f64", 2e-6, 2e-5)


def test_freq_xlating_against_fixtures_and_f64(lay):
    fx = fixtures()
    fc, fs, chunk = fx["xlat_meta"]
    x = fx["xlat_x"]
    for D in (1, 4):
        for tag in ("f", "c"):
            t = fx["xlat_taps_" + tag]
            xin = np.concatenate([np.zeros(len(t) - 1, np.complex64), x])
            got, ph = lay.freq_xlating_fir(D, t, fc, fs, xin, len(x) // D, chunk_out=int(chunk) // D)
            assert_parity(got, fx["xlat_cc%s_d%d" % (tag, D)], "fixture", 2e-6, 2e-5)
            assert abs(ph - fx["xlat_cc%s_d%d_phase" % (tag, D)]) < 1e-5
            # The rotator's float phasor recurrence (rotator.h:40-41 / VOLK rotator) is only renormalised every 512
            # items: between those points |phase| wanders, and its angle drifts off the increment's by up to an ulp per
            # item (systematically when fc/fs is a simple ratio).  Over these 4000 items the reference therefore sits
            # 6-8e-6 NRMS (max 4e-5 x RMS) from its own float64 closed form: inside north_star's 1e-5 / 1e-4, not by
            # much, and not for much longer runs (2e-5 at 400 000 items).
            assert_parity(got, po.f64_freq_xlating_fir(D, t, fc, fs, x), "f64", 1e-5, 1e-4)

def test_freq_xlating_qa_cases(lay):
    # gr-filter/python/filter/qa_freq_xlating_fir_filter.py:44-50,96-127 (ccf_001/002) and :129-160 (ccc_001/002):
    # expected = fir_filter(mix(sig_source_c(fs, -fc), src), taps, decim), 5 places.  The QA designs its taps with
    # firdes.low_pass / complex_band_pass (bw 0.1, transition bw/4); the same pass band comes from low_pass_2 here
    # and, for ccc, from those taps moved to +bw/8 -- the identity under test does not depend on the tap values.
    fs, fc, bw = 1.0, 0.3, 0.1
    lp = lay.low_pass_2(1, fs, bw, bw / 4, 53)
    src = np.exp(-2j * np.pi * fc / fs * (np.arange(100) / 100.0))
    lo = np.exp(-2j * np.pi * fc / fs * np.arange(100))
    for taps in (lp, (lp * np.exp(2j * np.pi * bw / 8 * np.arange(len(lp)))).astype(np.complex64)):
        for D in (1, 4):
            xin = np.concatenate([np.zeros(len(taps) - 1, np.complex64), src.astype(np.complex64)])
            got, _ = lay.freq_xlating_fir(D, taps, fc, fs, xin, 100 // D)
            exp = po.f64_fir(D, taps, lo * src)
            np.testing.assert_allclose(got, exp, atol=0.5e-5, rtol=0)


def test_freq_xlating_phase_carries_across_calls(lay):
    # rotator state (rotator.h:25-27) survives work() calls: two runs with the phase handed over equal one run
    rng = np.random.default_rng(5)
    taps = rng.uniform(-1, 1, 21).astype(np.float32)
    x = crand(rng, 2048 + 20)
    one, ph1 = lay.freq_xlating_fir(2, taps, 1234.5, 48000.0, x, 1024, chunk_out=512)
    a, pa = lay.freq_xlating_fir(2, taps, 1234.5, 48000.0, x, 512)
    b, pb = lay.freq_xlating_fir(2, taps, 1234.5, 48000.0, x[1024:], 512, phase=pa)
    # (set_phase renormalises the phase it is handed, rotator.h:30: last-bit differences are allowed)
    np.testing.assert_allclose(np.concatenate([a, b]), one, atol=2e-6, rtol=0)
    assert abs(ph1 - pb) < 1e-6


def test_fft_filter_chunked_state(lay):
    # the OLA tail (d_tail) carries between filter() calls: one call == many calls
    rng = np.random.default_rng(5)
    x, taps = crand(rng, 7000), crand(rng, 100)
    ns = lay.fft_filter_nsamples(100)
    whole = lay.fft_filter_ccc(1, taps, x)
    parts = lay.fft_filter_ccc(1, taps, x, chunk_out=ns)
    np.testing.assert_array_equal(whole, parts)


def test_fft_golden_vectors(lay):
    # gr-fft/python/fft/qa_fft.py: 32-point golden vectors (Octave), abs_eps 1e-9 rel_eps 4e-4
    q = qa_vectors()
    primes = q["primes"]
    src = np.array([complex(primes[2 * i], primes[2 * i + 1]) for i in range(32)], np.complex64)
    assert_tuples_almost_equal2(q["primes_transformed"], lay.fft_vcc(32, 1, None, 0, src), 1e-9, 4e-4)
    rev_in = (q["primes_transformed"] / 32).astype(np.complex64)
    assert_tuples_almost_equal2(src, lay.fft_vcc(32, 0, None, 0, rev_in), 1e-9, 4e-4)
    ham = lay.window(0, 32)
    assert_tuples_almost_equal2(q["test_window"], lay.fft_vcc(32, 1, ham, 0, src), 1e-9, 4e-4)
    assert_tuples_almost_equal2(q["test_reverse_window_shift"], lay.fft_vcc(32, 0, ham, 1, rev_in), 1e-9, 4e-4)


def test_fft_vcc_fixtures_and_f64(lay):
    fx = fixtures()
    w = fx["win_bh_64"]
    for fwd in (0, 1):
        for sh in (0, 1):
            assert_parity(lay.fft_vcc(64, fwd, w, sh, fx["vcc_x"]), fx["vcc64_f%d_s%d_w" % (fwd, sh)], "w", 2e-6, 2e-5)
            assert_parity(lay.fft_vcc(64, fwd, None, sh, fx["vcc_x"]), fx["vcc64_f%d_s%d" % (fwd, sh)], "", 2e-6, 2e-5)
            assert_parity(fx["vcc64_f%d_s%d_w" % (fwd, sh)], po.f64_fft_vcc(64, fwd, w, sh, fx["vcc_x"]), "f64", 2e-6, 2e-5)
        assert_parity(lay.fft_vcc(15, fwd, None, 1, fx["vcc_x15"]), fx["vcc15_f%d_s1" % fwd], "odd", 2e-6, 2e-5)
    with pytest.raises(RuntimeError):
        lay.fft_vcc(64, 1, np.ones(63, np.float32), 0, fx["vcc_x"])  # fft_vcc_fftw.cc:46-47


def test_windows_and_firdes(lay):
    fx = fixtures()
    np.testing.assert_allclose(lay.window(0, 32), fx["win_hamming_32"], atol=2e-7)
    np.testing.assert_allclose(lay.window(5, 64), fx["win_bh_64"], atol=5e-7)
    np.testing.assert_allclose(lay.window(5, 4096), fx["win_bh_4096"], atol=5e-7)
    taps = lay.low_pass_2(1, 25000, 2500, 500, 80, 5)
    assert len(taps) == len(fx["lp2_qa_taps"])
    np.testing.assert_allclose(taps, fx["lp2_qa_taps"], atol=1e-7)


def test_mag_squared(lay):
    q, fx = qa_vectors(), fixtures()
    np.testing.assert_allclose(lay.mag_squared(q["magsq_src"].astype(np.complex64)), q["magsq_expected"], rtol=1e-6)
    np.testing.assert_allclose(lay.mag_squared(fx["vcc_x"]), fx["magsq"], rtol=1e-6)


def test_channelizer_fixtures_and_closed_form(lay):
    fx = fixtures()
    for tag in "abc":
        M, osr, nout, cons, T, nouts = fx["pfb_%s_meta" % tag]
        M, nout, cons, T, nouts = int(M), int(nout), int(cons), int(T), int(nouts)
        taps, ins, cmap = fx["pfb_%s_taps" % tag], fx["pfb_%s_in" % tag], fx["pfb_%s_map" % tag]
        assert lay.pfb_taps_per_filter(M, len(taps)) == T
        got, c = lay.pfb_channelizer(M, taps, float(osr), ins, nout, channel_map=cmap, nouts=nouts)
        assert c == cons
        assert_parity(got, fx["pfb_%s_out" % tag], "fixture " + tag, 2e-6, 2e-5)
        # float64 closed form, channel map applied afterwards
        hp = np.concatenate([taps, np.zeros(M * T - len(taps), np.float32)])
        x = ins[:, T:].T.reshape(-1)
        f64 = po.f64_channelizer(M, hp, x, nout, float(osr))[cmap[:nouts]]
        assert_parity(got, f64, "f64 " + tag, 2e-6, 2e-5)
    with pytest.raises(ValueError):
        lay.pfb_channelizer(36, fx["pfb_a_taps"], 10.1334, np.zeros((36, 64), np.complex64), 4)  # :49-53


def test_channelizer_qa_tones(lay):
    # qa_pfb_channelizer.py:25-153: 5 tones, low_pass_2 taps, last 50 samples normalised, 3 places
    freqs, fs, N = [110., -513., 203., -230, 121], 5000, 1000
    M = len(freqs)
    ifs = M * fs
    taps = lay.low_pass_2(1, ifs, fs / 2, fs / 10, 80, 5)
    x = np.sum([sig_source_c(ifs, f + i * fs, 1, N) for i, f in enumerate(freqs)], axis=0).astype(np.complex64)
    T = lay.pfb_taps_per_filter(M, len(taps))
    rows = N // M
    ins = np.concatenate([np.zeros((M, T), np.complex64), x[: rows * M].reshape(rows, M).T], axis=1)
    got, _ = lay.pfb_channelizer(M, taps, 1.0, ins, rows)
    L = rows
    delay = int(-(np.ceil(len(taps) / float(M)) - 1.0) / 2.0)
    t = np.arange(delay, L + delay) / fs
    for i, f in enumerate(freqs):
        exp = np.exp(2j * np.pi * f * t)[-50:]
        rec = got[i][-50:]
        np.testing.assert_allclose(rec / rec[0], exp / exp[0], atol=0.5e-3 * 2)


def test_pfb_decimator_fixtures_and_closed_form(lay):
    fx = fixtures()
    taps, x = fx["lp2_qa_taps"], fx["pfbdec_x"]
    for ch in (0, 3):
        f64 = po.f64_pfb_decimator(5, taps, ch, x)
        for mode in range(4 if lay.kind == "reference" else 1):
            got = lay.pfb_decimator(5, taps, ch, x, mode)
            assert_parity(got, fx["pfbdec_c%d_m%d" % (ch, mode)], "fixture ch %d mode %d" % (ch, mode), 1e-6, 1e-5)
            assert_parity(got, f64[: len(got)], "f64 ch %d mode %d" % (ch, mode), 1e-6, 1e-5)
        # the four variants of the reference agree with each other: one device kernel may stand for all of them
        for mode in range(1, 4):
            a, b = fx["pfbdec_c%d_m0" % ch], fx["pfbdec_c%d_m%d" % (ch, mode)]
            assert_parity(b, a[: len(b)], "variant %d vs work_fir_exp" % mode, 1e-6, 1e-5)
    got = lay.pfb_decimator(8, fx["pfbdec_b_taps"], 5, fx["pfbdec_b_x"])
    assert_parity(got, fx["pfbdec_b_c5"], "ragged taps fixture", 1e-6, 1e-5)
    assert_parity(got, po.f64_pfb_decimator(8, fx["pfbdec_b_taps"], 5, fx["pfbdec_b_x"]), "ragged taps f64", 1e-6, 1e-5)


def pfb_decimator_qa_case(channel):
    """gr-filter/python/filter/qa_pfb_decimator.py:22-76 (run_test): five tones, one per channel, summed at
    ifs = M*fs and dealt round-robin onto M streams; returns (streams, taps_args, expected tone of the channel)."""
    N, M, fs = 1000, 5, 5000.0
    ifs = M * fs
    freqs = [-230., 121., 110., -513., 203.]
    Mch = ((len(freqs) - 1) // 2 + channel) % len(freqs)
    x = np.zeros(N, np.complex128)
    for i in range(M):
        f = freqs[i] + (M // 2 - M + i + 1) * fs
        x += np.array(sig_source_c(ifs, f, 1, N))
    streams = x.astype(np.complex64).reshape(N // M, M).T.copy()  # stream_to_streams
    phase = [0.11058476216852586, 4.5108246571401693, 3.9739891674564594, 2.2820531095511924,
             1.3782797467397869][channel]
    return streams, freqs[Mch], phase, fs


def pfb_decimator_qa_expected(ntaps, L, f, phase, fs, M=5):
    tpf = int(np.ceil(ntaps / float(M)))
    delay = int(-(tpf - 1.0) / 2.0)
    t = np.arange(delay, L + delay) / fs
    return np.cos(2. * np.pi * f * t + phase) + 1j * np.sin(2. * np.pi * f * t + phase)


@pytest.mark.parametrize("channel", [0, 1, 2, 3, 4])
def test_pfb_decimator_qa_tones(lay, channel):
    # qa_pfb_decimator.py:78-141 (test_000..004): last 50 outputs against the channel's tone, 4 places, all variants
    taps = fixtures()["lp2_qa_taps"]  # low_pass_2(1, 25000, 2500, 500, 80, BLACKMAN_hARRIS), qa_pfb_decimator.py:29-31
    streams, f, phase, fs = pfb_decimator_qa_case(channel)
    for mode in range(4 if lay.kind == "reference" else 1):
        got = lay.pfb_decimator(5, taps, channel, streams, mode)
        exp = pfb_decimator_qa_expected(len(taps), len(got), f, phase, fs)
        np.testing.assert_allclose(got[-50:], exp[-50:], atol=0.5e-4, rtol=0)


INTERP_QA_TAPS = [1, 10, 100, 1000, 10000]  # gr-filter/python/filter/qa_interp_fir_filter.py:25-35 (test_fff)
INTERP_QA_SRC = (0, 2, 3, 5, 7, 11, 13, 17)
INTERP_QA_EXPECTED = (0, 0, 0, 2, 20, 200, 2003, 20030, 300, 3005, 30050, 500, 5007, 50070, 700, 7011, 70110,
                      1100, 11013, 110130, 1300, 13017, 130170)


def test_interp_fir_qa_golden_and_fixtures(lay):
    # the QA's only case is the fff instantiation; the same integers through the complex-stream blocks must come out
    # exactly (every partial sum is an integer below 2^24), on the real and on the imaginary rail
    src = np.array(INTERP_QA_SRC, np.float32)
    exp = np.array(INTERP_QA_EXPECTED, np.float32)
    for taps in (np.array(INTERP_QA_TAPS, np.float32), np.array(INTERP_QA_TAPS, np.complex64)):
        got = lay.interp_fir(3, taps, (src + 1j * 2 * src).astype(np.complex64))
        assert np.array_equal(got[: len(exp)].real, exp) and np.array_equal(got[: len(exp)].imag, 2 * exp)
    fx = fixtures()
    for key, L, tk in (("interp_ccf_l4", 4, "interp_taps_f"), ("interp_ccc_l3", 3, "interp_taps_c")):
        got = lay.interp_fir(L, fx[tk], fx["interp_x"])
        assert_parity(got, fx[key], key, 1e-6, 1e-5)
        assert_parity(got, po.f64_interp_fir(L, fx[tk], fx["interp_x"]), key + " f64", 1e-6, 1e-5)


# gr-filter/python/filter/qa_rational_resampler.py:71-85 (test_000_1_to_1) and :87-109 (test_001_interp), fff
# instantiation; the block shadows history(), so the first output is the first full window (no zeros in front)
RESAMP_QA0_TAPS, RESAMP_QA0_SRC = (-4, 5), (234, -4, 23, -56, 45, 98, -23, -7)
RESAMP_QA0_EXPECTED = (1186, -112, 339, -460, -167, 582)
RESAMP_QA1_EXPECTED = (2, 20, 200, 2003, 20030, 300, 3005, 30050, 500, 5007, 50070, 700, 7011, 70110, 1100, 11013, 110130,
                       1300, 13017, 130170, 1700.0, 17000.0, 170000.0, 0.0)


def test_rational_resampler_qa_golden_and_fixtures(lay):
    # the same integers through the complex-stream blocks come out exactly (partial sums are integers below 2^24)
    for cplx in (False, True):
        dt = np.complex64 if cplx else np.float32
        src = np.array(RESAMP_QA0_SRC, np.float32)
        got, used, ctr = lay.rational_resampler(1, 1, np.array(RESAMP_QA0_TAPS, dt), (src + 2j * src).astype(np.complex64),
                                                noutput=6)
        exp = np.array(RESAMP_QA0_EXPECTED, np.float32)
        assert np.array_equal(got.real, exp) and np.array_equal(got.imag, 2 * exp) and (used, ctr) == (6, 0)
        # test_001: 24 outputs from 8 inputs -- the last three windows read one item past the stream, which the QA's
        # freshly zeroed buffer supplies (general_work tests count < ninput only, :160)
        src = np.array(INTERP_QA_SRC + (0,), np.float32)
        got, used, ctr = lay.rational_resampler(3, 1, np.array(INTERP_QA_TAPS, dt), (src + 2j * src).astype(np.complex64),
                                                noutput=24, nin=8)
        exp = np.array(RESAMP_QA1_EXPECTED, np.float32)
        assert np.array_equal(got.real, exp) and np.array_equal(got.imag, 2 * exp) and (used, ctr) == (8, 0)
    fx = fixtures()
    for key, L, D, tk in (("resamp_ccf_4_7", 4, 7, "resamp_taps_f"), ("resamp_ccc_3_2", 3, 2, "resamp_taps_c")):
        got, used, ctr = lay.rational_resampler(L, D, fx[tk], fx["resamp_x"])
        assert_parity(got, fx[key], key, 1e-6, 1e-5)
        assert_parity(got, po.f64_rational_resampler(L, D, fx[tk], fx["resamp_x"], len(got)), key + " f64", 1e-6, 1e-5)
        T = -(-len(fx[tk]) // L)
        nin = len(fx["resamp_x"]) - (T - 1)
        assert len(got) == -(-nin * L // D) and used == len(got) * D // L and ctr == len(got) * D % L


def test_rational_resampler_split_calls_carry_ctr(lay):
    # d_ctr and the consumed count carry a stream across general_work() calls (:156,171-172)
    rng = np.random.default_rng(11)
    x = crand(rng, 1200)
    taps = rng.uniform(-1, 1, 23).astype(np.float32)
    whole, _, _ = lay.rational_resampler(5, 3, taps, x)
    parts, pos, ctr = [], 0, 0
    for nout in (1, 7, 64, 333, 10 ** 6):
        got, used, ctr = lay.rational_resampler(5, 3, taps, x[pos:], noutput=nout, ctr=ctr)
        parts.append(got)
        pos += used
    assert_parity(np.concatenate(parts), whole, "split", 1e-6, 1e-5)


def test_port_matches_reference_layer_on_fresh_data():
    if not po.have_ref():
        pytest.skip("oracle/_ref not built here")
    R, P = po.Ref(), po.Port()
    rng = np.random.default_rng(99)
    x = crand(rng, 9000)
    for nt, D in ((1, 1), (7, 2), (64, 1), (1024, 8)):
        t = rng.uniform(-1, 1, nt).astype(np.float32)
        assert_parity(P.fir_stream(D, t, x), R.fir_stream(D, t, x), "fir %d/%d" % (nt, D), 1e-6, 1e-5)
    for N in (2, 4, 8, 16, 64, 100, 4096):
        v = crand(rng, N * 2)
        for fwd in (0, 1):
            assert_parity(P.fft_vcc(N, fwd, None, 1, v), R.fft_vcc(N, fwd, None, 1, v), "fft %d" % N, 2e-6, 2e-5)
