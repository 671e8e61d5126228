#!/usr/bin/env python
"""Regenerates the committed fixtures of tests/golden/ -- run in the BUILD container only (needs /root/reference).

  qa_vectors.json    golden vectors lifted (by ast, nothing retyped) from the reference's own QA files:
                     gr-fft/python/fft/qa_fft.py (32-point primes FFT, Hamming-windowed, reverse+window+shift),
                     gr-blocks/python/blocks/qa_type_conversions.py:129-137 (complex_to_mag_squared).
  ref_fixtures.npz   outputs of oracle/_ref/libgrref.so -- the reference's OWN fir_filter.cc, fft_filter.cc,
                     polyphase_filterbank.cc, firdes.cc, window.cc, fir_filter_with_buffer.cc, filterbank.cc (and the
                     header-only single_pole_iir.h) compiled unmodified -- on seeded inputs.
                     Inputs are stored next to the outputs so the fixtures stand alone on the GPU box.

    python tests/golden/make_golden.py
"""
import ast
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = "/root/reference"


def _c(v):
    return [complex(x).real for x in v], [complex(x).imag for x in v]


def lift_qa_vectors():
    out = {}
    tree = ast.parse(open(os.path.join(REF, "gr-fft/python/fft/qa_fft.py")).read())
    for node in ast.walk(tree):
        if isinstance(node, ast.Assign) and isinstance(node.targets[0], ast.Name):
            name = node.targets[0].id
            if name in ("primes", "primes_transformed"):
                out[name] = ast.literal_eval(node.value)
        if isinstance(node, ast.FunctionDef) and node.name in ("test_window", "test_reverse_window_shift"):
            for sub in ast.walk(node):
                if isinstance(sub, ast.Assign) and getattr(sub.targets[0], "id", "") == "expected_result":
                    out[node.name] = ast.literal_eval(sub.value)
    res = {"primes": list(out["primes"])}
    for k in ("primes_transformed", "test_window", "test_reverse_window_shift"):
        re, im = _c(out[k])
        res[k] = {"re": re, "im": im}
    tree = ast.parse(open(os.path.join(REF, "gr-blocks/python/blocks/qa_type_conversions.py")).read())
    for node in ast.walk(tree):
        if isinstance(node, ast.FunctionDef) and node.name == "test_complex_to_mag_squared":
            for sub in ast.walk(node):
                if isinstance(sub, ast.Assign) and getattr(sub.targets[0], "id", "") in ("src_data", "expected_data"):
                    v = ast.literal_eval(sub.value)
                    if sub.targets[0].id == "src_data":
                        re, im = _c(v)
                        res["magsq_src"] = {"re": re, "im": im}
                    else:
                        res["magsq_expected"] = [float(x) for x in v]
    return res


def crand(rng, n):
    return (rng.uniform(-1, 1, n) + 1j * rng.uniform(-1, 1, n)).astype(np.complex64)


def ref_fixtures():
    from oracle import pyoracle as po

    po.build()
    R = po.Ref()
    rng = np.random.default_rng(20261104)
    fx = {}
    # ---- fir_filter_ccf / ccc ----
    x = crand(rng, 3000)
    fx["fir_x"] = x
    fx["fir_taps_f"] = rng.uniform(-1, 1, 64).astype(np.float32)
    fx["fir_taps_c"] = crand(rng, 37)
    for D in (1, 4):
        fx["fir_ccf_d%d" % D] = R.fir_stream(D, fx["fir_taps_f"], x)
        fx["fir_ccc_d%d" % D] = R.fir_stream(D, fx["fir_taps_c"], x)
    # ---- fft_filter_ccc ----
    fx["fftfilt_x"] = crand(rng, 5000)
    fx["fftfilt_taps"] = crand(rng, 300)
    for D in (1, 3):
        fx["fftfilt_d%d" % D] = R.fft_filter_ccc(D, fx["fftfilt_taps"], fx["fftfilt_x"])
    fx["fftfilt_nsamples_300"] = np.int64(R.fft_filter_nsamples(300))
    # ---- firdes / window (host-side tap and window generators feeding the path) ----
    fx["lp2_qa_taps"] = R.low_pass_2(1, 25000, 2500, 500, 80, 5)  # qa_pfb_channelizer.py:37-40
    fx["win_hamming_32"] = R.window(0, 32)
    fx["win_bh_64"] = R.window(5, 64)
    fx["win_bh_4096"] = R.window(5, 4096)
    # ---- pfb_channelizer_ccf ----
    for tag, M, nt, osr in (("a", 5, len(fx["lp2_qa_taps"]), 1.0), ("b", 8, 50, 2.0), ("c", 6, 31, 1.5)):
        taps = fx["lp2_qa_taps"] if tag == "a" else rng.uniform(-1, 1, nt).astype(np.float32)
        T = R.pfb_taps_per_filter(M, len(taps))
        Rr = int(round(M / osr))
        om = 1
        while (om * Rr) % M:
            om += 1
        nout = om * 40
        cons = nout * Rr // M
        ins = np.concatenate([np.zeros((M, T), np.complex64), crand(rng, M * cons).reshape(M, cons)], axis=1)
        cmap = np.arange(M, dtype=np.int32) if tag == "a" else rng.permutation(M).astype(np.int32)
        nouts = M if tag != "c" else 4
        y, c = R.pfb_channelizer(M, taps, osr, ins, nout, channel_map=cmap, nouts=nouts)
        assert c == cons
        fx["pfb_%s_taps" % tag] = taps
        fx["pfb_%s_in" % tag] = ins
        fx["pfb_%s_map" % tag] = cmap
        fx["pfb_%s_out" % tag] = y
        fx["pfb_%s_meta" % tag] = np.array([M, osr, nout, cons, T, nouts], np.float64)
    # ---- fft_vcc ----
    fx["vcc_x"] = crand(rng, 64 * 6)
    w = fx["win_bh_64"]
    for fwd in (0, 1):
        for sh in (0, 1):
            fx["vcc64_f%d_s%d_w" % (fwd, sh)] = R.fft_vcc(64, fwd, w, sh, fx["vcc_x"])
            fx["vcc64_f%d_s%d" % (fwd, sh)] = R.fft_vcc(64, fwd, None, sh, fx["vcc_x"])
    fx["vcc_x15"] = crand(rng, 15 * 3)  # odd length: ceil/floor halves of the shift differ (fft_vcc_fftw.cc:79,107)
    for fwd in (0, 1):
        fx["vcc15_f%d_s1" % fwd] = R.fft_vcc(15, fwd, None, 1, fx["vcc_x15"])
    fx["magsq"] = R.mag_squared(fx["vcc_x"])
    # ---- freq_xlating_fir_filter_ccf / ccc (own generator: the arrays above stay bit-identical) ----
    rng2 = np.random.default_rng(20261105)
    fx["xlat_x"] = crand(rng2, 4000)
    fx["xlat_taps_f"] = rng2.uniform(-1, 1, 48).astype(np.float32)
    fx["xlat_taps_c"] = crand(rng2, 33)
    fx["xlat_meta"] = np.array([0.3, 1.0, 1100], np.float64)  # center_freq, sampling_freq, scheduler-like chunk
    for D in (1, 4):
        for tag in ("f", "c"):
            t = fx["xlat_taps_" + tag]
            xin = np.concatenate([np.zeros(len(t) - 1, np.complex64), fx["xlat_x"]])
            y, ph = R.freq_xlating_fir(D, t, 0.3, 1.0, xin, len(fx["xlat_x"]) // D, chunk_out=1100 // D)
            fx["xlat_cc%s_d%d" % (tag, D)] = y
            fx["xlat_cc%s_d%d_phase" % (tag, D)] = np.complex64(ph)
    # ---- pfb_decimator_ccf: the four work() variants of the reference on one input ----
    rng3 = np.random.default_rng(20261106)
    fx["pfbdec_x"] = crand(rng3, 5 * 920).reshape(5, 920)  # 920 = 4 output multiples of the FFT-filter variants
    for ch in (0, 3):
        for mode in range(4):
            fx["pfbdec_c%d_m%d" % (ch, mode)] = R.pfb_decimator(5, fx["lp2_qa_taps"], ch, fx["pfbdec_x"], mode)
    fx["pfbdec_b_x"] = crand(rng3, 8 * 500).reshape(8, 500)
    fx["pfbdec_b_taps"] = rng3.uniform(-1, 1, 50).astype(np.float32)  # 50 taps over 8 branches: ragged last column
    fx["pfbdec_b_c5"] = R.pfb_decimator(8, fx["pfbdec_b_taps"], 5, fx["pfbdec_b_x"], 0)
    # ---- interp_fir_filter_ccf / ccc ----
    rng4 = np.random.default_rng(20261107)
    fx["interp_x"] = crand(rng4, 1500)
    fx["interp_taps_f"] = rng4.uniform(-1, 1, 50).astype(np.float32)  # 50 taps over 4 phases: zero-padded to 52
    fx["interp_taps_c"] = crand(rng4, 33)
    fx["interp_ccf_l4"] = R.interp_fir(4, fx["interp_taps_f"], fx["interp_x"])
    fx["interp_ccc_l3"] = R.interp_fir(3, fx["interp_taps_c"], fx["interp_x"])
    # ---- rational_resampler_base_ccf / ccc (one general_work() over the whole stream, no history in front) ----
    rng5 = np.random.default_rng(20261108)
    fx["resamp_x"] = crand(rng5, 3000)
    fx["resamp_taps_f"] = rng5.uniform(-1, 1, 50).astype(np.float32)  # 50 taps over 4 phases: zero-padded to 52
    fx["resamp_taps_c"] = crand(rng5, 33)
    fx["resamp_ccf_4_7"] = R.rational_resampler(4, 7, fx["resamp_taps_f"], fx["resamp_x"])[0]
    fx["resamp_ccc_3_2"] = R.rational_resampler(3, 2, fx["resamp_taps_c"], fx["resamp_x"])[0]
    # ---- round 2: fir_filter_fff, fft_vfc, logpwrfft, pfb_synthesizer_ccf, filterbank_vcvcf (own generator again) ----
    rng6 = np.random.default_rng(20261109)
    fx["fff_x"] = rng6.uniform(-1, 1, 4000).astype(np.float32)
    fx["fff_taps"] = rng6.uniform(-1, 1, 101).astype(np.float32)
    for D in (1, 3):
        xin = np.concatenate([np.zeros(100, np.float32), fx["fff_x"]])
        fx["fff_d%d" % D] = R.fir_fff(D, fx["fff_taps"], xin, len(fx["fff_x"]) // D)
    fx["vfc_x"] = rng6.uniform(-1, 1, 64 * 5).astype(np.float32)
    fx["vfc64_f1_w"] = R.fft_vfc(64, True, fx["win_bh_64"], fx["vfc_x"])
    fx["vfc64_f0"] = R.fft_vfc(64, False, None, fx["vfc_x"])
    fx["logpwr_xc"] = crand(rng6, 64 * 37)
    fx["logpwr_xf"] = rng6.uniform(-1, 1, 64 * 37).astype(np.float32)
    fx["logpwr_c_d3_avg"] = R.logpwrfft(True, 64, 2.0, 3, 0.25, True, fx["win_bh_64"], fx["logpwr_xc"])[0]
    fx["logpwr_c_d1"] = R.logpwrfft(True, 64, 2.0, 1, 1.0, False, fx["win_bh_64"], fx["logpwr_xc"])[0]
    fx["logpwr_f_d2_avg"] = R.logpwrfft(False, 64, 1.0, 2, 0.1, True, fx["win_bh_64"], fx["logpwr_xf"])[0]
    fx["synth_taps"] = R.low_pass_2(5, 5000, 500, 100, 80, 5)  # qa_pfb_synthesizer.py:39-41
    fx["synth_in5"] = crand(rng6, 5 * 400).reshape(5, 400)
    fx["synth_5_1x"] = R.pfb_synthesizer(5, fx["synth_taps"], False, fx["synth_in5"])
    fx["synth_map"] = np.array([3, 0, 4], np.int32)
    fx["synth_5_1x_map3"] = R.pfb_synthesizer(5, fx["synth_taps"], False, fx["synth_in5"][:3], channel_map=fx["synth_map"])
    fx["synth_in8"] = crand(rng6, 8 * 300).reshape(8, 300)
    fx["synth_taps8"] = rng6.uniform(-1, 1, 8 * 7 - 3).astype(np.float32)
    fx["synth_8_2x"] = R.pfb_synthesizer(8, fx["synth_taps8"], True, fx["synth_in8"])
    fx["fbank_taps"] = rng6.uniform(-1, 1, (10, 5)).astype(np.float32)
    fx["fbank_taps"][3] = 0
    fx["fbank_taps"][7] = 0
    fx["fbank_x"] = crand(rng6, 10 * (200 + 5))
    fx["fbank_out"] = R.filterbank(fx["fbank_taps"], fx["fbank_x"], 200)
    return fx


if __name__ == "__main__":
    json.dump(lift_qa_vectors(), open(os.path.join(HERE, "qa_vectors.json"), "w"), indent=0)
    fx = ref_fixtures()
    path = os.path.join(HERE, "ref_fixtures.npz")
    if os.path.exists(path) and "--all" not in sys.argv:
        # arrays already committed are kept as they are (the reference's auto-vectorised float loops may differ in
        # the last bit between builds); only new keys are added.  --all regenerates everything.
        old = dict(np.load(path))
        fx = {k: old.get(k, v) for k, v in fx.items()}
    np.savez_compressed(os.path.join(HERE, "ref_fixtures.npz"), **fx)
    print("wrote", len(fx), "arrays,", os.path.getsize(os.path.join(HERE, "ref_fixtures.npz")), "bytes")
