#!/usr/bin/env python
"""bench.py -- complex Msps of the gr-filter / gr-fft hot path on B200, with roofline and CPU baseline.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl reference]

One "step" = one pass of the hot path over one batch of synthetic complex64 input that is already resident
in HBM (`value`), launched through the C-ABI device entry points on torch's current stream and timed with CUDA
events.  `e2e` is the same metric through the host-buffer C-ABI call (pinned host buffers, H2D + kernel + D2H
inside the timed region).  Headline workload (BASELINE.json metric "fir_filter_ccf/pfb_channelizer_ccf at 1-8
B200"): configs[2], pfb_channelizer_ccf 64 channels x 16 taps/branch over 2^30 samples per GPU; the other four
configs are measured at N=1 and attached under "other_configs".  Multi-GPU: one process per GPU (torchrun),
time-sharded stream, weak scaling; the only exchange is the taps_per_filter*numchans-sample history halo sent
to the right neighbour over NCCL.

`--impl reference` times the reference's own CPU kernels (oracle/_ref/libgrref.so = the reference sources
compiled unmodified; the C port when that file is absent) on the host cores, for the same workload.
"""
import argparse
import concurrent.futures as cf
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

HBM_FALLBACK_GBS = 6650.0  # /opt/skills/guides/B200_PROFILING.md fallback


def fp32_peak():
    """FP32 (CUDA-core, packed FFMA2) peak in TFLOP/s measured on this pool by tools/ubench.cu (profiles/r1_ubench.json)."""
    try:
        d = json.load(open(os.path.join(ROOT, "profiles", "r1_ubench.json")))
        return 2.0 * float(d["ffma2_tfma_s"]), "measured (tools/ubench.cu packed-FFMA2 loop, profiles/r1_ubench.json)"
    except Exception:
        return 66.6, "measured in round 1 (tools/ubench.cu): 66.6 TFLOP/s"


def roofline_entry(wl, kernel_ms, samples_per_launch, hbm_peak, hbm_src):
    """The roofline object of one kernel launch: achieved = algorithmic bytes (or flops) per launch / mean launch time."""
    gbs = wl.alg_bytes * samples_per_launch / (kernel_ms * 1e-3) / 1e9
    tfl = wl.flops * samples_per_launch / (kernel_ms * 1e-3) / 1e12
    fpk, fsrc = fp32_peak()
    r = {"kernel": wl.kernel, "alg_bytes_per_sample": wl.alg_bytes, "alg_flops_per_sample": wl.flops,
         "hbm_gbs": gbs, "hbm_frac": gbs / hbm_peak, "fp32_tflops": tfl, "fp32_frac": tfl / fpk}
    if getattr(wl, "bound", "hbm") == "fp32":
        r.update({"bound": "fp32", "achieved": tfl, "peak": fpk, "unit": "TFLOP/s", "frac": tfl / fpk, "peak_source": fsrc})
    else:
        r.update({"bound": "hbm", "achieved": gbs, "peak": hbm_peak, "unit": "GB/s", "frac": gbs / hbm_peak, "peak_source": hbm_src})
    return r


def measured_peaks():
    try:
        d = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, burst copy)"
    except Exception:
        return HBM_FALLBACK_GBS, "fallback (B200_PROFILING.md)"


# ---------------------------------------------------------------------------------------------------------
# synthetic taps / windows (host side, float64 formulas of gr-filter/lib/firdes.cc:77-119 and gr-fft/lib/window.cc:94-104)
# ---------------------------------------------------------------------------------------------------------
def coswindow(n, c0, c1, c2, c3=0.0):
    M = n - 1
    k = np.arange(n)
    return (c0 - c1 * np.cos(2 * np.pi * k / M) + c2 * np.cos(4 * np.pi * k / M) - c3 * np.cos(6 * np.pi * k / M))


def blackman_harris(n):
    return coswindow(n, 0.35875, 0.48829, 0.14128, 0.01168).astype(np.float32)


def hamming(n):
    return coswindow(n, 0.54, 0.46, 0.0).astype(np.float32)


def lowpass(ntaps, cutoff, window):
    n = np.arange(ntaps) - (ntaps - 1) / 2.0
    h = 2 * cutoff * np.sinc(2 * cutoff * n) * window.astype(np.float64)
    return (h / h.sum()).astype(np.float32)


# ---------------------------------------------------------------------------------------------------------
# workloads: each builds its device state lazily and exposes step() (device-resident) and, for the headline,
# e2e_step() (host buffers).  alg_bytes = SURVEY.md section 8(d) algorithmic bytes per input sample.
# ---------------------------------------------------------------------------------------------------------
class Workload:
    name = ""
    block = ""
    alg_bytes = 16.0
    flops = 0.0
    log2n = 30
    kernel = ""
    launches_per_step = 1
    seed = 0

    def __init__(self, torch, grb, device, log2n=None):
        self.torch, self.grb, self.device = torch, grb, device
        if log2n is not None:
            self.log2n = log2n
        self.n = 1 << self.log2n
        self.stream = torch.cuda.current_stream().cuda_stream

    def rand_c64(self, nitems, seed):
        g = self.torch.Generator(device="cuda")
        g.manual_seed(seed)
        x = self.torch.empty(2 * nitems, dtype=self.torch.float32, device="cuda")
        x.uniform_(-1.0, 1.0, generator=g)
        return x

    def describe(self):
        return self.name

    # -- parity OUTSIDE the timed region: first/last segment + 64 seeded interior windows of the last step's output against
    #    the CPU oracle and float64 (oracle/spotcheck.py -- the checker, never the thing measured) -------------------------
    def fetch_c64(self, t):
        return lambda a, b: t[2 * a: 2 * b].cpu().numpy().view(np.complex64)

    def parity(self, sc):
        return None

    def free(self):
        for k in list(self.__dict__):
            if k.startswith("d_"):
                delattr(self, k)
        self.torch.cuda.empty_cache()


class PfbChannelizer(Workload):
    """configs[2]: pfb_channelizer_ccf 64 channels x 16 taps/branch, 2^30 samples (per GPU)."""
    name = "pfb_channelizer_ccf 64ch x 16 taps/branch, 2^%d complex64 samples per GPU (BASELINE configs[2])"
    block = "pfb_channelizer_ccf"
    alg_bytes, flops, log2n, seed = 16.0, 94.0, 30, 3
    kernel = "pfb_stream_kernel"
    M, T = 64, 16
    # profiles/r1_pfb_stream_kernel_ncu.txt: 2.157 GB read + 2.103 GB written for a 2^28-sample launch
    ncu_dram_bytes_per_sample = (2.157e9 + 2.103e9) / (1 << 28)
    ncu_source = "profiles/r1_pfb_stream_kernel_ncu.txt (2^28-sample launch, scaled)"

    def setup(self):
        M, T = self.M, self.T
        self.taps = lowpass(M * T, 0.5 / M, blackman_harris(M * T)) * np.float32(M)
        self.blk = self.grb.filter.pfb_channelizer_ccf(M, self.taps, 1.0, device=self.device)
        self.nout = self.n // M
        self.d_in = self.rand_c64(self.n + T * M, self.seed)  # history rows in front
        self.d_out = self.torch.empty(2 * self.n, dtype=self.torch.float32, device="cuda")
        self.halo_items = T * M

    def step(self):
        self.blk.work_interleaved_dev(self.d_in.data_ptr(), self.d_out.data_ptr(), self.nout, self.M, self.nout,
                                      self.stream)

    def describe(self):
        return self.name % self.log2n

    def parity(self, sc):
        M, nout = self.M, self.nout
        xv, yv = self.d_in.view(-1, 2 * M), self.d_out.view(M, 2 * nout)
        return sc.pfb_channelizer(lambda a, b: xv[a:b].cpu().numpy().view(np.complex64),
                                  lambda a, b: np.ascontiguousarray(yv[:, 2 * a: 2 * b].cpu().numpy()).view(np.complex64),
                                  M, self.taps, nout, width=64, count=64, seed=self.seed)

    # -- multi-GPU: the segment's history comes from the left neighbour's tail ---------------------------
    def halo_views(self):
        h = 2 * self.halo_items
        return [(self.d_in[:h], self.d_in[-h:])]

    # -- e2e through host buffers ---------------------------------------------------------------------
    def e2e_setup(self, log2n):
        t, M, T = self.torch, self.M, self.T
        self.e2e_n = 1 << log2n
        self.e2e_nout = self.e2e_n // M
        self.h_in = t.empty(2 * (self.e2e_n + T * M), dtype=t.float32).pin_memory()
        self.h_in.uniform_(-1, 1)
        self.h_out = t.empty((M, 2 * self.e2e_nout), dtype=t.float32).pin_memory()
        self.h_in_np = self.h_in.numpy().view(np.complex64)
        self.h_out_np = [self.h_out[j].numpy().view(np.complex64) for j in range(M)]
        self.blk_e2e = self.grb.filter.pfb_channelizer_ccf(M, self.taps, 1.0, device=self.device)
        return 8 * (self.e2e_n + T * M), 8 * self.e2e_n

    def e2e_step(self):
        self.blk_e2e.work_interleaved(self.h_in_np, self.h_out_np, self.e2e_nout)

    # -- CPU leg: one bounded slice of the same workload per thread --------------------------------------
    def cpu_setup(self, lay, log2n):
        rng = np.random.default_rng(self.seed)
        M, T = self.M, self.T
        n = 1 << log2n
        self.cpu_nout = n // M
        self.cpu_ins = (rng.uniform(-1, 1, (M, self.cpu_nout + T)) + 1j * rng.uniform(-1, 1, (M, self.cpu_nout + T))
                        ).astype(np.complex64)
        self.cpu_taps = lowpass(M * T, 0.5 / M, blackman_harris(M * T)) * np.float32(M)
        self.cpu_lay = lay
        return n

    def cpu_call(self):
        self.cpu_lay.pfb_channelizer(self.M, self.cpu_taps, 1.0, self.cpu_ins, self.cpu_nout)


class FirCcf(Workload):
    """configs[0]: fir_filter_ccf 64 taps, decim 1, 2^24 samples (4 rotating buffer sets so no step re-reads L2)."""
    name = "fir_filter_ccf 64 taps decim 1, 2^%d samples x %d rotating buffer sets (BASELINE configs[0])"
    block = "fir_filter_ccf"
    alg_bytes, flops, log2n, seed = 16.0, 256.0, 24, 1
    kernel = "fir_tc_kernel"
    ntaps, decim, nsets = 64, 1, 4
    # profiles/r1_fir_tc_ncu.txt: 134.2 MB read + 84.3 MB written per 2^24-sample launch (part of the output is still in L2)
    ncu_dram_bytes_per_sample = (134.2e6 + 84.3e6) / (1 << 24)
    ncu_source = "profiles/r1_fir_tc_ncu.txt"

    def setup(self):
        self.taps = lowpass(self.ntaps, 0.1, hamming(self.ntaps))
        # one block per buffer set: on several GPUs every set has its own halo source (the left neighbour's set)
        self.blks = [self.grb.filter.fir_filter_ccf(self.decim, self.taps, device=self.device) for _ in range(self.nsets)]
        self.blk = self.blks[0]
        self.nout = self.n // self.decim
        self.d_in = [self.rand_c64(self.n + self.ntaps - 1, self.seed + i) for i in range(self.nsets)]
        self.d_out = [self.torch.empty(2 * self.nout, dtype=self.torch.float32, device="cuda") for _ in range(self.nsets)]
        self.launches_per_step = self.nsets
        self.samples_per_step = self.n * self.nsets

    def step(self):
        for i in range(len(self.d_in)):
            self.blks[i % len(self.blks)].work_dev(self.d_in[i].data_ptr(), self.d_out[i].data_ptr(), self.nout, self.stream)

    def parity(self, sc):
        k = len(self.d_in) - 1  # the last buffer set the step wrote
        return sc.fir(self.fetch_c64(self.d_in[k]), self.fetch_c64(self.d_out[k]), self.decim, self.taps, self.nout,
                      width=512, count=64, seed=self.seed)

    def halo_views(self):  # ntaps-1 history items in front of every segment
        h = 2 * (self.ntaps - 1)
        self.halo_items = self.ntaps - 1
        return [(d[:h], d[-h:]) for d in self.d_in]

    def describe(self):
        return self.name % (self.log2n, self.nsets)

    def cpu_setup(self, lay, log2n):
        rng = np.random.default_rng(self.seed)
        n = 1 << log2n
        self.cpu_x = (rng.uniform(-1, 1, n + self.ntaps - 1) + 1j * rng.uniform(-1, 1, n + self.ntaps - 1)).astype(np.complex64)
        self.cpu_lay, self.cpu_nout = lay, n // self.decim
        self.cpu_taps = lowpass(self.ntaps, 0.1, hamming(self.ntaps))
        return n

    def cpu_call(self):
        self.cpu_lay.fir(self.decim, self.cpu_taps, self.cpu_x, self.cpu_nout)


class FirLongDecim(FirCcf):
    """configs[3]: fir_filter_ccf 1024 taps, decim 8 (2^27 input samples per GPU)."""
    name = "fir_filter_ccf 1024 taps decim 8, 2^%d input samples per GPU x %d set(s) (BASELINE configs[3])"
    alg_bytes, flops, log2n, seed = 9.0, 512.0, 27, 4
    kernel = "fir_tc_kernel<M_DEC8>"
    ntaps, decim, nsets = 1024, 8, 1
    # profiles/r1_firdec_tc_ncu.txt: 1.074 GB read + 128 MB written per 2^27-input-sample launch
    ncu_dram_bytes_per_sample = (1.0738e9 + 128.07e6) / (1 << 27)
    ncu_source = "profiles/r1_firdec_tc_ncu.txt"

    def setup(self):
        self.taps = lowpass(self.ntaps, 1.0 / 32, blackman_harris(self.ntaps))
        self.blk = self.grb.filter.fir_filter_ccf(self.decim, self.taps, device=self.device)
        self.blks = [self.blk]
        self.nout = self.n // self.decim
        self.d_in = [self.rand_c64(self.n + self.ntaps - 1, self.seed)]
        self.d_out = [self.torch.empty(2 * self.nout, dtype=self.torch.float32, device="cuda")]
        self.launches_per_step = 1
        self.samples_per_step = self.n

    def cpu_setup(self, lay, log2n):
        n = FirCcf.cpu_setup(self, lay, log2n)
        self.cpu_taps = lowpass(self.ntaps, 1.0 / 32, blackman_harris(self.ntaps))
        return n


class FftFilter(Workload):
    """configs[1]: fft_filter_ccc 4096-tap low-pass, 2^28 samples on 1 B200."""
    name = "fft_filter_ccc 4096 taps, 2^%d samples (BASELINE configs[1])"
    block = "fft_filter_ccc"
    # two tap partitions on the 4096-point kernel: (2 x 5 N log2 N + 16 N) / (N/2) flops per output sample, N = 4096
    alg_bytes, flops, log2n, seed = 16.0, 272.0, 28, 2
    kernel = "fftfilt_upols2_kernel"
    bound = "fp32"  # ncu: DRAM 23 % busy, FMA pipe 65 %, shared-memory pipe 64 % -- the butterflies bound it, not HBM
    ncu_dram_bytes_per_sample = (2.164e9 + 2.105e9) / (1 << 28)
    ncu_source = "profiles/r2_fftfilt_upols2_ncu.txt"
    ntaps = 4096

    def setup(self):
        self.taps = lowpass(self.ntaps, 0.05, blackman_harris(self.ntaps)).astype(np.complex64)
        self.blk = self.grb.filter.fft_filter_ccc(1, self.taps, device=self.device)
        ns = self.blk.nsamples()
        self.nout = (self.n // ns) * ns  # 65520 x 4097: the block only emits whole multiples of nsamples
        self.d_in = self.rand_c64(self.n + self.ntaps - 1, self.seed)
        self.d_out = self.torch.empty(2 * self.nout, dtype=self.torch.float32, device="cuda")
        self.samples_per_step = self.nout

    def step(self):
        self.blk.work_dev(self.d_in.data_ptr(), self.d_out.data_ptr(), self.nout, self.stream)

    def describe(self):
        return self.name % self.log2n

    def parity(self, sc):
        return sc.fft_filter(self.fetch_c64(self.d_in), self.fetch_c64(self.d_out), self.taps, self.nout, count=64,
                             seed=self.seed)

    def cpu_setup(self, lay, log2n):
        rng = np.random.default_rng(self.seed)
        n = 1 << log2n
        self.cpu_x = (rng.uniform(-1, 1, n) + 1j * rng.uniform(-1, 1, n)).astype(np.complex64)
        self.cpu_taps = lowpass(self.ntaps, 0.05, blackman_harris(self.ntaps)).astype(np.complex64)
        self.cpu_lay = lay
        ns = lay.fft_filter_nsamples(self.ntaps)
        return (n // ns) * ns

    def cpu_call(self):
        self.cpu_lay.fft_filter_ccc(1, self.cpu_taps, self.cpu_x)


class FftVccMagSq(Workload):
    """configs[4]: fft_vcc 4096-point Blackman-Harris (shift) + complex_to_mag_squared, 2^30 samples, fused."""
    name = "fft_vcc 4096 Blackman-Harris fwd shift + complex_to_mag_squared fused, 2^%d samples (BASELINE configs[4])"
    block = "fft_vcc+complex_to_mag_squared"
    alg_bytes, flops, log2n, seed = 12.0, 65.0, 30, 5
    kernel = "fft_vcc_stream_kernel"
    # profiles/r2_fftvcc_stream_ncu.txt: 8.590 GB read + 4.267 GB written per 2^30-sample launch (algorithmic 8.59 + 4.29)
    ncu_dram_bytes_per_sample = (8.590198e9 + 4.267297e9) / (1 << 30)
    ncu_source = "profiles/r2_fftvcc_stream_ncu.txt"
    N = 4096

    def setup(self):
        self.win = blackman_harris(self.N)
        self.blk = self.grb.fft.fft_vcc(self.N, True, self.win, True, device=self.device)
        self.nvec = self.n // self.N
        self.d_in = self.rand_c64(self.n, self.seed)
        self.d_out = self.torch.empty(self.n, dtype=self.torch.float32, device="cuda")

    def step(self):
        self.blk.work_dev(self.d_in.data_ptr(), self.d_out.data_ptr(), self.nvec, self.stream, magsq=True)

    def describe(self):
        return self.name % self.log2n

    def parity(self, sc):
        N = self.N
        xv, yv = self.d_in.view(self.nvec, 2 * N), self.d_out.view(self.nvec, N)
        return sc.fft_vcc(lambda a, b: xv[a:b].cpu().numpy().view(np.complex64), lambda a, b: yv[a:b].cpu().numpy(), N, True,
                          self.win, True, self.nvec, magsq=True, count=64, seed=self.seed)

    def cpu_setup(self, lay, log2n):
        rng = np.random.default_rng(self.seed)
        n = 1 << log2n
        self.cpu_x = (rng.uniform(-1, 1, n) + 1j * rng.uniform(-1, 1, n)).astype(np.complex64)
        self.cpu_lay, self.cpu_win = lay, blackman_harris(self.N)
        return n

    def cpu_call(self):
        self.cpu_lay.mag_squared(self.cpu_lay.fft_vcc(self.N, True, self.cpu_win, True, self.cpu_x))


class FreqXlating(FirCcf):
    """freq_xlating_fir_filter_ccf, 128-tap channel filter, decimate by 8 (not a BASELINE config)."""
    name = "freq_xlating_fir_filter_ccf 128 taps decim 8, 2^%d input samples per GPU x %d set(s) (section 8(f) row)"
    block = "freq_xlating_fir_filter_ccf"
    # one read per input item, one write per output item; 128 complex composite taps at decimation 8 take the overlap-save
    # route (4096-point blocks, 3969 new items each: a 4096-point forward transform, the spectrum product, the fold to 512 bins
    # and a 512-point inverse per block), then the rotator
    alg_bytes, flops, log2n, seed = 8.0 + 8.0 / 8, (5 * 4096 * 12 + 5 * 512 * 9 + 6 * 4096 + 2 * 4096) / 3969.0, 28, 6
    kernel = "fftfilt_ols3_kernel<DEC 8> + rotate_kernel"
    bound = "fp32"
    ntaps, decim, nsets = 128, 8, 1

    def setup(self):
        self.taps = lowpass(self.ntaps, 0.05, hamming(self.ntaps))
        self.blk = self.grb.filter.freq_xlating_fir_filter_ccf(self.decim, self.taps, 0.21, 1.0, device=self.device)
        self.blks = [self.blk]
        self.nout = self.n // self.decim
        self.d_in = [self.rand_c64(self.n + self.ntaps - 1, self.seed)]
        self.d_out = [self.torch.empty(2 * self.nout, dtype=self.torch.float32, device="cuda")]
        self.launches_per_step = 2
        self.samples_per_step = self.n

    def parity(self, sc):
        from oracle import pyoracle as po

        # a fresh block: the rotator phase of the timed block has run on over every step
        blk = self.grb.filter.freq_xlating_fir_filter_ccf(self.decim, self.taps, 0.21, 1.0, device=self.device)
        blk.work_dev(self.d_in[0].data_ptr(), self.d_out[0].data_ptr(), self.nout, self.stream)
        self.torch.cuda.synchronize()
        ctaps = blk.composite_taps()
        delta = po.xlat_increment_angle(np.float32(2 * np.pi * 0.21 / 1.0), self.decim)
        return sc.fir(self.fetch_c64(self.d_in[0]), self.fetch_c64(self.d_out[0]), self.decim, ctaps, self.nout, width=512,
                      count=64, seed=self.seed, rotator=(0.0, delta))

    def cpu_call(self):
        self.cpu_lay.freq_xlating_fir(self.decim, self.cpu_taps, 0.21, 1.0, self.cpu_x, self.cpu_nout)


class PfbDecimator(Workload):
    """pfb_decimator_ccf 64 branches x 16 taps (not a BASELINE config)."""
    name = "pfb_decimator_ccf 64 branches x 16 taps, 2^%d input samples per GPU (section 8(f) row)"
    block = "pfb_decimator_ccf"
    alg_bytes, flops, log2n, seed = 8.0 * (1 + 1.0 / 64), 72.0, 28, 7
    kernel = "pfbdec_tile_kernel"
    M, T = 64, 16
    ncu_dram_bytes_per_sample = (2.148e9 + 34.1e6) / (1 << 28)
    ncu_source = "profiles/r1_pfbdec_tile_ncu.txt"

    def setup(self):
        M, T = self.M, self.T
        self.taps = lowpass(M * T, 0.5 / M, blackman_harris(M * T))
        self.blk = self.grb.filter.pfb_decimator_ccf(M, self.taps, 5, False, False, device=self.device)
        self.nout = self.n // M
        self.stride = (self.nout + T - 1 + 1) & ~1
        self.d_in = self.rand_c64(M * self.stride, self.seed)
        self.d_out = self.torch.empty(2 * self.nout, dtype=self.torch.float32, device="cuda")
        self.samples_per_step = self.n

    def step(self):
        self.blk.work_dev(self.d_in.data_ptr(), self.stride, self.d_out.data_ptr(), self.nout, self.stream)

    def describe(self):
        return self.name % self.log2n

    def parity(self, sc):
        xv = self.d_in.view(self.M, 2 * self.stride)
        return sc.pfb_decimator(lambda s_, a, b: xv[s_, 2 * a: 2 * b].cpu().numpy().view(np.complex64),
                                self.fetch_c64(self.d_out), self.M, self.taps, 5, self.nout, width=128, count=64, seed=self.seed)

    def cpu_setup(self, lay, log2n):
        rng = np.random.default_rng(self.seed)
        n = 1 << log2n
        self.cpu_lay = lay
        self.cpu_x = (rng.uniform(-1, 1, (self.M, n // self.M)) + 1j * rng.uniform(-1, 1, (self.M, n // self.M))).astype(np.complex64)
        self.cpu_taps = lowpass(self.M * self.T, 0.5 / self.M, blackman_harris(self.M * self.T))
        return n

    def cpu_call(self):
        self.cpu_lay.pfb_decimator(self.M, self.cpu_taps, 5, self.cpu_x)


class FirFff(Workload):
    """fir_filter_fff, 256 taps -- the one block the reference holds published numbers for
    (gnuradio-runtime/examples/mp-sched/perf-data/*.dat:2, BASELINE.md section 1: 14.2 - 18.4 Msps on 2007-8 CPUs)."""
    name = "fir_filter_fff 256 taps decim 1, 2^%d float samples per GPU (section 8(f) row; mp-sched perf-data shape)"
    block = "fir_filter_fff"
    # 256 taps run on the overlap-save kernel (two halves of the call as re / im of one 4096-point complex block, 3841 new
    # items each): (2 x 5 N log2 N + 6 N) / (2 x 3841) flops per real sample, N = 4096 (the time-domain sum would be 512)
    alg_bytes, flops, log2n, seed = 8.0, (2 * 5 * 4096 * 12 + 6 * 4096) / (2.0 * 3841), 28, 8
    kernel = "fftfilt_ols3_kernel<RP>"
    bound = "hbm"
    ntaps = 256

    def setup(self):
        self.taps = lowpass(self.ntaps, 0.1, blackman_harris(self.ntaps))
        self.blk = self.grb.filter.fir_filter_fff(1, self.taps, device=self.device)
        g = self.torch.Generator(device="cuda")
        g.manual_seed(self.seed)
        self.d_in = self.torch.empty(self.n + self.ntaps - 1, dtype=self.torch.float32, device="cuda")
        self.d_in.uniform_(-1.0, 1.0, generator=g)
        self.d_out = self.torch.empty(self.n, dtype=self.torch.float32, device="cuda")

    def step(self):
        self.blk.work_dev(self.d_in.data_ptr(), self.d_out.data_ptr(), self.n, self.stream)

    def describe(self):
        return self.name % self.log2n

    def parity(self, sc):
        f = lambda t: (lambda a, b: t[a:b].cpu().numpy())
        return sc.fir_fff(f(self.d_in), f(self.d_out), 1, self.taps, self.n, width=1024, count=64, seed=self.seed)

    def cpu_setup(self, lay, log2n):
        rng = np.random.default_rng(self.seed)
        n = 1 << log2n
        self.cpu_x = rng.uniform(-1, 1, n + self.ntaps - 1).astype(np.float32)
        self.cpu_lay, self.cpu_taps = lay, lowpass(self.ntaps, 0.1, blackman_harris(self.ntaps))
        self.cpu_nout = n
        return n

    def cpu_call(self):
        self.cpu_lay.fir_fff(1, self.cpu_taps, self.cpu_x, self.cpu_nout)


class PfbSynthesizer(Workload):
    """pfb_synthesizer_ccf 64 channels x 16 taps/branch (the channelizer transposed)."""
    name = "pfb_synthesizer_ccf 64ch x 16 taps/branch, 2^%d output samples per GPU (section 8(f) row)"
    block = "pfb_synthesizer_ccf"
    alg_bytes, flops, log2n, seed = 16.0, 94.0, 28, 9
    kernel = "pfb_synth_kernel"
    M, T = 64, 16

    def setup(self):
        M, T = self.M, self.T
        self.taps = lowpass(M * T, 0.5 / M, blackman_harris(M * T)) * np.float32(M)
        self.blk = self.grb.filter.pfb_synthesizer_ccf(M, self.taps, False, device=self.device)
        self.ntimes = self.n // M
        self.stride = (self.ntimes + T - 1 + 1) & ~1
        self.d_in = self.rand_c64(M * self.stride, self.seed)
        self.d_out = self.torch.empty(2 * self.n, dtype=self.torch.float32, device="cuda")

    def step(self):
        self.blk.work_dev(self.d_in.data_ptr(), self.stride, self.M, self.d_out.data_ptr(), self.n, 0, self.stream)

    def describe(self):
        return self.name % self.log2n

    def parity(self, sc):
        xv, M = self.d_in.view(self.M, 2 * self.stride), self.M
        return sc.pfb_synthesizer(lambda a, b: np.ascontiguousarray(xv[:, 2 * a: 2 * b].cpu().numpy()).view(np.complex64),
                                  lambda a, b: self.d_out[2 * a * M: 2 * b * M].cpu().numpy().view(np.complex64),
                                  M, self.taps, self.ntimes, width=64, count=64, seed=self.seed)

    def cpu_setup(self, lay, log2n):
        rng = np.random.default_rng(self.seed)
        n = 1 << log2n
        self.cpu_lay = lay
        self.cpu_ins = (rng.uniform(-1, 1, (self.M, n // self.M)) + 1j * rng.uniform(-1, 1, (self.M, n // self.M))).astype(np.complex64)
        self.cpu_taps = lowpass(self.M * self.T, 0.5 / self.M, blackman_harris(self.M * self.T)) * np.float32(self.M)
        return n

    def cpu_call(self):
        self.cpu_lay.pfb_synthesizer(self.M, self.cpu_taps, False, self.cpu_ins)


class Filterbank(Workload):
    """filterbank_vcvcf 64 filters x 16 taps (the branch-FIR half of channelizer_hier_ccf)."""
    name = "filterbank_vcvcf 64 filters x 16 taps, 2^%d samples per GPU (section 8(f) row)"
    block = "filterbank_vcvcf"
    alg_bytes, flops, log2n, seed = 16.0, 64.0, 28, 10
    kernel = "filterbank_stream_kernel"
    M, T = 64, 16

    def setup(self):
        M, T = self.M, self.T
        rng = np.random.default_rng(self.seed)
        self.taps = rng.uniform(-1, 1, (M, T)).astype(np.float32)
        self.blk = self.grb.filter.filterbank_vcvcf(self.taps, device=self.device)
        self.nvec = self.n // M
        self.d_in = self.rand_c64(M * (self.nvec + T), self.seed)
        self.d_out = self.torch.empty(2 * self.n, dtype=self.torch.float32, device="cuda")

    def step(self):
        self.blk.work_dev(self.d_in.data_ptr(), self.d_out.data_ptr(), self.nvec, self.stream)

    def describe(self):
        return self.name % self.log2n

    def parity(self, sc):
        M = self.M
        fi = lambda a, b: self.d_in[2 * a * M: 2 * b * M].cpu().numpy().view(np.complex64)
        fo = lambda a, b: self.d_out[2 * a * M: 2 * b * M].cpu().numpy().view(np.complex64).reshape(-1, M)
        return sc.filterbank(fi, fo, self.taps, self.nvec, width=64, count=64, seed=self.seed)

    def cpu_setup(self, lay, log2n):
        rng = np.random.default_rng(self.seed)
        n = 1 << log2n
        self.cpu_lay = lay
        self.cpu_taps = rng.uniform(-1, 1, (self.M, self.T)).astype(np.float32)
        self.cpu_nvec = n // self.M
        self.cpu_x = (rng.uniform(-1, 1, n + self.M * self.T) + 1j * rng.uniform(-1, 1, n + self.M * self.T)).astype(np.complex64)
        return n

    def cpu_call(self):
        self.cpu_lay.filterbank(self.cpu_taps, self.cpu_x, self.cpu_nvec)


class LogPwrFft(Workload):
    """logpwrfft_c 4096 bins, every frame kept (decimation 1), averaging on -- the worst case
    for the averaging tail (normally one frame in tens is kept and the block's cost is the frames it reads)."""
    name = "logpwrfft_c 4096 Blackman-Harris, decimation 1, avg_alpha 0.1, 2^%d samples per GPU (section 8(f) row)"
    block = "logpwrfft_c"
    alg_bytes, flops, log2n, seed = 12.0, 70.0, 28, 11
    kernel = "fft_vcc_stream_kernel<mode 3: run-local IIR> + logpwr_carry/fixup"
    launches_per_step = 3
    N = 4096

    def setup(self):
        N = self.N
        self.blk = self.grb.fft.logpwrfft_c(float(N), N, 2.0, 1.0, 0.1, True, device=self.device)
        self.nvec = self.n // N
        self.d_in = self.rand_c64(self.n, self.seed)
        self.d_out = self.torch.empty(self.n, dtype=self.torch.float32, device="cuda")

    def step(self):
        self.blk.work_dev(self.d_in.data_ptr(), self.n, self.d_out.data_ptr(), self.nvec, self.stream)

    def describe(self):
        return self.name % self.log2n


class Resampler(Workload):
    """rational_resampler_base_ccf 3/2, 16 taps per phase; rates in OUTPUT items."""
    name = "rational_resampler_base_ccf %d/%d, %d taps per phase, 2^%d output items per GPU (section 8(f) row; output items/s)"
    block = "rational_resampler_base_ccf"
    L, D, T = 3, 2, 16
    log2n, seed = 28, 12
    kernel = "resamp_stream_kernel"

    def __init__(self, *a, **k):
        super().__init__(*a, **k)
        self.alg_bytes = 8.0 * (1 + self.D / self.L)  # per OUTPUT item
        self.flops = 4.0 * self.T

    def make(self, taps):
        return self.grb.filter.rational_resampler_base_ccf(self.L, self.D, taps, device=self.device)

    def setup(self):
        L, D, T = self.L, self.D, self.T
        self.taps = lowpass(L * T, 0.4 / max(L, D), blackman_harris(L * T)) * L
        self.blk = self.make(self.taps)
        self.nout = self.n
        self.nin = ((self.nout - 1) * D) // L + 1
        self.d_in = self.rand_c64(self.nin + T, self.seed)
        self.d_out = self.torch.empty(2 * self.nout, dtype=self.torch.float32, device="cuda")
        self.samples_per_step = self.nout

    def step(self):
        self.blk.set_ctr(0)
        self.blk.work_dev(self.d_in.data_ptr(), self.nin, self.d_out.data_ptr(), self.nout, self.stream)

    def describe(self):
        return self.name % (self.L, self.D, self.T, self.log2n)

    def parity(self, sc):
        return sc.rational_resampler(self.fetch_c64(self.d_in), self.fetch_c64(self.d_out), self.L, self.D, self.taps, self.nout,
                                     seed=self.seed)

    def cpu_setup(self, lay, log2n):
        rng = np.random.default_rng(self.seed)
        n = 1 << log2n
        self.cpu_lay = lay
        self.cpu_nin = (n * self.D) // self.L
        self.cpu_x = (rng.uniform(-1, 1, self.cpu_nin + self.T) + 1j * rng.uniform(-1, 1, self.cpu_nin + self.T)).astype(np.complex64)
        self.cpu_taps = lowpass(self.L * self.T, 0.4 / max(self.L, self.D), blackman_harris(self.L * self.T)) * self.L
        return n

    def cpu_call(self):
        self.cpu_lay.rational_resampler(self.L, self.D, self.cpu_taps, self.cpu_x, nin=self.cpu_nin)


class Interpolator(Resampler):
    """interp_fir_filter_ccf x4, 16 taps per phase; rates in OUTPUT items."""
    name = "interp_fir_filter_ccf x%d (/%d), %d taps per phase, 2^%d output items per GPU (section 8(f) row; output items/s)"
    block = "interp_fir_filter_ccf"
    L, D, T = 4, 1, 16
    seed = 13

    def make(self, taps):
        return self.grb.filter.interp_fir_filter_ccf(self.L, taps, device=self.device)

    def step(self):
        self.blk.work_dev(self.d_in.data_ptr(), self.d_out.data_ptr(), self.nout, self.stream)


WORKLOADS = {"pfb": PfbChannelizer, "fir": FirCcf, "fftfilt": FftFilter, "firdec": FirLongDecim, "fftvcc": FftVccMagSq,
             "xlat": FreqXlating, "pfbdec": PfbDecimator, "firfff": FirFff, "pfbsynth": PfbSynthesizer, "fbank": Filterbank,
             "logpwr": LogPwrFft, "interp": Interpolator, "resamp": Resampler}


# ---------------------------------------------------------------------------------------------------------
# clocks during the timed region (NVML; the recipe's nvidia-smi line, sampled in-process)
# ---------------------------------------------------------------------------------------------------------
class ClockSampler:
    REASONS = {0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
               0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting"}

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.nv = None

    def _run(self):
        while not self._stop.is_set():
            try:
                self.samples.append(float(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM)))
                r = int(self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                for bit, name in self.REASONS.items():
                    if r & bit and name != "gpu_idle":
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.05)

    def __enter__(self):
        if self.nv is not None:
            self._thr = threading.Thread(target=self._run, daemon=True)
            self._thr.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._thr:
            self._thr.join()

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["nvml unavailable"]}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


# ---------------------------------------------------------------------------------------------------------
# CPU legs (the ONLY place bench.py touches oracle/)
# ---------------------------------------------------------------------------------------------------------
def cpu_layer():
    from oracle import pyoracle as po

    lay = po.best()
    return lay, lay.kind


def cpu_run(wl_cls, threads, log2n_per_thread, reps):
    """`threads` host threads each run the reference kernel over its own 2^log2n-sample slice, `reps` times."""
    lay, kind = cpu_layer()
    wl = wl_cls.__new__(wl_cls)
    n = wl.cpu_setup(lay, log2n_per_thread)
    wl.cpu_call()  # warm
    t0 = time.perf_counter()
    if threads == 1:
        for _ in range(reps):
            wl.cpu_call()
    else:
        with cf.ThreadPoolExecutor(max_workers=threads) as ex:
            list(ex.map(lambda _: [wl.cpu_call() for _ in range(reps)], range(threads)))
    dt = time.perf_counter() - t0
    return n * reps * threads / dt / 1e6, kind, n * reps * threads, dt


def run_reference_arm(args, wl_cls):
    """--impl reference: the reference's own CPU kernels on all host cores, same workload and metric."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    threads = os.cpu_count() or 1
    log2n = 22
    msps_list = []
    for _ in range(args.warmup):
        cpu_run(wl_cls, threads, log2n, 1)
    t_all = 0.0
    nsamp = 0
    for _ in range(args.steps):
        msps, kind, n, dt = cpu_run(wl_cls, threads, log2n, 1)
        msps_list.append(msps)
        t_all += dt
        nsamp += n
    value = nsamp / t_all / 1e6
    wl = wl_cls.__new__(wl_cls)
    sample = "%d host threads x 2^%d-sample slices of the workload per step" % (threads, log2n)
    line = {
        "impl": "reference", "metric": "complex Msps (%s)" % wl_cls.block, "value": value, "unit": "Msps",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_all / max(1, args.steps),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": wl_cls.name.replace("%d", "n") if "%" in wl_cls.name else wl_cls.name,
                   "cpu_impl": "oracle/_ref/libgrref.so (reference sources compiled unmodified against a VOLK/FFTW shim)"
                   if kind == "reference" else "oracle/libgrb_oracle.so (plain-C port of the reference)"},
        "cpu_baseline": {"value": value, "unit": "Msps", "cores": threads, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": "Msps", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


# ---------------------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------------------
def timed_steps(torch, dist, wl, steps, warmup, halo=None):
    def one():
        if halo is not None:
            halo()
        wl.step()

    for _ in range(warmup):
        one()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
        torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
    ev[0].record()
    for i in range(steps):
        one()
        ev[i + 1].record()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    per = [ev[i].elapsed_time(ev[i + 1]) for i in range(steps)]
    total = ev[0].elapsed_time(ev[steps])
    return total, per


def attach_peer_halo(torch, dist, wl, device, rank, world):
    """Multi-GPU: the only exchange of a time-sharded stream is the history halo.  Each rank maps the left neighbour's tail
    (CUDA IPC, once) and the block fetches it itself over NVLink inside work_dev -- no host-launched send/recv per step
    (round 1: a batch_isend_irecv group per step, ~0.25 ms of host time).  Ring: every rank does the same work."""
    if not hasattr(wl, "halo_views"):
        return None
    from gnuradio_b200 import sharding

    views = wl.halo_views()
    blks = getattr(wl, "blks", [wl.blk])
    if len(views) != len(blks):
        raise RuntimeError("one block per buffer set with a device-side halo")
    # once, through NCCL: the buffer's own history region holds the same items the block will read from the neighbour, so
    # the parity spot check (which reads this rank's buffers) sees exactly what the kernels saw
    sharding.exchange_halos(dist, views, rank, world, ring=True)
    torch.cuda.synchronize()
    ph = sharding.PeerHalo(dist, device, [v[1].data_ptr() for v in views], rank, world, ring=True)
    for b, q in zip(blks, ph.ptrs):
        b.set_halo_source(q)
    dist.barrier()
    return ph


def sharded_run(torch, dist, grb, cls, device, rank, world, steps, peak, log2n=None, **attrs):
    """One workload time-sharded over all ranks with the device-side halo: aggregate Msps (max over ranks) + parity."""
    o = cls(torch, grb, device, log2n)
    for k, v in attrs.items():
        setattr(o, k, v)
    o.setup()
    ph = attach_peer_halo(torch, dist, o, device, rank, world)
    sps = getattr(o, "samples_per_step", o.n)
    tot, pr = timed_steps(torch, dist, o, steps, 3)
    tt = torch.tensor([tot], dtype=torch.float64, device="cuda")
    dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    tot = float(tt.item())
    msps = sps * world * steps / (tot * 1e-3) / 1e6
    res = {"workload": o.describe(), "msps": msps, "msps_per_gpu": msps / world, "samples_per_gpu_per_step": sps,
           "hbm_frac_per_gpu": o.alg_bytes * msps / world * 1e6 / 1e9 / peak, "ms_per_step": tot / steps,
           "halo_items": getattr(o, "halo_items", 0), "halo": "fetched by the block from the left neighbour's memory (NVLink peer access)",
           "parity": run_parity(torch, o)}
    torch.cuda.synchronize()
    dist.barrier()
    for b in getattr(o, "blks", [o.blk]):
        b.set_halo_source(None)
    if ph is not None:
        ph.close()
    o.free()
    return res


def run_parity(torch, wl):
    """Spot-window parity of the workload's LAST step at its full size (outside any timed region): first + last segment and 64
    seeded interior windows against the CPU oracle and float64.  north_star tolerance: NRMS <= 1e-5, max-abs <= 1e-4 x RMS."""
    try:
        from oracle import spotcheck as sc  # the checker (test infrastructure), used as such

        torch.cuda.synchronize()
        res = wl.parity(sc)
        if res is None:
            return None
        res["ok"] = bool(sc.ok(res))
        res["tolerance"] = {"nrms": sc.NRMS_TOL, "max_over_rms": sc.MAXABS_TOL}
        return res
    except Exception as e:  # a checker failure must be visible in the line, not hide the measurement
        return {"error": repr(e)[:300], "ok": False}


# ---------------------------------------------------------------------------------------------------------
# End-to-end through the host-buffer C-ABI, shaped like a flowgraph: the stream is handed to the block in calls of
# `call_items` items -- what block_executor does with a block whose buffers hold that much (gnuradio-runtime/lib/
# block_executor.cc:61-84; 4096 items at the default 32 KiB buffers, 2^22 after INTEGRATION.md's set_min_output_buffer) --
# history carried in front of every call like the scheduler's read pointer does, out of host memory that is either page-locked
# (cudaHostAlloc), registered (plain memory + cudaHostRegister, what an integration can do once for its gr::buffers) or
# pageable (a gr::buffer as it comes).  mode "blocking": work() returns when its outputs are valid; "lag": work() returns the
# previous call's outputs (grb200_*_work_lag).  Calls go straight through ctypes with precomputed addresses.
# ---------------------------------------------------------------------------------------------------------
class HostMem:
    def __init__(self, grb, kind):
        self.grb, self.kind, self.keep = grb, kind, []

    def alloc(self, shape, dtype=np.complex64):
        import ctypes as C

        from gnuradio_b200 import _lib

        dtype = np.dtype(dtype)
        n = int(np.prod(shape))
        if self.kind == "pinned":
            p = C.c_void_p()
            _lib.check(_lib.lib().grb200_host_alloc(C.byref(p), max(n * dtype.itemsize, 64)))
            a = np.frombuffer((C.c_char * (n * dtype.itemsize)).from_address(p.value), dtype=dtype).reshape(shape)
            self.keep.append(("alloc", p))
        else:
            a = np.empty(shape, dtype)
            if self.kind == "registered":
                self.grb.host_register(a)
                self.keep.append(("reg", a))
        return a

    def free(self):
        from gnuradio_b200 import _lib

        for what, obj in self.keep:
            if what == "alloc":
                _lib.lib().grb200_host_free(obj)
            else:
                self.grb.host_unregister(obj)
        self.keep = []


class E2E:
    """One block + one kind of host memory, buffers allocated once for `max_items`; run() pushes a stream through it in calls
    of `call_items` items."""

    def __init__(self, grb, device, block, mem, max_items):
        self.grb, self.block, self.mem = grb, block, mem
        self.hm = HostMem(grb, mem)
        if block == "fir":
            self.nt = 64
            self.blk = grb.filter.fir_filter_ccf(1, lowpass(self.nt, 0.1, hamming(self.nt)), device=device)
            self.x = self.hm.alloc(max_items + self.nt - 1)
            self.y = self.hm.alloc(max_items)
            self.x.view(np.float32)[:] = 0.25
            self.y.view(np.float32)[:] = 0
        else:
            self.M, self.T = 64, 16
            M, T = self.M, self.T
            self.blk = grb.filter.pfb_channelizer_ccf(M, lowpass(M * T, 0.5 / M, blackman_harris(M * T)) * np.float32(M), 1.0,
                                                      device=device)
            self.nout_max = max_items // M
            self.outs = self.hm.alloc((M, self.nout_max))
            self.outs.view(np.float32)[:] = 0
            self.x = self.hm.alloc((self.nout_max + T) * M) if block == "pfb_interleaved" else self.hm.alloc((M, self.nout_max + T))
            self.x.view(np.float32)[:] = 0.25

    def close(self):
        self.blk = None
        self.hm.free()

    def run(self, call_items, total_items, mode, reps=1, dist=None, torch=None):
        """Returns (Msps, h2d_bytes_per_call, d2h_bytes_per_call, calls) or None when the call size does not apply."""
        import ctypes as C

        from gnuradio_b200 import _lib

        L = _lib.lib()
        got, used = C.c_longlong(0), C.c_longlong(0)
        h = self.blk._h
        if self.block == "fir":
            ncalls = max(1, total_items // call_items)
            total = ncalls * call_items
            xp, yp = self.x.ctypes.data, self.y.ctypes.data
            bi, bo = (call_items + self.nt - 1) * 8, call_items * 8

            def run_once():
                if mode == "lag":
                    done = 0
                    for c in range(ncalls):
                        rc = L.grb200_fir_work_lag(h, C.c_void_p(xp + c * call_items * 8), call_items, C.c_void_p(yp + done * 8),
                                                   total - done, C.byref(got))
                        if rc:
                            _lib.check(rc)
                        done += got.value
                    L.grb200_fir_flush(h, C.c_void_p(yp + done * 8), total - done, C.byref(got))
                    return done + got.value
                for c in range(ncalls):
                    rc = L.grb200_fir_work(h, C.c_void_p(xp + c * call_items * 8), C.c_void_p(yp + c * call_items * 8), call_items)
                    if rc:
                        _lib.check(rc)
                return total
        else:
            M, T = self.M, self.T
            nout_call = call_items // M
            if nout_call < 1:
                return None
            ncalls = max(1, (total_items // M) // nout_call)
            total = ncalls * nout_call * M
            nout_all = ncalls * nout_call
            inter = self.block == "pfb_interleaved"
            xp = self.x.ctypes.data
            rowp = None if inter else [self.x[j].ctypes.data for j in range(M)]
            orow = [self.outs[j].ctypes.data for j in range(M)]
            bi, bo = (nout_call + T) * M * 8, nout_call * M * 8
            VP = C.c_void_p * M

            def run_once():
                done = 0
                for c in range(ncalls):
                    op = VP(*[o + (done if mode == "lag" else c * nout_call) * 8 for o in orow])
                    if inter:
                        src = C.c_void_p(xp + c * nout_call * M * 8)
                        if mode == "lag":
                            rc = L.grb200_pfb_chan_work_interleaved_lag(h, src, nout_call, op, M, nout_all - done, C.byref(used), C.byref(got))
                        else:
                            rc = L.grb200_pfb_chan_work_interleaved(h, src, op, M, nout_call, C.byref(used))
                    else:
                        ip = VP(*[r + c * nout_call * 8 for r in rowp])
                        if mode == "lag":
                            rc = L.grb200_pfb_chan_work_lag(h, ip, nout_call, op, M, nout_all - done, C.byref(used), C.byref(got))
                        else:
                            rc = L.grb200_pfb_chan_work(h, ip, op, M, nout_call, C.byref(used))
                    if rc:
                        _lib.check(rc)
                    done += got.value if mode == "lag" else nout_call
                if mode == "lag":
                    op = VP(*[o + done * 8 for o in orow])
                    L.grb200_pfb_chan_flush(h, op, M, nout_all - done, C.byref(got))
                    done += got.value
                return done * M

        n = run_once()  # warm-up: page faults, staging buffers, the lag ring
        assert n == total, (n, total)
        if mode == "lag" and ncalls < 2:
            run_once()  # the two parking buffers of the lag ring alternate per call: touch both before timing
        if dist is not None:
            dist.barrier()
        t0 = time.perf_counter()
        for _ in range(reps):
            run_once()
        dt = time.perf_counter() - t0
        if dist is not None:
            t = torch.tensor([dt], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.item())
        return total * reps / dt / 1e6, bi, bo, ncalls


def pcie_probe(torch, dist, seconds=0.25):
    """Page-locked H2D and D2H copies running at once on this rank's GPU, every rank at the same time: the per-rank rate each
    way (GB/s) -- the ceiling of any host-staged path on this box at this rank count."""
    n = 1 << 28
    h_in = torch.empty(n, dtype=torch.uint8).pin_memory()
    h_out = torch.empty(n, dtype=torch.uint8).pin_memory()
    d_in = torch.empty(n, dtype=torch.uint8, device="cuda")
    d_out = torch.empty(n, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

    def go(reps):
        for _ in range(reps):
            with torch.cuda.stream(s1):
                d_in.copy_(h_in, non_blocking=True)
            with torch.cuda.stream(s2):
                h_out.copy_(d_out, non_blocking=True)
        torch.cuda.synchronize()

    go(1)
    if dist is not None:
        dist.barrier()
    t0 = time.perf_counter()
    reps = 4
    go(reps)
    dt = time.perf_counter() - t0
    if dist is not None:
        t = torch.tensor([dt], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
    return reps * n / dt / 1e9


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--workload", default="pfb", choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="grb200", choices=["grb200", "reference"])
    ap.add_argument("--log2n", type=int, default=None, help="override the workload's per-GPU sample count")
    ap.add_argument("--no-others", action="store_true", help="skip the other BASELINE configs at N=1")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-sweep", action="store_true", help="skip the e2e call-size sweep (N=1 only)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "grb200" else args.warmup
    wl_cls = WORKLOADS[args.workload]
    if args.impl == "reference":
        return run_reference_arm(args, wl_cls)

    # library chatter on fd 1 (e.g. NCCL's version banner) goes to stderr; stdout carries the ONE JSON line
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)

    import torch

    import gnuradio_b200 as grb

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; libgrb200 has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist_mod

        dist = dist_mod
        if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"  # the version banner goes to stdout and would precede the JSON line
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    launches0 = grb.launch_count()
    # a real (non-default) stream; the CUDA events below sit on the stream the kernels are launched on
    bench_stream = torch.cuda.Stream()
    torch.cuda.set_stream(bench_stream)

    wl = wl_cls(torch, grb, local, args.log2n)
    wl.setup()
    samples_per_step = getattr(wl, "samples_per_step", wl.n)

    halo = None
    peer = attach_peer_halo(torch, dist, wl, local, rank, world) if dist is not None else None

    with ClockSampler(local) as clk:
        total_ms, per = timed_steps(torch, dist, wl, args.steps, args.warmup, halo)
    t = torch.tensor([total_ms], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    launches = grb.launch_count() - launches0
    launches_timed = wl.launches_per_step * args.steps
    value = samples_per_step * world * args.steps / (total_ms * 1e-3) / 1e6

    peak, peak_src = measured_peaks()
    kernel_ms = float(np.mean(per)) / wl.launches_per_step  # one dominant kernel per launch, no other work in a step
    alg_bytes_launch = wl.alg_bytes * samples_per_step / wl.launches_per_step
    achieved = alg_bytes_launch / (kernel_ms * 1e-3) / 1e9
    # DRAM bytes of the dominant kernel per launch: dram__bytes_read.sum + dram__bytes_write.sum of one `ncu --set full`
    # capture (committed under profiles/), scaled from the captured launch size to this launch's sample count
    traffic = getattr(wl, "ncu_dram_bytes_per_sample", None)
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic * samples_per_step / wl.launches_per_step if traffic else None,
                "traffic_source": getattr(wl, "ncu_source", None),
                "kernel": wl.kernel, "peak_source": peak_src,
                "alg_bytes_per_sample": wl.alg_bytes, "alg_flops_per_sample": wl.flops,
                "fp32_tflops": wl.flops * samples_per_step / wl.launches_per_step / (kernel_ms * 1e-3) / 1e12}

    parity = run_parity(torch, wl)

    line = {
        "metric": "complex Msps (%s)" % wl.block, "value": value, "unit": "Msps", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic (uniform complex64 in [-1,1)^2, seeded torch Philox on device)",
        "config": {"workload": wl.describe(), "samples_per_step_per_gpu": samples_per_step,
                   "l2": "inputs larger than L2 (%.1f GiB read per step)" % (8.0 * samples_per_step / 2 ** 30),
                   "parallelism": "time-sharded x%d, weak scaling, %d-sample history halo read from the left neighbour's memory over NVLink (CUDA IPC peer access, inside the kernel)" % (
                       world, getattr(wl, "halo_items", 0)) if world > 1 else "single GPU"},
        "roofline": roofline, "parity": parity, "gpu_launches": launches_timed, "gpu_launches_total_process": launches,
    }

    if not args.no_e2e and args.workload == "pfb":
        # Headline e2e: the call a GNU Radio user's flowgraph makes after INTEGRATION.md's edits -- pfb.channelizer_ccf (one
        # interleaved stream in, 64 streams out) handed 2^22 items per work() (set_min_output_buffer), lag-by-one, out of the
        # scheduler's own buffers page-locked ONCE at start() (grb200_host_register: plain host memory made DMA-able where it
        # lies -- the base contract's "pinned host memory").  Beside it: the same calls out of untouched PAGEABLE memory (the
        # bounce path, bound by host memcpy) and the round-1 figure (one 2 GiB call).  Every rank pushes its own stream through
        # its own GPU; barrier on both sides, max over ranks.
        wl.free()
        probe = pcie_probe(torch, dist)
        items_call, total = 1 << 22, 1 << 27
        e = E2E(grb, local, "pfb_interleaved", "registered", total)
        msps, bi, bo, ncalls = e.run(items_call, total, "lag", reps=2, dist=dist, torch=torch)
        e.close()
        e = E2E(grb, local, "pfb_interleaved", "pageable", total)
        pageable, _, _, _ = e.run(items_call, total, "lag", reps=2, dist=dist, torch=torch)
        e.close()
        e = E2E(grb, local, "pfb_interleaved", "pinned", 1 << 28)
        best, _, _, _ = e.run(1 << 28, 1 << 28, "blocking", reps=2, dist=dist, torch=torch)
        e.close()
        gbs = msps * 1e6 * 8 / 1e9
        line["e2e"] = {"value": msps * world, "unit": "Msps", "h2d_bytes_per_step": bi * world, "d2h_bytes_per_step": bo * world,
                       "steps": 2 * ncalls, "step": "one work() call of 2^22 items per GPU",
                       "call": "grb200_pfb_chan_work_interleaved_lag: 2^22 items per call, lag-by-one, host buffers = plain "
                               "memory page-locked once with grb200_host_register (what the block does for its gr::buffers at start())",
                       "gbs_each_way_per_gpu": gbs, "pcie_probe_gbs_each_way_per_gpu": probe,
                       "frac_of_pcie_probe": gbs / probe,
                       "pageable": {"value": pageable * world, "unit": "Msps", "frac_of_pcie_probe": pageable * 8e-3 / probe,
                                    "call": "the same calls out of untouched pageable memory: every byte goes through the pinned "
                                            "bounce on the copy threads (bound by host memcpy bandwidth, not PCIe)"},
                       "best_case": {"value": best * world, "unit": "Msps", "frac_of_pcie_probe": best * 8e-3 / probe,
                                     "call": "grb200_pfb_chan_work_interleaved: one 2^28-item call, cudaHostAlloc buffers (the "
                                             "round-1 e2e figure)"}}
        line["gpu_launches_e2e"] = grb.launch_count() - launches0 - launches
        if rank == 0 and world == 1 and not args.no_sweep:
            sweep = {}
            for block in ("fir", "pfb_streams", "pfb_interleaved"):
                sweep[block] = {}
                for mem in ("pinned", "registered", "pageable"):
                    try:
                        e = E2E(grb, local, block, mem, 1 << 28)
                    except Exception as ex:
                        sweep[block][mem] = {"error": repr(ex)[:160]}
                        continue
                    for mode in ("blocking", "lag"):
                        row = {}
                        for lg in (12, 16, 20, 22, 28):
                            items = 1 << lg
                            tot = min(1 << 28, max(items, min(items * 256, 1 << 26)))
                            try:
                                r = e.run(items, tot, mode, reps=1)
                                row["2^%d" % lg] = None if r is None else {"msps": round(r[0], 1), "frac_of_pcie_probe":
                                                                           round(r[0] * 8e-3 / probe, 3), "calls": r[3]}
                            except Exception as ex:
                                row["2^%d" % lg] = {"error": repr(ex)[:120]}
                        sweep[block]["%s/%s" % (mem, mode)] = row
                    e.close()
            line["e2e_sweep"] = {"unit": "Msps of input through the host-buffer C-ABI, H2D + kernel + D2H per call",
                                 "items_per_call": "2^12 (default gr::buffer) .. 2^28", "pcie_probe_gbs_each_way": probe,
                                 "blocks": sweep}

    with ClockSampler(local) as _:
        pass
    line["clocks"] = clk.summary()

    if rank == 0 and world == 1 and not args.no_others:
        others = {}
        wl.free()
        for key, cls in WORKLOADS.items():
            if key == args.workload:
                continue
            try:
                o = cls(torch, grb, local)
                o.setup()
                sps = getattr(o, "samples_per_step", o.n)
                tot, pr = timed_steps(torch, None, o, max(5, args.steps // 2), 3)
                kms = float(np.mean(pr)) / o.launches_per_step
                ach = o.alg_bytes * sps / o.launches_per_step / (kms * 1e-3) / 1e9
                others[key] = {"workload": o.describe(), "msps": sps * len(pr) / (tot * 1e-3) / 1e6,
                               "hbm_gbs": ach, "hbm_frac": ach / peak, "kernel_ms": kms,
                               "fp32_tflops": o.flops * sps / o.launches_per_step / (kms * 1e-3) / 1e12,
                               "kernel": o.kernel, "roofline": roofline_entry(o, kms, sps / o.launches_per_step, peak, peak_src),
                               "parity": run_parity(torch, o)}
                o.free()
                del o
            except Exception as e:  # a failing side config must not hide the headline
                others[key] = {"error": repr(e)[:200]}
        line["other_configs"] = others

    if world > 1 and not args.no_others and args.workload == "pfb":
        # the FIR configs time-sharded the same way, exactly as at N = 1 -- same buffer sets, same launches per step, plus the halo fetch (weak: the BASELINE size per GPU), and the two
        # configs BASELINE.json words as a fixed total -- configs[2] "2^30 samples, sharded 1/2/4/8" and configs[3]
        # "time-sharded 8 x B200" -- as STRONG scaling: 2^30 input samples in total, 2^30 / N per GPU
        others = {}
        for b in getattr(wl, "blks", [wl.blk]):
            b.set_halo_source(None)
        if peer is not None:
            peer.close()
        wl.free()
        nst = max(5, args.steps // 2)
        lg = 30 - int(round(np.log2(world)))
        others["fir"] = sharded_run(torch, dist, grb, FirCcf, local, rank, world, nst, peak)
        others["firdec"] = sharded_run(torch, dist, grb, FirLongDecim, local, rank, world, nst, peak)
        strong = {"pfb": sharded_run(torch, dist, grb, PfbChannelizer, local, rank, world, nst, peak, log2n=lg),
                  "firdec": sharded_run(torch, dist, grb, FirLongDecim, local, rank, world, nst, peak, log2n=lg)}
        for v in strong.values():
            v["scaling"] = "strong"
            v["total_samples"] = 1 << 30
        line["other_configs"] = others
        line["strong_scaling"] = strong

    if rank == 0 and world == 1 and not args.no_cpu:
        cores = os.cpu_count() or 1
        # bounded sample: ~3 s on one core, then ~1.5 s wall on all cores (about 25-30 core-seconds in total)
        _, kind, n0, dt0 = cpu_run(wl_cls, 1, 22, 1)
        r1 = max(1, int(round(3.0 / dt0)))
        v1, kind, n1, dt1 = cpu_run(wl_cls, 1, 22, r1)
        _, kind, n0, dtp = cpu_run(wl_cls, cores, 22, 1)
        rN = max(1, int(round(1.5 / dtp)))
        vN, kind, nN, dtN = cpu_run(wl_cls, cores, 22, rN)
        line["cpu_baseline"] = {"value": vN, "unit": "Msps", "cores": cores, "kind": kind,
                                "sample": "%d passes over a 2^22-sample slice of the workload per thread (%d threads, %.1f s "
                                          "wall = %.0f core-seconds); 1 thread = %.2f Msps over %.1f s (what one GNU Radio block "
                                          "thread delivers)" % (rN, cores, dtN, dtN * cores, v1, dt1),
                                "value_1core": v1}
    if dist is not None:
        dist.destroy_process_group()
    sys.stdout.flush()
    os.dup2(saved_stdout, 1)
    if rank == 0:
        print(json.dumps(line), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
