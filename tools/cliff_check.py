"""Developer probe: device-resident throughput of the path's blocks across parameters the bench does not visit
(looking for cliffs between a fast path and its fallback).  usage: cliff_check.py [fir] [fft] [pfb] [fftfilt]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import gnuradio_b200 as g

rng = np.random.default_rng(0)
ts = torch.cuda.Stream()
what = [a for a in sys.argv[1:] if "=" not in a] or ["fir", "fft", "pfb", "fftfilt"]
opts = dict(a.split("=") for a in sys.argv[1:] if "=" in a)


def timed(fn, reps=5):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(ts)
    for _ in range(reps):
        fn()
    e1.record(ts)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


n = 1 << 27
xd = torch.randn(2 * (n + 70000), dtype=torch.float32, device="cuda")
yd = torch.empty(2 * n + 64, dtype=torch.float32, device="cuda")
if "fir" in what:
    for ntaps, D, cplx in ((16, 1, 0), (32, 1, 0), (64, 1, 0), (65, 1, 0), (128, 1, 0), (256, 1, 0), (64, 2, 0), (64, 4, 0), (128, 4, 0),
                           (256, 8, 0), (128, 8, 0), (512, 16, 0), (1024, 8, 0), (2048, 8, 0), (64, 1, 1), (128, 4, 1), (32, 2, 1)):
        taps = (rng.uniform(-1, 1, ntaps) + 1j * rng.uniform(-1, 1, ntaps)).astype(np.complex64) if cplx else rng.uniform(-1, 1, ntaps).astype(np.float32)
        blk = (g.filter.fir_filter_ccc if cplx else g.filter.fir_filter_ccf)(D, taps)
        ms = timed(lambda: blk.work_dev(xd.data_ptr(), yd.data_ptr(), n // D, ts.cuda_stream))
        print("fir_%s ntaps=%5d D=%2d  %.3f ms  %6.1f Gsps  %5.0f GB/s  %5.1f TFLOP/s" % ("ccc" if cplx else "ccf", ntaps, D, ms, n / ms / 1e6,
              n / ms / 1e6 * 8 * (1 + 1 / D), n / ms / 1e6 * ntaps / D * (8 if cplx else 4) / 1e3), flush=True)
if "fft" in what:
    for N in (16, 64, 256, 512, 1024, 2048, 4096, 8192, 16384, 32768, 65536, 262144, 1000, 1536):
        nv = n // N
        blk = g.fft.fft_vcc(N, True, np.ones(N, np.float32), False)
        ms = timed(lambda: blk.work_dev(xd.data_ptr(), yd.data_ptr(), nv, ts.cuda_stream))
        print("fft_vcc N=%5d  %.3f ms  %6.1f Gsps  %5.0f GB/s" % (N, ms, nv * N / ms / 1e6, nv * N / ms / 1e6 * 16), flush=True)
if "pfb" in what:
    for M, T, osr, remap in ((64, 16, 1.0, 0), (64, 16, 2.0, 0), (64, 16, 1.0, 1), (32, 16, 2.0, 0), (64, 12, 64 / 48, 0), (64, 40, 1.0, 0)):
        taps = rng.uniform(-1, 1, M * T).astype(np.float32)
        blk = g.filter.pfb_channelizer_ccf(M, taps, osr)
        if remap:
            blk.set_channel_map(list(reversed(range(M))))
        nout = int((n // M) / osr) // 4 * 4  # output items per channel so that M * nout fits the output buffer
        ms = timed(lambda: blk.work_interleaved_dev(xd.data_ptr(), yd.data_ptr(), nout, M, nout, ts.cuda_stream))
        nin = nout / osr * M
        print("pfb M=%3d T=%2d osr=%.2f map=%d  %.3f ms  %6.1f Gsps (input)" % (M, T, osr, remap, ms, nin / ms / 1e6), flush=True)
if "fftfilt" in what:
    for ntaps in [int(v) for v in opts.get("taps", "16,64,256,1024,2049,4096,6000,12000").split(",")]:
        taps = (rng.uniform(-1, 1, ntaps) + 1j * rng.uniform(-1, 1, ntaps)).astype(np.complex64)
        blk = g.filter.fft_filter_ccc(1, taps)
        ns = blk.nsamples()
        nout = (n // ns) * ns
        ms = timed(lambda: blk.work_dev(xd.data_ptr(), yd.data_ptr(), nout, ts.cuda_stream))
        print("fft_filter_ccc ntaps=%5d  %.3f ms  %6.1f Gsps" % (ntaps, ms, nout / ms / 1e6), flush=True)
