"""Developer probe: device-resident pfb_channelizer_ccf (interleaved entry) throughput for several channel counts."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import gnuradio_b200 as g

rng = np.random.default_rng(0)
ts = torch.cuda.Stream()
cases = [(int(a), int(b)) for a, b in (c.split("x") for c in sys.argv[1:])] or [(8, 16), (16, 16), (32, 16), (64, 16), (128, 16), (256, 16), (48, 16), (100, 12)]
for M, T in cases:
    n = 1 << 28
    nout = n // M
    taps = rng.uniform(-1, 1, M * T).astype(np.float32)
    blk = g.filter.pfb_channelizer_ccf(M, taps, 1.0)
    xd = torch.randn(2 * (n + M * T), dtype=torch.float32, device="cuda")
    yd = torch.empty(2 * n, dtype=torch.float32, device="cuda")
    for _ in range(3):
        blk.work_interleaved_dev(xd.data_ptr(), yd.data_ptr(), nout, M, nout, ts.cuda_stream)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(ts)
    for _ in range(5):
        blk.work_interleaved_dev(xd.data_ptr(), yd.data_ptr(), nout, M, nout, ts.cuda_stream)
    e1.record(ts)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print("M=%4d T=%2d  %.3f ms  %.1f Gsps  %.0f GB/s" % (M, T, ms, n / ms / 1e6, n / ms / 1e6 * 16), flush=True)
