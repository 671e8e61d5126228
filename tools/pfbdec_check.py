"""Developer probe: device-resident pfb_decimator_ccf throughput (input items/s)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import gnuradio_b200 as g

rng = np.random.default_rng(0)
ts = torch.cuda.Stream()
for M, T, log2n in ((64, 16, 28), (8, 16, 28), (16, 64, 28), (4, 8, 27)):
    n = (1 << log2n) // M
    stride = (n + T - 1 + 1) & ~1  # even: 16-byte aligned rows -> the pair kernel
    xd = torch.randn(2 * M * stride, dtype=torch.float32, device="cuda")
    yd = torch.empty(2 * n, dtype=torch.float32, device="cuda")
    blk = g.filter.pfb_decimator_ccf(M, rng.uniform(-1, 1, M * T).astype(np.float32), 1, False, False)
    for _ in range(3):
        blk.work_dev(xd.data_ptr(), stride, yd.data_ptr(), n, ts.cuda_stream)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(ts)
    for _ in range(5):
        blk.work_dev(xd.data_ptr(), stride, yd.data_ptr(), n, ts.cuda_stream)
    e1.record(ts)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    gs = n * M / ms / 1e6
    print("M=%3d T=%3d  %.3f ms  %.1f G input items/s  %.0f GB/s" % (M, T, ms, gs, gs * 8 * (1 + 1.0 / M)), flush=True)
