"""Developer probe: device-resident rational_resampler_base_ccf throughput (input + output items/s)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import gnuradio_b200 as g

rng = np.random.default_rng(0)
ts = torch.cuda.Stream()
for L, D, T, log2ni in ((3, 2, 16, 27), (2, 3, 16, 28), (160, 147, 10, 27), (147, 160, 10, 27), (1, 4, 64, 28), (4, 1, 16, 26)):
    ni = 1 << log2ni
    no = ni * L // D
    xd = torch.randn(2 * (ni + T), dtype=torch.float32, device="cuda")
    yd = torch.empty(2 * no, dtype=torch.float32, device="cuda")
    blk = g.filter.rational_resampler_base_ccf(L, D, rng.uniform(-1, 1, L * T).astype(np.float32))
    for _ in range(3):
        blk.set_ctr(0)
        blk.work_dev(xd.data_ptr(), ni, yd.data_ptr(), no, ts.cuda_stream)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(ts)
    for _ in range(5):
        blk.set_ctr(0)
        blk.work_dev(xd.data_ptr(), ni, yd.data_ptr(), no, ts.cuda_stream)
    e1.record(ts)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print("L=%3d D=%3d T=%2d  %.3f ms  %.1f G input items/s  %.1f G output items/s  %.0f GB/s" % (
        L, D, T, ms, ni / ms / 1e6, no / ms / 1e6, (ni + no) * 8 / ms / 1e6), flush=True)
