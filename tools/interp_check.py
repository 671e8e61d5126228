"""Developer probe: device-resident interp_fir_filter_ccf throughput (output items/s)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import gnuradio_b200 as g

rng = np.random.default_rng(0)
ts = torch.cuda.Stream()
for L, T, log2no in ((4, 16, 28), (2, 32, 28), (8, 8, 28), (4, 11, 28)):
    no = 1 << log2no
    ni = no // L
    xd = torch.randn(2 * (ni + T), dtype=torch.float32, device="cuda")
    yd = torch.empty(2 * no, dtype=torch.float32, device="cuda")
    blk = g.filter.interp_fir_filter_ccf(L, rng.uniform(-1, 1, L * T).astype(np.float32))
    for _ in range(3):
        blk.work_dev(xd.data_ptr(), yd.data_ptr(), no, ts.cuda_stream)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(ts)
    for _ in range(5):
        blk.work_dev(xd.data_ptr(), yd.data_ptr(), no, ts.cuda_stream)
    e1.record(ts)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    gs = no / ms / 1e6
    print("L=%2d T=%2d  %.3f ms  %.1f G output items/s  %.0f GB/s" % (L, T, ms, gs, gs * 8 * (1 + 1.0 / L)), flush=True)
