"""Developer probe: fir_filter_fff device-resident throughput vs tap count, FIR kernels against the overlap-save route."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import gnuradio_b200 as g

rng = np.random.default_rng(0)
ts = torch.cuda.Stream()
n = 1 << 28
xd = torch.randn(n + 20000, dtype=torch.float32, device="cuda")
yd = torch.empty(n, dtype=torch.float32, device="cuda")
for ntaps, D in ((32, 1), (64, 1), (96, 1), (128, 1), (192, 1), (256, 1), (1024, 1), (4097, 1), (8192, 1), (256, 4), (1024, 8)):
    taps = rng.uniform(-1, 1, ntaps).astype(np.float32)
    row = []
    for pref in ("0", "1"):
        os.environ["GRB200_FIR_FFT"] = pref
        blk = g.filter.fir_filter_fff(D, taps)
        for _ in range(2):
            blk.work_dev(xd.data_ptr(), yd.data_ptr(), n // D, ts.cuda_stream)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(ts)
        for _ in range(3):
            blk.work_dev(xd.data_ptr(), yd.data_ptr(), n // D, ts.cuda_stream)
        e1.record(ts)
        torch.cuda.synchronize()
        row.append(n / (e0.elapsed_time(e1) / 3) / 1e6)
    print("fir_filter_fff ntaps=%5d D=%d   FIR kernel %6.1f Gsps   overlap-save %6.1f Gsps" % (ntaps, D, row[0], row[1]), flush=True)
