"""Developer probe: device-resident fft_filter_ccc throughput per tap count, plain overlap-save vs two partitions."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import gnuradio_b200 as g

n = 1 << 27
rng = np.random.default_rng(0)
xd = torch.randn(2 * (n + 16384), dtype=torch.float32, device="cuda")
yd = torch.empty(2 * n, dtype=torch.float32, device="cuda")
ts = torch.cuda.Stream()
for nt in [int(a) for a in sys.argv[1:]] or [1026, 1500, 2049, 2050, 3000, 4096, 6000, 8192, 10240]:
    for mode in ("0", "1"):
        os.environ["GRB200_FFTFILT_UPOLS"] = mode
        taps = ((rng.uniform(-1, 1, nt) + 1j * rng.uniform(-1, 1, nt)) / nt).astype(np.complex64)
        blk = g.filter.fft_filter_ccc(1, taps)
        for _ in range(2):
            blk.work_dev(xd.data_ptr(), yd.data_ptr(), n, ts.cuda_stream)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(ts)
        for _ in range(5):
            blk.work_dev(xd.data_ptr(), yd.data_ptr(), n, ts.cuda_stream)
        e1.record(ts)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 5
        print("ntaps %5d  upols=%s  %.3f ms  %.1f Gsps" % (nt, mode, ms, n / ms / 1e6), flush=True)
