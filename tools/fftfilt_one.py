"""Developer probe: one device-resident fft_filter_ccc launch (for ncu).  argv: ntaps log2n"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import gnuradio_b200 as g

nt = int(sys.argv[1]); n = 1 << int(sys.argv[2])
rng = np.random.default_rng(0)
xd = torch.randn(2 * (n + 16384), dtype=torch.float32, device="cuda")
yd = torch.empty(2 * n, dtype=torch.float32, device="cuda")
taps = ((rng.uniform(-1, 1, nt) + 1j * rng.uniform(-1, 1, nt)) / nt).astype(np.complex64)
blk = g.filter.fft_filter_ccc(1, taps)
ts = torch.cuda.Stream()
for _ in range(2):
    blk.work_dev(xd.data_ptr(), yd.data_ptr(), n, ts.cuda_stream)
torch.cuda.synchronize()
