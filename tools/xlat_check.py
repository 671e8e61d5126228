"""Developer probe: device-resident throughput of freq_xlating_fir_filter (FIR + rotator) vs the plain FIR."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import gnuradio_b200 as g

n = 1 << 27
rng = np.random.default_rng(0)
xd = torch.randn(2 * (n + 1024), dtype=torch.float32, device="cuda")
yd = torch.empty(2 * n, dtype=torch.float32, device="cuda")
ts = torch.cuda.Stream()  # a null stream handle would select the library's own stream
st = ts.cuda_stream
for name, mk, D in (("fir_ccc 64", lambda: g.filter.fir_filter_ccc(1, (rng.uniform(-1, 1, 64) + 1j * rng.uniform(-1, 1, 64)).astype(np.complex64)), 1),
                    ("xlat_ccf 64", lambda: g.filter.freq_xlating_fir_filter_ccf(1, rng.uniform(-1, 1, 64).astype(np.float32), 0.1, 1.0), 1),
                    ("xlat_ccf 64 D=8", lambda: g.filter.freq_xlating_fir_filter_ccf(8, rng.uniform(-1, 1, 64).astype(np.float32), 0.1, 1.0), 8)):
    blk = mk()
    nout = n // D
    for _ in range(3):
        blk.work_dev(xd.data_ptr(), yd.data_ptr(), nout, st)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(ts)
    for _ in range(10):
        blk.work_dev(xd.data_ptr(), yd.data_ptr(), nout, st)
    e1.record(ts)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print("%-18s %.3f ms  %.1f G input samples/s" % (name, ms, nout * D / ms / 1e6))
