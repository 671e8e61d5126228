"""Developer check: accuracy of the tcgen05 FIR path vs float64, next to the CUDA-core path (GRB200_FIR_TC=0),
and device-resident throughput of both (2^24 samples, CUDA events)."""
import os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gnuradio_b200 as grb
from oracle import pyoracle as po

rng = np.random.default_rng(11)
n = 1 << 20
x = (rng.uniform(-1, 1, n) + 1j * rng.uniform(-1, 1, n)).astype(np.complex64)
def crand(k):
    return (rng.uniform(-1, 1, k) + 1j * rng.uniform(-1, 1, k)).astype(np.complex64)
cases = (("ccf uniform64", rng.uniform(-1, 1, 64).astype(np.float32)),
         ("ccf lowpass64", (np.sinc(0.2 * (np.arange(64) - 31.5)) * np.hamming(64)).astype(np.float32)),
         ("ccf uniform37", rng.uniform(-1, 1, 37).astype(np.float32)),
         ("ccc uniform64", crand(64)), ("ccc uniform21", crand(21)))
cases += (("dec8 uniform1024", rng.uniform(-1, 1, 1024).astype(np.float32)),
          ("dec8 lowpass1024", (np.sinc((np.arange(1024) - 511.5) / 16) * np.blackman(1024) / 16).astype(np.float32)),
          ("dec8 uniform968", rng.uniform(-1, 1, 968).astype(np.float32)),
          ("dec8 uniform1001", rng.uniform(-1, 1, 1001).astype(np.float32)),
          ("dec8 uniform300", rng.uniform(-1, 1, 300).astype(np.float32)))
st = torch.cuda.Stream()
for name, taps in cases:
    nt = len(taps)
    D = 8 if name.startswith("dec8") else 1
    cls = grb.filter.fir_filter_ccc if name.startswith("ccc") else grb.filter.fir_filter_ccf
    xin = np.concatenate([np.zeros(nt - 1, np.complex64), x])
    ref = po.f64_fir(D, taps, x)
    for tc in ("1", "0"):
        os.environ["GRB200_FIR_TC"] = tc
        blk = cls(D, taps)
        y = blk.process(xin, n // D)
        N = (1 << 24) if D == 1 else (1 << 27) // D  # outputs per launch
        d_in = torch.rand(2 * (N * D + nt - 1), device="cuda") - 0.5
        d_out = torch.empty(2 * N, device="cuda")
        with torch.cuda.stream(st):
            for _ in range(3):
                blk.work_dev(d_in.data_ptr(), d_out.data_ptr(), N, st.cuda_stream)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(10):
                blk.work_dev(d_in.data_ptr(), d_out.data_ptr(), N, st.cuda_stream)
            e1.record()
        torch.cuda.synchronize()
        print(name, "tc=" + tc, "nrms %.3g max/rms %.3g" % po.nrms(y, ref), "| %.0f G input samples/s" % (10 * N * D / e0.elapsed_time(e1) / 1e6))
