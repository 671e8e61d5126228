"""Developer check: accuracy of the tcgen05 FIR path vs float64, next to the CUDA-core path (GRB200_FIR_TC=0)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gnuradio_b200 as grb
from oracle import pyoracle as po

rng = np.random.default_rng(11)
n = 1 << 20
x = (rng.uniform(-1, 1, n) + 1j * rng.uniform(-1, 1, n)).astype(np.complex64)
for name, taps in (("uniform64", rng.uniform(-1, 1, 64).astype(np.float32)),
                   ("lowpass64", (np.sinc(0.2 * (np.arange(64) - 31.5)) * np.hamming(64)).astype(np.float32)),
                   ("uniform37", rng.uniform(-1, 1, 37).astype(np.float32)),
                   ("one", np.array([1.0], np.float32))):
    nt = len(taps)
    xin = np.concatenate([np.zeros(nt - 1, np.complex64), x])
    ref = po.f64_fir(1, taps, x)
    for tc in ("1", "0"):
        os.environ["GRB200_FIR_TC"] = tc
        y = grb.filter.fir_filter_ccf(1, taps).process(xin, n)
        print(name, "tc=" + tc, "nrms %.3g max/rms %.3g" % po.nrms(y, ref))
