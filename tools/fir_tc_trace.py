"""Developer aid: per-tile clock64() stamps of CTA 0 of fir_tc_kernel (GRB200_FIR_TC_TRACE)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gnuradio_b200 as grb
ntaps = int(sys.argv[1]) if len(sys.argv) > 1 else 64
D = int(sys.argv[2]) if len(sys.argv) > 2 else 1
n = 1 << 23
rng = np.random.default_rng(1)
x = (rng.uniform(-1, 1, n + ntaps - 1) + 1j * rng.uniform(-1, 1, n + ntaps - 1)).astype(np.complex64)
cplx = len(sys.argv) > 3 and sys.argv[3] == "ccc"
taps = (rng.uniform(-1, 1, ntaps) + 1j * rng.uniform(-1, 1, ntaps)).astype(np.complex64) if cplx else rng.uniform(-1, 1, ntaps).astype(np.float32)
blk = (grb.filter.fir_filter_ccc if cplx else grb.filter.fir_filter_ccf)(D, taps)
blk.process(x, n // D)
os.environ["GRB200_FIR_TC_TRACE"] = "gpurun_out/fir_tc_trace.txt"
blk.process(x, n // D)
rows = [list(map(int, l.split())) for l in open("gpurun_out/fir_tc_trace.txt") if l.strip()]
names = "P.start P.gotEmpty P.sts0 P.arrived | M.start M.gotB M.gotD M.issued | E.start E.gotD E.ldDone E.stored".replace("| ", "").split()
print(" it " + " ".join("%10s" % s for s in names))
for i, r in enumerate(rows[:20]):
    print("%3d " % i + " ".join("%10d" % v for v in r))
