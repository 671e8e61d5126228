"""Developer aid: per-tile clock64() stamps of CTA 0 of fir_tc_kernel (GRB200_FIR_TC_TRACE), one device-resident launch.
usage: fir_tc_trace.py ntaps decim [ccc] [log2n] [first_it]"""
import os, sys
os.makedirs("gpurun_out", exist_ok=True)
os.environ["GRB200_FIR_TC_TRACE"] = "gpurun_out/fir_tc_trace.txt"  # read once, at the first launch
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gnuradio_b200 as grb
ntaps = int(sys.argv[1]) if len(sys.argv) > 1 else 64
D = int(sys.argv[2]) if len(sys.argv) > 2 else 1
cplx = len(sys.argv) > 3 and sys.argv[3] == "ccc"
n = 1 << (int(sys.argv[4]) if len(sys.argv) > 4 else 26)
first = int(sys.argv[5]) if len(sys.argv) > 5 else 30
rng = np.random.default_rng(1)
taps = (rng.uniform(-1, 1, ntaps) + 1j * rng.uniform(-1, 1, ntaps)).astype(np.complex64) if cplx else rng.uniform(-1, 1, ntaps).astype(np.float32)
blk = (grb.filter.fir_filter_ccc if cplx else grb.filter.fir_filter_ccf)(D, taps)
x = torch.rand(2 * (n + ntaps), device="cuda") - 0.5
y = torch.empty(2 * (n // D), device="cuda")
for _ in range(2):
    blk.work_dev(x.data_ptr(), y.data_ptr(), n // D)
    torch.cuda.synchronize()
rows = [list(map(int, l.split())) for l in open("gpurun_out/fir_tc_trace.txt") if l.strip()]
names = "P.start P.gotRaw P.gotB P.arrived | M.start M.gotB M.gotD M.issued | E.start E.gotD E.ldDone E.stored".replace("| ", "").split()
print(" it " + " ".join("%9s" % s for s in names))
for i, r in enumerate(rows[first:first + 12]):
    print("%3d " % (i + first) + " ".join("%9d" % (v - rows[first][0]) for v in r))
per = [rows[i + 1][0] - rows[i][0] for i in range(first, min(first + 30, len(rows) - 1)) if rows[i + 1][0] > 0]
print("tile period (converter start to start):", sum(per) / max(len(per), 1), "clk")
