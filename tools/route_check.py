"""Developer probe: where the overlap-save route overtakes the direct FIR kernels (fir_filter_ccf / ccc / fff), per tap
count and decimation.  GRB200_FIR_FFT is read when a handle is made: 0 = direct kernels, 1 = overlap-save whenever a plan exists."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import gnuradio_b200 as g

rng = np.random.default_rng(0)
ts = torch.cuda.Stream()
n = 1 << 27
xd = torch.randn(2 * (n + 70000), dtype=torch.float32, device="cuda")
yd = torch.empty(2 * n + 64, dtype=torch.float32, device="cuda")


def timed(fn, reps=3):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(ts)
    for _ in range(reps):
        fn()
    e1.record(ts)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


kinds = [a for a in sys.argv[1:] if "=" not in a] or ["ccf", "ccc", "fff"]
for kind in kinds:
    for D in [int(a[2:]) for a in sys.argv[1:] if a.startswith('D=')] or (1, 2, 4, 8):
        for ntaps in (8, 16, 24, 32, 40, 48, 64, 96, 128, 192, 256, 384, 512):
            if kind == "ccc":
                taps = (rng.uniform(-1, 1, ntaps) + 1j * rng.uniform(-1, 1, ntaps)).astype(np.complex64)
            else:
                taps = rng.uniform(-1, 1, ntaps).astype(np.float32)
            row = []
            for pref in ("0", "1", None):
                if pref is None:
                    os.environ.pop("GRB200_FIR_FFT", None)
                else:
                    os.environ["GRB200_FIR_FFT"] = pref
                blk = getattr(g.filter, "fir_filter_" + kind)(D, taps)
                nn = (2 * n if kind == "fff" else n)
                ms = timed(lambda: blk.work_dev(xd.data_ptr(), yd.data_ptr(), nn // D, ts.cuda_stream))
                row.append(nn / ms / 1e6)
            print("fir_filter_%s D=%d ntaps=%4d   direct %6.1f   overlap-save %6.1f   auto %6.1f  %s" % (
                kind, D, ntaps, row[0], row[1], row[2], "<-- auto is not the best" if row[2] < 0.97 * max(row[0], row[1]) else ""), flush=True)
