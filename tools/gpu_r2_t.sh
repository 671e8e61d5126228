#!/bin/bash
mkdir -p gpurun_out
python - <<'PY'
import os, sys
sys.path.insert(0, os.getcwd())
import numpy as np, torch
import gnuradio_b200 as g
rng = np.random.default_rng(0); ts = torch.cuda.Stream(); n = 1 << 27
xd = torch.randn(2 * (n + 70000), dtype=torch.float32, device="cuda"); yd = torch.empty(2 * n + 64, dtype=torch.float32, device="cuda")
for ntaps, D in ((1024, 8), (700, 8), (224, 8), (2048, 8), (1024, 4), (2048, 4)):
    taps = rng.uniform(-1, 1, ntaps).astype(np.float32)
    row = []
    for pref in ("0", "1", None):
        if pref is None: os.environ.pop("GRB200_FIR_FFT", None)
        else: os.environ["GRB200_FIR_FFT"] = pref
        blk = g.filter.fir_filter_ccf(D, taps)
        f = lambda: blk.work_dev(xd.data_ptr(), yd.data_ptr(), n // D, ts.cuda_stream)
        f(); f(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(ts)
        for _ in range(5): f()
        e1.record(ts); torch.cuda.synchronize()
        row.append(n / (e0.elapsed_time(e1) / 5) / 1e6)
    print("ccf ntaps=%d D=%d direct %.1f routed %.1f auto %.1f" % (ntaps, D, *row), flush=True)
PY
for w in xlat firdec; do python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e --no-others --workload $w 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['config']['workload'][:40], d['value'], d['roofline']['frac'], (d.get('parity') or {}).get('ok'))"; done
