"""Child process of tests/test_gpu_staging.py: runs every host-staged work() once with the chunk size and the host-memory
kind given on the command line (GRB200_CHUNK_MB is read once per process, so each setting needs its own process) and saves
the outputs.   python staging_child.py <out.npz> <pageable|pinned>"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import gnuradio_b200 as grb  # noqa: E402
from gnuradio_b200 import _lib  # noqa: E402

out_path, kind = sys.argv[1], sys.argv[2]
keep = []


def buf(shape, dtype):
    """A host array of the requested kind: plain numpy (pageable) or page-locked through grb200_host_alloc."""
    dtype = np.dtype(dtype)
    n = int(np.prod(shape))
    if kind == "pageable":
        return np.zeros(shape, dtype)
    p = C.c_void_p()
    _lib.check(_lib.lib().grb200_host_alloc(C.byref(p), max(n * dtype.itemsize, 64)))
    keep.append(p)
    a = np.frombuffer((C.c_char * (n * dtype.itemsize)).from_address(p.value), dtype=dtype).reshape(shape)
    a[...] = 0
    return a


def put(a):
    b = buf(a.shape, a.dtype)
    b[...] = a
    return b


rng = np.random.default_rng(1234)
cr = lambda n: (rng.uniform(-1, 1, n) + 1j * rng.uniform(-1, 1, n)).astype(np.complex64)
res = {}
N = 1 << 20  # 8 MiB of complex64: 8 chunks at GRB200_CHUNK_MB=1

# fir_filter_ccf (CUDA-core path: bit-reproducible under chunking), ccc with decimation, fff
os.environ["GRB200_FIR_TC"] = "0"
os.environ["GRB200_FIR_FFT"] = "0"  # (the overlap-save route of long filters re-blocks per chunk: parity, not bits)
taps = rng.uniform(-1, 1, 37).astype(np.float32)
x = put(cr(N + 36))
y = buf(N, np.complex64)
grb.filter.fir_filter_ccf(1, taps).work(N, [x], [y])
res["fir_ccf"] = y.copy()
tc = cr(21)
y = buf(N // 4, np.complex64)
grb.filter.fir_filter_ccc(4, tc).work((N - 20 + 20) // 4 - 8, [x], [y])
res["fir_ccc_d4"] = y.copy()
xf = put(rng.uniform(-1, 1, 2 * N + 99).astype(np.float32))
y = buf(2 * N, np.float32)
grb.filter.fir_filter_fff(1, rng.uniform(-1, 1, 100).astype(np.float32)).work(2 * N, [xf], [y])
res["fir_fff"] = y.copy()
# freq_xlating: rotator phase per chunk
blk = grb.filter.freq_xlating_fir_filter_ccf(2, taps, 0.11, 1.0)
y = buf(N // 2, np.complex64)
blk.work(0, [x], [y])
blk.work(N // 2, [x], [y])
res["xlat"] = y.copy()
# fft_filter: history splice between chunks and calls (two calls)
ff = grb.filter.fft_filter_ccc(1, cr(300))
ns = ff.nsamples()
n1 = (N // 2 // ns) * ns
xa = put(cr(2 * n1))
y = buf(2 * n1, np.complex64)
ff.work(0, [xa], [y])
ff.work(n1, [xa[:n1]], [y[:n1]])
ff.work(n1, [xa[n1:]], [y[n1:]])
res["fftfilt"] = y.copy()
# channelizer: M separate streams (row-strided when pinned: the cudaMemcpy2DAsync branch) and the interleaved form
M, T = 64, 8
ptaps = rng.uniform(-1, 1, M * T).astype(np.float32)
nout = N // M
ins = put(cr(M * (nout + T)).reshape(M, nout + T))
outs = buf((M, nout), np.complex64)
ch = grb.filter.pfb_channelizer_ccf(M, ptaps, 1.0)
il, ol = [ins[j] for j in range(M)], [outs[j] for j in range(M)]
ch.general_work(0, [], il, ol)
ch.general_work(nout, [], il, ol)
res["pfb_streams"] = outs.copy()
xil = put(np.ascontiguousarray(ins.T).reshape(-1))
outs2 = buf((M, nout), np.complex64)
ch.work_interleaved(xil, [outs2[j] for j in range(M)], nout)
res["pfb_interleaved"] = outs2.copy()
# pfb_decimator, interp_fir, rational_resampler
dec = grb.filter.pfb_decimator_ccf(8, rng.uniform(-1, 1, 8 * 6).astype(np.float32), 3, False, False)
dins = put(cr(8 * (N // 8 + 5)).reshape(8, N // 8 + 5))
y = buf(N // 8, np.complex64)
dl = [dins[j] for j in range(8)]
dec.work(0, dl, [y])
dec.work(N // 8, dl, [y])
res["pfbdec"] = y.copy()
it = grb.filter.interp_fir_filter_ccf(3, rng.uniform(-1, 1, 3 * 7).astype(np.float32))
xi = put(cr(N // 2 + 6))
y = buf(3 * (N // 2), np.complex64)
it.work(0, [xi], [y])
it.work(3 * (N // 2), [xi], [y])
res["interp"] = y.copy()
rs = grb.filter.rational_resampler_base_ccf(3, 5, rng.uniform(-1, 1, 3 * 8).astype(np.float32))
y = buf(N, np.complex64)
rs.general_work(0, [N], [x], [y])
n = rs.general_work(N * 3 // 5 - 16, [N], [x], [y])
res["resamp"] = y[:n].copy()
# fft_vcc (+magsq), standalone magsq, logpwrfft (averaging: filter memory across chunks), fft_vfc
f = grb.fft.fft_vcc(1024, True, rng.uniform(0.1, 1, 1024).astype(np.float32), True)
xv = put(cr(N))
y = buf(N, np.complex64)
f.work(N // 1024, [xv], [y])
res["fft_vcc"] = y.copy()
y = buf(N, np.float32)
f.work_magsq(N // 1024, [xv], [y])
res["fft_vcc_magsq"] = y.copy()
y = buf(N, np.float32)
grb.blocks.complex_to_mag_squared().work(N, [xv], [y])
res["magsq"] = y.copy()
lp = grb.fft.logpwrfft_c(1024.0 * 2, 1024, 2.0, 1.0, 0.2, True)
y = buf(N // 2, np.float32)
k = lp.general_work(N // 2048, [N], [xv], [y])
res["logpwr"] = y[: k * 1024].copy()
y = buf(2 * N, np.complex64)
grb.fft.fft_vfc(1024, True, []).work(2 * N // 1024, [xf], [y])
res["fft_vfc"] = y.copy()
# pfb_synthesizer (state across chunks AND calls), filterbank, vector_to_streams
sy = grb.filter.pfb_synthesizer_ccf(64, ptaps)
sins = put(cr(64 * (N // 64)).reshape(64, N // 64))
y = buf(N, np.complex64)
h = N // 128
sy.work(0, [sins[j] for j in range(64)], [y])
sy.work(h * 64, [sins[j][:h] for j in range(64)], [y[: h * 64]])
sy.work(N - h * 64, [sins[j][h:] for j in range(64)], [y[h * 64:]])
res["synth"] = y.copy()
fb = grb.filter.filterbank_vcvcf(rng.uniform(-1, 1, (32, 7)).astype(np.float32))
xfb = put(cr(32 * (N // 32 + 7)))
y = buf(N, np.complex64)
fb.general_work(N // 32, [N // 32 + 7], [xfb], [y])
res["fbank"] = y.copy()
vouts = buf((16, N // 16), np.complex64)
grb.blocks.vector_to_streams(8, 16).work(N // 16, [xv], [vouts[j] for j in range(16)])
res["v2ss"] = vouts.copy()
# inputs the parent needs for the oracle comparison
res["in_x"], res["in_taps"], res["in_ptaps"], res["in_ins"] = np.asarray(x).copy(), taps, ptaps, np.asarray(ins).copy()
np.savez(out_path, **res)
print("ok", len(res), "arrays,", grb.launch_count(), "launches")
