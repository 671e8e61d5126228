"""GPU: the host staging path (csrc/staging.cu).

* every host work() with GRB200_CHUNK_MB=1 (8 chunks per call: slot reuse and retire, bounce regrow, per-chunk geometry --
  resampler phase, fft_filter history splice, rotator phase, synthesizer delay lines, logpwrfft filter memory), once from
  pageable and once from page-locked buffers (direct copies; the cudaMemcpy2DAsync branch for row-strided stream sets), must
  equal the same calls made with one big chunk, and the oracle;
* lag-by-one mode: results arrive one call late, flush() returns the tail, and they equal the blocking results;
* NULL stream = the legacy default stream: producer -> work_dev -> consumer on one stream need no device-wide sync."""
import os
import subprocess
import sys

import numpy as np
import pytest

from common import assert_parity, crand
from conftest import ROOT

pytestmark = pytest.mark.gpu
CHILD = os.path.join(ROOT, "tests", "staging_child.py")


def run_child(tmp_path, tag, chunk_mb, kind):
    out = str(tmp_path / ("staging_%s.npz" % tag))
    env = dict(os.environ, GRB200_CHUNK_MB=str(chunk_mb))
    r = subprocess.run([sys.executable, CHILD, out, kind], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    return dict(np.load(out))


def test_chunked_pageable_and_pinned_equal_single_chunk(tmp_path, orc):
    one = run_child(tmp_path, "one", 1024, "pageable")   # at most the 4-way split of calls above 1 MiB
    page = run_child(tmp_path, "page", 1, "pageable")    # 1 MiB chunks through the pinned bounce
    pin = run_child(tmp_path, "pin", 1, "pinned")        # 1 MiB chunks, direct copies from / to page-locked memory
    keys = [k for k in one if not k.startswith("in_")]
    assert len(keys) >= 18
    for k in keys:
        assert one[k].shape == page[k].shape == pin[k].shape, k
        assert np.isfinite(one[k].view(np.float32)).all(), k
        # chunk boundaries change tile boundaries, never the arithmetic of an output: equal to rounding (the FFT-based
        # blocks re-block their transforms per chunk; the FIR family is bit-reproducible)
        assert_parity(page[k], one[k], "pageable chunks: " + k, 2e-6, 2e-5)
        assert_parity(pin[k], one[k], "pinned chunks: " + k, 2e-6, 2e-5)
        np.testing.assert_array_equal(pin[k], page[k], err_msg="pinned vs pageable: " + k)
    # (freq_xlating: the rotator angle of a chunk's first item is reduced mod 2 pi in float64 and handed to sincosf as a
    # float -- a different chunk start rounds that angle differently in the last bit; parity above covers it)
    for k in ("fir_ccf", "fir_fff", "pfb_streams", "pfb_interleaved", "pfbdec", "interp", "resamp", "magsq", "fbank", "v2ss"):
        np.testing.assert_array_equal(page[k], one[k], err_msg="bit-reproducible under chunking: " + k)
    # and against the oracle (spot: the blocks whose per-chunk geometry is the intricate one)
    x, taps, ptaps, ins = one["in_x"], one["in_taps"], one["in_ptaps"], one["in_ins"]
    n = 1 << 20
    assert_parity(page["fir_ccf"][-50000:], orc.fir(1, taps, x[n - 50000:], 50000), "fir tail vs oracle")
    M, T = 64, 8
    nout = n // M
    ref, _ = orc.pfb_channelizer(M, ptaps, 1.0, ins[:, nout - 2000:], 2000 - 0)
    assert_parity(page["pfb_streams"][:, nout - 2000:], ref, "channelizer tail vs oracle")
    np.testing.assert_array_equal(page["pfb_streams"], page["pfb_interleaved"])


def test_lag_by_one_fir_and_channelizer(orc):
    import gnuradio_b200 as grb

    rng = np.random.default_rng(7)
    nt, D, n, calls = 64, 2, 1 << 18, 5
    taps = rng.uniform(-1, 1, nt).astype(np.float32)
    x = crand(rng, calls * n * D + nt - 1)
    ref = grb.filter.fir_filter_ccf(D, taps).process(x, calls * n)
    blk = grb.filter.fir_filter_ccf(D, taps)
    out = np.zeros(n, np.complex64)
    got = []
    for c in range(calls):
        seg = x[c * n * D: (c + 1) * n * D + nt - 1]
        r = blk.general_work_lag(n, [seg], [out])
        assert r == (0 if c == 0 else n) and blk._consumed == n * D  # one call late
        if r:
            got.append(out[:r].copy())
    r = blk.flush([out])
    assert r == n
    got.append(out[:r].copy())
    assert blk.flush([out]) == 0
    assert_parity(np.concatenate(got), ref, "lag-by-one fir vs blocking", 2e-6, 2e-5)
    assert_parity(np.concatenate(got)[:20000], orc.fir(D, taps, x[: 20000 * D + nt - 1], 20000), "lag-by-one fir vs oracle")
    # a result that does not fit the room offered now is refused, not truncated
    blk.general_work_lag(n, [x[: n * D + nt - 1]], [out])
    with pytest.raises(ValueError):
        blk.general_work_lag(n, [x[: n * D + nt - 1]], [out[: n // 2]])

    M, T = 64, 16
    ptaps = rng.uniform(-1, 1, M * T).astype(np.float32)
    nout, calls = 1 << 12, 4
    xin = crand(rng, (calls * nout + T) * M)
    whole = grb.filter.pfb_channelizer_ccf(M, ptaps, 1.0)
    ref = np.zeros((M, calls * nout), np.complex64)
    whole.work_interleaved(xin, [ref[j] for j in range(M)], calls * nout)
    for interleaved in (True, False):
        ch = grb.filter.pfb_channelizer_ccf(M, ptaps, 1.0)
        outs = np.zeros((M, nout), np.complex64)
        ol = [outs[j] for j in range(M)]
        got = []
        for c in range(calls):
            seg = xin[c * nout * M: ((c + 1) * nout + T) * M]
            if interleaved:
                r = ch.work_lag(seg, ol, nout, interleaved=True)
            else:
                st = np.ascontiguousarray(seg.reshape(-1, M).T)
                r = ch.work_lag([st[j] for j in range(M)], ol, nout)
            assert r == (0 if c == 0 else nout)
            if r:
                got.append(outs[:, :r].copy())
        assert ch.flush(ol) == nout
        got.append(outs.copy())
        np.testing.assert_array_equal(np.concatenate(got, axis=1), ref)


def test_lag_by_one_delivers_in_place_into_page_locked_output(orc):
    """Page-locked output with room behind the items being returned: the new results are DMA'd straight to where the next call
    looks for them (no CPU copy).  The caller moves its output pointer like a gr::buffer write pointer: by what was returned."""
    import ctypes as C

    import gnuradio_b200 as grb
    from gnuradio_b200 import _lib

    rng = np.random.default_rng(17)
    nt, n, calls = 48, 1 << 17, 6
    taps = rng.uniform(-1, 1, nt).astype(np.float32)
    x = crand(rng, calls * n + nt - 1)
    ref = grb.filter.fir_filter_ccf(1, taps).process(x, calls * n)
    y = np.zeros(calls * n, np.complex64)
    grb.host_register(y)
    try:
        blk = grb.filter.fir_filter_ccf(1, taps)
        L, got, done = _lib.lib(), C.c_longlong(0), 0
        for c in range(calls):
            seg = x[c * n: (c + 1) * n + nt - 1]
            _lib.check(L.grb200_fir_work_lag(blk._h, C.c_void_p(seg.ctypes.data), n, C.c_void_p(y.ctypes.data + 8 * done),
                                             len(y) - done, C.byref(got)))
            assert got.value == (0 if c == 0 else n)
            done += got.value
        _lib.check(L.grb200_fir_flush(blk._h, C.c_void_p(y.ctypes.data + 8 * done), len(y) - done, C.byref(got)))
        done += got.value
        assert done == calls * n
        assert_parity(y, ref, "in-place lag delivery", 2e-6, 2e-5)
    finally:
        grb.host_unregister(y)
    # channelizer, 64 page-locked output streams
    M, T, nout, calls = 64, 16, 1 << 11, 5
    ptaps = rng.uniform(-1, 1, M * T).astype(np.float32)
    xin = crand(rng, (calls * nout + T) * M)
    refo = np.zeros((M, calls * nout), np.complex64)
    grb.filter.pfb_channelizer_ccf(M, ptaps, 1.0).work_interleaved(xin, [refo[j] for j in range(M)], calls * nout)
    outs = np.zeros((M, calls * nout), np.complex64)
    grb.host_register(outs)
    try:
        ch = grb.filter.pfb_channelizer_ccf(M, ptaps, 1.0)
        done = 0
        for c in range(calls):
            seg = xin[c * nout * M: ((c + 1) * nout + T) * M]
            done += ch.work_lag(seg, [outs[j][done:] for j in range(M)], nout, interleaved=True)
        done += ch.flush([outs[j][done:] for j in range(M)])
        assert done == calls * nout
        np.testing.assert_array_equal(outs, refo)
    finally:
        grb.host_unregister(outs)


def test_null_stream_is_the_legacy_default_stream(orc):
    """producer -> work_dev(stream 0) -> consumer, all on torch's default stream, with NO device-wide synchronisation: the
    result is right only if the library's launch is ordered on that stream (stream 0 used to mean a private
    non-blocking stream)."""
    import torch

    import gnuradio_b200 as grb

    rng = np.random.default_rng(3)
    nt, n = 64, 1 << 22
    taps = rng.uniform(-1, 1, nt).astype(np.float32)
    blk = grb.filter.fir_filter_ccf(1, taps)
    base = torch.from_numpy(crand(rng, n + nt - 1).view(np.float32)).cuda()
    torch.cuda.synchronize()
    for rep in range(3):
        x = base * float(rep + 1)                 # producer: a kernel on the default stream, not waited for
        y = torch.empty(2 * n, dtype=torch.float32, device="cuda")
        blk.work_dev(x.data_ptr(), y.data_ptr(), n, 0)
        z = (y * 0.5).cpu().numpy().view(np.complex64)  # consumer on the default stream; .cpu() waits for that stream only
        xin = (base.cpu().numpy() * float(rep + 1)).view(np.complex64)
        assert_parity(z[:30000], 0.5 * orc.fir(1, taps, xin[: 30000 + nt - 1], 30000), "default-stream ordering, rep %d" % rep)
        assert_parity(z[-30000:], 0.5 * orc.fir(1, taps, xin[n - 30000:], 30000), "default-stream ordering (tail), rep %d" % rep)
    # the handle's own stream is still available through the sentinel
    y = torch.empty(2 * n, dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    blk.work_dev(base.data_ptr(), y.data_ptr(), n, (1 << 64) - 1)  # GRB200_STREAM_OWN
    torch.cuda.synchronize()
    assert_parity(y.cpu().numpy().view(np.complex64)[:30000], orc.fir(1, taps, base.cpu().numpy().view(np.complex64)[: 30000 + nt - 1], 30000), "own stream")


def test_device_scope_restores_the_callers_device():
    import torch

    import gnuradio_b200 as grb

    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    torch.cuda.set_device(1)
    blk = grb.filter.fir_filter_ccf(1, [1.0, 2.0], device=0)
    out = blk.process(np.ones(101, np.complex64), 100)
    assert torch.cuda.current_device() == 1 and np.allclose(out, 3.0)
    torch.cuda.set_device(0)


def test_host_register_makes_numpy_memory_direct(orc):
    import gnuradio_b200 as grb

    rng = np.random.default_rng(5)
    nt, n = 32, 1 << 20
    taps = rng.uniform(-1, 1, nt).astype(np.float32)
    x, y = crand(rng, n + nt - 1), np.zeros(n, np.complex64)
    blk = grb.filter.fir_filter_ccf(1, taps)
    ref = blk.process(x, n)
    grb.host_register(x)
    grb.host_register(y)
    try:
        blk.work(n, [x], [y])
        np.testing.assert_array_equal(y, ref)
    finally:
        grb.host_unregister(x)
        grb.host_unregister(y)
    y[:] = 0
    blk.work(n, [x], [y])  # pageable again: the pinned-ness cache was invalidated
    np.testing.assert_array_equal(y, ref)
