"""GPU (2 or more): the device-side halo of a time-sharded stream (include/grb200.h "peer memory").  One process, two GPUs:
grb200_enable_peer + set_halo_source; the right segment's history comes out of the left GPU's memory, and the two segments
together equal the unsharded stream.  (The multi-process form -- CUDA IPC handles through torch.distributed -- is what
bench.py runs under torchrun; its host logic is covered on gloo in tests/test_multi_gloo.py.)"""
import ctypes as C

import numpy as np
import pytest

from common import assert_parity, crand

pytestmark = pytest.mark.gpu


def test_two_gpu_segments_with_peer_halo(orc):
    import torch

    import gnuradio_b200 as grb
    from gnuradio_b200 import _lib, sharding

    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    L = _lib.lib()
    _lib.check(L.grb200_enable_peer(1, 0))
    rng = np.random.default_rng(1)
    # fir_filter_ccf: two segments, the right one's ntaps-1 history items fetched from GPU 0
    nt, n = 64, 1 << 20
    taps = rng.uniform(-1, 1, nt).astype(np.float32)
    x = np.concatenate([np.zeros(nt - 1, np.complex64), crand(rng, 2 * n)])
    ref = orc.fir(1, taps, x, 2 * n)
    seg = [sharding.segment(r, 2, 2 * n, 1, nt) for r in range(2)]
    d_in, d_out, blk = [], [], []
    for r in range(2):
        with torch.cuda.device(r):
            xi = x[seg[r].in_begin: seg[r].in_end + nt - 1].copy()
            if r == 1:
                xi[: nt - 1] = 0  # NOT provided: the block must fetch it from the neighbour
            d_in.append(torch.from_numpy(xi.view(np.float32)).cuda(r))
            d_out.append(torch.empty(2 * n, dtype=torch.float32, device="cuda:%d" % r))
            blk.append(grb.filter.fir_filter_ccf(1, taps, device=r))
    tail0 = d_in[0].data_ptr() + 8 * n  # GPU 0's newest ntaps-1 items
    blk[1].set_halo_source(tail0)
    for r in range(2):
        with torch.cuda.device(r):
            blk[r].work_dev(d_in[r].data_ptr(), d_out[r].data_ptr(), n)
    for r in range(2):
        torch.cuda.synchronize(r)
    got = np.concatenate([d_out[r].cpu().numpy().view(np.complex64) for r in range(2)])
    assert_parity(got, ref, "two FIR segments, peer halo")
    # pfb_channelizer_ccf streaming kernel: the T history rows are read in place from GPU 0
    M, T, nout = 64, 16, 1 << 14
    ptaps = rng.uniform(-1, 1, M * T).astype(np.float32)
    xs = np.concatenate([np.zeros(T * M, np.complex64), crand(rng, 2 * nout * M)])
    whole = np.zeros((M, 2 * nout), np.complex64)
    grb.filter.pfb_channelizer_ccf(M, ptaps, 1.0).work_interleaved(xs, [whole[j] for j in range(M)], 2 * nout)
    outs = []
    ins = []
    for r in range(2):
        with torch.cuda.device(r):
            xi = xs[r * nout * M: ((r + 1) * nout + T) * M].copy()
            if r == 1:
                xi[: T * M] = 0
            ins.append(torch.from_numpy(xi.view(np.float32)).cuda(r))
            o = torch.empty(2 * M * nout, dtype=torch.float32, device="cuda:%d" % r)
            ch = grb.filter.pfb_channelizer_ccf(M, ptaps, 1.0, device=r)
            if r == 1:
                ch.set_halo_source(ins[0].data_ptr() + 8 * nout * M)
            ch.work_interleaved_dev(ins[r].data_ptr(), o.data_ptr(), nout, M, nout)
            outs.append(o)
    for r in range(2):
        torch.cuda.synchronize(r)
    got = np.concatenate([outs[r].cpu().numpy().view(np.complex64).reshape(M, nout) for r in range(2)], axis=1)
    np.testing.assert_array_equal(got, whole)


@pytest.mark.parametrize("case", ["fir_overlap_save", "pfb_oversampled", "fir_fff"])
def test_halo_source_on_the_paths_that_copy_it_in_first(orc, case):
    """One GPU is enough to check the halo plumbing of the kernels that do NOT read the neighbour's memory in place (the
    overlap-save route of long FIR filters, the oversampled channelizer, fir_filter_fff): the history of the right segment is
    left zero in its own buffer, the halo source points at the left segment's tail (same device here), and the two segments
    together must equal the unsharded stream."""
    import torch

    import gnuradio_b200 as grb

    rng = np.random.default_rng(9)
    st = torch.cuda.current_stream().cuda_stream
    if case in ("fir_overlap_save", "fir_fff"):
        real = case == "fir_fff"
        nt, n = 301, 1 << 17
        taps = (rng.uniform(-1, 1, nt) / nt).astype(np.float32)
        data = rng.uniform(-1, 1, 2 * n).astype(np.float32) if real else crand(rng, 2 * n)
        x = np.concatenate([np.zeros(nt - 1, data.dtype), data])
        cls = grb.filter.fir_filter_fff if real else grb.filter.fir_filter_ccf
        isz = 4 if real else 8
        whole = cls(1, taps).process(x, 2 * n)
        seg0 = torch.from_numpy(x[: n + nt - 1].copy()).cuda()
        x1 = x[n: 2 * n + nt - 1].copy()
        x1[: nt - 1] = 0
        seg1 = torch.from_numpy(x1).cuda()
        outs = [torch.empty(n, dtype=seg0.dtype, device="cuda") for _ in range(2)]
        b0, b1 = cls(1, taps), cls(1, taps)
        b1.set_halo_source(seg0.data_ptr() + isz * n)
        b0.work_dev(seg0.data_ptr(), outs[0].data_ptr(), n, st)
        b1.work_dev(seg1.data_ptr(), outs[1].data_ptr(), n, st)
        torch.cuda.synchronize()
        got = np.concatenate([o.cpu().numpy() for o in outs])
        assert_parity(got, whole, case)
    else:
        M, T, nout = 32, 12, 1 << 13  # oversample rate 2: nout outputs consume nout / 2 rows
        ptaps = rng.uniform(-1, 1, M * T).astype(np.float32)
        rows = nout  # two segments of nout outputs
        xs = np.concatenate([np.zeros(T * M, np.complex64), crand(rng, rows * M)])
        whole = np.zeros((M, 2 * nout), np.complex64)
        grb.filter.pfb_channelizer_ccf(M, ptaps, 2.0).work_interleaved(xs, [whole[j] for j in range(M)], 2 * nout)
        half = rows // 2
        i0 = torch.from_numpy(xs[: (half + T) * M].copy().view(np.float32)).cuda()
        x1 = xs[half * M: (rows + T) * M].copy()
        x1[: T * M] = 0
        i1 = torch.from_numpy(x1.view(np.float32)).cuda()
        outs = [torch.empty(2 * M * nout, dtype=torch.float32, device="cuda") for _ in range(2)]
        c0, c1 = grb.filter.pfb_channelizer_ccf(M, ptaps, 2.0), grb.filter.pfb_channelizer_ccf(M, ptaps, 2.0)
        c1.set_halo_source(i0.data_ptr() + 8 * half * M)
        c0.work_interleaved_dev(i0.data_ptr(), outs[0].data_ptr(), nout, M, nout, st)
        c1.work_interleaved_dev(i1.data_ptr(), outs[1].data_ptr(), nout, M, nout, st)
        torch.cuda.synchronize()
        got = np.concatenate([o.cpu().numpy().view(np.complex64).reshape(M, nout) for o in outs], axis=1)
        assert_parity(got, whole, case)
