"""N>1 host logic on CPU: time-segment sharding + history halo exchange (gnuradio_b200/sharding.py),
world_size 2 and 3 over gloo.  The per-rank "block" is the CPU oracle (tests may call it): what is under test is
that segments + halo reproduce the single-stream result item for item, i.e. the arithmetic bench.py and a
multi-GPU flowgraph rely on."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from gnuradio_b200 import sharding  # noqa: E402


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _c64_to_t(a):
    return torch.from_numpy(np.ascontiguousarray(a).view(np.float32).copy())


def _worker(rank, world, port, case, outdir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import pyoracle as po

        orc = po.Port()
        rng = np.random.default_rng(1234)  # every rank draws the same whole stream, then keeps its segment only
        if case == "fir":
            ntaps, D, nout_total = 37, 3, 6000
            taps = rng.uniform(-1, 1, ntaps).astype(np.float32)
            x = (rng.uniform(-1, 1, nout_total * D) + 1j * rng.uniform(-1, 1, nout_total * D)).astype(np.complex64)
            seg = sharding.segment(rank, world, nout_total, decimation=D, history=ntaps)
            buf = np.zeros(seg.halo + seg.ninput, np.complex64)  # rank 0: zeros = preloaded history
            buf[seg.halo:] = x[seg.in_begin:seg.in_end]
            tbuf = _c64_to_t(buf)
            head, tail = tbuf[: 2 * seg.halo], tbuf[-2 * seg.halo:].clone()
            sharding.exchange_halo(dist, head, tail, rank, world)
            y = orc.fir(D, taps, tbuf.numpy().view(np.complex64), seg.noutput)
            full = orc.fir_stream(D, taps, x)
        elif case == "pfb":
            M, T, nout_total = 8, 5, 900
            taps = rng.uniform(-1, 1, M * T).astype(np.float32)
            x = (rng.uniform(-1, 1, nout_total * M) + 1j * rng.uniform(-1, 1, nout_total * M)).astype(np.complex64)
            seg = sharding.segment(rank, world, nout_total, history=T + 1, items_per_output=M)
            buf = np.zeros(seg.halo + seg.ninput, np.complex64)
            buf[seg.halo:] = x[seg.in_begin:seg.in_end]
            tbuf = _c64_to_t(buf)
            head, tail = tbuf[: 2 * seg.halo], tbuf[-2 * seg.halo:].clone()
            # batched variant, the halo cut in two pairs (what bench.py does for several buffer sets per step)
            h = seg.halo
            sharding.exchange_halos(dist, [(head[:h], tail[:h]), (head[h:], tail[h:])], rank, world)
            ins = np.ascontiguousarray(tbuf.numpy().view(np.complex64).reshape(-1, M).T)
            y, consumed = orc.pfb_channelizer(M, taps, 1.0, ins, seg.noutput)
            assert consumed == seg.noutput
            xf = np.concatenate([np.zeros(T * M, np.complex64), x])
            full, _ = orc.pfb_channelizer(M, taps, 1.0, np.ascontiguousarray(xf.reshape(-1, M).T), nout_total)
            full = full[:, seg.out_begin:seg.out_end]
            np.save(os.path.join(outdir, "ok_%d.npy" % rank), np.array([np.array_equal(y, full)]))
            return
        elif case == "xlat":
            # freq_xlating_fir_filter: FIR halo as above + the rotator taken over at the segment's first output
            # (grb200_freq_xlating_fir_advance on the device; the phase handed to the oracle here)
            ntaps, D, nout_total, fc, fs = 29, 4, 3000, 1.5e3, 48e3
            taps = rng.uniform(-1, 1, ntaps).astype(np.float32)
            x = (rng.uniform(-1, 1, nout_total * D) + 1j * rng.uniform(-1, 1, nout_total * D)).astype(np.complex64)
            seg = sharding.segment(rank, world, nout_total, decimation=D, history=ntaps)
            buf = np.zeros(seg.halo + seg.ninput, np.complex64)
            buf[seg.halo:] = x[seg.in_begin:seg.in_end]
            tbuf = _c64_to_t(buf)
            head, tail = tbuf[: 2 * seg.halo], tbuf[-2 * seg.halo:].clone()
            sharding.exchange_halo(dist, head, tail, rank, world)
            delta = po.xlat_increment_angle(np.float32(2 * np.pi * fc / fs), D)
            y, _ = orc.freq_xlating_fir(D, taps, fc, fs, tbuf.numpy().view(np.complex64), seg.noutput,
                                        phase=np.exp(1j * delta * seg.out_begin))
            xin = np.concatenate([np.zeros(ntaps - 1, np.complex64), x])
            full, _ = orc.freq_xlating_fir(D, taps, fc, fs, xin, nout_total)
            # the float phasor recurrence restarted mid-stream rounds differently: equal to a few ulp, not bit for bit
            ok = np.allclose(y, full[seg.out_begin:seg.out_end], atol=2e-5, rtol=0)
            np.save(os.path.join(outdir, "ok_%d.npy" % rank), np.array([ok]))
            return
        elif case == "pfbdec":
            # pfb_decimator_ccf: M input streams, each sharded in time with its own T-1 item halo
            M, T, nout_total = 6, 7, 1200
            taps = rng.uniform(-1, 1, M * T - 3).astype(np.float32)
            xs = (rng.uniform(-1, 1, (M, nout_total)) + 1j * rng.uniform(-1, 1, (M, nout_total))).astype(np.complex64)
            seg = sharding.segment(rank, world, nout_total, history=T)
            bufs = []
            for m in range(M):
                b = np.zeros(seg.halo + seg.ninput, np.complex64)
                b[seg.halo:] = xs[m, seg.in_begin:seg.in_end]
                bufs.append(_c64_to_t(b))
            pairs = [(t[: 2 * seg.halo], t[-2 * seg.halo:].clone()) for t in bufs]
            sharding.exchange_halos(dist, pairs, rank, world)
            streams = np.stack([t.numpy().view(np.complex64) for t in bufs])
            # the oracle wrapper prepends T-1 zeros itself: drop the outputs those produce, keep the segment's
            y = orc.pfb_decimator(M, taps, 2, streams)[seg.halo:]
            full = orc.pfb_decimator(M, taps, 2, xs)
        elif case == "interp":
            # interp_fir_filter: sharded on INPUT positions (T-1 item halo); rank r owns outputs [in_begin*L, in_end*L)
            L, ntaps, nin_total = 4, 30, 2000
            taps = rng.uniform(-1, 1, ntaps).astype(np.float32)
            T = -(-ntaps // L)
            x = (rng.uniform(-1, 1, nin_total) + 1j * rng.uniform(-1, 1, nin_total)).astype(np.complex64)
            seg = sharding.segment(rank, world, nin_total, history=T)
            buf = np.zeros(seg.halo + seg.ninput, np.complex64)
            buf[seg.halo:] = x[seg.in_begin:seg.in_end]
            tbuf = _c64_to_t(buf)
            head, tail = tbuf[: 2 * seg.halo], tbuf[-2 * seg.halo:].clone()
            sharding.exchange_halo(dist, head, tail, rank, world)
            y = orc.interp_fir(L, taps, tbuf.numpy().view(np.complex64))[seg.halo * L:]
            full = orc.interp_fir(L, taps, x)
            np.save(os.path.join(outdir, "ok_%d.npy" % rank),
                    np.array([np.array_equal(y, full[seg.out_begin * L:seg.out_end * L])]))
            return
        else:
            raise AssertionError(case)
        np.save(os.path.join(outdir, "ok_%d.npy" % rank),
                np.array([np.array_equal(y, full[seg.out_begin:seg.out_end])]))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
@pytest.mark.parametrize("case", ["fir", "pfb", "xlat", "pfbdec", "interp"])
def test_segments_plus_halo_equal_single_stream(tmp_path, world, case):
    mp.spawn(_worker, args=(world, _free_port(), case, str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        assert bool(np.load(tmp_path / ("ok_%d.npy" % r))[0]), "rank %d differs from the single-stream result" % r


def test_split_outputs_respects_multiples():
    b = sharding.split_outputs(4097 * 10, 4, multiple=4097)
    assert b == [0, 3 * 4097, 6 * 4097, 8 * 4097, 10 * 4097]
    with pytest.raises(ValueError):
        sharding.split_outputs(100, 3, multiple=7)
    s = sharding.segment(1, 2, 1000, decimation=8, history=1024)
    assert (s.out_begin, s.out_end, s.in_begin, s.in_end, s.halo) == (500, 1000, 4000, 8000, 1023)
    s = sharding.segment(0, 8, 1 << 24, history=17, items_per_output=64)
    assert s.halo == 16 * 64 and s.ninput == (1 << 27)


@pytest.mark.parametrize("world", [2, 3, 8])
@pytest.mark.parametrize("L,D", [(3, 2), (2, 3), (160, 147), (1, 5), (7, 1)])
def test_resampler_segments_equal_single_stream(orc, world, L, D):
    # rational_resampler_base shards on outputs with no exchange: window start and d_ctr of a rank's first output are
    # closed-form (sharding.resampler_segment); the ranks' outputs concatenate to the single-stream result bit for bit
    rng = np.random.default_rng(world * 1000 + L * 10 + D)
    taps = rng.uniform(-1, 1, 61).astype(np.float32)
    T = -(-61 // L)
    x = (rng.uniform(-1, 1, 3000) + 1j * rng.uniform(-1, 1, 3000)).astype(np.complex64)
    port = __import__("oracle.pyoracle", fromlist=["Port"]).Port()  # the C port sums in one order whatever the alignment
    full, _, _ = port.rational_resampler(L, D, taps, x)
    parts = []
    for r in range(world):
        seg = sharding.resampler_segment(r, world, len(full), L, D, T)
        assert seg.in_end <= len(x) and 0 <= seg.ctr < L
        y, used, _ = port.rational_resampler(L, D, taps, x[seg.in_begin:seg.in_end], noutput=seg.noutput, ctr=seg.ctr,
                                             nin=seg.ninput)
        assert len(y) == seg.noutput
        parts.append(y)
    assert np.array_equal(np.concatenate(parts), full)
