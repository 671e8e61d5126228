"""Time-segment sharding of the hot-path blocks across the GPUs of one box (host logic, no compute).

SURVEY.md section 8(e): every block on the path shards by time segment.  Rank g owns a contiguous run of
OUTPUT items; the input it needs is its own segment plus a history halo that is the left neighbour's newest
items (fir: ntaps-1, gr-filter/lib/fir_filter_blk_impl.cc:41; channelizer: taps_per_filter rows of numchans
items, gr-filter/lib/pfb_channelizer_ccf_impl.cc:101; fft_filter: ntaps-1, the d_tail of
gr-filter/lib/fft_filter.cc:262-267; fft_vcc: none).  The halo is the only exchange step the path has:
one send/recv pair per rank and step over NCCL (gloo in the CPU tests).
"""
from dataclasses import dataclass


@dataclass(frozen=True)
class Segment:
    rank: int
    world: int
    out_begin: int   # first output item owned
    out_end: int     # one past the last output item owned
    in_begin: int    # first NEW input item consumed (halo excluded)
    in_end: int
    halo: int        # history items in front of in_begin that come from the left neighbour (zeros for rank 0)

    @property
    def noutput(self):
        return self.out_end - self.out_begin

    @property
    def ninput(self):
        return self.in_end - self.in_begin


def split_outputs(noutput_total, world, multiple=1):
    """Cut [0, noutput_total) into `world` contiguous runs whose lengths are multiples of `multiple`
    (the block's output_multiple); the remainder units go to the lowest ranks."""
    if world < 1 or multiple < 1:
        raise ValueError("split_outputs: world and multiple must be >= 1")
    if noutput_total % multiple:
        raise ValueError("split_outputs: %d outputs is not a multiple of %d" % (noutput_total, multiple))
    units = noutput_total // multiple
    base, extra = divmod(units, world)
    bounds = [0]
    for r in range(world):
        bounds.append(bounds[-1] + (base + (1 if r < extra else 0)) * multiple)
    return bounds


def segment(rank, world, noutput_total, decimation=1, history=1, multiple=1, items_per_output=1):
    """Segment of a sync_decimator-like block: output i consumes input items
    [i*decimation*items_per_output, ...) plus (history-1)*items_per_output items of history.
    items_per_output = numchans for the interleaved channelizer stream, fft_size is NOT used here
    (fft_vcc items are whole vectors)."""
    b = split_outputs(noutput_total, world, multiple)
    step = decimation * items_per_output
    return Segment(rank, world, b[rank], b[rank + 1], b[rank] * step, b[rank + 1] * step,
                   (history - 1) * items_per_output)


def exchange_halos(dist, pairs, rank, world, ring=False):
    """exchange_halo for several (head, tail) pairs of one step in ONE batched send/recv group (one NCCL group launch
    instead of len(pairs): the host-side cost of a halo exchange is per group, not per byte)."""
    if world == 1:
        return
    right, left = rank + 1, rank - 1
    if ring:
        right, left = right % world, left % world
    ops = []
    for head, tail in pairs:
        if head.numel() == 0:
            continue
        if right < world:
            ops.append(dist.P2POp(dist.isend, tail, right))
        if left >= 0:
            ops.append(dist.P2POp(dist.irecv, head, left))
    if ops:
        for r in dist.batch_isend_irecv(ops):
            r.wait()


def exchange_halo(dist, head, tail, rank, world, ring=False):
    """Fill `head` (this rank's halo region, in front of its segment) with the left neighbour's `tail`
    (its newest halo-many items).  Rank 0 keeps what it has (zeros = the scheduler's preloaded history)
    unless ring=True (bench: every rank does the same work).  Tensors may be CUDA (nccl) or CPU (gloo)."""
    if world == 1 or head.numel() == 0:
        return
    right, left = rank + 1, rank - 1
    ops = []
    if ring:
        right, left = right % world, left % world
    if right < world:
        ops.append(dist.P2POp(dist.isend, tail, right))
    if left >= 0:
        ops.append(dist.P2POp(dist.irecv, head, left))
    if ops:
        for r in dist.batch_isend_irecv(ops):
            r.wait()


@dataclass(frozen=True)
class ResamplerSegment:
    rank: int
    world: int
    out_begin: int   # first output item owned
    out_end: int
    in_begin: int    # input item the first owned output's window starts at
    in_end: int      # one past the last input item any owned window reads (taps_per_filter-1 items of look-ahead included)
    ctr: int         # d_ctr at out_begin: the rank sets it with set_ctr() before its first call
    ninput: int      # ninput_items[0] for the rank's call: in_end - in_begin - (taps_per_filter-1)

    @property
    def noutput(self):
        return self.out_end - self.out_begin


def resampler_segment(rank, world, noutput_total, interpolation, decimation, taps_per_filter, ctr0=0):
    """Segment of rational_resampler_base (gr-filter/lib/rational_resampler_base_impl.cc:156-172): output i filters the
    window that starts at input (ctr0 + i*decimation) / interpolation with phase (ctr0 + i*decimation) % interpolation,
    so a rank needs no neighbour data -- it reads its own span of the source, look-ahead included, and takes the phase
    counter over in closed form.  No exchange step."""
    b = split_outputs(noutput_total, world)
    o0, o1 = b[rank], b[rank + 1]
    s0 = ctr0 + o0 * decimation
    q0 = s0 // interpolation
    if o1 == o0:
        return ResamplerSegment(rank, world, o0, o1, q0, q0, s0 % interpolation, 0)
    q_last = (ctr0 + (o1 - 1) * decimation) // interpolation
    return ResamplerSegment(rank, world, o0, o1, q0, q_last + taps_per_filter, s0 % interpolation, q_last + 1 - q0)


class PeerHalo:
    """The halo without a host round trip per step (include/grb200.h "peer memory"): every rank exports the address of its
    segment's TAIL (its newest `halo` items) once, opens the left neighbour's, and hands that pointer to its block
    (set_halo_source) -- from then on the block fetches its history over NVLink itself, on its own stream.

    One process per GPU (torchrun): CUDA IPC handles travel through torch.distributed.all_gather_object (72 bytes each).
    ring=True lets rank 0 read the last rank's tail (bench: every rank does the same work)."""

    def __init__(self, dist, device, tail_ptrs, rank, world, ring=False):
        import ctypes as C

        from . import _lib

        self._device, self._open = int(device), []
        self.ptrs = [None] * len(tail_ptrs)
        if world == 1:
            return
        L = _lib.lib()
        mine = []
        for p in tail_ptrs:
            blob = (C.c_ubyte * 72)()
            _lib.check(L.grb200_ipc_export(self._device, C.c_void_p(int(p)), blob))
            mine.append(bytes(blob))
        everyone = [None] * world
        dist.all_gather_object(everyone, mine)
        left = rank - 1
        if ring:
            left %= world
        if left < 0:
            return
        for k, blob in enumerate(everyone[left]):
            q = C.c_void_p()
            _lib.check(L.grb200_ipc_open(self._device, (C.c_ubyte * 72).from_buffer_copy(blob), C.byref(q)))
            self._open.append(q)
            self.ptrs[k] = q.value

    def close(self):
        from . import _lib

        for q in self._open:
            _lib.lib().grb200_ipc_close(self._device, q)
        self._open = []
