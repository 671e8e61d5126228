"""A ~100-line stand-in for block_executor::run_one_iteration (gnuradio-runtime/lib/block_executor.cc:234-575)
driving ONE block between a vector_source and vector_sink(s): history-1 zeros preloaded
(lib/buffer.cc:184-195), chunks capped at `max_noutput` and rounded to the output multiple (:61-84),
fixed-rate blocks sized from the items available (:373-387), general blocks through forecast() (:423-450),
a 0 return retried (the d_updated protocol).  Test infrastructure: it calls the block's general_work only.
"""
import numpy as np


def run_sync(blk, src, max_noutput=4096, item_in=1, item_out=1, out_dtype=np.complex64):
    """vector_source -> sync block/decimator -> vector_sink.  item_in/out: scalars per item (vlen)."""
    src = np.ascontiguousarray(src)
    hist = blk.history()
    buf = np.concatenate([np.zeros((hist - 1) * item_in, src.dtype), src])
    n_items = len(buf) // item_in
    out = []
    consumed = 0
    zero_returns = 0
    while True:
        avail = n_items - consumed
        nout = blk.fixed_rate_ninput_to_noutput(avail)
        nout = min(nout, max_noutput)
        nout = nout // blk.output_multiple() * blk.output_multiple()
        if nout <= 0:
            break
        ob = np.empty(nout * item_out, out_dtype)
        r = blk.general_work(nout, [avail], [buf[consumed * item_in:]], [ob])
        if r < 0:
            break
        if r == 0:
            zero_returns += 1
            assert zero_returns < 8, "block keeps returning 0"
            if blk.history() != hist:  # re-declared history: keep the stream position, re-pad the front
                newh = blk.history()
                pos = consumed + hist - 1  # stream position of the next new item
                pad = max(0, (newh - 1) - pos)
                buf = np.concatenate([np.zeros(pad * item_in, src.dtype), buf])
                consumed = pos + pad - (newh - 1)
                n_items = len(buf) // item_in
                hist = newh
            continue
        out.append(ob[: r * item_out])
        consumed += blk._consumed
    return np.concatenate(out) if out else np.empty(0, out_dtype)


def run_general(blk, srcs, nouts, max_noutput=4096):
    """M vector_sources -> general block (pfb_channelizer_ccf) -> nouts vector_sinks."""
    M = len(srcs)
    hist = blk.history()
    bufs = [np.concatenate([np.zeros(hist - 1, np.complex64), np.ascontiguousarray(s, np.complex64)]) for s in srcs]
    n_items = min(len(b) for b in bufs)
    outs = [[] for _ in range(nouts)]
    consumed = 0
    zero_returns = 0
    while True:
        avail = n_items - consumed
        nout = max_noutput // blk.output_multiple() * blk.output_multiple()
        while nout > 0 and max(blk.forecast(nout, M)) > avail:  # block_executor.cc:444-450, stepped down
            nout -= blk.output_multiple()
        if nout <= 0:
            break
        obs = [np.empty(nout, np.complex64) for _ in range(nouts)]
        r = blk.general_work(nout, [avail] * M, [b[consumed:] for b in bufs], obs)
        if r < 0:
            break
        if r == 0:
            zero_returns += 1
            assert zero_returns < 8
            continue
        for k in range(nouts):
            outs[k].append(obs[k][:r])
        consumed += blk._consumed
    return [np.concatenate(o) if o else np.empty(0, np.complex64) for o in outs]
