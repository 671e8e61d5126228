// capi_pfb.cu -- C-ABI of pfb_channelizer_ccf (include/grb200.h): tap partition, channel map, LUT-free
// closed form of the branch rotation, host-staged and device-resident work().
#include "capi_common.cuh"

using namespace grb;

struct grb200_pfb {
    int device = 0;
    int M = 0;       // d_nfilts
    int T = 0;       // d_taps_per_filter
    int R = 0;       // d_rate_ratio
    int om = 1;      // d_output_multiple
    float osr = 1.f; // d_oversample_rate
    std::vector<float> padded; // prototype zero-padded to T*M; d_taps[i][j] = padded[i + j*M]
    std::vector<int> cmap;     // d_channel_map
    bool identity = true;
    const void* halo_src = nullptr; // left neighbour's newest T rows (peer memory), see grb200_pfb_chan_set_halo_source
    DevBuf h, wM, dmap, dinv; // dinv: output stream of DFT bin c (-1 = none), when no two outputs share a channel
    bool invertible = false;
    Stager sg;
    std::mutex mu;
};

static int pfb_upload_taps(grb200_pfb* p, const float* taps, int ntaps)
{
    // polyphase_filterbank::set_taps, gr-filter/lib/polyphase_filterbank.cc:53-81
    const int M = p->M;
    const int T = (int)ceil((double)ntaps / (double)M);
    std::vector<float> padded((size_t)T * M, 0.f);
    for (int i = 0; i < ntaps; i++)
        padded[i] = taps[i];
    int rc = p->h.upload(padded.data(), padded.size() * sizeof(float));
    if (rc)
        return rc;
    p->padded.swap(padded);
    p->T = T;
    return 0;
}

static int pfb_upload_map(grb200_pfb* p)
{
    p->identity = true;
    for (size_t i = 0; i < p->cmap.size(); i++)
        if (p->cmap[i] != (int)i)
            p->identity = false;
    if ((int)p->cmap.size() != p->M)
        p->identity = false;
    std::vector<int> inv((size_t)p->M, -1);
    p->invertible = true;
    for (size_t n = 0; n < p->cmap.size(); n++) {
        const int c = p->cmap[n];
        if (c < 0 || c >= p->M || inv[c] != -1)
            p->invertible = false;
        else
            inv[c] = (int)n;
    }
    int rc = p->dinv.upload(inv.data(), inv.size() * sizeof(int));
    if (rc)
        return rc;
    return p->dmap.upload(p->cmap.data(), p->cmap.size() * sizeof(int));
}

// Device map for a call writing `nouts` streams; a user map shorter than nouts cannot be honoured.
static int pfb_check_nouts(const grb200_pfb* p, int nouts)
{
    if (nouts < 1 || nouts > p->M)
        return fail(GRB200_EINVAL, "pfb_channelizer_ccf: %d output streams (1..%d allowed)", nouts, p->M);
    if (nouts > (int)p->cmap.size())
        return fail(GRB200_EINVAL, "pfb_channelizer_ccf: channel map has %zu entries for %d output streams",
                    p->cmap.size(), nouts);
    return 0;
}

extern "C" {

int grb200_pfb_chan_ccf_create(unsigned int numchans, const float* taps, int ntaps, float oversample_rate,
                               int device, grb200_pfb** out)
{
    if (!out)
        return fail(GRB200_EINVAL, "pfb_channelizer_ccf: null output handle");
    *out = nullptr;
    if (numchans < 1 || numchans > 65536)
        return fail(GRB200_EINVAL, "pfb_channelizer_ccf: bad channel count %u", numchans);
    if (ntaps < 1 || !taps)
        return fail(GRB200_EINVAL, "pfb_channelizer_ccf: at least one tap is required");
    if (!(oversample_rate > 0.f))
        return fail(GRB200_EINVAL, "pfb_channelizer: oversample rate must be N/i for i in [1, N]");
    // gr-filter/lib/pfb_channelizer_ccf_impl.cc:49-53
    const double srate = numchans / oversample_rate;
    const double rsrate = round(srate);
    if (fabs(srate - rsrate) > 0.00001)
        return fail(GRB200_EINVAL, "pfb_channelizer: oversample rate must be N/i for i in [1, N]");
    std::unique_ptr<grb200_pfb> p(new grb200_pfb);
    p->device = device;
    p->M = (int)numchans;
    p->osr = oversample_rate;
    p->R = (int)rintf(numchans / oversample_rate); // :73
    if (p->R < 1 || p->R > p->M)
        return fail(GRB200_EINVAL, "pfb_channelizer: oversample rate must be N/i for i in [1, N]");
    p->om = 1; // :81-84
    while ((p->om * p->R) % p->M != 0)
        p->om++;
    p->cmap.resize(numchans); // :59-63
    for (unsigned int i = 0; i < numchans; i++)
        p->cmap[i] = (int)i;
    grb::DeviceScope _ds_create(device); // the caller keeps its own current device
    int rc = p->sg.init(device);
    if (rc)
        return rc;
    std::vector<float2> w = make_roots(p->M, +1);
    if ((rc = p->wM.upload(w.data(), w.size() * sizeof(float2))) || (rc = pfb_upload_taps(p.get(), taps, ntaps)) ||
        (rc = pfb_upload_map(p.get()))) {
        grb200_pfb_chan_destroy(p.release());
        return rc;
    }
    *out = p.release();
    return 0;
}

int grb200_pfb_chan_set_taps(grb200_pfb* p, const float* taps, int ntaps)
{
    if (!p || !taps || ntaps < 1)
        return fail(GRB200_EINVAL, "pfb_channelizer_ccf::set_taps: bad arguments");
    std::lock_guard<std::mutex> lk(p->mu);
    grb::DeviceScope _ds(p->device);
    int rc = _ds.rc;
    if (rc)
        return rc;
    if ((rc = p->sg.quiesce())) // launches through this handle may still read the old state
        return rc;
    return pfb_upload_taps(p, taps, ntaps);
}

int grb200_pfb_chan_set_channel_map(grb200_pfb* p, const int* map, int n)
{
    if (!p || n < 0 || (n > 0 && !map))
        return fail(GRB200_EINVAL, "pfb_channelizer_ccf::set_channel_map: bad arguments");
    if (n == 0)
        return 0; // an empty map leaves the current one in place (:116)
    for (int i = 0; i < n; i++) // :117-121 (negative entries read out of bounds there; refused here)
        if (map[i] < 0 || map[i] >= p->M)
            return fail(GRB200_EINVAL, "pfb_channelizer_ccf_impl::set_channel_map: map range out of bounds.");
    std::lock_guard<std::mutex> lk(p->mu);
    grb::DeviceScope _ds(p->device);
    int rc = _ds.rc;
    if (rc)
        return rc;
    if ((rc = p->sg.quiesce())) // launches through this handle may still read the old state
        return rc;
    p->cmap.assign(map, map + n);
    return pfb_upload_map(p);
}

int grb200_pfb_chan_get_channel_map(const grb200_pfb* p, int* map, int n)
{
    if (!p)
        return fail(GRB200_EINVAL, "pfb_channelizer_ccf::channel_map: null handle");
    const int have = (int)p->cmap.size();
    for (int i = 0; i < n && i < have; i++)
        map[i] = p->cmap[i];
    return have;
}

int grb200_pfb_chan_set_halo_source(grb200_pfb* p, const void* peer_tail)
{
    if (!p)
        return fail(GRB200_EINVAL, "pfb_channelizer_ccf::set_halo_source: null handle");
    std::lock_guard<std::mutex> lk(p->mu);
    p->halo_src = peer_tail;
    return 0;
}

int grb200_pfb_chan_taps_per_filter(const grb200_pfb* p) { return p ? p->T : 0; }
int grb200_pfb_chan_output_multiple(const grb200_pfb* p) { return p ? p->om : 0; }

int grb200_pfb_chan_branch_taps(const grb200_pfb* p, float* out)
{
    if (!p || !out)
        return fail(GRB200_EINVAL, "pfb_channelizer_ccf::taps: bad arguments");
    for (int i = 0; i < p->M; i++) // polyphase_filterbank.cc:70-77
        for (int j = 0; j < p->T; j++)
            out[(size_t)i * p->T + j] = p->padded[(size_t)i + (size_t)j * p->M];
    return 0;
}

int grb200_pfb_chan_work_interleaved_dev(grb200_pfb* p, const void* in_dev, void* out_dev, long long out_stride,
                                         int nout_streams, long long noutput, void* stream)
{
    if (!p || noutput < 0 || (noutput > 0 && (!in_dev || !out_dev)) || out_stride < noutput)
        return fail(GRB200_EINVAL, "pfb_channelizer_ccf::work_interleaved_dev: bad arguments");
    std::lock_guard<std::mutex> lk(p->mu);
    int rc = pfb_check_nouts(p, nout_streams);
    if (rc)
        return rc;
    if (noutput % p->om != 0)
        return fail(GRB200_EINVAL, "pfb_channelizer_ccf: noutput (%lld) must be a multiple of %d", noutput, p->om);
    grb::DeviceScope _ds(p->device);
    if ((rc = _ds.rc))
        return rc;
    cudaStream_t st_ = pick_stream(stream, p->sg);
    return dev_done(pfb_launch(p->M, p->T, p->R, in_dev, out_dev, out_stride, nout_streams, (const int*)p->dmap.p,
                      p->identity ? 1 : 0, noutput, (const float*)p->h.p, (const float2*)p->wM.p,
                      sm_count(p->device), true, st_, p->halo_src, p->invertible ? (const int*)p->dinv.p : nullptr), p->sg, st_);
}

// Shared host-staged driver.  `interleaved` selects the input form:
//   0: M separate streams (the block's io_signature), de-interleaved data is re-interleaved on the device
//   1: one interleaved stream (what pfb.channelizer_ccf's stream_to_streams is fed with)
// lag != 0: queue this call, hand the PREVIOUS call's results to out_streams (room for out_cap items each); the count goes
// to *produced.  A zero-length lag call hands out what is still parked (end of stream).
static int pfb_work_host(grb200_pfb* p, int interleaved, const void* const* in_streams, const void* in_il,
                         void* const* out_streams, int nouts, long long noutput, long long* consumed, int lag = 0,
                         long long out_cap = 0, long long* produced = nullptr)
{
    if (!p || noutput < 0 || !out_streams || (noutput > 0 && (interleaved ? !in_il : !in_streams)))
        return fail(GRB200_EINVAL, "pfb_channelizer_ccf::work: bad arguments");
    std::lock_guard<std::mutex> lk(p->mu);
    int rc = pfb_check_nouts(p, nouts);
    if (rc)
        return rc;
    const int M = p->M, T = p->T, R = p->R, om = p->om;
    if (noutput % om != 0)
        return fail(GRB200_EINVAL, "pfb_channelizer_ccf: noutput (%lld) must be a multiple of %d", noutput, om);
    const long long rows_total = noutput * R / M; // toconsume, :158
    if (consumed)
        *consumed = rows_total;
    if (produced)
        *produced = lag ? 0 : noutput;
    if (noutput == 0 && !lag)
        return 0;
    grb::DeviceScope _ds(p->device);
    if ((rc = _ds.rc))
        return rc;
    if (lag) {
        const Stager::LagResult& newest = p->sg.lag[p->sg.lag_cur ^ 1];
        if (newest.pending && newest.nitems > out_cap)
            return fail(GRB200_EINVAL, "pfb_channelizer_ccf::work_lag: %lld items are waiting, room for %lld offered",
                        newest.nitems, out_cap);
        if (noutput == 0)
            return p->sg.lag_flush(out_streams, nouts, produced);
    }
    // chunk: outputs per chunk, a multiple of the output multiple so every chunk starts on a whole input row
    long long chunk = (long long)(chunk_bytes_for((size_t)noutput * R * 8) / 8) / R;
    chunk = std::max<long long>(1, chunk / om) * om;
    if (chunk > noutput)
        chunk = noutput;
    const long long nchunks = (noutput + chunk - 1) / chunk;
    const long long rows_chunk = T + chunk * R / M;
    Stager& sg = p->sg;
    if ((rc = sg.reserve_dev(0, (size_t)rows_chunk * M * 8)) || (rc = sg.reserve_dev(1, (size_t)chunk * nouts * 8)) ||
        (!interleaved && (rc = sg.reserve_dev(2, (size_t)rows_chunk * M * 8))))
        return rc;
    if (lag) {
        const Stager::LagResult& newest = sg.lag[sg.lag_cur ^ 1];
        const long long nprev = newest.pending ? newest.nitems : 0;
        if ((rc = sg.lag_begin((size_t)noutput * nouts * 8, nchunks * nouts, noutput)))
            return rc;
        if (nprev + noutput <= out_cap) // room behind the previous call's items in every output stream: deliver in place
            sg.lag_direct(out_streams, nouts, (size_t)nprev * 8);
    }
    rc = sg.run(
        nchunks,
        [&](long long c, int slot, Stager::Chunk& ch) {
            const long long o0 = c * chunk, n = std::min(chunk, noutput - o0);
            const long long row0 = o0 * R / M, rows = T + n * R / M;
            if (interleaved) {
                ch.h2d.push_back({ (const float2*)in_il + row0 * M, sg.dev[slot][0], (size_t)rows * M * 8 });
            } else {
                for (int j = 0; j < M; j++)
                    ch.h2d.push_back({ (const float2*)in_streams[j] + row0,
                                       (float2*)sg.dev[slot][2] + (size_t)j * rows_chunk, (size_t)rows * 8 });
            }
            for (int k = 0; k < nouts; k++)
                ch.d2h.push_back(Piece((const float2*)sg.dev[slot][1] + (size_t)k * chunk, (float2*)out_streams[k] + o0,
                                       (size_t)n * 8, k, (size_t)o0 * 8));
            return 0;
        },
        [&](long long c, int slot, cudaStream_t st) {
            const long long o0 = c * chunk, n = std::min(chunk, noutput - o0);
            const long long rows = T + n * R / M;
            if (!interleaved) {
                int r2 = streams_to_interleaved_launch(sg.dev[slot][2], rows_chunk, sg.dev[slot][0], M, rows, st);
                if (r2)
                    return r2;
            }
            return pfb_launch(M, T, R, sg.dev[slot][0], sg.dev[slot][1], chunk, nouts, (const int*)p->dmap.p,
                              p->identity ? 1 : 0, n, (const float*)p->h.p, (const float2*)p->wM.p,
                              sm_count(p->device), true, st, nullptr, p->invertible ? (const int*)p->dinv.p : nullptr);
        });
    if (!rc && lag)
        rc = sg.lag_collect(out_streams, nouts, produced);
    return rc;
}

int grb200_pfb_chan_work_lag(grb200_pfb* p, const void* const* in_streams, long long noutput, void* const* out_streams,
                             int nout_streams, long long out_capacity, long long* consumed, long long* produced)
{
    if (!produced || out_capacity < 0)
        return fail(GRB200_EINVAL, "pfb_channelizer_ccf::work_lag: bad arguments");
    return pfb_work_host(p, 0, in_streams, nullptr, out_streams, nout_streams, noutput, consumed, 1, out_capacity, produced);
}

int grb200_pfb_chan_work_interleaved_lag(grb200_pfb* p, const void* in_host, long long noutput, void* const* out_streams,
                                         int nout_streams, long long out_capacity, long long* consumed, long long* produced)
{
    if (!produced || out_capacity < 0)
        return fail(GRB200_EINVAL, "pfb_channelizer_ccf::work_lag: bad arguments");
    return pfb_work_host(p, 1, nullptr, in_host, out_streams, nout_streams, noutput, consumed, 1, out_capacity, produced);
}

int grb200_pfb_chan_flush(grb200_pfb* p, void* const* out_streams, int nout_streams, long long out_capacity,
                          long long* produced)
{
    if (!produced || out_capacity < 0)
        return fail(GRB200_EINVAL, "pfb_channelizer_ccf::flush: bad arguments");
    return pfb_work_host(p, 1, nullptr, nullptr, out_streams, nout_streams, 0, nullptr, 1, out_capacity, produced);
}

int grb200_pfb_chan_work(grb200_pfb* p, const void* const* in_streams, void* const* out_streams, int nout_streams,
                         long long noutput, long long* consumed)
{
    return pfb_work_host(p, 0, in_streams, nullptr, out_streams, nout_streams, noutput, consumed);
}

int grb200_pfb_chan_work_interleaved(grb200_pfb* p, const void* in_host, void* const* out_streams, int nout_streams,
                                     long long noutput, long long* consumed)
{
    return pfb_work_host(p, 1, nullptr, in_host, out_streams, nout_streams, noutput, consumed);
}

void grb200_pfb_chan_destroy(grb200_pfb* p)
{
    if (!p)
        return;
    grb::DeviceScope _ds(p->device);
    p->sg.destroy();
    p->h.release();
    p->wM.release();
    p->dmap.release();
    p->dinv.release();
    delete p;
}

} // extern "C"
