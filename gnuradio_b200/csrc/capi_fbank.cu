// capi_fbank.cu -- C-ABI of filterbank_vcvcf (include/grb200.h).
#include "capi_common.cuh"

using namespace grb;

struct grb200_fbank {
    int device = 0;
    int nf = 0, T = 0;
    DevBuf hrev, active;
    Stager sg;
    std::mutex mu;
};

// kernel::filterbank::set_taps, gr-filter/lib/filterbank.cc:50-79: d_active[i] = any tap != 0; taps stored reversed
// (fir_filter::set_taps, fir_filter.cc:62-64); device layout [t][nfilts] so a warp's tap loads are coalesced
static int fbank_upload(grb200_fbank* h, int nf, int T, const float* taps)
{
    std::vector<float> hrev((size_t)T * nf);
    std::vector<unsigned char> act((size_t)nf, 0);
    for (int i = 0; i < nf; i++)
        for (int t = 0; t < T; t++) {
            const float v = taps[(size_t)i * T + (T - 1 - t)];
            hrev[(size_t)t * nf + i] = v;
            if (v != 0)
                act[i] = 1;
        }
    int rc = h->hrev.upload(hrev.data(), hrev.size() * sizeof(float));
    if (rc)
        return rc;
    if ((rc = h->active.upload(act.data(), act.size())))
        return rc;
    h->nf = nf;
    h->T = T;
    return 0;
}

extern "C" {

int grb200_filterbank_vcvcf_create(int nfilts, int ntaps, const float* taps, int device, grb200_fbank** out)
{
    if (!out)
        return fail(GRB200_EINVAL, "filterbank_vcvcf: null output handle");
    *out = nullptr;
    if (nfilts < 1 || !taps) // filterbank.cc:30-32
        return fail(GRB200_EINVAL, "The taps vector may not be empty.");
    if (ntaps < 1)
        return fail(GRB200_EINVAL, "filterbank_vcvcf: at least one tap per filter is required");
    std::unique_ptr<grb200_fbank> h(new grb200_fbank);
    h->device = device;
    grb::DeviceScope _ds_create(device); // the caller keeps its own current device
    int rc = h->sg.init(device);
    if (rc)
        return rc;
    if ((rc = fbank_upload(h.get(), nfilts, ntaps, taps)))
        return rc;
    *out = h.release();
    return 0;
}

int grb200_filterbank_vcvcf_set_taps(grb200_fbank* h, int nfilts, int ntaps, const float* taps)
{
    if (!h || !taps || ntaps < 1)
        return fail(GRB200_EINVAL, "filterbank_vcvcf::set_taps: bad arguments");
    if (nfilts != h->nf) // filterbank.cc:54-56 (std::runtime_error)
        return fail(GRB200_ERUNTIME, "The number of filters is incorrect.");
    std::lock_guard<std::mutex> lk(h->mu);
    grb::DeviceScope _ds(h->device);
    int rc = _ds.rc;
    if (rc)
        return rc;
    if ((rc = h->sg.quiesce()))
        return rc;
    return fbank_upload(h, nfilts, ntaps, taps);
}

int grb200_filterbank_vcvcf_nfilts(const grb200_fbank* h) { return h ? h->nf : 0; }
int grb200_filterbank_vcvcf_ntaps(const grb200_fbank* h) { return h ? h->T : 0; }

int grb200_filterbank_vcvcf_work_dev(grb200_fbank* h, const void* in_dev, void* out_dev, long long noutput, void* stream)
{
    if (!h || noutput < 0 || (noutput > 0 && (!in_dev || !out_dev)))
        return fail(GRB200_EINVAL, "filterbank_vcvcf::work_dev: bad arguments");
    std::lock_guard<std::mutex> lk(h->mu);
    grb::DeviceScope _ds(h->device);
    int rc = _ds.rc;
    if (rc)
        return rc;
    cudaStream_t st_ = pick_stream(stream, h->sg);
    return dev_done(filterbank_launch(in_dev, out_dev, noutput, h->nf, h->T, (const float*)h->hrev.p,
                                      (const unsigned char*)h->active.p, sm_count(h->device), st_),
                    h->sg, st_);
}

// in_host: (noutput + ntaps) vectors, the first ntaps being history (set_history(ntaps + 1), filterbank_vcvcf_impl.cc:34)
int grb200_filterbank_vcvcf_work(grb200_fbank* h, const void* in_host, void* out_host, long long noutput)
{
    if (!h || noutput < 0 || (noutput > 0 && (!in_host || !out_host)))
        return fail(GRB200_EINVAL, "filterbank_vcvcf::work: bad arguments");
    if (noutput == 0)
        return 0;
    std::lock_guard<std::mutex> lk(h->mu);
    grb::DeviceScope _ds(h->device);
    int rc = _ds.rc;
    if (rc)
        return rc;
    const size_t vsz = (size_t)h->nf * 8;
    const int T = h->T;
    long long chunk = std::max<long long>(1, (long long)(chunk_bytes_for((size_t)noutput * vsz) / vsz));
    if (chunk > noutput)
        chunk = noutput;
    const long long nchunks = (noutput + chunk - 1) / chunk;
    if ((rc = h->sg.reserve_dev(0, (size_t)(chunk + T) * vsz)) || (rc = h->sg.reserve_dev(1, (size_t)chunk * vsz)))
        return rc;
    const char* in = (const char*)in_host;
    char* out = (char*)out_host;
    Stager& sg = h->sg;
    return sg.run(
        nchunks,
        [&](long long c, int slot, Stager::Chunk& ch) {
            const long long j0 = c * chunk, n = std::min(chunk, noutput - j0);
            ch.h2d.push_back({ in + (size_t)j0 * vsz, sg.dev[slot][0], (size_t)(n + T - 1) * vsz });
            ch.d2h.push_back({ sg.dev[slot][1], out + (size_t)j0 * vsz, (size_t)n * vsz });
            return 0;
        },
        [&](long long c, int slot, cudaStream_t st) {
            const long long j0 = c * chunk, n = std::min(chunk, noutput - j0);
            return filterbank_launch(sg.dev[slot][0], sg.dev[slot][1], n, h->nf, T, (const float*)h->hrev.p,
                                     (const unsigned char*)h->active.p, sm_count(h->device), st);
        });
}

void grb200_filterbank_vcvcf_destroy(grb200_fbank* h)
{
    if (!h)
        return;
    grb::DeviceScope _ds(h->device);
    h->sg.destroy();
    h->hrev.release();
    h->active.release();
    delete h;
}

} // extern "C"
