// capi_pfbsynth.cu -- C-ABI of pfb_synthesizer_ccf (include/grb200.h): tap partitions, channel map, the filters' delay-line
// state as the last T-1 items of every input stream, host-staged and device-resident work().
#include "capi_common.cuh"

using namespace grb;

struct grb200_pfbsynth {
    int device = 0;
    int M = 0;        // d_numchans
    int twox = 1;     // d_twox
    int T = 0;        // d_taps_per_filter
    std::vector<float> padded;   // prototype zero padded to T*M
    std::vector<int> cmap;       // d_channel_map (twox*M entries by default)
    int par = 0;                 // d_state: parity of the input items consumed so far (2x mode)
    int hist_nin = 0;            // input streams whose history is valid
    DevBuf h, wM, twNF, dinv, hist, scratch;
    long long hist_stride = 0;   // items between streams in `hist`
    Stager sg;
    std::mutex mu;
    int NF() const { return twox * M; }
};

static int synth_upload_taps(grb200_pfbsynth* p, const float* taps, int ntaps)
{
    // set_taps1 / set_taps2, gr-filter/lib/pfb_synthesizer_ccf_impl.cc:95-184: both partitions are h[i + j*M] of the
    // zero padded prototype (2x splits even / odd j over the filter pair, which the kernels fold into the bin choice)
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
    // set_taps re-creates every filter's delay line zeroed (fir_filter_with_buffer.cc:79-86): the history starts over
    p->hist_stride = ((long long)std::max(T - 1, 1) + 1) & ~1LL;
    std::vector<float2> z((size_t)p->NF() * p->hist_stride, make_float2(0.f, 0.f));
    if ((rc = p->hist.upload(z.data(), z.size() * sizeof(float2))))
        return rc;
    p->hist_nin = 0;
    return 0;
}

static int synth_upload_map(grb200_pfbsynth* p)
{
    // work(): d_fft->get_inbuf()[d_channel_map[i]] = in_i[n] for i < ninputs in order (:236-239) -- a later stream wins
    std::vector<int> inv((size_t)p->NF(), -1);
    for (size_t i = 0; i < p->cmap.size() && i < (size_t)p->NF(); i++)
        inv[p->cmap[i]] = (int)i;
    return p->dinv.upload(inv.data(), inv.size() * sizeof(int));
}

// inverse map restricted to the `nin` connected inputs
static int synth_upload_map_nin(grb200_pfbsynth* p, int nin)
{
    std::vector<int> inv((size_t)p->NF(), -1);
    for (int i = 0; i < nin && i < (int)p->cmap.size(); i++)
        inv[p->cmap[i]] = i;
    return p->dinv.upload(inv.data(), inv.size() * sizeof(int));
}

static int synth_run_dev(grb200_pfbsynth* p, const void* x, long long stride, int nin, void* out, long long ntimes, int par0,
                         cudaStream_t st)
{
    const int nsm = sm_count(p->device);
    if (pfb_synth_fast_supported(p->M, p->T, p->twox, x, stride, nin))
        return pfb_synth_fast_launch(p->M, p->T, x, stride, nin, out, ntimes, (const float*)p->h.p, (const float2*)p->wM.p,
                                     (const int*)p->dinv.p, nsm, st);
    const int NF = p->NF();
    if (!fft_size_is_fast(NF) && NF > fft_direct_max())
        return fail(GRB200_ERANGE, "pfb_synthesizer_ccf: %d channels are not supported by the device transform", NF);
    long long slab = pfb_synth_slab(NF, p->T);
    if (slab > ntimes + p->T - 1)
        slab = ntimes + p->T - 1;
    const size_t need = (size_t)2 * slab * NF * 8;
    if (need > p->scratch.cap) {
        GRB_CUDA(cudaStreamSynchronize(st));
        p->scratch.release();
        void* q = nullptr;
        cudaError_t e = cudaMalloc(&q, need);
        if (e != cudaSuccess)
            return fail(GRB200_ENOMEM, "cudaMalloc(%zu) failed: %s", need, cudaGetErrorString(e));
        p->scratch.p = q;
        p->scratch.cap = need;
    }
    return pfb_synth_general_launch(p->M, p->twox, p->T, x, stride, out, ntimes, (const float*)p->h.p,
                                    (const float2*)p->twNF.p, (const int*)p->dinv.p, par0, p->scratch.p, slab, nsm, st);
}

extern "C" {

int grb200_pfb_synth_ccf_create(unsigned int numchans, const float* taps, int ntaps, int twox, int device,
                                grb200_pfbsynth** out)
{
    if (!out)
        return fail(GRB200_EINVAL, "pfb_synthesizer_ccf: null output handle");
    *out = nullptr;
    if (numchans < 1 || numchans > 65536)
        return fail(GRB200_EINVAL, "pfb_synthesizer_ccf: bad channel count %u", numchans);
    if (ntaps < 1 || !taps)
        return fail(GRB200_EINVAL, "pfb_synthesizer_ccf: at least one tap is required");
    const int tw = twox ? 2 : 1;
    if (numchans % tw != 0) // :44-47
        return fail(GRB200_EINVAL, "pfb_synthesizer_ccf_impl: number of channels must be even for 2x oversampling.");
    std::unique_ptr<grb200_pfbsynth> p(new grb200_pfbsynth);
    p->device = device;
    p->M = (int)numchans;
    p->twox = tw;
    p->cmap.resize((size_t)tw * numchans); // :50,58
    for (size_t i = 0; i < p->cmap.size(); i++)
        p->cmap[i] = (int)i;
    grb::DeviceScope _ds_create(device); // the caller keeps its own current device
    int rc = p->sg.init(device);
    if (rc)
        return rc;
    std::vector<float2> w = make_roots(p->M, +1), wf = make_roots(p->NF(), -1);
    if ((rc = p->wM.upload(w.data(), w.size() * sizeof(float2))) || (rc = p->twNF.upload(wf.data(), wf.size() * sizeof(float2))) ||
        (rc = synth_upload_taps(p.get(), taps, ntaps)) || (rc = synth_upload_map(p.get())))
        return rc;
    *out = p.release();
    return 0;
}

int grb200_pfb_synth_set_taps(grb200_pfbsynth* p, const float* taps, int ntaps)
{
    if (!p || !taps || ntaps < 1)
        return fail(GRB200_EINVAL, "pfb_synthesizer_ccf::set_taps: bad arguments");
    std::lock_guard<std::mutex> lk(p->mu);
    grb::DeviceScope _ds(p->device);
    int rc = _ds.rc;
    if (rc)
        return rc;
    if ((rc = p->sg.quiesce()))
        return rc;
    return synth_upload_taps(p, taps, ntaps);
}

int grb200_pfb_synth_set_channel_map(grb200_pfbsynth* p, const int* map, int n)
{
    if (!p || n < 0 || (n > 0 && !map))
        return fail(GRB200_EINVAL, "pfb_synthesizer_ccf::set_channel_map: bad arguments");
    if (n == 0) // an empty map is ignored (:206)
        return 0;
    for (int i = 0; i < n; i++) // :207-212
        if (map[i] < 0 || map[i] >= p->NF())
            return fail(GRB200_EINVAL, "pfb_synthesizer_ccf_impl::set_channel_map: map range out of bounds.");
    std::lock_guard<std::mutex> lk(p->mu);
    grb::DeviceScope _ds(p->device);
    int rc = _ds.rc;
    if (rc)
        return rc;
    if ((rc = p->sg.quiesce()))
        return rc;
    p->cmap.assign(map, map + n);
    return synth_upload_map(p);
}

int grb200_pfb_synth_get_channel_map(const grb200_pfbsynth* p, int* map, int n)
{
    if (!p || !map || n < (int)p->cmap.size())
        return fail(GRB200_EINVAL, "pfb_synthesizer_ccf::channel_map: bad arguments");
    for (size_t i = 0; i < p->cmap.size(); i++)
        map[i] = p->cmap[i];
    return (int)p->cmap.size();
}

int grb200_pfb_synth_taps_per_filter(const grb200_pfbsynth* p) { return p ? p->T : 0; }
int grb200_pfb_synth_numchans(const grb200_pfbsynth* p) { return p ? p->M : 0; }
int grb200_pfb_synth_twox(const grb200_pfbsynth* p) { return p ? p->twox : 0; }

// d_taps as the reference's taps() returns them: twox*numchans rows of taps_per_filter (:95-184)
int grb200_pfb_synth_branch_taps(const grb200_pfbsynth* p, float* out)
{
    if (!p || !out)
        return fail(GRB200_EINVAL, "pfb_synthesizer_ccf::taps: bad arguments");
    const int M = p->M, T = p->T;
    for (int i = 0; i < M; i++)
        for (int j = 0; j < T; j++) {
            const float v = p->padded[(size_t)i + (size_t)j * M];
            if (p->twox == 1) {
                out[(size_t)i * T + j] = v;
            } else { // state toggles with j: even j -> filter M+i, odd j -> filter i
                out[(size_t)i * T + j] = (j & 1) ? v : 0.f;
                out[(size_t)(M + i) * T + j] = (j & 1) ? 0.f : v;
            }
        }
    return 0;
}

// Stateless device form: nin streams, stream i at x_dev + i*stride items, each holding noutput/numchans + T-1 items with the
// T-1 history items in front.  `parity` is d_state at the first output time (2x mode; ignored in 1x).
int grb200_pfb_synth_work_dev(grb200_pfbsynth* p, const void* x_dev, long long stride, int nin, void* out_dev,
                              long long noutput, int parity, void* stream)
{
    if (!p || noutput < 0 || nin < 1 || nin > p->NF() || (noutput > 0 && (!x_dev || !out_dev)))
        return fail(GRB200_EINVAL, "pfb_synthesizer_ccf::work_dev: bad arguments");
    if (noutput % p->M != 0)
        return fail(GRB200_EINVAL, "pfb_synthesizer_ccf: noutput (%lld) must be a multiple of %d", noutput, p->M);
    std::lock_guard<std::mutex> lk(p->mu);
    grb::DeviceScope _ds(p->device);
    int rc = _ds.rc;
    if (rc)
        return rc;
    if ((rc = synth_upload_map_nin(p, nin)))
        return rc;
    cudaStream_t st_ = pick_stream(stream, p->sg);
    return dev_done(synth_run_dev(p, x_dev, stride, nin, out_dev, noutput / p->M, parity & 1, st_), p->sg, st_);
}

// Host form = the block's work(): in_streams[i] (i < nin) holds noutput/numchans NEW items (history 1); the delay lines
// of the reference's filters live in the handle as the last T-1 items of every stream.
int grb200_pfb_synth_work(grb200_pfbsynth* p, const void* const* in_streams, int nin, void* out_host, long long noutput)
{
    if (!p || noutput < 0 || nin < 1 || nin > p->NF() || !in_streams || (noutput > 0 && !out_host))
        return fail(GRB200_EINVAL, "pfb_synthesizer_ccf::work: bad arguments");
    if (noutput % p->M != 0)
        return fail(GRB200_EINVAL, "pfb_synthesizer_ccf: noutput (%lld) must be a multiple of %d", noutput, p->M);
    if (noutput == 0)
        return 0;
    std::lock_guard<std::mutex> lk(p->mu);
    grb::DeviceScope _ds(p->device);
    int rc = _ds.rc;
    if (rc)
        return rc;
    if ((rc = synth_upload_map_nin(p, nin)))
        return rc;
    const int M = p->M, T = p->T, H = T - 1;
    const long long ntimes = noutput / M;
    long long chunk = std::max<long long>(1, (long long)(chunk_bytes_for((size_t)noutput * 8) / 8) / M); // times per chunk
    if (chunk > ntimes)
        chunk = ntimes;
    const long long nchunks = (ntimes + chunk - 1) / chunk;
    const long long stride = (chunk + H + 1) & ~1LL; // items between streams in the slot (even: 16-byte aligned rows)
    Stager& sg = p->sg;
    if ((rc = sg.reserve_dev(0, (size_t)stride * nin * 8)) || (rc = sg.reserve_dev(1, (size_t)chunk * M * 8)))
        return rc;
    if (nin > p->hist_nin) { // streams connected for the first time start from a zero delay line
        if (H > 0)
            GRB_CUDA(cudaMemset2DAsync((float2*)p->hist.p + (size_t)p->hist_nin * p->hist_stride, p->hist_stride * 8, 0,
                                       (size_t)H * 8, nin - p->hist_nin, sg.s_exec));
        p->hist_nin = nin;
    }
    const int par_start = p->par;
    rc = sg.run(
        nchunks,
        [&](long long c, int slot, Stager::Chunk& ch) {
            const long long t0 = c * chunk, n = std::min(chunk, ntimes - t0);
            for (int i = 0; i < nin; i++)
                ch.h2d.push_back({ (const float2*)in_streams[i] + t0, (float2*)sg.dev[slot][0] + (size_t)i * stride + H,
                                   (size_t)n * 8 });
            ch.d2h.push_back({ sg.dev[slot][1], (float2*)out_host + (size_t)t0 * M, (size_t)n * M * 8 });
            return 0;
        },
        [&](long long c, int slot, cudaStream_t st) {
            const long long t0 = c * chunk, n = std::min(chunk, ntimes - t0);
            float2* x = (float2*)sg.dev[slot][0];
            if (H > 0) // history in front of every stream
                GRB_CUDA(cudaMemcpy2DAsync(x, stride * 8, p->hist.p, p->hist_stride * 8, (size_t)H * 8, nin,
                                           cudaMemcpyDeviceToDevice, st));
            int r2 = synth_run_dev(p, x, stride, nin, sg.dev[slot][1], n, (par_start + (int)(t0 & 1)) & 1, st);
            if (r2)
                return r2;
            if (H > 0) // the newest T-1 items of (history ++ chunk) are the next call's history
                GRB_CUDA(cudaMemcpy2DAsync(p->hist.p, p->hist_stride * 8, x + n, stride * 8, (size_t)H * 8, nin,
                                           cudaMemcpyDeviceToDevice, st));
            return 0;
        });
    if (rc)
        return rc;
    p->par = (par_start + (int)(ntimes & 1)) & 1;
    return 0;
}

void grb200_pfb_synth_destroy(grb200_pfbsynth* p)
{
    if (!p)
        return;
    grb::DeviceScope _ds(p->device);
    p->sg.destroy();
    p->h.release();
    p->wM.release();
    p->twNF.release();
    p->dinv.release();
    p->hist.release();
    p->scratch.release();
    delete p;
}

} // extern "C"
