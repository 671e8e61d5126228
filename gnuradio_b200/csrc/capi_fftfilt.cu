// capi_fftfilt.cu -- C-ABI of fft_filter_ccc (include/grb200.h).  Host side only prepares taps and carries
// the ntaps-1 item overlap between work() calls; the convolution itself runs in fftfilt.cu.
#include "capi_common.cuh"

using namespace grb;

struct grb200_fftfilt {
    int device = 0;
    int decim = 1;
    int ntaps = 0;
    int nsamples = 0; // the reference's block length = the block's output_multiple (fft_filter.cc:208-209)
    int L = 0;        // device FFT length
    int parts = 1;    // tap partitions (fftfilt_device_plan)
    DevBuf H, tw;
    std::vector<float2> hist; // last ntaps-1 input items of the previous work() call (the d_tail equivalent)
    Stager sg;
    std::mutex mu;
};

// reference geometry, gr-filter/lib/fft_filter.cc:204-209
static int ref_nsamples(int ntaps)
{
    const int fftsize = (int)(2 * pow(2.0, ceil(log((double)ntaps) / log(2.0))));
    return fftsize - ntaps + 1;
}

namespace grb {
// Device plan + filter spectra for `ntaps` complex taps (also used by the FIR handle when it routes a long filter here):
// H_p = FFT_L(taps of partition p / L), scale on the host like fft_filter.cc:184-189, transform on the device.
// *L_io: the length `tw` currently holds roots for (0 = none); updated.  Returns GRB200_ERANGE when no plan exists.
int fftfilt_plan_build(const float* taps_re_im, int ntaps, int device, cudaStream_t st, DevBuf& H, DevBuf& tw, int* L_io,
                       int* K_out, int force_L)
{
    int K = 1;
    const int L = force_L > 0 ? force_L : fftfilt_device_plan(ntaps, &K); // (a forced length is a one-partition plan)
    if (L < 0)
        return fail(GRB200_ERANGE, "fft_filter_ccc: %d taps exceed the single-CTA transform (max %d)", ntaps,
                    16384 * 3 / 4 + 1);
    int rc;
    if (L != *L_io) {
        std::vector<float2> roots = make_roots(L, -1);
        if ((rc = tw.upload(roots.data(), roots.size() * sizeof(float2))))
            return rc;
        *L_io = L;
    }
    // One partition holds all taps; K hold taps [p L/2, (p+1) L/2).
    const int P = K == 1 ? ntaps : L / 2;
    std::vector<float2> padded((size_t)L * K, make_float2(0.f, 0.f));
    const float scale = 1.0f / (float)L;
    for (int i = 0; i < ntaps; i++)
        padded[(size_t)(i / P) * L + i % P] = make_float2(taps_re_im[2 * i] * scale, taps_re_im[2 * i + 1] * scale);
    DevBuf tmp;
    if ((rc = tmp.upload(padded.data(), padded.size() * sizeof(float2))))
        return rc;
    if ((rc = H.upload(padded.data(), padded.size() * sizeof(float2)))) { // sized; overwritten below
        tmp.release();
        return rc;
    }
    rc = fft_vcc_launch(L, true, tmp.p, H.p, K, nullptr, 0, 0, (const float2*)tw.p, sm_count(device), st);
    if (!rc) {
        cudaError_t e = cudaStreamSynchronize(st);
        if (e != cudaSuccess)
            rc = fail(GRB200_ECUDA, "fft_filter_ccc::set_taps: %s", cudaGetErrorString(e));
    }
    tmp.release();
    if (rc)
        return rc;
    *K_out = K;
    return 0;
}
} // namespace grb

static int fftfilt_prepare(grb200_fftfilt* h, const float* taps_re_im, int ntaps)
{
    int K = 1;
    int rc = fftfilt_plan_build(taps_re_im, ntaps, h->device, h->sg.s_exec, h->H, h->tw, &h->L, &K);
    if (rc)
        return rc;
    h->parts = K;
    h->ntaps = ntaps;
    h->nsamples = ref_nsamples(ntaps);
    h->hist.assign((size_t)(ntaps - 1), make_float2(0.f, 0.f)); // set_taps zeroes the overlap state (:177-179)
    return 0;
}

extern "C" {

int grb200_fftfilt_ccc_create(int decim, const float* taps_re_im, int ntaps, int device, grb200_fftfilt** out)
{
    if (!out)
        return fail(GRB200_EINVAL, "fft_filter_ccc: null output handle");
    *out = nullptr;
    if (decim < 1)
        return fail(GRB200_EINVAL, "fft_filter_ccc: decimation must be >= 1 (got %d)", decim);
    if (ntaps < 1 || !taps_re_im)
        return fail(GRB200_EINVAL, "fft_filter_ccc: at least one tap is required");
    std::unique_ptr<grb200_fftfilt> h(new grb200_fftfilt);
    h->device = device;
    h->decim = decim;
    grb::DeviceScope _ds_create(device); // the caller keeps its own current device
    int rc = h->sg.init(device);
    if (rc)
        return rc;
    if ((rc = fftfilt_prepare(h.get(), taps_re_im, ntaps))) {
        grb200_fftfilt_destroy(h.release());
        return rc;
    }
    *out = h.release();
    return 0;
}

int grb200_fftfilt_set_taps(grb200_fftfilt* h, const float* taps_re_im, int ntaps)
{
    if (!h || !taps_re_im || ntaps < 1)
        return fail(GRB200_EINVAL, "fft_filter_ccc::set_taps: bad arguments");
    std::lock_guard<std::mutex> lk(h->mu);
    grb::DeviceScope _ds(h->device);
    int rc = _ds.rc;
    if (rc)
        return rc;
    if ((rc = h->sg.quiesce())) // launches through this handle may still read the old state
        return rc;
    return fftfilt_prepare(h, taps_re_im, ntaps);
}

static std::vector<float> real_taps_as_complex(const float* taps, int ntaps)
{
    std::vector<float> t((size_t)2 * (ntaps > 0 ? ntaps : 0), 0.f);
    for (int i = 0; i < ntaps; i++)
        t[2 * (size_t)i] = taps[i]; // (t, 0): fft_filter.cc:317-320
    return t;
}

int grb200_fftfilt_ccf_create(int decim, const float* taps, int ntaps, int device, grb200_fftfilt** out)
{
    if (ntaps < 1 || !taps) {
        if (out)
            *out = nullptr;
        return fail(GRB200_EINVAL, "fft_filter_ccf: at least one tap is required");
    }
    const std::vector<float> t = real_taps_as_complex(taps, ntaps);
    return grb200_fftfilt_ccc_create(decim, t.data(), ntaps, device, out);
}

int grb200_fftfilt_set_taps_real(grb200_fftfilt* h, const float* taps, int ntaps)
{
    if (!h || !taps || ntaps < 1)
        return fail(GRB200_EINVAL, "fft_filter_ccf::set_taps: bad arguments");
    const std::vector<float> t = real_taps_as_complex(taps, ntaps);
    return grb200_fftfilt_set_taps(h, t.data(), ntaps);
}

int grb200_fftfilt_nsamples(const grb200_fftfilt* h) { return h ? h->nsamples : 0; }
int grb200_fftfilt_ntaps(const grb200_fftfilt* h) { return h ? h->ntaps : 0; }

int grb200_fftfilt_work_dev(grb200_fftfilt* h, const void* in_dev, void* out_dev, long long noutput, void* stream)
{
    if (!h || noutput < 0 || (noutput > 0 && (!in_dev || !out_dev)))
        return fail(GRB200_EINVAL, "fft_filter_ccc::work_dev: bad arguments");
    std::lock_guard<std::mutex> lk(h->mu);
    grb::DeviceScope _ds(h->device);
    int rc = _ds.rc;
    if (rc)
        return rc;
    cudaStream_t st_ = pick_stream(stream, h->sg);
    return dev_done(fftfilt_launch(h->L, h->parts, in_dev, out_dev, noutput, h->ntaps, h->decim, (const float2*)h->H.p,
                          (const float2*)h->tw.p, sm_count(h->device), st_), h->sg, st_);
}

int grb200_fftfilt_work(grb200_fftfilt* h, const void* in_host, void* out_host, long long noutput)
{
    if (!h || noutput < 0 || (noutput > 0 && (!in_host || !out_host)))
        return fail(GRB200_EINVAL, "fft_filter_ccc::work: bad arguments");
    if (noutput == 0)
        return 0;
    std::lock_guard<std::mutex> lk(h->mu);
    if (noutput % h->nsamples != 0)
        return fail(GRB200_EINVAL, "fft_filter_ccc::work: noutput (%lld) must be a multiple of nsamples (%d)", noutput,
                    h->nsamples);
    grb::DeviceScope _ds(h->device);
    int rc = _ds.rc;
    if (rc)
        return rc;
    const int D = h->decim, nh = h->ntaps - 1;
    const long long V = h->parts > 1 ? h->L / 2 : h->L - nh; // the device kernel's hop
    // chunk = whole device blocks, and whole decimation periods so that every chunk starts on phase 0
    long long per = (long long)(chunk_bytes_for((size_t)noutput * D * 8) / 8) / D; // outputs per chunk (target)
    long long unit = V;                               // outputs*D per chunk must be a multiple of V for no waste
    long long chunk = std::max<long long>(1, per / unit) * unit; // outputs per chunk
    if (chunk > noutput)
        chunk = noutput;
    const long long nchunks = (noutput + chunk - 1) / chunk;
    if ((rc = h->sg.reserve_dev(0, (size_t)(chunk * D + nh) * 8)) || (rc = h->sg.reserve_dev(1, (size_t)chunk * 8)))
        return rc;
    const float2* in = (const float2*)in_host;
    float2* out = (float2*)out_host;
    Stager& sg = h->sg;
    const float2* hist = h->hist.data();
    rc = sg.run(
        nchunks,
        [&](long long c, int slot, Stager::Chunk& ch) {
            const long long o0 = c * chunk, n = std::min(chunk, noutput - o0);
            const long long i0 = o0 * D; // first new input item of this chunk
            float2* d = (float2*)sg.dev[slot][0];
            if (i0 >= nh) {
                ch.h2d.push_back({ in + (i0 - nh), d, (size_t)(n * D + nh) * 8 });
            } else {
                const long long from_hist = nh - i0; // items still coming from the carried history
                ch.h2d.push_back({ hist + i0, d, (size_t)from_hist * 8 });
                ch.h2d.push_back({ in, d + from_hist, (size_t)(n * D + i0) * 8 });
            }
            ch.d2h.push_back({ sg.dev[slot][1], out + o0, (size_t)n * 8 });
            return 0;
        },
        [&](long long c, int slot, cudaStream_t st) {
            const long long o0 = c * chunk, n = std::min(chunk, noutput - o0);
            return fftfilt_launch(h->L, h->parts, sg.dev[slot][0], sg.dev[slot][1], n, h->ntaps, D, (const float2*)h->H.p,
                                  (const float2*)h->tw.p, sm_count(h->device), st);
        });
    if (rc)
        return rc;
    // carry the last ntaps-1 items of [hist | in] into the next call
    const long long nin = noutput * D;
    if (nh > 0) {
        if (nin >= nh) {
            memcpy(h->hist.data(), in + (nin - nh), (size_t)nh * 8);
        } else {
            memmove(h->hist.data(), h->hist.data() + nin, (size_t)(nh - nin) * 8);
            memcpy(h->hist.data() + (nh - nin), in, (size_t)nin * 8);
        }
    }
    return 0;
}

void grb200_fftfilt_destroy(grb200_fftfilt* h)
{
    if (!h)
        return;
    grb::DeviceScope _ds(h->device);
    h->sg.destroy();
    h->H.release();
    h->tw.release();
    delete h;
}

} // extern "C"
