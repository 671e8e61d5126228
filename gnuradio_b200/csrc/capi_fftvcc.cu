// capi_fftvcc.cu -- C-ABI of fft_vcc (+ fused / standalone complex_to_mag_squared), include/grb200.h.
#include "capi_common.cuh"

using namespace grb;

struct grb200_fftvcc {
    int device = 0;
    int N = 0;
    bool forward = true;
    bool shift = false;
    bool has_window = false;
    bool real_in = false; // fft_vfc: float32 input vectors
    DevBuf window, tw;
    Stager sg;
    std::mutex mu;
    size_t in_item() const { return real_in ? 4 : 8; }
};

static int vcc_set_window(grb200_fftvcc* h, const float* window, int nwindow)
{
    // fft_vcc_fftw::set_window, gr-fft/lib/fft_vcc_fftw.cc:54-61: empty or exactly fft_size, else refuse
    if (nwindow == 0) {
        h->has_window = false;
        return 0;
    }
    if (nwindow != h->N || !window)
        return fail(GRB200_EWINDOW, "fft_vcc: window not the same length as fft_size");
    int rc = h->window.upload(window, (size_t)nwindow * sizeof(float));
    if (rc)
        return rc;
    h->has_window = true;
    return 0;
}

static int vcc_run_dev(grb200_fftvcc* h, const void* in, void* out, long long nvec, int magsq, cudaStream_t st)
{
    VccArgs a{ in, out, nvec, h->has_window ? (const float*)h->window.p : nullptr, (const float2*)h->tw.p, h->shift ? 1 : 0,
               magsq ? 1 : 0, 1, 0, 0.f, 0.f, nullptr };
    return fft_vcc_launch_ex(h->N, h->forward, h->real_in, a, sm_count(h->device), st);
}

static int vcc_work_host(grb200_fftvcc* h, const void* in_host, void* out_host, long long nvectors, int magsq)
{
    if (!h || nvectors < 0 || (nvectors > 0 && (!in_host || !out_host)))
        return fail(GRB200_EINVAL, "fft_vcc::work: bad arguments");
    if (nvectors == 0)
        return 0;
    std::lock_guard<std::mutex> lk(h->mu);
    grb::DeviceScope _ds(h->device);
    int rc = _ds.rc;
    if (rc)
        return rc;
    const size_t vin = (size_t)h->N * h->in_item(), vout = (size_t)h->N * (magsq ? 4 : 8);
    long long chunk = std::max<long long>(1, (long long)(chunk_bytes_for((size_t)nvectors * vin) / vin));
    if (chunk > nvectors)
        chunk = nvectors;
    const long long nchunks = (nvectors + chunk - 1) / chunk;
    if ((rc = h->sg.reserve_dev(0, (size_t)chunk * vin)) || (rc = h->sg.reserve_dev(1, (size_t)chunk * vout)))
        return rc;
    const char* in = (const char*)in_host;
    char* out = (char*)out_host;
    Stager& sg = h->sg;
    return sg.run(
        nchunks,
        [&](long long c, int slot, Stager::Chunk& ch) {
            const long long v0 = c * chunk, n = std::min(chunk, nvectors - v0);
            ch.h2d.push_back({ in + (size_t)v0 * vin, sg.dev[slot][0], (size_t)n * vin });
            ch.d2h.push_back({ sg.dev[slot][1], out + (size_t)v0 * vout, (size_t)n * vout });
            return 0;
        },
        [&](long long c, int slot, cudaStream_t st) {
            const long long v0 = c * chunk, n = std::min(chunk, nvectors - v0);
            return vcc_run_dev(h, sg.dev[slot][0], sg.dev[slot][1], n, magsq, st);
        });
}

extern "C" void grb200_fft_vcc_destroy(grb200_fftvcc* h);

static int vcc_create(int fft_size, int forward, const float* window, int nwindow, int shift, int real_in, int device,
                      grb200_fftvcc** out)
{
    if (!out)
        return fail(GRB200_EINVAL, "fft_vcc: null output handle");
    *out = nullptr;
    if (fft_size <= 0) // fft_complex ctor, gr-fft/lib/fft.cc:161-163 -> std::out_of_range
        return fail(GRB200_ERANGE, "fft_impl_fftw: invalid fft_size");
    if (!fft_size_is_fast(fft_size) && !fft_size_is_big(fft_size) && fft_size > fft_direct_max())
        return fail(GRB200_ERANGE, "fft_vcc: fft_size %d is not supported by the device transform", fft_size);
    if (nwindow < 0)
        return fail(GRB200_EINVAL, "fft_vcc: negative window length");
    std::unique_ptr<grb200_fftvcc> h(new grb200_fftvcc);
    h->device = device;
    h->N = fft_size;
    h->forward = forward != 0;
    h->shift = shift != 0;
    h->real_in = real_in != 0;
    grb::DeviceScope _ds_create(device); // the caller keeps its own current device
    int rc = h->sg.init(device);
    if (rc)
        return rc;
    std::vector<float2> roots = make_roots(fft_size, -1);
    if ((rc = h->tw.upload(roots.data(), roots.size() * sizeof(float2))) ||
        (rc = vcc_set_window(h.get(), window, nwindow))) {
        grb200_fft_vcc_destroy(h.release());
        return rc;
    }
    *out = h.release();
    return 0;
}

extern "C" {

int grb200_fft_vcc_create(int fft_size, int forward, const float* window, int nwindow, int shift, int device,
                          grb200_fftvcc** out)
{
    return vcc_create(fft_size, forward, window, nwindow, shift, 0, device, out);
}
int grb200_fft_vfc_create(int fft_size, int forward, const float* window, int nwindow, int device, grb200_fftvcc** out)
{
    return vcc_create(fft_size, forward, window, nwindow, 0, 1, device, out);
}

int grb200_fft_vcc_set_window(grb200_fftvcc* h, const float* window, int nwindow)
{
    if (!h || nwindow < 0)
        return fail(GRB200_EINVAL, "fft_vcc::set_window: bad arguments");
    std::lock_guard<std::mutex> lk(h->mu);
    grb::DeviceScope _ds(h->device);
    int rc = _ds.rc;
    if (rc)
        return rc;
    if (nwindow != 0 && nwindow != h->N)
        return fail(GRB200_EWINDOW, "fft_vcc: window not the same length as fft_size");
    if ((rc = h->sg.quiesce())) // launches through this handle may still read the old state
        return rc;
    return vcc_set_window(h, window, nwindow);
}

int grb200_fft_vcc_fft_size(const grb200_fftvcc* h) { return h ? h->N : 0; }

int grb200_fft_vcc_work(grb200_fftvcc* h, const void* in_host, void* out_host, long long nvectors)
{
    return vcc_work_host(h, in_host, out_host, nvectors, 0);
}
int grb200_fft_vcc_work_magsq(grb200_fftvcc* h, const void* in_host, float* out_host, long long nvectors)
{
    return vcc_work_host(h, in_host, out_host, nvectors, 1);
}

int grb200_fft_vcc_work_dev(grb200_fftvcc* h, const void* in_dev, void* out_dev, long long nvectors, void* stream)
{
    if (!h || nvectors < 0 || (nvectors > 0 && (!in_dev || !out_dev)))
        return fail(GRB200_EINVAL, "fft_vcc::work_dev: bad arguments");
    std::lock_guard<std::mutex> lk(h->mu);
    grb::DeviceScope _ds(h->device);
    int rc = _ds.rc;
    if (rc)
        return rc;
    cudaStream_t st_ = pick_stream(stream, h->sg);
    return dev_done(vcc_run_dev(h, in_dev, out_dev, nvectors, 0, st_), h->sg, st_);
}
int grb200_fft_vcc_work_magsq_dev(grb200_fftvcc* h, const void* in_dev, float* out_dev, long long nvectors,
                                  void* stream)
{
    if (!h || nvectors < 0 || (nvectors > 0 && (!in_dev || !out_dev)))
        return fail(GRB200_EINVAL, "fft_vcc::work_magsq_dev: bad arguments");
    std::lock_guard<std::mutex> lk(h->mu);
    grb::DeviceScope _ds(h->device);
    int rc = _ds.rc;
    if (rc)
        return rc;
    cudaStream_t st_ = pick_stream(stream, h->sg);
    return dev_done(vcc_run_dev(h, in_dev, out_dev, nvectors, 1, st_), h->sg, st_);
}

void grb200_fft_vcc_destroy(grb200_fftvcc* h)
{
    if (!h)
        return;
    grb::DeviceScope _ds(h->device);
    h->sg.destroy();
    h->window.release();
    h->tw.release();
    delete h;
}

// ---- complex_to_mag_squared standalone (gr-blocks/lib/complex_to_mag_squared_impl.cc:37-48) ----------
int grb200_magsq_work_dev(int device, const void* in_dev, float* out_dev, long long nitems, void* stream)
{
    if (nitems < 0 || (nitems > 0 && (!in_dev || !out_dev)))
        return fail(GRB200_EINVAL, "complex_to_mag_squared::work_dev: bad arguments");
    Stager* sg = nullptr;
    int rc = shared_stager(device, &sg);
    if (rc)
        return rc;
    cudaStream_t st_ = pick_stream(stream, *sg);
    return dev_done(magsq_launch(in_dev, out_dev, nitems, sm_count(device), st_), *sg, st_);
}

int grb200_magsq_work(int device, const void* in_host, float* out_host, long long nitems)
{
    if (nitems < 0 || (nitems > 0 && (!in_host || !out_host)))
        return fail(GRB200_EINVAL, "complex_to_mag_squared::work: bad arguments");
    if (nitems == 0)
        return 0;
    Stager* sgp = nullptr;
    int rc = shared_stager(device, &sgp);
    if (rc)
        return rc;
    std::lock_guard<std::mutex> lk(shared_stager_mutex(device)); // the per-device stager is shared: one host-staged call at a time
    grb::DeviceScope _ds(device);
    Stager& sg = *sgp;
    long long chunk = std::max<long long>(1, (long long)(chunk_bytes_for((size_t)nitems * 8) / 8));
    if (chunk > nitems)
        chunk = nitems;
    const long long nchunks = (nitems + chunk - 1) / chunk;
    if ((rc = sg.reserve_dev(0, (size_t)chunk * 8)) || (rc = sg.reserve_dev(1, (size_t)chunk * 4)))
        return rc;
    const char* in = (const char*)in_host;
    char* out = (char*)out_host;
    return sg.run(
        nchunks,
        [&](long long c, int slot, Stager::Chunk& ch) {
            const long long i0 = c * chunk, n = std::min(chunk, nitems - i0);
            ch.h2d.push_back({ in + (size_t)i0 * 8, sg.dev[slot][0], (size_t)n * 8 });
            ch.d2h.push_back({ sg.dev[slot][1], out + (size_t)i0 * 4, (size_t)n * 4 });
            return 0;
        },
        [&](long long c, int slot, cudaStream_t st) {
            const long long i0 = c * chunk, n = std::min(chunk, nitems - i0);
            return magsq_launch(sg.dev[slot][0], (float*)sg.dev[slot][1], n, sm_count(device), st);
        });
}

} // extern "C"

// =====================================================================================================
// logpwrfft_c / logpwrfft_f  (gr-fft/python/fft/logpwrfft.py:29-64)
// =====================================================================================================
// The hier block stream_to_vector_decimator -> fft_vcc|fft_vfc(window) -> complex_to_mag_squared -> single_pole_iir_filter_ff
// -> nlog10_ff as ONE handle: the kept vectors are read straight out of the stream by the FFT kernel (vector stride =
// decimation), |X|^2 is its epilogue, and -- without averaging -- so is the logarithm.
struct grb200_logpwr {
    int device = 0;
    int N = 0;
    bool complex_input = true;
    double sample_rate = 1.0, frame_rate = 1.0, ref_scale = 2.0, avg_alpha = 1.0;
    bool average = false;
    int decim = 1;
    int count = 1; // keep_one_in_n's d_count (keep_one_in_n_impl.cc:45-53,66-74)
    float log_mul = 0.f, log_scale = 1.f;
    DevBuf window, tw, ystate, scratch;
    Stager sg;
    std::mutex mu;
    size_t in_item() const { return complex_input ? 8 : 4; }
};

static void logpwr_set_decim(grb200_logpwr* h, double decim)
{
    // stream_to_vector_decimator.set_decimation (gr-blocks/python/blocks/stream_to_vector_decimator.py:64-72): max(1, round)
    // -- Python 3 round(): half to even, which nearbyint() does under the default rounding mode -- then keep_one_in_n::set_n
    // (keep_one_in_n_impl.cc:41-53), which restarts the counter
    long long d = (long long)std::nearbyint(decim);
    h->decim = (int)std::max<long long>(1, std::min<long long>(d, 0x7fffffff));
    h->count = h->decim;
}

extern "C" {

int grb200_logpwrfft_create(int complex_input, double sample_rate, int fft_size, double ref_scale, double frame_rate,
                            double avg_alpha, int average, const float* window, int nwindow, int device, grb200_logpwr** out)
{
    if (!out)
        return fail(GRB200_EINVAL, "logpwrfft: null output handle");
    *out = nullptr;
    if (fft_size <= 0)
        return fail(GRB200_ERANGE, "fft_impl_fftw: invalid fft_size");
    if (!fft_size_is_fast(fft_size) && !fft_size_is_big(fft_size) && fft_size > fft_direct_max())
        return fail(GRB200_ERANGE, "logpwrfft: fft_size %d is not supported by the device transform", fft_size);
    if (nwindow != fft_size || !window)
        return fail(GRB200_EWINDOW, "fft_vcc: window not the same length as fft_size");
    if (!(avg_alpha >= 0.0 && avg_alpha <= 1.0)) // single_pole_iir::set_taps, single_pole_iir.h:56-57 (std::out_of_range)
        return fail(GRB200_ERANGE, "Alpha must be in [0, 1]");
    if (!(frame_rate > 0.0) || !(sample_rate > 0.0) || !(ref_scale > 0.0))
        return fail(GRB200_EINVAL, "logpwrfft: sample_rate, frame_rate and ref_scale must be positive");
    std::unique_ptr<grb200_logpwr> h(new grb200_logpwr);
    h->device = device;
    h->N = fft_size;
    h->complex_input = complex_input != 0;
    h->sample_rate = sample_rate;
    h->frame_rate = frame_rate;
    h->ref_scale = ref_scale;
    h->avg_alpha = avg_alpha;
    h->average = average != 0;
    logpwr_set_decim(h.get(), sample_rate / fft_size / frame_rate); // _update_decimator, stream_to_vector_decimator.py:74-75
    // logpwrfft.py:55-60 in double, then nlog10_ff_impl's float members (nlog10_ff_impl.cc:31-33)
    double wp = 0.0;
    for (int i = 0; i < fft_size; i++)
        wp += (double)window[i] * (double)window[i];
    const double kd = -20 * std::log10((double)fft_size) - 10 * std::log10(wp / fft_size) - 20 * std::log10(ref_scale / 2);
    const float n = 10.0f, k = (float)kd;
    h->log_mul = n / log2f(10.0f);
    h->log_scale = std::pow(10.0f, k / n);
    grb::DeviceScope _ds_create(device); // the caller keeps its own current device
    int rc = h->sg.init(device);
    if (rc)
        return rc;
    std::vector<float2> roots = make_roots(fft_size, -1);
    std::vector<float> zeros((size_t)fft_size, 0.f);
    if ((rc = h->tw.upload(roots.data(), roots.size() * sizeof(float2))) ||
        (rc = h->window.upload(window, (size_t)fft_size * sizeof(float))) ||
        (rc = h->ystate.upload(zeros.data(), zeros.size() * sizeof(float))))
        return rc;
    *out = h.release();
    return 0;
}

int grb200_logpwrfft_set_decimation(grb200_logpwr* h, double decim)
{
    if (!h)
        return fail(GRB200_EINVAL, "logpwrfft: null handle");
    std::lock_guard<std::mutex> lk(h->mu);
    logpwr_set_decim(h, decim);
    return 0;
}
int grb200_logpwrfft_set_vec_rate(grb200_logpwr* h, double vec_rate)
{
    if (!h || !(vec_rate > 0.0))
        return fail(GRB200_EINVAL, "logpwrfft::set_vec_rate: bad arguments");
    std::lock_guard<std::mutex> lk(h->mu);
    h->frame_rate = vec_rate;
    logpwr_set_decim(h, h->sample_rate / h->N / h->frame_rate);
    return 0;
}
int grb200_logpwrfft_set_sample_rate(grb200_logpwr* h, double sample_rate)
{
    if (!h || !(sample_rate > 0.0))
        return fail(GRB200_EINVAL, "logpwrfft::set_sample_rate: bad arguments");
    std::lock_guard<std::mutex> lk(h->mu);
    h->sample_rate = sample_rate;
    logpwr_set_decim(h, h->sample_rate / h->N / h->frame_rate);
    return 0;
}
int grb200_logpwrfft_set_average(grb200_logpwr* h, int average)
{
    if (!h)
        return fail(GRB200_EINVAL, "logpwrfft: null handle");
    std::lock_guard<std::mutex> lk(h->mu);
    h->average = average != 0;
    return 0;
}
int grb200_logpwrfft_set_avg_alpha(grb200_logpwr* h, double avg_alpha)
{
    if (!h)
        return fail(GRB200_EINVAL, "logpwrfft: null handle");
    if (!(avg_alpha >= 0.0 && avg_alpha <= 1.0))
        return fail(GRB200_ERANGE, "Alpha must be in [0, 1]");
    std::lock_guard<std::mutex> lk(h->mu);
    h->avg_alpha = avg_alpha;
    return 0;
}
int grb200_logpwrfft_decimation(const grb200_logpwr* h) { return h ? h->decim : 0; }
double grb200_logpwrfft_sample_rate(const grb200_logpwr* h) { return h ? h->sample_rate : 0.0; }
double grb200_logpwrfft_frame_rate(const grb200_logpwr* h)
{
    return h ? h->sample_rate / h->N / h->decim : 0.0; // stream_to_vector_decimator.frame_rate()
}
int grb200_logpwrfft_average(const grb200_logpwr* h) { return h ? (h->average ? 1 : 0) : 0; }
double grb200_logpwrfft_avg_alpha(const grb200_logpwr* h) { return h ? h->avg_alpha : 0.0; }
int grb200_logpwrfft_fft_size(const grb200_logpwr* h) { return h ? h->N : 0; }

} // extern "C"

// Which of `nvec_in` input vectors the decimator keeps, given its counter: the first kept vector and the count afterwards.
static void logpwr_plan(const grb200_logpwr* h, long long nvec_in, long long max_out, long long* first, long long* nkeep,
                        long long* consumed_vecs, int* count_after)
{
    // keep_one_in_n general_work (:66-74): d_count--, emit when it reaches 0, stop when inputs or outputs run out
    const long long n = h->decim, c = h->count;
    const long long f = c - 1; // index of the first vector that is kept
    long long k = nvec_in > f ? (nvec_in - f + n - 1) / n : 0;
    long long used = nvec_in;
    int cnt;
    if (k > max_out) {
        k = max_out;
        used = k > 0 ? f + (k - 1) * n + 1 : 0; // the loop stops right after the max_out-th emission
        cnt = k > 0 ? (int)n : (int)c;
    } else {
        const long long after_last = k > 0 ? nvec_in - (f + (k - 1) * n + 1) : nvec_in;
        cnt = k > 0 ? (int)(n - after_last) : (int)(c - nvec_in);
    }
    *first = f;
    *nkeep = k;
    *consumed_vecs = used;
    *count_after = cnt;
}

// the handle's averaging scratch, grown on demand (grb200_logpwrfft_work_dev_static exists for callers that must not allocate)
static int logpwr_reserve_scratch(grb200_logpwr* h, long long floats, cudaStream_t st)
{
    if ((size_t)floats * 4 <= h->scratch.cap)
        return 0;
    GRB_CUDA(cudaStreamSynchronize(st)); // an earlier call on this stream may still use the old scratch
    h->scratch.release();
    void* pnew = nullptr;
    cudaError_t e = cudaMalloc(&pnew, (size_t)floats * 4 + 256);
    if (e != cudaSuccess)
        return fail(GRB200_ENOMEM, "cudaMalloc(%lld) failed: %s", floats * 4, cudaGetErrorString(e));
    h->scratch.p = pnew;
    h->scratch.cap = (size_t)floats * 4 + 256;
    return 0;
}

// output frame j transforms input vector first + j*vstride of in_dev
static int logpwr_run_dev(grb200_logpwr* h, const void* in_dev, long long first, long long vstride, long long nkeep,
                          float* out_dev, cudaStream_t st)
{
    if (nkeep <= 0)
        return 0;
    const int nsm = sm_count(h->device);
    const bool avg = h->average && h->avg_alpha != 1.0;
    VccArgs a{ in_dev, out_dev, nkeep, (const float*)h->window.p, (const float2*)h->tw.p, 0, avg ? 1 : 2, vstride, first,
               h->log_mul, h->log_scale, (float*)h->ystate.p };
    int rc;
    // Long averaging calls at 2048 / 4096 / 8192 bins: the FFT kernel filters its own runs of frames (mode 3), two small
    // kernels finish.  (Only past what grb200_logpwrfft_work_dev_static accepts: up to there both entry points run the same
    // kernels and give the same bits.)
    const long long fused_need = logpwr_fused_scratch_floats(h->N, nsm);
    if (avg && (h->N == 2048 || h->N == 4096 || h->N == 8192) && nkeep > 8 * (logpwr_static_scratch_floats() / h->N)) {
        if ((rc = logpwr_reserve_scratch(h, fused_need, st)))
            return rc;
        long long run_len = 0, nruns = 0;
        bool done = false;
        if ((rc = fft_vcc_stream_iir_launch(h->N, !h->complex_input, a, h->avg_alpha, (float*)h->scratch.p, nsm, st, &run_len,
                                            &nruns, &done)))
            return rc;
        if (done)
            return logpwr_carry_fixup_launch(out_dev, nkeep, h->N, run_len, nruns, h->avg_alpha, (float*)h->scratch.p,
                                             (float*)h->ystate.p, h->log_mul, h->log_scale, st);
    }
    rc = fft_vcc_launch_ex(h->N, true, !h->complex_input, a, nsm, st);
    if (rc || !avg)
        return rc;
    const long long need = logpwr_scratch_floats(nkeep, h->N, nsm);
    if ((rc = logpwr_reserve_scratch(h, need, st)))
        return rc;
    return logpwr_iir_log_launch(out_dev, nkeep, h->N, h->avg_alpha, (float*)h->ystate.p, need ? (float*)h->scratch.p : nullptr,
                                 h->log_mul, h->log_scale, nsm, st);
}


extern "C" {

// in_dev: nitems stream items (float32 or complex64); whole vectors are consumed.  out_dev: room for max_out frames of
// fft_size floats.  *consumed: stream items used; *produced: frames written.
int grb200_logpwrfft_work_dev(grb200_logpwr* h, const void* in_dev, long long nitems, float* out_dev, long long max_out,
                              long long* consumed, long long* produced, void* stream)
{
    if (!h || nitems < 0 || max_out < 0 || !consumed || !produced)
        return fail(GRB200_EINVAL, "logpwrfft::work_dev: bad arguments");
    std::lock_guard<std::mutex> lk(h->mu);
    long long first, nkeep, used;
    int cnt;
    logpwr_plan(h, nitems / h->N, max_out, &first, &nkeep, &used, &cnt);
    if (nkeep > 0 && (!in_dev || !out_dev))
        return fail(GRB200_EINVAL, "logpwrfft::work_dev: null buffer");
    grb::DeviceScope _ds(h->device);
    int rc = _ds.rc;
    if (rc)
        return rc;
    cudaStream_t st_ = pick_stream(stream, h->sg);
    rc = dev_done(logpwr_run_dev(h, in_dev, first, h->decim, nkeep, out_dev, st_), h->sg, st_);
    if (rc)
        return rc;
    h->count = cnt;
    *consumed = used * h->N;
    *produced = nkeep;
    return 0;
}

// Host form: only the KEPT vectors cross PCIe (one frame in `decimation`), each as its own piece of the staged chunk.
int grb200_logpwrfft_work(grb200_logpwr* h, const void* in_host, long long nitems, float* out_host, long long max_out,
                          long long* consumed, long long* produced)
{
    if (!h || nitems < 0 || max_out < 0 || !consumed || !produced)
        return fail(GRB200_EINVAL, "logpwrfft::work: bad arguments");
    std::lock_guard<std::mutex> lk(h->mu);
    long long first, nkeep, used;
    int cnt;
    logpwr_plan(h, nitems / h->N, max_out, &first, &nkeep, &used, &cnt);
    *consumed = used * h->N;
    *produced = nkeep;
    if (nkeep > 0) {
        if (!in_host || !out_host)
            return fail(GRB200_EINVAL, "logpwrfft::work: null buffer");
        grb::DeviceScope _ds(h->device);
        int rc = _ds.rc;
        if (rc)
            return rc;
        const size_t vin = (size_t)h->N * h->in_item(), vout = (size_t)h->N * 4;
        long long chunk = std::max<long long>(1, (long long)(chunk_bytes_for((size_t)nkeep * vin) / vin));
        if (chunk > nkeep)
            chunk = nkeep;
        const long long nchunks = (nkeep + chunk - 1) / chunk;
        if ((rc = h->sg.reserve_dev(0, (size_t)chunk * vin)) || (rc = h->sg.reserve_dev(1, (size_t)chunk * vout)))
            return rc;
        const char* in = (const char*)in_host;
        char* out = (char*)out_host;
        Stager& sg = h->sg;
        const long long dec = h->decim;
        rc = sg.run(
            nchunks,
            [&](long long c, int slot, Stager::Chunk& ch) {
                const long long k0 = c * chunk, n = std::min(chunk, nkeep - k0);
                if (dec == 1) {
                    ch.h2d.push_back({ in + (size_t)(first + k0) * vin, sg.dev[slot][0], (size_t)n * vin });
                } else {
                    for (long long k = 0; k < n; k++)
                        ch.h2d.push_back({ in + (size_t)(first + (k0 + k) * dec) * vin, (char*)sg.dev[slot][0] + (size_t)k * vin, vin });
                }
                ch.d2h.push_back({ sg.dev[slot][1], out + (size_t)k0 * vout, (size_t)n * vout });
                return 0;
            },
            [&](long long c, int slot, cudaStream_t st) {
                const long long k0 = c * chunk, n = std::min(chunk, nkeep - k0);
                // the staged slot holds the kept vectors back to back
                return logpwr_run_dev(h, sg.dev[slot][0], 0, 1, n, (float*)sg.dev[slot][1], st);
            });
        if (rc)
            return rc;
    }
    h->count = cnt;
    return 0;
}

void grb200_logpwrfft_destroy(grb200_logpwr* h)
{
    if (!h)
        return;
    grb::DeviceScope _ds(h->device);
    h->sg.destroy();
    h->window.release();
    h->tw.release();
    h->ystate.release();
    h->scratch.release();
    delete h;
}

} // extern "C"

// ---- grb200_logpwrfft_work_dev_static: grb200_logpwrfft_work_dev that allocates nothing -------------------------------
// No counterpart in the reference: the averaging tail's scratch is the
// library's file-scope device array (logpwr.cu) instead of the handle's lazily cudaMalloc'ed buffer, so the call is legal
// where cudaMalloc is not.  A call that does not fit is refused BEFORE anything is launched: the handle's decimator counter
// and filter memory stay as they were.  Same kernels, same run lengths, same bits as grb200_logpwrfft_work_dev.
extern "C" int grb200_logpwrfft_work_dev_static(grb200_logpwr* h, const void* in_dev, long long nitems, float* out_dev,
                                                long long max_out, long long* consumed, long long* produced, void* stream)
{
    if (!h || nitems < 0 || max_out < 0 || !consumed || !produced)
        return fail(GRB200_EINVAL, "logpwrfft::work_dev_static: bad arguments");
    std::lock_guard<std::mutex> lk(h->mu);
    long long first, nkeep, used;
    int cnt;
    logpwr_plan(h, nitems / h->N, max_out, &first, &nkeep, &used, &cnt);
    if (nkeep > 0 && (!in_dev || !out_dev))
        return fail(GRB200_EINVAL, "logpwrfft::work_dev_static: null buffer");
    grb::DeviceScope _ds(h->device);
    int rc = _ds.rc;
    if (rc)
        return rc;
    const int nsm = sm_count(h->device);
    const bool avg = h->average && h->avg_alpha != 1.0;
    if (avg && nkeep > 0 && (rc = logpwr_static_check(nkeep, h->N, nsm)))
        return rc;
    cudaStream_t st_ = pick_stream(stream, h->sg);
    if (nkeep > 0) {
        VccArgs a{ in_dev, out_dev, nkeep, (const float*)h->window.p, (const float2*)h->tw.p, 0, avg ? 1 : 2, h->decim, first,
                   h->log_mul, h->log_scale, (float*)h->ystate.p };
        rc = fft_vcc_launch_ex(h->N, true, !h->complex_input, a, nsm, st_);
        if (!rc && avg)
            rc = logpwr_iir_log_launch_static(out_dev, nkeep, h->N, h->avg_alpha, (float*)h->ystate.p, h->log_mul,
                                              h->log_scale, nsm, h->device, st_);
    }
    rc = dev_done(rc, h->sg, st_);
    if (rc)
        return rc;
    h->count = cnt;
    *consumed = used * h->N;
    *produced = nkeep;
    return 0;
}
