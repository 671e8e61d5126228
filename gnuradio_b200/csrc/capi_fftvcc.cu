// capi_fftvcc.cu -- C-ABI of fft_vcc (+ fused / standalone complex_to_mag_squared), include/grb200.h.
#include "capi_common.cuh"

using namespace grb;

struct grb200_fftvcc {
    int device = 0;
    int N = 0;
    bool forward = true;
    bool shift = false;
    bool has_window = false;
    DevBuf window, tw;
    Stager sg;
    std::mutex mu;
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
    return fft_vcc_launch(h->N, h->forward, in, out, nvec, h->has_window ? (const float*)h->window.p : nullptr,
                          h->shift ? 1 : 0, magsq, (const float2*)h->tw.p, sm_count(h->device), st);
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
    const size_t vin = (size_t)h->N * 8, vout = (size_t)h->N * (magsq ? 4 : 8);
    long long chunk = std::max<long long>(1, (long long)(kChunkBytes / vin));
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

extern "C" {

int grb200_fft_vcc_create(int fft_size, int forward, const float* window, int nwindow, int shift, int device,
                          grb200_fftvcc** out)
{
    if (!out)
        return fail(GRB200_EINVAL, "fft_vcc: null output handle");
    *out = nullptr;
    if (fft_size <= 0) // fft_complex ctor, gr-fft/lib/fft.cc:161-163 -> std::out_of_range
        return fail(GRB200_ERANGE, "fft_impl_fftw: invalid fft_size");
    if (!fft_size_is_fast(fft_size) && fft_size > fft_direct_max())
        return fail(GRB200_ERANGE, "fft_vcc: fft_size %d is not supported by the device transform", fft_size);
    if (nwindow < 0)
        return fail(GRB200_EINVAL, "fft_vcc: negative window length");
    std::unique_ptr<grb200_fftvcc> h(new grb200_fftvcc);
    h->device = device;
    h->N = fft_size;
    h->forward = forward != 0;
    h->shift = shift != 0;
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
static std::mutex g_magsq_mu;
static Stager* g_magsq_sg[64];

static int magsq_stager(int device, Stager** out)
{
    grb::DeviceScope _ds(device);
    int rc = _ds.rc;
    if (rc)
        return rc;
    std::lock_guard<std::mutex> lk(g_magsq_mu);
    if (!g_magsq_sg[device]) {
        Stager* s = new Stager;
        if ((rc = s->init(device))) {
            delete s;
            return rc;
        }
        g_magsq_sg[device] = s;
    }
    *out = g_magsq_sg[device];
    return 0;
}

int grb200_magsq_work_dev(int device, const void* in_dev, float* out_dev, long long nitems, void* stream)
{
    if (nitems < 0 || (nitems > 0 && (!in_dev || !out_dev)))
        return fail(GRB200_EINVAL, "complex_to_mag_squared::work_dev: bad arguments");
    Stager* sg = nullptr;
    int rc = magsq_stager(device, &sg);
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
    int rc = magsq_stager(device, &sgp);
    if (rc)
        return rc;
    static std::mutex work_mu; // the per-device stager is shared: one host-staged call at a time
    std::lock_guard<std::mutex> lk(work_mu);
    Stager& sg = *sgp;
    long long chunk = (long long)(kChunkBytes / 8);
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
