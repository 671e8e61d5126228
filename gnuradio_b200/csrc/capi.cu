// capi.cu -- the C-ABI of libgrb200 (include/grb200.h): handles, tap/window preparation, host-staged and
// device-resident work() entry points.  No compute happens on the host beyond O(ntaps) tap re-layout.
#include "capi_common.cuh"

using namespace grb;

// =====================================================================================================
// library / device
// =====================================================================================================
extern "C" {

const char* grb200_last_error(void) { return g_err; }
const char* grb200_version(void) { return "grb200 0.1 (sm_100a)"; }
int grb200_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}
unsigned long long grb200_launch_count(void) { return g_launches.load(); }

int grb200_host_alloc(void** ptr, unsigned long long bytes)
{
    if (!ptr)
        return fail(GRB200_EINVAL, "null pointer");
    GRB_CUDA(cudaHostAlloc(ptr, (size_t)bytes, cudaHostAllocDefault));
    return 0;
}
int grb200_host_free(void* ptr)
{
    GRB_CUDA(cudaFreeHost(ptr));
    return 0;
}
int grb200_dev_alloc(int device, void** ptr, unsigned long long bytes)
{
    int rc = use_device(device);
    if (rc)
        return rc;
    cudaError_t e = cudaMalloc(ptr, (size_t)bytes);
    if (e != cudaSuccess)
        return fail(GRB200_ENOMEM, "cudaMalloc(%llu) failed: %s", bytes, cudaGetErrorString(e));
    return 0;
}
int grb200_dev_free(int device, void* ptr)
{
    int rc = use_device(device);
    if (rc)
        return rc;
    GRB_CUDA(cudaFree(ptr));
    return 0;
}
int grb200_memcpy_h2d(int device, void* dst, const void* src, unsigned long long bytes)
{
    int rc = use_device(device);
    if (rc)
        return rc;
    GRB_CUDA(cudaMemcpy(dst, src, (size_t)bytes, cudaMemcpyHostToDevice));
    return 0;
}
int grb200_memcpy_d2h(int device, void* dst, const void* src, unsigned long long bytes)
{
    int rc = use_device(device);
    if (rc)
        return rc;
    GRB_CUDA(cudaMemcpy(dst, src, (size_t)bytes, cudaMemcpyDeviceToHost));
    return 0;
}
int grb200_dev_sync(int device)
{
    int rc = use_device(device);
    if (rc)
        return rc;
    GRB_CUDA(cudaDeviceSynchronize());
    return 0;
}

} // extern "C"

// =====================================================================================================
// fir_filter_ccf / ccc
// =====================================================================================================
struct grb200_fir {
    int device = 0;
    bool ct = false;
    int decim = 1;
    int ntaps = 0;
    DevBuf gtaps, rev;
    Stager sg;
    std::mutex mu;
};

static int fir_upload_taps(grb200_fir* h, const float* taps, int ntaps)
{
    // reversed copy (fir_filter.cc:62-64) and its phase-major re-layout g_p[u] = rev[D*u + p]
    const int D = h->decim, es = h->ct ? 2 : 1;
    const int Upad = fir_upad(D, ntaps);
    std::vector<float> rev((size_t)ntaps * es), g((size_t)D * Upad * es, 0.f);
    for (int t = 0; t < ntaps; t++)
        for (int e = 0; e < es; e++)
            rev[(size_t)t * es + e] = taps[(size_t)(ntaps - 1 - t) * es + e];
    for (int t = 0; t < ntaps; t++) {
        const int p = t % D, u = t / D;
        for (int e = 0; e < es; e++)
            g[((size_t)p * Upad + u) * es + e] = rev[(size_t)t * es + e];
    }
    int rc = h->rev.upload(rev.data(), rev.size() * sizeof(float));
    if (rc)
        return rc;
    rc = h->gtaps.upload(g.data(), g.size() * sizeof(float));
    if (rc)
        return rc;
    h->ntaps = ntaps;
    return 0;
}

static int fir_create(bool ct, int decim, const float* taps, int ntaps, int device, grb200_fir** out)
{
    if (!out)
        return fail(GRB200_EINVAL, "fir: null output handle");
    *out = nullptr;
    if (decim < 1)
        return fail(GRB200_EINVAL, "fir_filter: decimation must be >= 1 (got %d)", decim);
    if (ntaps < 1 || !taps)
        return fail(GRB200_EINVAL, "fir_filter: at least one tap is required");
    std::unique_ptr<grb200_fir> h(new grb200_fir);
    h->device = device;
    h->ct = ct;
    h->decim = decim;
    int rc = h->sg.init(device);
    if (rc)
        return rc;
    rc = fir_upload_taps(h.get(), taps, ntaps);
    if (rc)
        return rc;
    *out = h.release();
    return 0;
}

extern "C" {

int grb200_fir_ccf_create(int decim, const float* taps, int ntaps, int device, grb200_fir** out)
{
    return fir_create(false, decim, taps, ntaps, device, out);
}
int grb200_fir_ccc_create(int decim, const float* taps, int ntaps, int device, grb200_fir** out)
{
    return fir_create(true, decim, taps, ntaps, device, out);
}

int grb200_fir_set_taps(grb200_fir* h, const float* taps, int ntaps)
{
    if (!h || !taps || ntaps < 1)
        return fail(GRB200_EINVAL, "fir_filter::set_taps: bad arguments");
    std::lock_guard<std::mutex> lk(h->mu);
    int rc = use_device(h->device);
    if (rc)
        return rc;
    GRB_CUDA(cudaDeviceSynchronize()); // kernels in flight still read the old taps
    return fir_upload_taps(h, taps, ntaps);
}

int grb200_fir_ntaps(const grb200_fir* h) { return h ? h->ntaps : 0; }
int grb200_fir_decimation(const grb200_fir* h) { return h ? h->decim : 0; }

int grb200_fir_work_dev(grb200_fir* h, const void* in_dev, void* out_dev, long long noutput, void* stream)
{
    if (!h || noutput < 0 || (noutput > 0 && (!in_dev || !out_dev)))
        return fail(GRB200_EINVAL, "fir_filter::work_dev: bad arguments");
    std::lock_guard<std::mutex> lk(h->mu);
    int rc = use_device(h->device);
    if (rc)
        return rc;
    return fir_launch(h->ct, h->decim, h->ntaps, h->gtaps.p, h->rev.p, in_dev, out_dev, noutput,
                      pick_stream(stream, h->sg));
}

int grb200_fir_work(grb200_fir* h, const void* in_host, void* out_host, long long noutput)
{
    if (!h || noutput < 0 || (noutput > 0 && (!in_host || !out_host)))
        return fail(GRB200_EINVAL, "fir_filter::work: bad arguments");
    if (noutput == 0)
        return 0;
    std::lock_guard<std::mutex> lk(h->mu);
    int rc = use_device(h->device);
    if (rc)
        return rc;
    const int D = h->decim, nt = h->ntaps;
    long long chunk = (long long)(kChunkBytes / 8) / D; // outputs per chunk
    if (chunk < 1)
        chunk = 1;
    if (chunk > noutput)
        chunk = noutput;
    const long long nchunks = (noutput + chunk - 1) / chunk;
    if ((rc = h->sg.reserve_dev(0, (size_t)(chunk * D + nt - 1) * 8)) || (rc = h->sg.reserve_dev(1, (size_t)chunk * 8)))
        return rc;
    const char* in = (const char*)in_host;
    char* out = (char*)out_host;
    Stager& sg = h->sg;
    return sg.run(
        nchunks,
        [&](long long c, int slot, Stager::Chunk& ch) {
            const long long o0 = c * chunk, n = std::min(chunk, noutput - o0);
            ch.h2d.push_back({ in + (size_t)o0 * D * 8, sg.dev[slot][0], (size_t)(n * D + nt - 1) * 8 });
            ch.d2h.push_back({ sg.dev[slot][1], out + (size_t)o0 * 8, (size_t)n * 8 });
            return 0;
        },
        [&](long long c, int slot, cudaStream_t st) {
            const long long o0 = c * chunk, n = std::min(chunk, noutput - o0);
            return fir_launch(h->ct, D, nt, h->gtaps.p, h->rev.p, sg.dev[slot][0], sg.dev[slot][1], n, st);
        });
}

void grb200_fir_destroy(grb200_fir* h)
{
    if (!h)
        return;
    cudaSetDevice(h->device);
    h->sg.destroy();
    h->gtaps.release();
    h->rev.release();
    delete h;
}

} // extern "C"
