// capi.cu -- the C-ABI of libgrb200 (include/grb200.h): handles, tap/window preparation, host-staged and
// device-resident work() entry points.  No compute happens on the host beyond O(ntaps) tap re-layout.
#include "capi_common.cuh"
#include <complex>
#include <cstring>

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
    pin_cache_invalidate();
    GRB_CUDA(cudaFreeHost(ptr));
    return 0;
}
int grb200_host_register(void* ptr, unsigned long long bytes)
{
    if (!ptr || !bytes)
        return fail(GRB200_EINVAL, "host_register: bad arguments");
    cudaError_t e = cudaHostRegister(ptr, (size_t)bytes, cudaHostRegisterDefault);
    if (e == cudaErrorHostMemoryAlreadyRegistered) {
        cudaGetLastError();
        e = cudaSuccess;
    }
    if (e != cudaSuccess)
        return fail(GRB200_ECUDA, "cudaHostRegister(%llu bytes) failed: %s", bytes, cudaGetErrorString(e));
    pin_cache_invalidate();
    return 0;
}
int grb200_host_unregister(void* ptr)
{
    cudaError_t e = cudaHostUnregister(ptr);
    pin_cache_invalidate();
    if (e != cudaSuccess)
        return fail(GRB200_ECUDA, "cudaHostUnregister failed: %s", cudaGetErrorString(e));
    return 0;
}
int grb200_dev_alloc(int device, void** ptr, unsigned long long bytes)
{
    grb::DeviceScope _ds(device);
    int rc = _ds.rc;
    if (rc)
        return rc;
    cudaError_t e = cudaMalloc(ptr, (size_t)bytes);
    if (e != cudaSuccess)
        return fail(GRB200_ENOMEM, "cudaMalloc(%llu) failed: %s", bytes, cudaGetErrorString(e));
    return 0;
}
int grb200_dev_free(int device, void* ptr)
{
    grb::DeviceScope _ds(device);
    int rc = _ds.rc;
    if (rc)
        return rc;
    GRB_CUDA(cudaFree(ptr));
    return 0;
}
int grb200_memcpy_h2d(int device, void* dst, const void* src, unsigned long long bytes)
{
    grb::DeviceScope _ds(device);
    int rc = _ds.rc;
    if (rc)
        return rc;
    GRB_CUDA(cudaMemcpy(dst, src, (size_t)bytes, cudaMemcpyHostToDevice));
    return 0;
}
int grb200_memcpy_d2h(int device, void* dst, const void* src, unsigned long long bytes)
{
    grb::DeviceScope _ds(device);
    int rc = _ds.rc;
    if (rc)
        return rc;
    GRB_CUDA(cudaMemcpy(dst, src, (size_t)bytes, cudaMemcpyDeviceToHost));
    return 0;
}
int grb200_dev_sync(int device)
{
    grb::DeviceScope _ds(device);
    int rc = _ds.rc;
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
    // freq_xlating_fir_filter (capi_xlat.cu) drives this handle with a rotator behind the filter: out[n] *= e^{j (theta + n delta)}
    bool rot = false;
    double theta = 0.0, delta = 0.0;
    int tc_pref = -1; // GRB200_FIR_TC as it stood when the handle was made
    const void* halo_src = nullptr; // left neighbour's newest ntaps-1 items (peer memory), see grb200_fir_set_halo_source
    bool real = false; // fir_filter_fff: float stream, float taps (fir_fff.cu)
    int isz() const { return real ? 4 : 8; }
    // Long filters run on the overlap-save kernels of fft_filter (fftfilt.cu) -- same linear convolution, ~3e-7 NRMS from the
    // time-domain sum: past ~48 real (16 complex) taps per decimation step the CUDA-core FIR is FP32-bound below their rate.
    int fft_pref = -1; // GRB200_FIR_FFT as it stood when the handle was made: 0 never, 1 whenever a plan exists, unset = auto
    bool fft_route = false;
    DevBuf fH, ftw;
    int fL = 0, fK = 1;
    Stager sg;
    std::mutex mu;
};

static int fir_upload_taps(grb200_fir* h, const float* taps, int ntaps)
{
    if (h->real) { // reversed taps zero padded to what fir_fff_kernel unrolls by
        std::vector<float> rev((size_t)fir_fff_pad(ntaps), 0.f);
        for (int t = 0; t < ntaps; t++)
            rev[t] = taps[ntaps - 1 - t];
        int rc = h->rev.upload(rev.data(), rev.size() * sizeof(float));
        if (rc)
            return rc;
        h->ntaps = ntaps;
        // long real filters: the overlap-save kernel with the two halves of a call as re / im of one complex stream
        h->fft_route = false;
        const int L = fftfilt_realpair_L(ntaps);
        // (measured, tools/route_check.py: ~470 G samples/s at any tap count up to 2049, ~590 / 635 / 645 at decimation 2 / 4 / 8
        // (folded inverse transform), ~450 at other decimations that divide 256, ~380 at the rest; the FIR kernel falls below
        // that past 16 taps, at decimation 4 and 8 past 8, at decimation 2 it is below at 8 taps already)
        // At decimations that are not powers of two the one-output-per-thread FIR kernel reads ntaps floats per output whatever
        // the decimation: ~15 - 27 T floats/s, so it holds against the overlap-save kernels (430 - 530 G samples/s at even
        // decimations, 425 at odd ones) up to 24 - 64 taps (profiles/r2_route_check.txt).
        const int D = h->decim;
        const bool dp2 = (D & (D - 1)) == 0;
        const int least = dp2 ? (D == 1 ? 16 : (D == 2 ? 4 : 8))
                              : ((D & 1) == 0 ? (D < 16 ? 40 : 48) : (D < 8 ? 24 : (D < 16 ? 44 : 64)));
        const bool want = h->fft_pref == 1 || (h->fft_pref < 0 && ntaps > least);
        if (want && L > 0) {
            std::vector<float> tc((size_t)2 * ntaps, 0.f);
            for (int t = 0; t < ntaps; t++)
                tc[2 * (size_t)t] = taps[t];
            if ((rc = fftfilt_plan_build(tc.data(), ntaps, h->device, h->sg.s_exec, h->fH, h->ftw, &h->fL, &h->fK, L)))
                return rc;
            h->fft_route = true;
        }
        return 0;
    }
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
    // route to the overlap-save kernels?  Measured crossovers (tools/route_check.py; the three-group kernel runs at ~245 G
    // items/s, ~200 at decimations that do not divide 256): complex taps past 16 at decimation 1 -- that includes the
    // tensor-core FIR of 17..64 complex taps (3 x 4 MMAs per tile: 205 G items/s) unless GRB200_FIR_TC=1 asks for it --, any
    // complex filter at decimation 2, 4, 8.., past 16 per step at other decimations; real taps past 64 at decimation 1, past 32
    // per step at decimation 2, 4, .., 48 per step otherwise; never where the real-tap tensor-core modes apply.  (Short filters
    // take the route on large calls only: fir_run_dev.)
    h->fft_route = false;
    // (decimation 2, 4, 8 and 16 .. 256: the overlap-save kernel folds the spectrum and inverts L/D (L/8) points -- 320 - 360 G
    // items/s, against 319 of the direct kernel up to 256 real taps at decimation 16 and 149 at decimation 32; the tensor-core
    // decimate-by-8 mode holds 345 - 357 from 256 taps up, 292 below)
    const bool tc_real = h->tc_pref != 0 && !h->ct && ((D == 1 && ntaps <= 64) || (D == 8 && ntaps >= 256 && ntaps <= 1024));
    const bool tc_forced = h->tc_pref == 1 && ((h->ct && D == 1 && ntaps <= 64) || (!h->ct && D == 8 && ntaps <= 1024));
    const bool p2 = D > 1 && D <= 256 && (D & (D - 1)) == 0;
    // (even decimations fold by their largest factor 2 / 4 / 8: 270 - 290 G items/s when the rest is not a power of two; the
    // direct kernels drop to 110 - 150 from decimation 24 on)
    (void)p2;
    const bool even = (D & 1) == 0;
    const long long least =
        h->ct ? ((even || D >= 24) ? 0 : 16LL * D)
              : (D == 1 ? 64
                        : (D == 2 ? 32 : (D == 4 ? 64 : (D == 8 || D == 16 ? 0 : (D >= 24 ? 0 : (even ? 13LL : 40LL) * D)))));
    const bool want = h->fft_pref == 1 || (h->fft_pref < 0 && !tc_real && !tc_forced && ntaps > least);
    if (want && fftfilt_device_plan(ntaps, nullptr) > 0) {
        std::vector<float> tc((size_t)2 * ntaps, 0.f);
        for (int t = 0; t < ntaps; t++) {
            tc[2 * (size_t)t] = taps[(size_t)t * es];
            if (h->ct)
                tc[2 * (size_t)t + 1] = taps[(size_t)t * es + 1];
        }
        if ((rc = fftfilt_plan_build(tc.data(), ntaps, h->device, h->sg.s_exec, h->fH, h->ftw, &h->fL, &h->fK)))
            return rc;
        h->fft_route = true;
    }
    return 0;
}

static int fir_fff_run_dev(grb200_fir* h, const void* in_dev, void* out_dev, long long noutput, cudaStream_t st)
{
    if (h->fft_route && noutput >= 65536)
        return fftfilt_realpair_launch(h->fL, in_dev, out_dev, noutput, h->ntaps, h->decim, (const float2*)h->fH.p,
                                       (const float2*)h->ftw.p, sm_count(h->device), st);
    return fir_fff_launch(h->decim, h->ntaps, (const float*)h->rev.p, in_dev, out_dev, noutput, sm_count(h->device), st);
}

// the FIR of one device-resident call: overlap-save when the handle routes there, else the FIR kernels
static int fir_run_dev(grb200_fir* h, const void* in_dev, void* out_dev, long long noutput, cudaStream_t st, const void* halo)
{
    // (a short filter on a scheduler-sized call: the direct kernels start faster than a 768-thread CTA per SM does)
    if (h->fft_route && (h->fft_pref == 1 || h->ntaps > 64LL * h->decim || noutput * h->decim >= 32768)) {
        if (halo) { // the overlap-save kernels read their history in place: bring the neighbour's items over first
            int rc = halo_fetch_launch(const_cast<void*>(in_dev), halo, (size_t)(h->ntaps - 1) * 8, st);
            if (rc)
                return rc;
        }
        return fftfilt_launch(h->fL, h->fK, in_dev, out_dev, noutput, h->ntaps, h->decim, (const float2*)h->fH.p,
                              (const float2*)h->ftw.p, sm_count(h->device), st);
    }
    return fir_launch(h->ct, h->decim, h->ntaps, h->gtaps.p, h->rev.p, in_dev, out_dev, noutput, h->tc_pref, st, halo);
}

static int fir_create(bool ct, int decim, const float* taps, int ntaps, int device, grb200_fir** out, bool real = false)
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
    h->tc_pref = fir_tc_env();
    if (const char* e = getenv("GRB200_FIR_FFT")) // (per handle, at create: tests flip it between handles)
        h->fft_pref = atoi(e) ? 1 : 0;
    h->real = real;
    grb::DeviceScope _ds_create(device); // the caller keeps its own current device
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
int grb200_fir_fff_create(int decim, const float* taps, int ntaps, int device, grb200_fir** out)
{
    return fir_create(false, decim, taps, ntaps, device, out, true);
}

int grb200_fir_set_taps(grb200_fir* h, const float* taps, int ntaps)
{
    if (!h || !taps || ntaps < 1)
        return fail(GRB200_EINVAL, "fir_filter::set_taps: bad arguments");
    std::lock_guard<std::mutex> lk(h->mu);
    grb::DeviceScope _ds(h->device);
    int rc = _ds.rc;
    if (rc)
        return rc;
    if ((rc = h->sg.quiesce())) // launches through this handle may still read the old state
        return rc;
    return fir_upload_taps(h, taps, ntaps);
}

int grb200_fir_set_halo_source(grb200_fir* h, const void* peer_tail)
{
    if (!h)
        return fail(GRB200_EINVAL, "fir_filter::set_halo_source: null handle");
    std::lock_guard<std::mutex> lk(h->mu);
    h->halo_src = peer_tail;
    return 0;
}

int grb200_fir_ntaps(const grb200_fir* h) { return h ? h->ntaps : 0; }
int grb200_fir_decimation(const grb200_fir* h) { return h ? h->decim : 0; }

int grb200_fir_work_dev(grb200_fir* h, const void* in_dev, void* out_dev, long long noutput, void* stream)
{
    if (!h || noutput < 0 || (noutput > 0 && (!in_dev || !out_dev)))
        return fail(GRB200_EINVAL, "fir_filter::work_dev: bad arguments");
    std::lock_guard<std::mutex> lk(h->mu);
    grb::DeviceScope _ds(h->device);
    int rc = _ds.rc;
    if (rc)
        return rc;
    cudaStream_t st = pick_stream(stream, h->sg);
    const size_t isz = (size_t)h->isz();
    const void* halo = (h->halo_src && h->ntaps > 1) ? h->halo_src : nullptr; // see grb200_fir_set_halo_source
    if (halo && h->real) { // fir_filter_fff stages through cp.async: bring the halo over first (one-CTA kernel, plain loads)
        if ((rc = halo_fetch_launch(const_cast<void*>(in_dev), halo, (size_t)(h->ntaps - 1) * isz, st)))
            return rc;
    }
    rc = h->real ? fir_fff_run_dev(h, in_dev, out_dev, noutput, st) : fir_run_dev(h, in_dev, out_dev, noutput, st, halo);
    if (!rc && h->rot) { // (a separate pass on purpose: rotating in the overlap-save kernel's stores was measured slower -- a
        // decimating block keeps one lane in D, and the float64 angle sequence issues for the whole warp: 159 vs 221 G items/s)
        rc = rotate_launch(out_dev, noutput, h->theta, h->delta, st);
        h->theta = rotate_advance(h->theta, h->delta, noutput);
    }
    return dev_done(rc, h->sg, st);
}

} // extern "C"

// lag != 0: queue this call, hand the PREVIOUS call's results to out_host (at most out_cap items), count in *produced
static int fir_work_host(grb200_fir* h, const void* in_host, void* out_host, long long noutput, int lag, long long out_cap,
                         long long* produced)
{
    if (!h || noutput < 0 || (noutput > 0 && (!in_host || (!lag && !out_host))))
        return fail(GRB200_EINVAL, "fir_filter::work: bad arguments");
    if (produced)
        *produced = lag ? 0 : noutput;
    if (noutput == 0 && !lag)
        return 0;
    std::lock_guard<std::mutex> lk(h->mu);
    grb::DeviceScope _ds(h->device);
    int rc = _ds.rc;
    if (rc)
        return rc;
    const int D = h->decim, nt = h->ntaps;
    const size_t isz = (size_t)h->isz();
    if (noutput == 0) { // lag mode, nothing to queue: hand out what is still parked (end of stream)
        void* outs[1] = { out_host };
        const Stager::LagResult& newest = h->sg.lag[h->sg.lag_cur ^ 1];
        if (newest.pending && newest.nitems > out_cap)
            return fail(GRB200_EINVAL, "fir_filter::flush: %lld items are waiting, room for %lld offered", newest.nitems, out_cap);
        return h->sg.lag_flush(outs, 1, produced);
    }
    long long chunk = (long long)(chunk_bytes_for((size_t)noutput * D * isz) / isz) / D; // outputs per chunk
    if (chunk < 1)
        chunk = 1;
    if (chunk > noutput)
        chunk = noutput;
    const long long nchunks = (noutput + chunk - 1) / chunk;
    if ((rc = h->sg.reserve_dev(0, (size_t)(chunk * D + nt - 1) * isz)) || (rc = h->sg.reserve_dev(1, (size_t)chunk * isz)))
        return rc;
    const char* in = (const char*)in_host;
    char* out = (char*)out_host;
    Stager& sg = h->sg;
    void* lag_out1[1] = { nullptr };
    if (lag) {
        // the previous call's result must fit what the caller offers now (the block mirrors submit equal-sized calls)
        const Stager::LagResult& prev = sg.lag[sg.lag_cur ^ 1];
        if (prev.pending && prev.nitems > out_cap)
            return fail(GRB200_EINVAL, "fir_filter::work_lag: %lld items are waiting, room for %lld offered", prev.nitems, out_cap);
        if ((rc = sg.lag_begin((size_t)noutput * isz, nchunks, noutput)))
            return rc;
        const long long nprev = prev.pending ? prev.nitems : 0;
        if (out_host && nprev + noutput <= out_cap) { // room behind the previous call's items: deliver in place
            lag_out1[0] = out_host;
            sg.lag_direct(lag_out1, 1, (size_t)nprev * isz);
        }
    }
    rc = sg.run(
        nchunks,
        [&](long long c, int slot, Stager::Chunk& ch) {
            const long long o0 = c * chunk, n = std::min(chunk, noutput - o0);
            ch.h2d.push_back({ in + (size_t)o0 * D * isz, sg.dev[slot][0], (size_t)(n * D + nt - 1) * isz });
            ch.d2h.push_back(Piece(sg.dev[slot][1], out + (size_t)o0 * isz, (size_t)n * isz, 0, (size_t)o0 * isz));
            return 0;
        },
        [&](long long c, int slot, cudaStream_t st) {
            const long long o0 = c * chunk, n = std::min(chunk, noutput - o0);
            int r2 = h->real ? fir_fff_run_dev(h, sg.dev[slot][0], sg.dev[slot][1], n, st)
                             : fir_run_dev(h, sg.dev[slot][0], sg.dev[slot][1], n, st, nullptr);
            if (!r2 && h->rot)
                r2 = rotate_launch(sg.dev[slot][1], n, rotate_advance(h->theta, h->delta, o0), h->delta, st);
            return r2;
        });
    if (!rc && h->rot)
        h->theta = rotate_advance(h->theta, h->delta, noutput);
    if (!rc && lag) {
        void* outs[1] = { out_host };
        long long got = 0;
        rc = sg.lag_collect(outs, 1, &got);
        if (produced)
            *produced = got;
    }
    return rc;
}

extern "C" {

int grb200_fir_work(grb200_fir* h, const void* in_host, void* out_host, long long noutput)
{
    return fir_work_host(h, in_host, out_host, noutput, 0, noutput, nullptr);
}

int grb200_fir_work_lag(grb200_fir* h, const void* in_host, long long noutput, void* out_host, long long out_capacity,
                        long long* produced)
{
    if (!produced || out_capacity < 0 || (out_capacity > 0 && !out_host))
        return fail(GRB200_EINVAL, "fir_filter::work_lag: bad arguments");
    return fir_work_host(h, in_host, out_host, noutput, 1, out_capacity, produced);
}

int grb200_fir_flush(grb200_fir* h, void* out_host, long long out_capacity, long long* produced)
{
    if (!produced || out_capacity < 0 || (out_capacity > 0 && !out_host))
        return fail(GRB200_EINVAL, "fir_filter::flush: bad arguments");
    return fir_work_host(h, nullptr, out_host, 0, 1, out_capacity, produced);
}


void grb200_fir_destroy(grb200_fir* h)
{
    if (!h)
        return;
    grb::DeviceScope _ds(h->device);
    h->sg.destroy();
    h->gtaps.release();
    h->rev.release();
    h->fH.release();
    h->ftw.release();
    delete h;
}

} // extern "C"

// =====================================================================================================
// freq_xlating_fir_filter_ccf / ccc  (gr-filter/lib/freq_xlating_fir_filter_impl.cc)
// =====================================================================================================
// The reference is fir_filter<gr_complex,gr_complex,gr_complex> over composite taps followed by blocks::rotator.
// Here it is the same FIR handle (so the tcgen05 complex-tap kernel serves it) with the rotator hook switched on.
struct grb200_xlat {
    grb200_fir* fir = nullptr;
    bool cplx_taps = false;
    std::vector<std::complex<float>> proto;
    double center_freq = 0.0, sampling_freq = 1.0;
    std::mutex mu; // serialises the setters (the FIR handle's own mutex orders them against work())
};

// build_composite_fir(), freq_xlating_fir_filter_impl.cc:74-93, float arithmetic as written there
static void xlat_composite(const grb200_xlat* x, int decim, std::vector<std::complex<float>>& ctaps, double* delta)
{
    const float fwT0 = 2 * M_PI * x->center_freq / x->sampling_freq;
    ctaps.resize(x->proto.size());
    for (unsigned int i = 0; i < x->proto.size(); i++) {
        const std::complex<float> e = std::exp(std::complex<float>(0, i * fwT0));
        ctaps[i] = x->cplx_taps ? x->proto[i] * e : x->proto[i].real() * e;
    }
    // rotator::set_phase_incr (rotator.h:31) normalises the float phasor; the device rotates by its angle
    std::complex<float> incr = std::exp(std::complex<float>(0, -fwT0 * decim));
    incr = incr / std::abs(incr);
    *delta = std::atan2((double)incr.imag(), (double)incr.real());
}

static void xlat_store_proto(grb200_xlat* x, const float* taps, int ntaps)
{
    x->proto.resize(ntaps);
    for (int i = 0; i < ntaps; i++)
        x->proto[i] = x->cplx_taps ? std::complex<float>(taps[2 * i], taps[2 * i + 1]) : std::complex<float>(taps[i], 0.f);
}

static int xlat_create(bool cplx_taps, int decim, const float* taps, int ntaps, double center_freq, double sampling_freq,
                       int device, grb200_xlat** out)
{
    if (!out)
        return fail(GRB200_EINVAL, "freq_xlating_fir_filter: null output handle");
    *out = nullptr;
    if (decim < 1)
        return fail(GRB200_EINVAL, "freq_xlating_fir_filter: decimation must be >= 1 (got %d)", decim);
    if (ntaps < 1 || !taps)
        return fail(GRB200_EINVAL, "freq_xlating_fir_filter: at least one tap is required");
    if (sampling_freq == 0.0)
        return fail(GRB200_EINVAL, "freq_xlating_fir_filter: sampling_freq must not be 0");
    std::unique_ptr<grb200_xlat> x(new grb200_xlat);
    x->cplx_taps = cplx_taps;
    x->center_freq = center_freq;
    x->sampling_freq = sampling_freq;
    xlat_store_proto(x.get(), taps, ntaps);
    std::vector<std::complex<float>> ctaps;
    double delta = 0.0;
    xlat_composite(x.get(), decim, ctaps, &delta);
    int rc = fir_create(true, decim, reinterpret_cast<const float*>(ctaps.data()), ntaps, device, &x->fir);
    if (rc)
        return rc;
    x->fir->rot = true;
    x->fir->delta = delta;
    *out = x.release();
    return 0;
}

static int xlat_rebuild(grb200_xlat* x)
{
    std::vector<std::complex<float>> ctaps;
    double delta = 0.0;
    xlat_composite(x, x->fir->decim, ctaps, &delta);
    int rc = grb200_fir_set_taps(x->fir, reinterpret_cast<const float*>(ctaps.data()), (int)ctaps.size());
    if (rc)
        return rc;
    std::lock_guard<std::mutex> lk(x->fir->mu);
    x->fir->delta = delta; // the phase itself carries on (rotator::set_phase_incr leaves d_phase alone)
    return 0;
}

extern "C" {

int grb200_freq_xlating_fir_ccf_create(int decim, const float* taps, int ntaps, double center_freq, double sampling_freq,
                                       int device, grb200_xlat** out)
{
    return xlat_create(false, decim, taps, ntaps, center_freq, sampling_freq, device, out);
}
int grb200_freq_xlating_fir_ccc_create(int decim, const float* taps_re_im, int ntaps, double center_freq,
                                       double sampling_freq, int device, grb200_xlat** out)
{
    return xlat_create(true, decim, taps_re_im, ntaps, center_freq, sampling_freq, device, out);
}

int grb200_freq_xlating_fir_set_taps(grb200_xlat* x, const float* taps, int ntaps)
{
    if (!x || !taps || ntaps < 1)
        return fail(GRB200_EINVAL, "freq_xlating_fir_filter::set_taps: bad arguments");
    std::lock_guard<std::mutex> lk(x->mu);
    xlat_store_proto(x, taps, ntaps);
    return xlat_rebuild(x);
}

int grb200_freq_xlating_fir_set_center_freq(grb200_xlat* x, double center_freq)
{
    if (!x)
        return fail(GRB200_EINVAL, "freq_xlating_fir_filter::set_center_freq: null handle");
    std::lock_guard<std::mutex> lk(x->mu);
    x->center_freq = center_freq;
    return xlat_rebuild(x);
}

int grb200_freq_xlating_fir_advance(grb200_xlat* x, long long noutput)
{
    if (!x)
        return fail(GRB200_EINVAL, "freq_xlating_fir_filter::advance: null handle");
    std::lock_guard<std::mutex> lk(x->fir->mu);
    x->fir->theta = rotate_advance(x->fir->theta, x->fir->delta, noutput);
    return 0;
}

double grb200_freq_xlating_fir_center_freq(const grb200_xlat* x) { return x ? x->center_freq : 0.0; }
int grb200_freq_xlating_fir_ntaps(const grb200_xlat* x) { return x ? grb200_fir_ntaps(x->fir) : 0; }
int grb200_freq_xlating_fir_decimation(const grb200_xlat* x) { return x ? grb200_fir_decimation(x->fir) : 0; }

int grb200_freq_xlating_fir_composite_taps(const grb200_xlat* x, float* taps_re_im_out, int capacity)
{
    if (!x || !taps_re_im_out || capacity < (int)x->proto.size())
        return fail(GRB200_EINVAL, "freq_xlating_fir_filter::composite_taps: bad arguments");
    std::vector<std::complex<float>> ctaps;
    double delta;
    xlat_composite(x, x->fir->decim, ctaps, &delta);
    std::memcpy(taps_re_im_out, ctaps.data(), ctaps.size() * sizeof(ctaps[0]));
    return 0;
}

int grb200_freq_xlating_fir_work(grb200_xlat* x, const void* in_host, void* out_host, long long noutput)
{
    if (!x)
        return fail(GRB200_EINVAL, "freq_xlating_fir_filter::work: null handle");
    return grb200_fir_work(x->fir, in_host, out_host, noutput);
}

int grb200_freq_xlating_fir_work_dev(grb200_xlat* x, const void* in_dev, void* out_dev, long long noutput, void* stream)
{
    if (!x)
        return fail(GRB200_EINVAL, "freq_xlating_fir_filter::work_dev: null handle");
    return grb200_fir_work_dev(x->fir, in_dev, out_dev, noutput, stream);
}

void grb200_freq_xlating_fir_destroy(grb200_xlat* x)
{
    if (!x)
        return;
    grb200_fir_destroy(x->fir);
    delete x;
}

} // extern "C"
