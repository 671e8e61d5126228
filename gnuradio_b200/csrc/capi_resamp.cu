// capi_resamp.cu -- rational_resampler_base_ccf / _ccc (gr-filter/lib/rational_resampler_base_impl.cc): kernel + C-ABI.
//
// The reference keeps `interpolation` polyphase FIRs, xtaps[n][k] = taps[n + k*interpolation] (install_taps :97-117), and
// walks a phase counter: out[i] = firs[ctr].filter(in); ctr += decimation; every time ctr wraps past the interpolation
// the input pointer advances (general_work :151-173).  In closed form, with s_i = ctr0 + i*decimation:
//     out[i] = sum_t rev[s_i % L][t] * in[s_i / L + t],   consumed = s_n / L,   ctr' = s_n % L
// so every output is independent.  A thread owns one output item: phase and window start from one divide (32-bit when
// the launch's largest s_i fits), phase taps from shared memory (tap-major: the phases a warp holds at one tap index sit
// in different banks), window through the read-only path (neighbouring outputs share all but D/L of their window, so
// L1 serves the re-reads).  8*D/L + 8 algorithmic bytes per output item.
#include <cstdint>
#include <type_traits>
#include "capi_common.cuh"

using namespace grb;

namespace grb {

// rev[(t*L + p) * ES + e]: reversed taps of phase p (fir_filter.cc:62-64), tap-major so the phases a warp holds at one
// tap index sit in different banks; ES = 1 (real) or 2 (complex).  W32: every s_i fits 32 bits (one 32-bit divide).
template <bool CT, bool W32>
__global__ void __launch_bounds__(256) resamp_kernel(const float2* __restrict__ in, float2* __restrict__ out,
                                                     long long nout, unsigned ctr0, unsigned L, unsigned D, int T,
                                                     const float* __restrict__ rev)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* sh = reinterpret_cast<float*>(smem_raw);
    constexpr int ES = CT ? 2 : 1;
    for (int k = threadIdx.x; k < (int)L * T * ES; k += blockDim.x)
        sh[k] = rev[k];
    __syncthreads();
    const long long step = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nout; i += step) {
        unsigned long long q;
        unsigned p;
        if (W32) {
            const unsigned s = ctr0 + (unsigned)i * D;
            const unsigned q32 = s / L;
            p = s - q32 * L;
            q = q32;
        } else {
            const unsigned long long s = (unsigned long long)ctr0 + (unsigned long long)i * D;
            q = s / L;
            p = (unsigned)(s - q * L);
        }
        const float2* x = in + q;
        float2 a = make_float2(0.f, 0.f);
        if (CT) {
            const float2* h = reinterpret_cast<const float2*>(sh) + p;
#pragma unroll 8
            for (int t = 0; t < T; t++) {
                const float2 v = __ldg(x + t);
                const float2 c = h[(size_t)t * L];
                a.x = fmaf(v.x, c.x, fmaf(-v.y, c.y, a.x));
                a.y = fmaf(v.x, c.y, fmaf(v.y, c.x, a.y));
            }
        } else {
            const float* h = sh + p;
#pragma unroll 8
            for (int t = 0; t < T; t++) {
                const float2 v = __ldg(x + t);
                const float c = h[(size_t)t * L];
                a.x = fmaf(v.x, c, a.x);
                a.y = fmaf(v.y, c, a.y);
            }
        }
        out[i] = a;
    }
}

// Phase-fixed variant: with G = (a multiple of L) active threads and a grid stride of G outputs, thread g keeps the phase
// (ctr0 + g*D) mod L for all of its outputs and its window start advances by the integer G*D/L per step -- the TT phase
// taps live in registers (no shared-memory reads, no divide in the loop) and the loop is TT/2 16-byte read-only loads + FMAs.
// PAIR: the window is read in 16-byte pairs (pays when windows of neighbouring threads overlap little, D/L >= 0.8);
// otherwise item by item (neighbouring threads hit the same L1 lines).  Same FMA order either way.
template <bool CT, int TT, bool PAIR>
__global__ void __launch_bounds__(256) resamp_phase_kernel(const float2* __restrict__ in, float2* __restrict__ out,
                                                           long long nout, unsigned ctr0, unsigned L, unsigned D,
                                                           long long G, long long qstep, const float* __restrict__ rev)
{
    const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= G)
        return;
    const unsigned long long s = (unsigned long long)ctr0 + (unsigned long long)g * D;
    unsigned long long q = s / L;
    const unsigned p = (unsigned)(s - q * L);
    // 16-byte window loads: the window is read from the even item at or below q; an odd q shifts the taps up by one
    // (zero in front) and needs one more item at the end.  G/L * D is even, so a thread keeps its parity.
    const unsigned odd = PAIR ? (unsigned)(q & 1) : 0u;
    q -= odd;
    float hr[TT + 1], hi[CT ? TT + 1 : 1];
#pragma unroll
    for (int j = 0; j <= TT; j++) {
        const int t = j - (int)odd;
        const bool ok = t >= 0 && t < TT;
        if (CT) {
            const float2 c = ok ? __ldg(reinterpret_cast<const float2*>(rev) + (size_t)t * L + p) : make_float2(0.f, 0.f);
            hr[j] = c.x;
            hi[j] = c.y;
        } else {
            hr[j] = ok ? __ldg(rev + (size_t)t * L + p) : 0.f;
        }
    }
    auto mac = [&](float2& a, const float2 v, const int j) {
        if (CT) {
            a.x = fmaf(v.x, hr[j], fmaf(-v.y, hi[j], a.x));
            a.y = fmaf(v.x, hi[j], fmaf(v.y, hr[j], a.y));
        } else {
            a.x = fmaf(v.x, hr[j], a.x);
            a.y = fmaf(v.y, hr[j], a.y);
        }
    };
    for (long long i = g; i < nout; i += G, q += qstep) {
        float2 a = make_float2(0.f, 0.f);
        if (PAIR) {
            const float4* x = reinterpret_cast<const float4*>(in + q);
#pragma unroll
            for (int j = 0; j < TT / 2; j++) {
                const float4 v = __ldg(x + j);
                // an odd window starts one item late: in[q] lies OUTSIDE the reference's window, so it must not enter the
                // sum at all (0 * Inf/NaN would poison an output the reference keeps finite)
                if (!(j == 0 && odd))
                    mac(a, make_float2(v.x, v.y), 2 * j);
                mac(a, make_float2(v.z, v.w), 2 * j + 1);
            }
            if (odd)
                mac(a, __ldg(in + q + TT), TT);
        } else {
#pragma unroll
            for (int t = 0; t < TT; t++)
                mac(a, __ldg(in + q + t), t);
        }
        out[i] = a;
    }
}

template <bool CT>
static int resamp_phase_launch(const void* in, void* out, long long nout, unsigned ctr0, unsigned L, unsigned D, int T,
                               const float* rev, long long nblk, cudaStream_t st, bool* done)
{
    *done = false;
    // 16-byte pairs when neighbouring windows overlap little and the stream is 16-byte aligned
    const bool pair = 5ull * D >= 4ull * L && (reinterpret_cast<uintptr_t>(in) & 15) == 0;
    const long long unit = (pair && (D & 1)) ? 2ll * L : (long long)L; // pairs: G/L * D even keeps a thread's parity
    const long long G = nblk * 256 / unit * unit;
    if (G == 0)
        return 0;
    const long long qstep = G / L * D;
    auto launch = [&](auto kern) -> int {
        kern<<<(unsigned)nblk, 256, 0, st>>>((const float2*)in, (float2*)out, nout, ctr0, L, D, G, qstep, rev);
        GRB_LAUNCH_CHECK();
        *done = true;
        return 0;
    };
    auto pick = [&](auto tt) -> int {
        constexpr int TTv = decltype(tt)::value;
        return pair ? launch(resamp_phase_kernel<CT, TTv, true>) : launch(resamp_phase_kernel<CT, TTv, false>);
    };
    switch (T) {
    case 4:
        return pick(std::integral_constant<int, 4>{});
    case 8:
        return pick(std::integral_constant<int, 8>{});
    case 10:
        return pick(std::integral_constant<int, 10>{});
    case 12:
        return pick(std::integral_constant<int, 12>{});
    case 16:
        return pick(std::integral_constant<int, 16>{});
    case 24:
        return pick(std::integral_constant<int, 24>{});
    case 32:
        return pick(std::integral_constant<int, 32>{});
    case 64:
        return pick(std::integral_constant<int, 64>{});
    }
    return 0;
}

static int resamp_launch(bool ct, const void* in, void* out, long long nout, unsigned ctr0, unsigned L, unsigned D, int T,
                         const float* rev, int nsm, cudaStream_t st)
{
    if (nout <= 0)
        return 0;
    const size_t smem = (size_t)L * T * (ct ? 8 : 4);
    if (smem > 200 * 1024)
        return fail(GRB200_ERANGE, "rational_resampler_base: %u x %d taps do not fit in shared memory", L, T);
    long long nblk = (nout + 255) / 256;
    const long long cap = (long long)nsm * 8;
    if (nblk > cap)
        nblk = cap;
    auto launch = [&](auto kern) -> int {
        if (smem > 48 * 1024)
            GRB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        kern<<<(unsigned)nblk, 256, smem, st>>>((const float2*)in, (float2*)out, nout, ctr0, L, D, T, rev);
        GRB_LAUNCH_CHECK();
        return 0;
    };
    bool done = false;
    int rc = ct ? resamp_phase_launch<true>(in, out, nout, ctr0, L, D, T, rev, nblk, st, &done)
                : resamp_phase_launch<false>(in, out, nout, ctr0, L, D, T, rev, nblk, st, &done);
    if (rc || done)
        return rc;
    const bool w32 = (unsigned long long)ctr0 + (unsigned long long)nout * D < (1ull << 32);
    if (w32)
        return ct ? launch(resamp_kernel<true, true>) : launch(resamp_kernel<false, true>);
    return ct ? launch(resamp_kernel<true, false>) : launch(resamp_kernel<false, false>);
}

} // namespace grb

struct grb200_resamp {
    int device = 0;
    bool ct = false;
    unsigned L = 1, D = 1, ctr = 0; // interpolation, decimation, d_ctr
    int T = 0, ntaps = 0;           // taps per phase, taps as given
    DevBuf rev;
    Stager sg;
    std::mutex mu;
};

static int resamp_upload(grb200_resamp* h, const float* taps, int ntaps)
{
    // set_taps (:77-93): pad with zeros to a multiple of the interpolation; install_taps (:97-117): deal round-robin
    const int L = (int)h->L, es = h->ct ? 2 : 1;
    const int T = (ntaps + L - 1) / L;
    std::vector<float> rev((size_t)L * T * es, 0.f);
    for (int i = 0; i < ntaps; i++) {
        const int p = i % L, k = i / L; // xtaps[p][k]
        for (int e = 0; e < es; e++)
            rev[((size_t)(T - 1 - k) * L + p) * es + e] = taps[(size_t)i * es + e];
    }
    int rc = h->rev.upload(rev.data(), rev.size() * sizeof(float));
    if (rc)
        return rc;
    h->T = T;
    h->ntaps = ntaps;
    return 0;
}

static int resamp_create(bool ct, unsigned interpolation, unsigned decimation, const float* taps, int ntaps, int device,
                         grb200_resamp** out)
{
    if (!out)
        return fail(GRB200_EINVAL, "rational_resampler_base: null output handle");
    *out = nullptr;
    if (interpolation == 0) // std::out_of_range in the reference (:46-48)
        return fail(GRB200_ERANGE, "rational_resampler_base_impl: interpolation must be > 0");
    if (decimation == 0) // std::out_of_range in the reference (:49-51)
        return fail(GRB200_ERANGE, "rational_resampler_base_impl: decimation must be > 0");
    if (ntaps < 1 || !taps)
        return fail(GRB200_EINVAL, "rational_resampler_base_impl: no filter taps provided.");
    std::unique_ptr<grb200_resamp> h(new grb200_resamp);
    h->device = device;
    h->ct = ct;
    h->L = interpolation;
    h->D = decimation;
    grb::DeviceScope _ds_create(device); // the caller keeps its own current device
    int rc = h->sg.init(device);
    if (rc)
        return rc;
    if ((rc = resamp_upload(h.get(), taps, ntaps)))
        return rc;
    *out = h.release();
    return 0;
}

// the loop of general_work (:160-169) in closed form: outputs i = 0.. while i < noutput and s_i / L < nin
static void resamp_counts(const grb200_resamp* h, long long nin, long long noutput, long long* produced,
                          long long* consumed, unsigned* ctr_after)
{
    const unsigned long long L = h->L, D = h->D, c0 = h->ctr;
    long long n = 0;
    if (nin > 0 && noutput > 0) {
        const unsigned long long room = (unsigned long long)nin * L - c0; // c0 < L <= nin*L
        const unsigned long long fit = (room + D - 1) / D;
        n = (long long)std::min<unsigned long long>(fit, (unsigned long long)noutput);
    }
    const unsigned long long s = c0 + (unsigned long long)n * D;
    *produced = n;
    *consumed = (long long)(s / L);
    *ctr_after = (unsigned)(s % L);
}

extern "C" {

int grb200_rational_resampler_ccf_create(unsigned int interpolation, unsigned int decimation, const float* taps, int ntaps,
                                         int device, grb200_resamp** out)
{
    return resamp_create(false, interpolation, decimation, taps, ntaps, device, out);
}
int grb200_rational_resampler_ccc_create(unsigned int interpolation, unsigned int decimation, const float* taps_re_im,
                                         int ntaps, int device, grb200_resamp** out)
{
    return resamp_create(true, interpolation, decimation, taps_re_im, ntaps, device, out);
}

int grb200_rational_resampler_set_taps(grb200_resamp* h, const float* taps, int ntaps)
{
    if (!h || !taps || ntaps < 1)
        return fail(GRB200_EINVAL, "rational_resampler_base::set_taps: bad arguments");
    std::lock_guard<std::mutex> lk(h->mu);
    grb::DeviceScope _ds(h->device);
    int rc = _ds.rc;
    if (rc)
        return rc;
    if ((rc = h->sg.quiesce())) // launches through this handle may still read the old state
        return rc;
    return resamp_upload(h, taps, ntaps);
}

int grb200_rational_resampler_taps_per_filter(const grb200_resamp* h) { return h ? h->T : 0; }
unsigned int grb200_rational_resampler_interpolation(const grb200_resamp* h) { return h ? h->L : 0; }
unsigned int grb200_rational_resampler_decimation(const grb200_resamp* h) { return h ? h->D : 0; }
unsigned int grb200_rational_resampler_ctr(const grb200_resamp* h) { return h ? h->ctr : 0; }

int grb200_rational_resampler_set_ctr(grb200_resamp* h, unsigned int ctr)
{
    if (!h || ctr >= h->L)
        return fail(GRB200_EINVAL, "rational_resampler_base::set_ctr: the phase counter must be below the interpolation");
    std::lock_guard<std::mutex> lk(h->mu);
    h->ctr = ctr;
    return 0;
}

// in: nin + taps_per_filter-1 readable items; out: room for noutput items
int grb200_rational_resampler_work_dev(grb200_resamp* h, const void* in_dev, long long nin, void* out_dev,
                                       long long noutput, long long* consumed, long long* produced, void* stream)
{
    if (!h || nin < 0 || noutput < 0 || !consumed || !produced)
        return fail(GRB200_EINVAL, "rational_resampler_base::work_dev: bad arguments");
    std::lock_guard<std::mutex> lk(h->mu);
    unsigned ctr_after;
    long long n, used;
    resamp_counts(h, nin, noutput, &n, &used, &ctr_after);
    if (n > 0 && (!in_dev || !out_dev))
        return fail(GRB200_EINVAL, "rational_resampler_base::work_dev: null buffer");
    grb::DeviceScope _ds(h->device);
    int rc = _ds.rc;
    if (rc)
        return rc;
    cudaStream_t st_ = pick_stream(stream, h->sg);
    rc = dev_done(resamp_launch(h->ct, in_dev, out_dev, n, h->ctr, h->L, h->D, h->T, (const float*)h->rev.p,
                                sm_count(h->device), st_),
                  h->sg, st_);
    if (rc)
        return rc;
    h->ctr = ctr_after;
    *consumed = used;
    *produced = n;
    return 0;
}

int grb200_rational_resampler_work(grb200_resamp* h, const void* in_host, long long nin, void* out_host, long long noutput,
                                   long long* consumed, long long* produced)
{
    if (!h || nin < 0 || noutput < 0 || !consumed || !produced)
        return fail(GRB200_EINVAL, "rational_resampler_base::work: bad arguments");
    std::lock_guard<std::mutex> lk(h->mu);
    unsigned ctr_after;
    long long n, used;
    resamp_counts(h, nin, noutput, &n, &used, &ctr_after);
    *consumed = 0;
    *produced = 0;
    if (n > 0) {
        if (!in_host || !out_host)
            return fail(GRB200_EINVAL, "rational_resampler_base::work: null buffer");
        grb::DeviceScope _ds(h->device);
        int rc = _ds.rc;
        if (rc)
            return rc;
        const unsigned long long L = h->L, D = h->D, c0 = h->ctr;
        const int T = h->T;
        // outputs per chunk: bound both the output bytes and the input span (n*D/L items) of a chunk
        long long chunk = (long long)(chunk_bytes_for((size_t)std::max<unsigned long long>(n, (unsigned long long)n * D / L) * 8) / 8);
        if (D > L)
            chunk = std::max<long long>(1, (long long)((unsigned long long)chunk * L / D));
        if (chunk > n)
            chunk = n;
        const long long nchunks = (n + chunk - 1) / chunk;
        const long long span = (long long)(((unsigned long long)chunk * D + L - 1) / L) + 1 + T; // input items per chunk
        Stager& sg = h->sg;
        if ((rc = sg.reserve_dev(0, (size_t)span * 8)) || (rc = sg.reserve_dev(1, (size_t)chunk * 8)))
            return rc;
        const float2* in = (const float2*)in_host;
        float2* out = (float2*)out_host;
        auto geom = [&](long long c, long long* o0, long long* no, unsigned long long* q0, unsigned* cc, long long* nitems) {
            *o0 = c * chunk;
            *no = std::min(chunk, n - *o0);
            const unsigned long long s0 = c0 + (unsigned long long)*o0 * D;
            const unsigned long long sl = c0 + (unsigned long long)(*o0 + *no - 1) * D;
            *q0 = s0 / L;
            *cc = (unsigned)(s0 % L);
            *nitems = (long long)(sl / L - *q0) + T; // first window start .. last window end
        };
        rc = sg.run(
            nchunks,
            [&](long long c, int slot, Stager::Chunk& ch) {
                long long o0, no, nitems;
                unsigned long long q0;
                unsigned cc;
                geom(c, &o0, &no, &q0, &cc, &nitems);
                ch.h2d.push_back({ in + q0, sg.dev[slot][0], (size_t)nitems * 8 });
                ch.d2h.push_back({ sg.dev[slot][1], out + o0, (size_t)no * 8 });
                return 0;
            },
            [&](long long c, int slot, cudaStream_t st) {
                long long o0, no, nitems;
                unsigned long long q0;
                unsigned cc;
                geom(c, &o0, &no, &q0, &cc, &nitems);
                return resamp_launch(h->ct, sg.dev[slot][0], sg.dev[slot][1], no, cc, h->L, h->D, T, (const float*)h->rev.p,
                                     sm_count(h->device), st);
            });
        if (rc)
            return rc;
    }
    h->ctr = ctr_after;
    *consumed = used;
    *produced = n;
    return 0;
}

void grb200_rational_resampler_destroy(grb200_resamp* h)
{
    if (!h)
        return;
    grb::DeviceScope _ds(h->device);
    h->sg.destroy();
    h->rev.release();
    delete h;
}

} // extern "C"
