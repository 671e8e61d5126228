// capi_interp.cu -- interp_fir_filter_ccf / interp_fir_filter_ccc (gr-filter/lib/interp_fir_filter_impl.cc): kernel + C-ABI.
//
// The reference keeps `interpolation` polyphase FIRs, xtaps[n][k] = taps[n + k*interpolation] (install_taps :97-117), and
// for every input item i writes out[i*L + n] = firs[n].filter(&in[i]) (work :139-147).  On the device a thread owns one
// input position: it loads the T-item window once into registers and runs all L phase filters against it (phase taps in
// shared memory, broadcast reads), so an input item is fetched T times per L*T multiply-adds; the L results of a position
// are contiguous in the output.  8 + 8L algorithmic bytes per input item: the block is bound by its own output stream.
#include <cstdint>
#include "capi_common.cuh"

using namespace grb;

namespace grb {

// rev[(p*T + t) * ES + e]: reversed taps of phase p (fir_filter.cc:62-64), ES = 1 (real) or 2 (complex)
template <int TT, bool CT>
__global__ void __launch_bounds__(256) interp_fir_kernel(const float2* __restrict__ in, float2* __restrict__ out,
                                                         long long nin_pos, int L, int T, const float* __restrict__ rev)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* sh = reinterpret_cast<float*>(smem_raw);
    constexpr int ES = CT ? 2 : 1;
    for (int k = threadIdx.x; k < L * T * ES; k += blockDim.x)
        sh[k] = rev[k];
    __syncthreads();
    const long long step = (long long)gridDim.x * blockDim.x;
    const bool pair_ok = (reinterpret_cast<uintptr_t>(out) & 15) == 0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nin_pos; i += step) {
        const float2* x = in + i;
        float2* y = out + i * L;
        if (TT > 0) {
            float2 w[TT > 0 ? TT : 1];
#pragma unroll
            for (int t = 0; t < TT; t++)
                w[t] = __ldg(x + t);
            auto phase = [&](int p) {
                const float* h = sh + (size_t)p * TT * ES;
                float2 a = make_float2(0.f, 0.f);
#pragma unroll
                for (int t = 0; t < TT; t++) {
                    if (CT) {
                        const float hr = h[2 * t], hi = h[2 * t + 1];
                        a.x = fmaf(w[t].x, hr, fmaf(-w[t].y, hi, a.x));
                        a.y = fmaf(w[t].x, hi, fmaf(w[t].y, hr, a.y));
                    } else {
                        a.x = fmaf(w[t].x, h[t], a.x);
                        a.y = fmaf(w[t].y, h[t], a.y);
                    }
                }
                return a;
            };
            if ((L & 1) == 0 && pair_ok) { // two phases per 16-byte store: item i*L + p is even
                for (int p = 0; p < L; p += 2) {
                    const float2 a = phase(p), b = phase(p + 1);
                    *reinterpret_cast<float4*>(y + p) = make_float4(a.x, a.y, b.x, b.y);
                }
            } else {
                for (int p = 0; p < L; p++)
                    y[p] = phase(p);
            }
        } else {
            for (int p = 0; p < L; p++) {
                const float* h = sh + (size_t)p * T * ES;
                float2 a = make_float2(0.f, 0.f);
                for (int t = 0; t < T; t++) {
                    const float2 v = __ldg(x + t);
                    if (CT) {
                        const float hr = h[2 * t], hi = h[2 * t + 1];
                        a.x = fmaf(v.x, hr, fmaf(-v.y, hi, a.x));
                        a.y = fmaf(v.x, hi, fmaf(v.y, hr, a.y));
                    } else {
                        a.x = fmaf(v.x, h[t], a.x);
                        a.y = fmaf(v.y, h[t], a.y);
                    }
                }
                y[p] = a;
            }
        }
    }
}

template <bool CT>
static int interp_launch_t(const void* in, void* out, long long nin_pos, int L, int T, const float* rev, int nsm,
                           cudaStream_t st)
{
    const size_t smem = (size_t)L * T * (CT ? 8 : 4);
    if (smem > 200 * 1024)
        return fail(GRB200_ERANGE, "interp_fir_filter: %d x %d taps do not fit in shared memory", L, T);
    long long nblk = (nin_pos + 255) / 256;
    const long long cap = (long long)nsm * 8;
    if (nblk > cap)
        nblk = cap;
    auto launch = [&](auto kern) -> int {
        if (smem > 48 * 1024)
            GRB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        kern<<<(unsigned)nblk, 256, smem, st>>>((const float2*)in, (float2*)out, nin_pos, L, T, rev);
        GRB_LAUNCH_CHECK();
        return 0;
    };
    switch (T) {
    case 1:
        return launch(interp_fir_kernel<1, CT>);
    case 2:
        return launch(interp_fir_kernel<2, CT>);
    case 4:
        return launch(interp_fir_kernel<4, CT>);
    case 8:
        return launch(interp_fir_kernel<8, CT>);
    case 11:
        return launch(interp_fir_kernel<11, CT>);
    case 16:
        return launch(interp_fir_kernel<16, CT>);
    case 32:
        return launch(interp_fir_kernel<32, CT>);
    }
    return launch(interp_fir_kernel<0, CT>);
}

static int interp_launch(bool ct, const void* in, void* out, long long nin_pos, int L, int T, const float* rev, int nsm,
                         cudaStream_t st)
{
    if (nin_pos <= 0)
        return 0;
    return ct ? interp_launch_t<true>(in, out, nin_pos, L, T, rev, nsm, st)
              : interp_launch_t<false>(in, out, nin_pos, L, T, rev, nsm, st);
}

} // namespace grb

struct grb200_interp {
    int device = 0;
    bool ct = false;
    int L = 1, T = 0, ntaps = 0; // interpolation, taps per phase, taps as given
    DevBuf rev;
    Stager sg;
    std::mutex mu;
};

static int interp_upload(grb200_interp* h, const float* taps, int ntaps)
{
    // set_taps (:72-90): pad with zeros to a multiple of the interpolation; install_taps (:97-112): deal round-robin
    const int L = h->L, es = h->ct ? 2 : 1;
    const int T = (ntaps + L - 1) / L;
    std::vector<float> rev((size_t)L * T * es, 0.f);
    for (int i = 0; i < ntaps; i++) {
        const int p = i % L, k = i / L; // xtaps[p][k]
        for (int e = 0; e < es; e++)
            rev[((size_t)p * T + (T - 1 - k)) * es + e] = taps[(size_t)i * es + e];
    }
    int rc = h->rev.upload(rev.data(), rev.size() * sizeof(float));
    if (rc)
        return rc;
    h->T = T;
    h->ntaps = ntaps;
    return 0;
}

static int interp_create(bool ct, unsigned interpolation, const float* taps, int ntaps, int device, grb200_interp** out)
{
    if (!out)
        return fail(GRB200_EINVAL, "interp_fir_filter: null output handle");
    *out = nullptr;
    if (interpolation == 0) // std::out_of_range in the reference (:46-48)
        return fail(GRB200_ERANGE, "interp_fir_filter_impl: interpolation must be > 0");
    if (ntaps < 1 || !taps) // std::runtime_error in the reference (:50-52)
        return fail(GRB200_EINVAL, "interp_fir_filter_impl: no filter taps provided.");
    std::unique_ptr<grb200_interp> h(new grb200_interp);
    h->device = device;
    h->ct = ct;
    h->L = (int)interpolation;
    grb::DeviceScope _ds_create(device); // the caller keeps its own current device
    int rc = h->sg.init(device);
    if (rc)
        return rc;
    if ((rc = interp_upload(h.get(), taps, ntaps)))
        return rc;
    *out = h.release();
    return 0;
}

extern "C" {

int grb200_interp_fir_ccf_create(unsigned int interpolation, const float* taps, int ntaps, int device, grb200_interp** out)
{
    return interp_create(false, interpolation, taps, ntaps, device, out);
}
int grb200_interp_fir_ccc_create(unsigned int interpolation, const float* taps_re_im, int ntaps, int device,
                                 grb200_interp** out)
{
    return interp_create(true, interpolation, taps_re_im, ntaps, device, out);
}

int grb200_interp_fir_set_taps(grb200_interp* h, const float* taps, int ntaps)
{
    if (!h || !taps || ntaps < 1)
        return fail(GRB200_EINVAL, "interp_fir_filter::set_taps: bad arguments");
    std::lock_guard<std::mutex> lk(h->mu);
    grb::DeviceScope _ds(h->device);
    int rc = _ds.rc;
    if (rc)
        return rc;
    if ((rc = h->sg.quiesce())) // launches through this handle may still read the old state
        return rc;
    return interp_upload(h, taps, ntaps);
}

int grb200_interp_fir_taps_per_filter(const grb200_interp* h) { return h ? h->T : 0; }
int grb200_interp_fir_interpolation(const grb200_interp* h) { return h ? h->L : 0; }

// in: noutput/interpolation + taps_per_filter-1 items, history first; out: noutput items
int grb200_interp_fir_work_dev(grb200_interp* h, const void* in_dev, void* out_dev, long long noutput, void* stream)
{
    if (!h || noutput < 0 || (noutput > 0 && (!in_dev || !out_dev)))
        return fail(GRB200_EINVAL, "interp_fir_filter::work_dev: bad arguments");
    if (noutput % h->L != 0)
        return fail(GRB200_EINVAL, "interp_fir_filter: noutput (%lld) must be a multiple of the interpolation %d", noutput,
                    h->L);
    std::lock_guard<std::mutex> lk(h->mu);
    grb::DeviceScope _ds(h->device);
    int rc = _ds.rc;
    if (rc)
        return rc;
    cudaStream_t st_ = pick_stream(stream, h->sg);
    return dev_done(interp_launch(h->ct, in_dev, out_dev, noutput / h->L, h->L, h->T, (const float*)h->rev.p, sm_count(h->device),
                         st_), h->sg, st_);
}

int grb200_interp_fir_work(grb200_interp* h, const void* in_host, void* out_host, long long noutput)
{
    if (!h || noutput < 0 || (noutput > 0 && (!in_host || !out_host)))
        return fail(GRB200_EINVAL, "interp_fir_filter::work: bad arguments");
    if (noutput == 0)
        return 0;
    if (noutput % h->L != 0)
        return fail(GRB200_EINVAL, "interp_fir_filter: noutput (%lld) must be a multiple of the interpolation %d", noutput,
                    h->L);
    std::lock_guard<std::mutex> lk(h->mu);
    grb::DeviceScope _ds(h->device);
    int rc = _ds.rc;
    if (rc)
        return rc;
    const int L = h->L, T = h->T;
    const long long ni = noutput / L;
    long long chunk = std::max<long long>(1, (long long)(chunk_bytes_for((size_t)noutput * 8) / 8) / L); // input positions per chunk
    if (chunk > ni)
        chunk = ni;
    const long long nchunks = (ni + chunk - 1) / chunk;
    Stager& sg = h->sg;
    if ((rc = sg.reserve_dev(0, (size_t)(chunk + T - 1) * 8)) || (rc = sg.reserve_dev(1, (size_t)chunk * L * 8)))
        return rc;
    const float2* in = (const float2*)in_host;
    float2* out = (float2*)out_host;
    return sg.run(
        nchunks,
        [&](long long c, int slot, Stager::Chunk& ch) {
            const long long i0 = c * chunk, n = std::min(chunk, ni - i0);
            ch.h2d.push_back({ in + i0, sg.dev[slot][0], (size_t)(n + T - 1) * 8 });
            ch.d2h.push_back({ sg.dev[slot][1], out + i0 * L, (size_t)n * L * 8 });
            return 0;
        },
        [&](long long c, int slot, cudaStream_t st) {
            const long long i0 = c * chunk, n = std::min(chunk, ni - i0);
            return interp_launch(h->ct, sg.dev[slot][0], sg.dev[slot][1], n, L, T, (const float*)h->rev.p,
                                 sm_count(h->device), st);
        });
}

void grb200_interp_fir_destroy(grb200_interp* h)
{
    if (!h)
        return;
    grb::DeviceScope _ds(h->device);
    h->sg.destroy();
    h->rev.release();
    delete h;
}

} // extern "C"
