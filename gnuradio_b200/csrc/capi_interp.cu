// capi_interp.cu -- interp_fir_filter_ccf / interp_fir_filter_ccc (gr-filter/lib/interp_fir_filter_impl.cc): kernel + C-ABI.
//
// The reference keeps `interpolation` polyphase FIRs, xtaps[n][k] = taps[n + k*interpolation] (install_taps :97-117), and
// for every input item i writes out[i*L + n] = firs[n].filter(&in[i]) (work :139-147).  On the device a thread owns one
// input position: it loads the T-item window once into registers and runs all L phase filters against it (phase taps in
// shared memory, broadcast reads), so an input item is fetched T times per L*T multiply-adds; the L results of a position
// are contiguous in the output.  8 + 8L algorithmic bytes per input item: the block is bound by its own output stream.
#include <cstdint>
#include <type_traits>
#include "capi_common.cuh"
#include "tma.cuh"

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


// Tile form (EXPERIMENT, off by default; taps per phase 4 / 8 / 16 / 32, interpolation <= 16): ncu on the kernel above (profiles/r2_interp_ncu.txt) shows
// the load/store unit, not HBM, as the limit -- L1/TEX 78 % busy, long-scoreboard stalls 11 per issue: every FMA pair costs a
// scalar shared-memory tap load, the window comes through L1 one item at a time, and a warp's output stores are 16-byte
// pieces 8L bytes apart.  Here a CTA stages the 256 + T-1 input items of a tile ONCE (coalesced), a thread reads its window
// from shared memory (consecutive lanes, consecutive items), the taps of a phase arrive four per 16-byte broadcast load, and
// the 256 L results go through shared memory so that the tile's output leaves as one contiguous run of 16-byte stores.
template <int TT, bool CT>
__global__ void __launch_bounds__(256) interp_fir_tile_kernel(const float2* __restrict__ in, float2* __restrict__ out,
                                                              long long nin_pos, int L, const float* __restrict__ rev)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int ES = CT ? 2 : 1, PT = 256;
    float* sh = reinterpret_cast<float*>(smem_raw);                       // [L][TT][ES] taps
    float2* xs = reinterpret_cast<float2*>(sh + ((L * TT * ES + 3) & ~3)); // [PT + TT - 1] inputs (+1 pad)
    float2* ys = xs + ((PT + TT + 1) & ~1);                               // [PT * L] outputs, rows skewed by one pair per 8 positions
    const int tid = threadIdx.x;
    for (int k = tid; k < L * TT * ES; k += PT)
        sh[k] = rev[k];
    const long long ntiles = (nin_pos + PT - 1) / PT;
    const bool out16 = (reinterpret_cast<uintptr_t>(out) & 15) == 0;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const long long i0 = tile * PT;
        const int npos = (int)(nin_pos - i0 < PT ? nin_pos - i0 : PT);
        __syncthreads(); // the previous tile's outputs have left ys, its windows have been read
        for (int k = tid; k < npos + TT - 1; k += PT)
            xs[k] = __ldg(in + i0 + k);
        __syncthreads();
        if (tid < npos) {
            float2 w[TT];
#pragma unroll
            for (int t = 0; t < TT; t++)
                w[t] = xs[tid + t];
            float2* y = ys + (size_t)tid * L;
            for (int p = 0; p < L; p++) {
                const float4* h4 = reinterpret_cast<const float4*>(sh + (size_t)p * TT * ES);
                float2 a = make_float2(0.f, 0.f);
                if (CT) {
#pragma unroll
                    for (int t = 0; t < TT; t += 2) {
                        const float4 h = h4[t / 2]; // (hr, hi) of taps t, t+1
                        a.x = fmaf(w[t].x, h.x, fmaf(-w[t].y, h.y, a.x));
                        a.y = fmaf(w[t].x, h.y, fmaf(w[t].y, h.x, a.y));
                        a.x = fmaf(w[t + 1].x, h.z, fmaf(-w[t + 1].y, h.w, a.x));
                        a.y = fmaf(w[t + 1].x, h.w, fmaf(w[t + 1].y, h.z, a.y));
                    }
                } else {
#pragma unroll
                    for (int t = 0; t < TT; t += 4) {
                        const float4 h = h4[t / 4];
                        a = __ffma2_rn(make_float2(h.x, h.x), w[t], a);
                        a = __ffma2_rn(make_float2(h.y, h.y), w[t + 1], a);
                        a = __ffma2_rn(make_float2(h.z, h.z), w[t + 2], a);
                        a = __ffma2_rn(make_float2(h.w, h.w), w[t + 3], a);
                    }
                }
                y[p] = a;
            }
        }
        __syncthreads();
        // the tile's npos * L outputs are contiguous in the stream
        const long long o0 = i0 * L;
        const int nout = npos * L;
        if (out16 && (o0 & 1) == 0) {
            const float4* y4 = reinterpret_cast<const float4*>(ys);
            float4* o4 = reinterpret_cast<float4*>(out + o0);
            for (int k = tid; k < nout / 2; k += PT)
                o4[k] = y4[k];
            if ((nout & 1) && tid == 0)
                out[o0 + nout - 1] = ys[nout - 1];
        } else {
            for (int k = tid; k < nout; k += PT)
                out[o0 + k] = ys[k];
        }
    }
}

// ---- streaming form (round 2): the resampler's streaming kernel (capi_resamp.cu, resamp_stream_kernel) with decimation 1 ----
// What bounds the kernels above is the load/store unit: every multiply-add pair drags a shared-memory tap load along and the
// window comes through L1 item by item.  In the streaming form nothing but the stream itself moves: thread (phase l, run g)
// keeps the taps of its phase AND the last TP input items of its run in registers and walks down the run -- per output one
// 8-byte shared load (the new item, broadcast to the L lanes of the run), TP packed FMAs and one store; the L lanes of a run
// store L consecutive outputs (a whole 32-byte sector at L = 4); a producer warp streams the runs' new items with one bulk
// copy per run and iteration.  Same summation order as interp_fir_kernel (one accumulator, oldest tap first): a stream cut
// into calls of any length gives the same bits whichever kernel a call takes.
template <bool CT>
static int interp_launch_t(const void* in, void* out, long long nin_pos, int L, int T, const float* rev, const float* rev_tm,
                           int nsm, cudaStream_t st)
{
    const size_t smem = (size_t)L * T * (CT ? 8 : 4);
    if (smem > 200 * 1024)
        return fail(GRB200_ERANGE, "interp_fir_filter: %d x %d taps do not fit in shared memory", L, T);
    static const bool use_tile = [] {
        // measured on B200 (gpurun_out/r2_interp_check_v2.txt): 217 / 313 G output items/s at 4 x 16 / 8 x 8 against 355 / 439
        // for the one-position-per-thread kernel -- three barriers per 256-position tile cost more than the load/store
        // unit saves.  Kept as an experiment behind GRB200_INTERP_TILE=1.
        const char* e = getenv("GRB200_INTERP_TILE");
        return e && e[0] == '1';
    }();
    static const bool use_stream = [] {
        const char* e = GRB_GETENV_ONCE("GRB200_INTERP_STREAM"); // 0 = the one-position-per-thread kernel only
        return !(e && e[0] == '0');
    }();
    if (use_stream && rev_tm && nin_pos >= 8192) {
        bool done = false;
        const int rc = resamp_stream_try(CT, in, out, nin_pos * L, 0u, (unsigned)L, 1u, T, rev_tm, nsm, st, &done);
        if (rc || done)
            return rc;
    }
    if (use_tile && L <= 16 && (T == 4 || T == 8 || T == 16 || T == 32) && nin_pos >= 1024) {
        const size_t tsm = (((size_t)L * T * (CT ? 2 : 1) + 3) & ~(size_t)3) * 4 + (size_t)((256 + T + 1) & ~1) * 8 +
                           (size_t)256 * L * 8;
        long long ntiles = (nin_pos + 255) / 256, grid = (long long)nsm * 6;
        if (grid > ntiles)
            grid = ntiles;
        auto tl = [&](auto kern) -> int {
            if (tsm > 48 * 1024)
                GRB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tsm));
            kern<<<(unsigned)grid, 256, tsm, st>>>((const float2*)in, (float2*)out, nin_pos, L, rev);
            GRB_LAUNCH_CHECK();
            return 0;
        };
        switch (T) {
        case 4:
            return tl(interp_fir_tile_kernel<4, CT>);
        case 8:
            return tl(interp_fir_tile_kernel<8, CT>);
        case 16:
            return tl(interp_fir_tile_kernel<16, CT>);
        default:
            return tl(interp_fir_tile_kernel<32, CT>);
        }
    }
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

static int interp_launch(bool ct, const void* in, void* out, long long nin_pos, int L, int T, const float* rev,
                         const float* rev_tm, int nsm, cudaStream_t st)
{
    if (nin_pos <= 0)
        return 0;
    return ct ? interp_launch_t<true>(in, out, nin_pos, L, T, rev, rev_tm, nsm, st)
              : interp_launch_t<false>(in, out, nin_pos, L, T, rev, rev_tm, nsm, st);
}

} // namespace grb

struct grb200_interp {
    int device = 0;
    bool ct = false;
    int L = 1, T = 0, ntaps = 0; // interpolation, taps per phase, taps as given
    DevBuf rev, rev_tm; // phase-major reversed taps; the same tap-major (the streaming kernel's table)
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
    std::vector<float> tm(rev.size());
    for (int p = 0; p < L; p++)
        for (int t = 0; t < T; t++)
            for (int e = 0; e < es; e++)
                tm[((size_t)t * L + p) * es + e] = rev[((size_t)p * T + t) * es + e];
    if ((rc = h->rev_tm.upload(tm.data(), tm.size() * sizeof(float))))
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
    return dev_done(interp_launch(h->ct, in_dev, out_dev, noutput / h->L, h->L, h->T, (const float*)h->rev.p, (const float*)h->rev_tm.p, sm_count(h->device),
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
                                 (const float*)h->rev_tm.p, sm_count(h->device), st);
        });
}

void grb200_interp_fir_destroy(grb200_interp* h)
{
    if (!h)
        return;
    grb::DeviceScope _ds(h->device);
    h->sg.destroy();
    h->rev.release();
    h->rev_tm.release();
    delete h;
}

} // extern "C"
