// capi_pfbdec.cu -- pfb_decimator_ccf (gr-filter/lib/pfb_decimator_ccf_impl.cc): kernel + C-ABI.
//
// The reference offers four work() variants (FIR or FFT branch filters x rotator table or M-point FFT,
// pfb_decimator_ccf_impl.cc:135-257); all four compute
//     out[i] = sum_j  rot[j] * sum_t taps_j[t] * in_{M-1-j}[i - t],   rot[j] = e^{+j 2 pi j chan / M},
// and differ only in rounding.  The device evaluates that sum directly: one pass over the M input streams, real-tap
// branch dot products, one complex multiply per branch: 4T + 8 flops per INPUT item against its 8 bytes.
#include <cstdint>
#include <cstdlib>
#include "capi_common.cuh"
#include "tma.cuh"

using namespace grb;

namespace grb {

// X: M streams, stream s at X + s*stride, nout + T-1 items each (T-1 history items first).
// crev[s*T + t] = taps_{M-1-s}[T-1-t] (reversed branch taps, indexed by input stream), rot[s] = rot of branch M-1-s.
template <int TT>
__global__ void __launch_bounds__(256) pfbdec_kernel(const float2* __restrict__ X, long long stride,
                                                     float2* __restrict__ out, long long nout, int M, int T,
                                                     const float* __restrict__ crev, const float2* __restrict__ rot)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float2* srot = reinterpret_cast<float2*>(smem_raw);
    float* srev = reinterpret_cast<float*>(smem_raw + (size_t)M * 8);
    for (int k = threadIdx.x; k < M; k += blockDim.x)
        srot[k] = rot[k];
    for (int k = threadIdx.x; k < M * T; k += blockDim.x)
        srev[k] = crev[k];
    __syncthreads();
    const long long step = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nout; i += step) {
        float2 acc = make_float2(0.f, 0.f);
        const float2* x = X + i;
        // branch order of work_fir_exp (:146-150): j = M-1 .. 0, i.e. input streams 0 .. M-1
        for (int s = 0; s < M; s++, x += stride) {
            float2 a = make_float2(0.f, 0.f);
            const float* h = srev + s * T;
            if (TT > 0) {
#pragma unroll
                for (int t = 0; t < TT; t++) {
                    const float2 v = __ldg(x + t);
                    a.x = fmaf(v.x, h[t], a.x);
                    a.y = fmaf(v.y, h[t], a.y);
                }
            } else {
                for (int t = 0; t < T; t++) {
                    const float2 v = __ldg(x + t);
                    a.x = fmaf(v.x, h[t], a.x);
                    a.y = fmaf(v.y, h[t], a.y);
                }
            }
            const float2 r = srot[s];
            acc.x = fmaf(a.x, r.x, fmaf(-a.y, r.y, acc.x));
            acc.y = fmaf(a.x, r.y, fmaf(a.y, r.x, acc.y));
        }
        out[i] = acc;
    }
}

// Two outputs per thread (i = 2p, 2p+1) for even T, 16-byte aligned streams: the pair's T+1-item window comes in as T/2
// aligned 16-byte loads + one 8-byte load, 2.1x fewer load instructions and 44 % less L1 traffic per output than the
// kernel above -- which is bound by exactly that (T loads of 8 bytes per output and stream).
template <int TT>
__global__ void __launch_bounds__(256) pfbdec_pair_kernel(const float2* __restrict__ X, long long stride,
                                                          float2* __restrict__ out, long long nout, int M,
                                                          const float* __restrict__ crev, const float2* __restrict__ rot)
{
    static_assert(TT % 2 == 0, "even taps per branch");
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float2* srot = reinterpret_cast<float2*>(smem_raw);
    float* srev = reinterpret_cast<float*>(smem_raw + (size_t)M * 8);
    for (int k = threadIdx.x; k < M; k += blockDim.x)
        srot[k] = rot[k];
    for (int k = threadIdx.x; k < M * TT; k += blockDim.x)
        srev[k] = crev[k];
    __syncthreads();
    const long long npair = (nout + 1) / 2;
    const long long step = (long long)gridDim.x * blockDim.x;
    for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < npair; p += step) {
        const long long i = 2 * p;
        const bool two = i + 1 < nout;
        float2 acc0 = make_float2(0.f, 0.f), acc1 = make_float2(0.f, 0.f);
        const float2* x = X + i;
        for (int s = 0; s < M; s++, x += stride) {
            const float* h = srev + s * TT;
            const float4* x4 = reinterpret_cast<const float4*>(x);
            float2 a0 = make_float2(0.f, 0.f), a1 = make_float2(0.f, 0.f);
            float4 w = __ldg(x4);
#pragma unroll
            for (int t = 0; t < TT; t += 2) {
                // items i+t = (w.x, w.y), i+t+1 = (w.z, w.w), i+t+2 = first item of the next load
                float4 nx;
                if (t + 2 < TT) {
                    nx = __ldg(x4 + t / 2 + 1);
                } else {
                    const float2 last = two ? __ldg(x + TT) : make_float2(0.f, 0.f);
                    nx = make_float4(last.x, last.y, 0.f, 0.f);
                }
                const float h0 = h[t], h1 = h[t + 1];
                a0.x = fmaf(w.x, h0, a0.x);
                a0.y = fmaf(w.y, h0, a0.y);
                a0.x = fmaf(w.z, h1, a0.x);
                a0.y = fmaf(w.w, h1, a0.y);
                a1.x = fmaf(w.z, h0, a1.x);
                a1.y = fmaf(w.w, h0, a1.y);
                a1.x = fmaf(nx.x, h1, a1.x);
                a1.y = fmaf(nx.y, h1, a1.y);
                w = nx;
            }
            const float2 r = srot[s];
            acc0.x = fmaf(a0.x, r.x, fmaf(-a0.y, r.y, acc0.x));
            acc0.y = fmaf(a0.x, r.y, fmaf(a0.y, r.x, acc0.y));
            acc1.x = fmaf(a1.x, r.x, fmaf(-a1.y, r.y, acc1.x));
            acc1.y = fmaf(a1.x, r.y, fmaf(a1.y, r.x, acc1.y));
        }
        out[i] = acc0;
        if (two)
            out[i + 1] = acc1;
    }
}

// Tiled kernel (T = 8, 16, 32, 64; aligned rows): a CTA walks the M streams of one tile of 256*R outputs.  Each stream's
// NT+T-1 item row segment arrives by ONE TMA bulk copy into a two-deep shared-memory ring (the copy of stream s+2 is in
// flight while s is multiplied); a thread owns R = 9 CONSECUTIVE outputs, so every item it loads (one 8-byte LDS) feeds
// up to min(R, T) packed FMAs (fma.rn.f32x2: re and im together) -- (T+R-1)/R = 2.7 shared loads per output and stream
// instead of T.  R is odd on purpose: lanes sit 2R = 18 words apart, which spreads a half-warp's 8-byte loads over all
// 32 banks without padding, so the linear layout a bulk copy produces is already conflict-free.
__device__ __forceinline__ void ffma2(float2& d, float2 a, float h)
{
    unsigned long long dd = *reinterpret_cast<unsigned long long*>(&d);
    const unsigned long long aa = *reinterpret_cast<const unsigned long long*>(&a);
    const float2 hh2 = make_float2(h, h);
    const unsigned long long hh = *reinterpret_cast<const unsigned long long*>(&hh2);
    asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(dd) : "l"(aa), "l"(hh));
    d = *reinterpret_cast<float2*>(&dd);
}

template <int TT>
struct PfbDecTile {
    static constexpr int R = 9, NTHR = 256, NT = R * NTHR;
    static constexpr int SEG = (NT + TT + 1) & ~1; // items per ring slot (even: slots stay 16-byte aligned)
    static constexpr int RING_BYTES = 2 * SEG * 8;
};

template <int TT>
__global__ void __launch_bounds__(256) pfbdec_tile_kernel(const float2* __restrict__ X, long long stride,
                                                          float2* __restrict__ out, long long nout, int M,
                                                          const float* __restrict__ crev, const float2* __restrict__ rot)
{
    using G = PfbDecTile<TT>;
    constexpr int R = G::R, NT = G::NT, SEG = G::SEG;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float2* ring = reinterpret_cast<float2*>(smem_raw);
    unsigned long long* bar = reinterpret_cast<unsigned long long*>(smem_raw + G::RING_BYTES);
    float2* srot = reinterpret_cast<float2*>(smem_raw + G::RING_BYTES + 16);
    float* srev = reinterpret_cast<float*>(smem_raw + G::RING_BYTES + 16 + (size_t)M * 8);
    const int t = threadIdx.x;
    for (int k = t; k < M; k += blockDim.x)
        srot[k] = rot[k];
    for (int k = t; k < M * TT; k += blockDim.x)
        srev[k] = crev[k];
    if (t == 0) {
        mbar_init(&bar[0], 1);
        mbar_init(&bar[1], 1);
    }
    __syncthreads();
    const long long ntiles = (nout + NT - 1) / NT;
    unsigned ph = 0; // bit b = parity the next wait on bar[b] expects
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const long long i0 = tile * NT;
        long long nit = stride - i0; // items of the row segment the copy may touch
        nit = nit > SEG ? SEG : nit;
        const unsigned bytes = (unsigned)(nit * 8);
        auto issue = [&](int s) {
            mbar_expect_tx(&bar[s & 1], bytes);
            bulk_g2s(ring + (s & 1) * SEG, X + (long long)s * stride + i0, bytes, &bar[s & 1]);
        };
        if (t == 0) {
            issue(0);
            if (M > 1)
                issue(1);
        }
        float2 acc[R];
#pragma unroll
        for (int r = 0; r < R; r++)
            acc[r] = make_float2(0.f, 0.f);
        for (int s = 0; s < M; s++) {
            const int b = s & 1;
            float h[TT];
#pragma unroll
            for (int q = 0; q < TT; q++)
                h[q] = srev[s * TT + q];
            mbar_wait(&bar[b], (ph >> b) & 1);
            ph ^= 1u << b;
            const float2* w = ring + b * SEG + t * R;
            float2 a[R];
#pragma unroll
            for (int r = 0; r < R; r++)
                a[r] = make_float2(0.f, 0.f);
#pragma unroll
            for (int j = 0; j < TT + R - 1; j++) {
                const float2 v = w[j]; // item i0 + t*R + j: output r = j - q takes it with tap q
#pragma unroll
                for (int r = 0; r < R; r++) {
                    const int q = j - r;
                    if (q >= 0 && q < TT)
                        ffma2(a[r], v, h[q]);
                }
            }
            const float2 rr = srot[s];
#pragma unroll
            for (int r = 0; r < R; r++) {
                acc[r].x = fmaf(a[r].x, rr.x, fmaf(-a[r].y, rr.y, acc[r].x));
                acc[r].y = fmaf(a[r].x, rr.y, fmaf(a[r].y, rr.x, acc[r].y));
            }
            __syncthreads(); // every thread has read slot b
            if (t == 0 && s + 2 < M)
                issue(s + 2);
        }
#pragma unroll
        for (int r = 0; r < R; r++) {
            const long long i = i0 + (long long)t * R + r;
            if (i < nout)
                out[i] = acc[r];
        }
    }
}

static int pfbdec_launch(const void* X, long long stride, void* out, long long nout, int M, int T, const float* crev,
                         const float2* rot, int nsm, cudaStream_t st)
{
    if (nout <= 0)
        return 0;
    const size_t smem = (size_t)M * 8 + (size_t)M * T * 4;
    if (smem > 200 * 1024)
        return fail(GRB200_ERANGE, "pfb_decimator_ccf: %d x %d taps do not fit in shared memory", M, T);
    const long long cap = (long long)nsm * 8;
    auto launch = [&](auto kern, long long nthreads, bool takes_T) -> int {
        long long nblk = (nthreads + 255) / 256;
        if (nblk > cap)
            nblk = cap;
        if (smem > 48 * 1024)
            GRB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        (void)takes_T;
        kern<<<(unsigned)nblk, 256, smem, st>>>((const float2*)X, stride, (float2*)out, nout, M, T, crev, rot);
        GRB_LAUNCH_CHECK();
        return 0;
    };
    auto launch_pair = [&](auto kern) -> int {
        long long nblk = ((nout + 1) / 2 + 255) / 256;
        if (nblk > cap)
            nblk = cap;
        if (smem > 48 * 1024)
            GRB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        kern<<<(unsigned)nblk, 256, smem, st>>>((const float2*)X, stride, (float2*)out, nout, M, crev, rot);
        GRB_LAUNCH_CHECK();
        return 0;
    };
    const bool aligned = ((((uintptr_t)X) & 15) == 0) && (stride % 2 == 0);
    static const bool use_tile = [] {
        const char* e = getenv("GRB200_PFBDEC_TILE");
        return !(e && e[0] == '0');
    }();
    auto launch_tile = [&](auto kern, int ring_bytes, int NT) -> int {
        const size_t sm = (size_t)ring_bytes + 16 + smem;
        GRB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
        int occ = 1;
        GRB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, 256, sm));
        long long nblk = (nout + NT - 1) / NT;
        const long long fit = (long long)nsm * (occ < 1 ? 1 : occ);
        if (nblk > fit)
            nblk = fit;
        kern<<<(unsigned)nblk, 256, sm, st>>>((const float2*)X, stride, (float2*)out, nout, M, crev, rot);
        GRB_LAUNCH_CHECK();
        return 0;
    };
    if (aligned && use_tile && nout >= 4 * PfbDecTile<16>::NT && smem <= 64 * 1024) {
        switch (T) {
        case 8:
            return launch_tile(pfbdec_tile_kernel<8>, PfbDecTile<8>::RING_BYTES, PfbDecTile<8>::NT);
        case 16:
            return launch_tile(pfbdec_tile_kernel<16>, PfbDecTile<16>::RING_BYTES, PfbDecTile<16>::NT);
        case 32:
            return launch_tile(pfbdec_tile_kernel<32>, PfbDecTile<32>::RING_BYTES, PfbDecTile<32>::NT);
        case 64:
            return launch_tile(pfbdec_tile_kernel<64>, PfbDecTile<64>::RING_BYTES, PfbDecTile<64>::NT);
        }
    }
    if (aligned) {
        switch (T) {
        case 4:
            return launch_pair(pfbdec_pair_kernel<4>);
        case 8:
            return launch_pair(pfbdec_pair_kernel<8>);
        case 16:
            return launch_pair(pfbdec_pair_kernel<16>);
        case 32:
            return launch_pair(pfbdec_pair_kernel<32>);
        case 64:
            return launch_pair(pfbdec_pair_kernel<64>);
        }
    }
    switch (T) {
    case 4:
        return launch(pfbdec_kernel<4>, nout, true);
    case 8:
        return launch(pfbdec_kernel<8>, nout, true);
    case 16:
        return launch(pfbdec_kernel<16>, nout, true);
    case 32:
        return launch(pfbdec_kernel<32>, nout, true);
    }
    return launch(pfbdec_kernel<0>, nout, true);
}

} // namespace grb

struct grb200_pfbdec {
    int device = 0;
    int M = 0, T = 0, chan = 0;
    std::vector<float> proto;      // prototype taps as given
    std::vector<float2> tail;      // carry mode: last T-1 items of every input stream, stream-major
    DevBuf crev, rot;
    Stager sg;
    std::mutex mu;
};

static int pfbdec_upload(grb200_pfbdec* p)
{
    // polyphase_filterbank::set_taps (gr-filter/lib/polyphase_filterbank.cc:53-81): taps_j[t] = padded[j + t*M];
    // the branch FIR keeps them reversed (fir_filter.cc:62-64).  Indexed here by INPUT stream s = M-1-j.
    const int M = p->M, ntaps = (int)p->proto.size();
    const int T = (int)ceil((double)ntaps / (double)M);
    std::vector<float> padded((size_t)T * M, 0.f);
    for (int i = 0; i < ntaps; i++)
        padded[i] = p->proto[i];
    std::vector<float> crev((size_t)M * T);
    std::vector<float2> rot((size_t)M);
    for (int s = 0; s < M; s++) {
        const int j = M - 1 - s;
        for (int t = 0; t < T; t++)
            crev[(size_t)s * T + t] = padded[j + (size_t)(T - 1 - t) * M];
        // d_rotator[i] = gr_expj(i * d_chan * 2 * GR_M_PI / d_rate), pfb_decimator_ccf_impl.cc:49-51 (float phase)
        const float ph = (float)((unsigned)j * (unsigned)p->chan * 2 * M_PI / (unsigned)M);
        rot[s] = make_float2(cosf(ph), sinf(ph));
    }
    int rc = p->crev.upload(crev.data(), crev.size() * sizeof(float));
    if (!rc)
        rc = p->rot.upload(rot.data(), rot.size() * sizeof(float2));
    if (rc)
        return rc;
    if (T != p->T || p->tail.empty())
        p->tail.assign((size_t)M * (T > 1 ? T - 1 : 0), make_float2(0.f, 0.f));
    p->T = T;
    return 0;
}

extern "C" {

int grb200_pfb_dec_ccf_create(unsigned int decim, const float* taps, int ntaps, unsigned int channel, int device,
                              grb200_pfbdec** out)
{
    if (!out)
        return fail(GRB200_EINVAL, "pfb_decimator_ccf: null output handle");
    *out = nullptr;
    if (decim < 1)
        return fail(GRB200_EINVAL, "pfb_decimator_ccf: decimation must be >= 1");
    if (ntaps < 1 || !taps)
        return fail(GRB200_EINVAL, "pfb_decimator_ccf: at least one tap is required");
    if (channel >= decim)
        return fail(GRB200_EINVAL, "pfb_decimator_ccf: channel %u out of range (decimation %u)", channel, decim);
    std::unique_ptr<grb200_pfbdec> p(new grb200_pfbdec);
    p->device = device;
    p->M = (int)decim;
    p->chan = (int)channel;
    p->proto.assign(taps, taps + ntaps);
    grb::DeviceScope _ds_create(device); // the caller keeps its own current device
    int rc = p->sg.init(device);
    if (rc)
        return rc;
    if ((rc = pfbdec_upload(p.get())))
        return rc;
    *out = p.release();
    return 0;
}

int grb200_pfb_dec_set_taps(grb200_pfbdec* p, const float* taps, int ntaps)
{
    if (!p || !taps || ntaps < 1)
        return fail(GRB200_EINVAL, "pfb_decimator_ccf::set_taps: bad arguments");
    std::lock_guard<std::mutex> lk(p->mu);
    grb::DeviceScope _ds(p->device);
    int rc = _ds.rc;
    if (rc)
        return rc;
    if ((rc = p->sg.quiesce())) // launches through this handle may still read the old state
        return rc;
    p->proto.assign(taps, taps + ntaps);
    return pfbdec_upload(p);
}

int grb200_pfb_dec_set_channel(grb200_pfbdec* p, unsigned int channel)
{
    if (!p)
        return fail(GRB200_EINVAL, "pfb_decimator_ccf::set_channel: null handle");
    if ((int)channel >= p->M)
        return fail(GRB200_EINVAL, "pfb_decimator_ccf: channel %u out of range (decimation %d)", channel, p->M);
    std::lock_guard<std::mutex> lk(p->mu);
    grb::DeviceScope _ds(p->device);
    int rc = _ds.rc;
    if (rc)
        return rc;
    if ((rc = p->sg.quiesce())) // launches through this handle may still read the old state
        return rc;
    p->chan = (int)channel;
    return pfbdec_upload(p);
}

int grb200_pfb_dec_taps_per_filter(const grb200_pfbdec* p) { return p ? p->T : 0; }
int grb200_pfb_dec_decimation(const grb200_pfbdec* p) { return p ? p->M : 0; }

int grb200_pfb_dec_work_dev(grb200_pfbdec* p, const void* x_dev, long long stride, void* out_dev, long long noutput,
                            void* stream)
{
    if (!p || noutput < 0 || (noutput > 0 && (!x_dev || !out_dev)) || stride < noutput + p->T - 1)
        return fail(GRB200_EINVAL, "pfb_decimator_ccf::work_dev: bad arguments");
    std::lock_guard<std::mutex> lk(p->mu);
    grb::DeviceScope _ds(p->device);
    int rc = _ds.rc;
    if (rc)
        return rc;
    cudaStream_t st_ = pick_stream(stream, p->sg);
    return dev_done(pfbdec_launch(x_dev, stride, out_dev, noutput, p->M, p->T, (const float*)p->crev.p, (const float2*)p->rot.p,
                         sm_count(p->device), st_), p->sg, st_);
}

// carry_state = 0: in_streams[s] holds noutput + T-1 items, history first (the block's set_history(T) mode, :61).
// carry_state = 1: in_streams[s] holds noutput items; the T-1 items before them are the ones the previous call ended
//                  with (zeros at first) -- the set_history(1) mode of the FFT-filter variants (:57-59), whose
//                  fft_filter kernels carry that overlap themselves.
int grb200_pfb_dec_work(grb200_pfbdec* p, const void* const* in_streams, void* out_host, long long noutput,
                        int carry_state)
{
    if (!p || noutput < 0 || !in_streams || (noutput > 0 && !out_host))
        return fail(GRB200_EINVAL, "pfb_decimator_ccf::work: bad arguments");
    if (noutput == 0)
        return 0;
    std::lock_guard<std::mutex> lk(p->mu);
    grb::DeviceScope _ds(p->device);
    int rc = _ds.rc;
    if (rc)
        return rc;
    const int M = p->M, T = p->T, Hn = T - 1;
    for (int s = 0; s < M; s++)
        if (!in_streams[s])
            return fail(GRB200_EINVAL, "pfb_decimator_ccf::work: input stream %d is null", s);
    long long chunk = std::max<long long>(1, (long long)(chunk_bytes_for((size_t)noutput * M * 8) / 8) / M);
    if (chunk > noutput)
        chunk = noutput;
    const long long nchunks = (noutput + chunk - 1) / chunk;
    const long long stride = (chunk + Hn + 1) & ~1LL; // even: every stream row starts 16-byte aligned (pair kernel)
    Stager& sg = p->sg;
    if ((rc = sg.reserve_dev(0, (size_t)stride * M * 8)) || (rc = sg.reserve_dev(1, (size_t)chunk * 8)))
        return rc;
    // In carry mode stream item k of this call sits at host index k; history item -h (h = 1..Hn) comes from p->tail.
    const long long hoff = carry_state ? 0 : Hn; // host index of the call's first NEW item
    rc = sg.run(
        nchunks,
        [&](long long c, int slot, Stager::Chunk& ch) {
            const long long o0 = c * chunk, n = std::min(chunk, noutput - o0);
            for (int s = 0; s < M; s++) {
                const float2* src = (const float2*)in_streams[s];
                float2* dst = (float2*)sg.dev[slot][0] + (size_t)s * stride;
                // the chunk needs stream items o0-Hn .. o0+n-1
                long long first = o0 - Hn;
                if (carry_state && first < 0) { // the leading -first items come from the saved tail
                    const long long nt = std::min<long long>(-first, n + Hn);
                    ch.h2d.push_back({ p->tail.data() + (size_t)s * Hn + (Hn + first), dst, (size_t)nt * 8 });
                    if (n + Hn > nt)
                        ch.h2d.push_back({ src, dst + nt, (size_t)(n + Hn - nt) * 8 });
                } else {
                    ch.h2d.push_back({ src + hoff + first, dst, (size_t)(n + Hn) * 8 });
                }
            }
            ch.d2h.push_back({ sg.dev[slot][1], (float2*)out_host + o0, (size_t)n * 8 });
            return 0;
        },
        [&](long long c, int slot, cudaStream_t st) {
            const long long o0 = c * chunk, n = std::min(chunk, noutput - o0);
            return pfbdec_launch(sg.dev[slot][0], stride, sg.dev[slot][1], n, M, T, (const float*)p->crev.p,
                                 (const float2*)p->rot.p, sm_count(p->device), st);
        });
    if (!rc && carry_state && Hn > 0) { // new tail = last Hn items of (old tail ++ this call's items), per stream
        std::vector<float2> nt((size_t)M * Hn);
        for (int s = 0; s < M; s++) {
            const float2* src = (const float2*)in_streams[s];
            for (int h = 0; h < Hn; h++) {
                const long long k = noutput - Hn + h; // stream index within this call
                nt[(size_t)s * Hn + h] = k >= 0 ? src[k] : p->tail[(size_t)s * Hn + (Hn + k)];
            }
        }
        p->tail.swap(nt);
    }
    return rc;
}

void grb200_pfb_dec_destroy(grb200_pfbdec* p)
{
    if (!p)
        return;
    grb::DeviceScope _ds(p->device);
    p->sg.destroy();
    p->crev.release();
    p->rot.release();
    delete p;
}

} // extern "C"
