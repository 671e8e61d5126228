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
#include "tma.cuh"

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

// ---- streaming form (round 2): interpolation <= 32, decimation <= 4 (and <= taps per phase), <= 32 taps per phase ----------
// The kernels above read a T-item window per output through L1: load/store-unit bound (0.34 of HBM at 3/2 x 16).  Outputs of
// the SAME phase are L apart and their windows D items apart, so for small D a thread that keeps one phase can keep its
// window in a register ring and fetch D new items per output -- interp_stream_kernel (capi_interp.cu) with a step of D:
// lanes = (output offset l < L, run g); lane l owns outputs n L + l of its run, phase (ctr0 + l D) mod L, window start
// n D + (ctr0 + l D) / L; a producer warp streams each run's new items (one bulk copy of KT D + D + 2 items per run and
// iteration).  Summation order = the kernels above (one accumulator, oldest tap first): a stream cut into calls of any
// length gives the same bits whichever kernel a call takes.
namespace rstream {
constexpr int NCW = 8, MAXNS = 4; // consumer warps, stages <=  (KT = groups of L outputs per run and iteration: 32, or 16 when
                                  // the stages of a long row -- small interpolation, decimation 3 or 4 -- would not fit)
// row stride: even (16-byte copies) with an odd half -- 4 x odd words: the runs of a warp start in different banks
__host__ __device__ constexpr int row_items(int D, int KT) { return (((KT * D + D + 3) / 2) | 1) * 2; }
}

template <int TP, int D, bool CT, int KT>
__global__ void __launch_bounds__(288)
resamp_stream_kernel(const float2* __restrict__ in /* 16-byte aligned; the stream starts at in[gs] */, float2* __restrict__ out,
                     long long nout, unsigned ctr0, int L, int T, const float* __restrict__ rev, long long nitems, int iters,
                     int gs, int NS)
{
    using namespace rstream;
    constexpr int ROW = row_items(D, KT);
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int RW = 32 / L, NG = NCW * RW;
    float2* xs = reinterpret_cast<float2*>(smem_raw); // [NS][NG][ROW]
    unsigned long long* full = reinterpret_cast<unsigned long long*>(smem_raw + (size_t)NS * NG * ROW * 8);
    unsigned long long* empty = full + MAXNS;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const long long region = (long long)iters * KT; // groups per run
    const long long cta0 = (long long)blockIdx.x * NG * region;
    const long long nitems_even = nitems & ~1LL; // (nitems counts from the aligned base: gs included)
    if (tid == 0)
        for (int s = 0; s < NS; s++) {
            mbar_init(&full[s], 32);
            mbar_init(&empty[s], NCW);
        }
    __syncthreads();
    // the first new item of a run's group m (for a lane at offset 0) is item (n0 + m) D + T - D + gs of the aligned buffer
    if (warp == NCW) {
        for (int it = 0; it < iters; it++) {
            const int s = it % NS, use = it / NS;
            if (use > 0)
                mbar_wait(&empty[s], (unsigned)((use - 1) & 1));
            unsigned bytes = 0;
            for (int g = lane; g < NG; g += 32) {
                const long long a0 = ((cta0 + (long long)g * region + (long long)it * KT) * D + T - D + gs) & ~1LL;
                long long n = nitems_even - a0;
                n = n < 0 ? 0 : (n > ROW ? ROW : n);
                bytes += (unsigned)(n * 8);
            }
            mbar_expect_tx(&full[s], bytes);
            for (int g = lane; g < NG; g += 32) {
                const long long a0 = ((cta0 + (long long)g * region + (long long)it * KT) * D + T - D + gs) & ~1LL;
                long long n = nitems_even - a0;
                n = n < 0 ? 0 : (n > ROW ? ROW : n);
                if (n)
                    bulk_g2s(xs + ((size_t)s * NG + g) * ROW, in + a0, (unsigned)(n * 8), &full[s]);
            }
        }
        return;
    }
    const int l = lane % L, gw = lane / L;
    const bool idle = gw >= RW;
    const int g = warp * RW + (idle ? 0 : gw);
    const long long n0 = cta0 + (long long)g * region; // first group of this run
    const unsigned sl = ctr0 + (unsigned)l * (unsigned)D;
    const int ql = (int)(sl / (unsigned)L), ph = (int)(sl - (unsigned)ql * (unsigned)L); // window offset and phase of this lane
    const long long b0 = n0 * D + T - D + gs; // item of (group 0, new item 0) for a lane at offset 0
    const int off = (int)(b0 & 1) + ql;       // (KT D is even: the parity holds for every iteration)
    // tap u multiplies the item u behind the newest one of a window: rev index T - 1 - u (tap-major table)
    float hr[TP], hi[CT ? TP : 1];
#pragma unroll
    for (int u = 0; u < TP; u++) {
        const bool ok = u < T && !idle;
        if (CT) {
            const float2 c = ok ? __ldg(reinterpret_cast<const float2*>(rev) + (size_t)(T - 1 - u) * L + ph) : make_float2(0.f, 0.f);
            hr[u] = c.x;
            hi[u] = c.y;
        } else {
            hr[u] = ok ? __ldg(rev + (size_t)(T - 1 - u) * L + ph) : 0.f;
        }
    }
    // ring slot of the item r = m D + d (group m of the run, new item d): r mod TP; the T - D items in front of group 0 warm it up
    float2 ring[TP];
#pragma unroll
    for (int u = 0; u < TP; u++)
        ring[u] = make_float2(0.f, 0.f);
#pragma unroll
    for (int r = 1; r < TP; r++) { // item b0 + ql - r
        const long long x = b0 + ql - r;
        if (r <= T - D && x >= gs && x < nitems)
            ring[(TP - r) % TP] = in[x];
    }
    const long long last_odd = (nitems & 1) ? nitems - 1 : -1; // the one item the 16-byte copies cannot bring
    static_assert(KT % TP == 0, "the ring position returns to 0 every iteration");
    for (int it = 0; it < iters; it++) {
        const int s = it % NS;
        mbar_wait(&full[s], (unsigned)((it / NS) & 1));
        const float2* src = xs + ((size_t)s * NG + g) * ROW + off;
        const long long nb = n0 + (long long)it * KT; // first group of this block
        const long long ib = b0 + ql + (long long)it * KT * D; // item index of src[0]
        float2* o = out + nb * L + l;
        const long long left = nout - (nb * L + l); // this lane's outputs of the block that exist: k L < left
        auto block = [&](auto patch) {
            constexpr bool PATCH = decltype(patch)::value; // this block holds the odd last item: it is read where it lies
#pragma unroll
            for (int k = 0; k < KT; k++) {
#pragma unroll
                for (int d = 0; d < D; d++) {
                    float2 x = src[k * D + d];
                    if (PATCH && ib + k * D + d == last_odd)
                        x = in[last_odd];
                    ring[(k * D + d) % TP] = x;
                }
                float2 a = make_float2(0.f, 0.f);
#pragma unroll
                for (int u = TP - 1; u >= 0; u--) { // oldest tap first, one accumulator: the order of resamp_kernel (two
                                                    // chains by tap parity were measured: no faster here, slower above)
                    const float2 w = ring[(k * D + D - 1 - u + 4 * TP) % TP];
                    if (CT) {
                        a = __ffma2_rn(make_float2(-hi[u], hr[u]), make_float2(w.y, w.y), a);
                        a = __ffma2_rn(make_float2(hr[u], hi[u]), make_float2(w.x, w.x), a);
                    } else {
                        a = __ffma2_rn(make_float2(hr[u], hr[u]), w, a);
                    }
                }
                if ((long long)k * L < left && !idle)
                    o[(long long)k * L] = a;
            }
        };
        if (last_odd >= ib && last_odd < ib + KT * D)
            block(std::true_type{});
        else
            block(std::false_type{});
        __syncwarp();
        if (lane == 0)
            mbar_arrive_one(&empty[s]);
    }
}

template <int TP, int D, bool CT, int KT>
static int resamp_stream_launch_k(const void* in, void* out, long long nout, unsigned ctr0, int L, int T, const float* rev,
                                  int nsm, cudaStream_t st, bool* done)
{
    using namespace rstream;
    const int NG = NCW * (32 / L);
    const size_t stage = (size_t)NG * row_items(D, KT) * 8;
    const int NS = 3 * stage <= 80 * 1024 ? 3 : 2;
    const size_t smem = (size_t)NS * stage + 2 * MAXNS * 8 + 16;
    if (smem > 160 * 1024)
        return 0; // not done: the caller falls back
    auto kern = resamp_stream_kernel<TP, D, CT, KT>;
    GRB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int occ = 1;
    GRB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, 288, smem));
    if (occ < 1)
        occ = 1;
    const long long ngroups = (nout + L - 1) / L;
    const long long per_iter = (long long)NG * KT;
    long long grid = (long long)nsm * occ;
    const long long need = (ngroups + per_iter - 1) / per_iter;
    if (grid > need)
        grid = need;
    long long iters = (ngroups + grid * per_iter - 1) / (grid * per_iter);
    grid = (ngroups + iters * per_iter - 1) / (iters * per_iter);
    if (iters > 0x7fffffffLL / (KT * D))
        return 0;
    const int gs = (int)((((uintptr_t)in) >> 3) & 1);
    // items the call may read: the last output's window ends at (ctr0 + (nout-1) D) / L + T
    const long long nitems = (long long)(((unsigned long long)ctr0 + (unsigned long long)(nout - 1) * D) / (unsigned)L) + T + gs;
    kern<<<(unsigned)grid, 288, smem, st>>>((const float2*)in - gs, (float2*)out, nout, ctr0, L, T, rev, nitems, (int)iters, gs,
                                            NS);
    GRB_LAUNCH_CHECK();
    *done = true;
    return 0;
}

template <int TP, int D, bool CT>
static int resamp_stream_launch_t(const void* in, void* out, long long nout, unsigned ctr0, int L, int T, const float* rev,
                                  int nsm, cudaStream_t st, bool* done)
{
    using namespace rstream;
    const size_t stage32 = (size_t)NCW * (32 / L) * row_items(D, 32) * 8;
    if (2 * stage32 + 128 <= 160 * 1024 || TP > 16)
        return resamp_stream_launch_k<TP, D, CT, 32>(in, out, nout, ctr0, L, T, rev, nsm, st, done);
    if constexpr (TP <= 16)
        return resamp_stream_launch_k<TP, D, CT, 16>(in, out, nout, ctr0, L, T, rev, nsm, st, done);
    return 0;
}

template <bool CT>
static int resamp_stream_launch(const void* in, void* out, long long nout, unsigned ctr0, unsigned L, unsigned D, int T,
                                const float* rev, int nsm, cudaStream_t st, bool* done)
{
    *done = false;
    static const bool use_stream = [] {
        const char* e = GRB_GETENV_ONCE("GRB200_RESAMP_STREAM"); // 0 = the window-through-L1 kernels only
        return !(e && e[0] == '0');
    }();
    if (!use_stream || L < 2 || L > 32 || D > 4 || (int)D > T || T > (CT ? 16 : 32) || nout < 32768)
        return 0;
    auto byd = [&](auto tp) -> int {
        constexpr int TPv = decltype(tp)::value;
        switch (D) {
        case 1:
            return resamp_stream_launch_t<TPv, 1, CT>(in, out, nout, ctr0, (int)L, T, rev, nsm, st, done);
        case 2:
            return resamp_stream_launch_t<TPv, 2, CT>(in, out, nout, ctr0, (int)L, T, rev, nsm, st, done);
        case 3:
            return resamp_stream_launch_t<TPv, 3, CT>(in, out, nout, ctr0, (int)L, T, rev, nsm, st, done);
        default:
            return resamp_stream_launch_t<TPv, 4, CT>(in, out, nout, ctr0, (int)L, T, rev, nsm, st, done);
        }
    };
    if (T <= 8)
        return byd(std::integral_constant<int, 8>{});
    if (T <= 16)
        return byd(std::integral_constant<int, 16>{});
    if constexpr (!CT)
        return byd(std::integral_constant<int, 32>{});
    return 0;
}

// (also the streaming kernel of interp_fir_filter: an interpolator is the resampler with decimation 1 and counter 0)
int resamp_stream_try(bool ct, const void* in, void* out, long long nout, unsigned ctr0, unsigned L, unsigned D, int T,
                      const float* rev_tap_major, int nsm, cudaStream_t st, bool* done)
{
    return ct ? resamp_stream_launch<true>(in, out, nout, ctr0, L, D, T, rev_tap_major, nsm, st, done)
              : resamp_stream_launch<false>(in, out, nout, ctr0, L, D, T, rev_tap_major, nsm, st, done);
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
    int rc = ct ? resamp_stream_launch<true>(in, out, nout, ctr0, L, D, T, rev, nsm, st, &done)
                : resamp_stream_launch<false>(in, out, nout, ctr0, L, D, T, rev, nsm, st, &done);
    if (rc || done)
        return rc;
    rc = ct ? resamp_phase_launch<true>(in, out, nout, ctr0, L, D, T, rev, nblk, st, &done)
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
    // interpolation 1: the block is a decimating FIR (one phase, the input advances D per output) -- its calls go to a FIR
    // handle, whose kernels (tensor cores up to 64 taps, overlap-save for long filters) are 2 - 8 x faster than one output per
    // thread.  Same window, same taps, same counts.
    grb200_fir* fir = nullptr;
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
    if (L == 1) {
        if (h->fir)
            return grb200_fir_set_taps(h->fir, taps, ntaps);
        return h->ct ? grb200_fir_ccc_create((int)h->D, taps, ntaps, h->device, &h->fir)
                     : grb200_fir_ccf_create((int)h->D, taps, ntaps, h->device, &h->fir);
    }
    return 0;
}

// one device-resident call: n outputs from the phase counter cc, `in` = the first window's first item
static int resamp_run(grb200_resamp* h, const void* in, void* out, long long n, unsigned cc, cudaStream_t st)
{
    if (h->fir && n > 0 && h->D <= 0x7fffffffu)
        return grb200_fir_work_dev(h->fir, in, out, n, (void*)st);
    return resamp_launch(h->ct, in, out, n, cc, h->L, h->D, h->T, (const float*)h->rev.p, sm_count(h->device), st);
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
    if ((rc = resamp_upload(h.get(), taps, ntaps))) {
        grb200_rational_resampler_destroy(h.release());
        return rc;
    }
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
    rc = dev_done(resamp_run(h, in_dev, out_dev, n, h->ctr, st_), h->sg, st_);
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
                return resamp_run(h, sg.dev[slot][0], sg.dev[slot][1], no, cc, st);
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
    if (h->fir)
        grb200_fir_destroy(h->fir);
    delete h;
}

} // extern "C"
