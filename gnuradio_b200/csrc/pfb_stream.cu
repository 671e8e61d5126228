// pfb_stream.cu -- pfb_channelizer_ccf streaming kernel (K4, second generation; oversample_rate 1, identity map).
//
// Same closed form as pfb.cu (reference gr-filter/lib/pfb_channelizer_ccf_impl.cc:128-193 over
// gr-filter/lib/polyphase_filterbank.cc:53-81):  v_i[o] = sum_u h[i + u*M] * x[(o + T - u)*M + (M-1-i)],
// out_c[o] = sum_i v_i[o] e^{+j 2 pi i c / M}.
//
// What changed against pfb_fast_kernel, driven by its ncu capture (profiles/r1_pfb_fast_kernel_v1_ncu.txt:
// 2x redundant L1 loads, 25 % occupancy, long-scoreboard + barrier stalls):
//  * persistent CTAs that own a CONTIGUOUS run of output times, split into NG group runs; thread (branch i,
//    group g) walks down its run with the last TP input samples of its branch in a register ring, so every
//    input sample is fetched exactly once and the FIR is TP packed FFMA2 per output;
//  * the KT x M input rows each group needs per iteration are one contiguous block of the interleaved
//    stream: a single TMA bulk copy (cp.async.bulk, mbarrier completion) per group, issued one iteration
//    ahead right after the FIR phase has drained the staging buffer, so HBM latency hides behind the FFT;
//  * the M-point inverse DFT runs in place in ONE shared buffer (pass 1 writes back to the slots it read),
//    lanes over time, so channel stores are contiguous 128-byte runs.
// HBM traffic stays the compulsory 8 B in + 8 B out per input sample (16 B/sample, the roofline figure).
#include <cstdlib>
#include <type_traits>
#include "fft.cuh"
#include "tma.cuh"

namespace grb {

template <int M, int NT>
struct PfbStreamGeom {
    // M = R1 * R2: pass 1 runs radix-R1 transforms over the branches j + R2 q, pass 2 radix-R2 ones over j (R2 = 1: none)
    static constexpr int R1 = M >= 128 ? 16 : (M >= 32 ? 8 : (M == 16 ? 4 : M));
    static constexpr int R2 = M / R1;
    static constexpr int NG = NT / M;                            // group runs per CTA
    static constexpr int KT = 16;                                // output times per group per iteration
    static constexpr int OT = NG * KT;                           // output times per CTA per iteration
    static constexpr int LD = M + 1;                             // padded row of the [time][branch] buffer
    static constexpr int NI1 = OT * R2 / NT, NI2 = OT * R1 / NT; // DFT items per thread in pass 1 / pass 2
    static constexpr int XS_BYTES = NG * KT * M * 8;
    static constexpr int VB_BYTES = OT * LD * 8;
    static constexpr int SMEM = XS_BYTES + VB_BYTES + M * 8 + 16 + M * 4;
    static_assert(NT % M == 0 && (OT * R2) % NT == 0 && (OT * R1) % NT == 0 && R1 * R2 == M && R2 <= 16, "geometry");
};

// OS = 2: oversample_rate 2 (R = M/2).  Output time t = 2 m + p uses the samples x[(m - u) M + (p + 1) R - 1 - i] and the
// rotation q = (i - p R) mod M (pfb.cu's closed form with (t + 2) R = (m + 1) M + p R): the odd times are the plain channelizer
// on x, the even ones the plain channelizer on x delayed by R samples, and the rotation by M/2 of the odd ones is a factor
// (-1)^c on channel c.  So half of a CTA's groups walk the stream and half walk it R samples earlier; lanes 0-15 / 16-31 of a
// warp's channel store hold the even / odd times of the same 32 consecutive output times.
// ACCUM: the results are ADDED to what `out` holds -- the second launch of a filter with 33 .. 64 taps per branch, whose older
// taps run as a second channelizer on the same stream (the DFT is linear: pfb_stream_split_launch).
template <int M, int TP, int NT, int OS, bool MAPPED, bool ACCUM = false>
__global__ void __launch_bounds__(NT)
pfb_stream_kernel(const float2* __restrict__ x, float2* __restrict__ out, long long out_stride, int nouts,
                  long long noutput, int T, long long nrows, const float* __restrict__ h,
                  const float2* __restrict__ wM, int iters, const float2* __restrict__ halo,
                  const int* __restrict__ inv /* output stream of DFT bin c (-1: none), or null = identity */)
{
    using Geo = PfbStreamGeom<M, NT>;
    constexpr int NG = Geo::NG, KT = Geo::KT, OT = Geo::OT, LD = Geo::LD, R1 = Geo::R1, R2 = Geo::R2;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float2* xs = reinterpret_cast<float2*>(smem_raw);                     // [NG][KT][M]
    float2* vbuf = reinterpret_cast<float2*>(smem_raw + Geo::XS_BYTES);   // [OT][LD]
    float2* wtab = vbuf + OT * LD;                                        // [M] e^{+j 2 pi k / M}
    unsigned long long* bar = reinterpret_cast<unsigned long long*>(wtab + M);
    int* sinv = reinterpret_cast<int*>(bar + 2); // [M] when a channel map is in force

    const int tid = threadIdx.x;
    const int bi = tid % M, bg = tid / M;
    const int col = M - 1 - bi;                       // input stream feeding branch bi
    const long long region = (long long)iters * KT;   // output times (OS = 2: pairs of them) per group run
    const long long cta0 = (long long)blockIdx.x * (NG / OS) * region;
    const long long gs = cta0 + (long long)(bg / OS) * region; // this thread's first output time (pair)
    static_assert(NG % OS == 0, "an oversampled CTA needs both parities");
    // OS = 2: the groups of even output times read the stream R = M/2 samples earlier; their last pair may lack its odd time
    const int delay = (OS == 2 && (bg % OS) == 0) ? M / 2 : 0;
    x -= delay;
    if (OS == 2)
        nrows = T + (noutput + 1 - (bg % OS)) / 2;

    for (int k = tid; k < M; k += NT)
        wtab[k] = wM[k];
    if (MAPPED)
        for (int k = tid; k < M; k += NT)
            sinv[k] = inv[k];
    if (tid == 0)
        mbar_init(bar, 1);

    float taps[TP];
#pragma unroll
    for (int u = 0; u < TP; u++)
        taps[u] = (u < T) ? __ldg(h + u * M + bi) : 0.f;

    // register ring: after the row of output o is in, ring[(n - u) mod TP] = row (o + T - u), n = o - gs
    float2 ring[TP];
#pragma unroll
    for (int u = 1; u < TP; u++) {
        const long long r = gs + T - u;
        // the T history rows of a time-sharded segment are the left neighbour's newest rows: with a halo source they are
        // read where they are -- the neighbour GPU's memory, plain loads over NVLink -- instead of being copied in first
        const float2* src = (halo != nullptr && r < T) ? halo : x;
        ring[(TP - u) % TP] = (r >= (delay ? 1 : 0) && r < nrows) ? src[r * M + col] : make_float2(0.f, 0.f);
    }
    ring[0] = make_float2(0.f, 0.f);
    __syncthreads(); // barrier initialised, twiddle table written

    // one elected thread streams the input: NG contiguous blocks of KT rows per iteration
    auto issue = [&](int it) {
        // (the issuing thread belongs to group 0: its x is the EVEN groups' delayed one -- the odd groups add the delay back --
        // and the row count is taken per parity here, not from its own nrows)
        auto rows_of = [&](int g, long long& r0) {
            r0 = cta0 + (long long)(g / OS) * region + (long long)it * KT + T;
            const long long nr = OS == 2 ? T + (noutput + 1 - (g % OS)) / 2 : nrows;
            const long long nv = nr - r0;
            return (unsigned)((nv < 0 ? 0 : (nv > KT ? KT : nv)) * M * 8);
        };
        unsigned total = 0;
        long long r0;
        constexpr int UNR = NG <= 16 ? NG : 1; // (M = 8 has 32 groups: rolled there, straight-line code elsewhere)
#pragma unroll UNR
        for (int g = 0; g < NG; g++)
            total += rows_of(g, r0);
        mbar_expect_tx(bar, total);
#pragma unroll UNR
        for (int g = 0; g < NG; g++) {
            const unsigned bytes = rows_of(g, r0);
            if (bytes)
                bulk_g2s(xs + g * KT * M, x + r0 * M + ((OS == 2 && (g % OS)) ? M / 2 : 0), bytes, bar);
        }
    };
    if (tid == 0)
        issue(0);

    auto iteration = [&](int it, auto phc) {
        constexpr int PH = decltype(phc)::value; // ring position of this iteration's first row
        mbar_wait(bar, it & 1);
        // ---------------- FIR: one new row per output, TP packed FMAs ----------------
#pragma unroll
        for (int k = 0; k < KT; k++) {
            ring[(PH + k) % TP] = xs[(bg * KT + k) * M + col];
            float2 a0 = make_float2(0.f, 0.f), a1 = make_float2(0.f, 0.f);
#pragma unroll
            for (int u = 0; u < TP; u += 2) {
                a0 = __ffma2_rn(make_float2(taps[u], taps[u]), ring[(PH + k - u + 2 * TP) % TP], a0);
                a1 = __ffma2_rn(make_float2(taps[u + 1], taps[u + 1]), ring[(PH + k - u - 1 + 2 * TP) % TP], a1);
            }
            vbuf[(bg * KT + k) * LD + bi] = __fadd2_rn(a0, a1);
        }
        __syncthreads(); // vbuf complete; xs drained
        if (tid == 0 && it + 1 < iters)
            issue(it + 1);
        // ---------------- inverse DFT pass 1 (in place): y[j][k1] -> slot j + R2*k1 ----------------
#pragma unroll
        for (int n = 0; n < Geo::NI1; n++) {
            const int idx = tid + n * NT;
            const int o = idx % OT, j = idx / OT;
            float2 v[R1];
#pragma unroll
            for (int q = 0; q < R1; q++)
                v[q] = vbuf[o * LD + j + q * R2];
            dft_r<R1, false>(v);
#pragma unroll
            for (int q = 0; q < R1; q++)
                vbuf[o * LD + j + q * R2] = v[q];
        }
        __syncthreads();
        // ---------------- pass 2: twiddle, DFT over j, channel-major stores ----------------
#pragma unroll
        for (int n = 0; n < Geo::NI2; n++) {
            const int idx = tid + n * NT;
            const int o = idx % OT, k1 = idx / OT;
            const int g = o / KT, k = o % KT;
            const long long t = OS * (cta0 + (long long)(g / OS) * region + (long long)it * KT + k) + (g % OS);
            float2 prev[ACCUM ? R2 : 1]; // what the first launch left: all R2 loads in flight behind the transform below
            if constexpr (ACCUM) {
#pragma unroll
                for (int k2 = 0; k2 < R2; k2++) {
                    const int c = k1 + R1 * k2;
                    prev[k2] = (t < noutput && c < nouts) ? out[(long long)c * out_stride + t] : make_float2(0.f, 0.f);
                }
            }
            float2 a[R2];
#pragma unroll
            for (int j = 0; j < R2; j++) {
                a[j] = vbuf[o * LD + k1 * R2 + j];
                if (j > 0)
                    a[j] = cmul(a[j], wtab[j * k1]);
            }
            dft_r<R2, false>(a);
            if (t < noutput) {
#pragma unroll
                for (int k2 = 0; k2 < R2; k2++) {
                    const int c = k1 + R1 * k2;
                    float2 y = a[k2];
                    if (OS == 2 && (g % OS) && (c & 1)) // the rotation by M/2 of the odd output times
                        y = make_float2(-y.x, -y.y);
                    const int n = MAPPED ? sinv[c] : c; // the channel map, inverted on the host: a scatter here
                    if (n >= 0 && n < nouts) {
                        if (ACCUM) // (identity map only)
                            y = make_float2(prev[k2].x + y.x, prev[k2].y + y.y);
                        out[(long long)n * out_stride + t] = y;
                    }
                }
            }
        }
        __syncthreads(); // pass-2 reads of vbuf are done before the next FIR phase overwrites it
    };

    if constexpr (TP <= KT) {
        for (int it = 0; it < iters; it++)
            iteration(it, std::integral_constant<int, 0>{});
    } else { // TP == 2*KT: the ring position alternates between 0 and KT
        for (int it = 0; it < iters; it += 2) {
            iteration(it, std::integral_constant<int, 0>{});
            if (it + 1 < iters)
                iteration(it + 1, std::integral_constant<int, KT>{});
        }
    }
}

// ---- host launcher ------------------------------------------------------------------------------------
template <int M, int TP, int NT, int OS = 1>
static int launch_stream(const void* x, void* out, long long out_stride, int nouts, long long noutput, int T,
                         long long nrows, const float* h, const float2* wM, int nsm, cudaStream_t st, const void* halo,
                         const int* inv, bool accum = false)
{
    using Geo = PfbStreamGeom<M, NT>;
    auto kern = inv != nullptr ? pfb_stream_kernel<M, TP, NT, OS, true> : pfb_stream_kernel<M, TP, NT, OS, false>;
    if constexpr (OS == 1) {
        if (accum) {
            if (inv != nullptr)
                return fail(GRB200_EINVAL, "pfb_channelizer: no accumulating kernel under a channel map");
            kern = pfb_stream_kernel<M, TP, NT, 1, false, true>;
        }
    }
    static thread_local bool configured[3][16] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    const int which = accum ? 2 : (inv != nullptr ? 1 : 0);
    if (!configured[which][dev & 15]) {
        GRB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Geo::SMEM));
        configured[which][dev & 15] = true;
    }
    int occ = 1;
    GRB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, NT, Geo::SMEM));
    if (occ < 1)
        occ = 1;
    long long grid = (long long)nsm * occ;
    const long long steps = (noutput + OS - 1) / OS; // output times (OS = 2: pairs) ...
    constexpr int OTS = Geo::OT / OS;                // ... and how many a CTA covers per iteration
    const long long need = (steps + OTS - 1) / OTS;
    if (grid > need)
        grid = need;
    long long iters = (steps + grid * OTS - 1) / (grid * OTS);
    if (TP > Geo::KT && (iters & 1))
        iters++;
    grid = (steps + iters * OTS - 1) / (iters * OTS); // drop CTAs the rounding left without work
    if (iters > 0x7fffffffLL / Geo::KT)
        return fail(GRB200_ERANGE, "pfb_channelizer: work() call too long for one launch");
    kern<<<(unsigned)grid, NT, Geo::SMEM, st>>>((const float2*)x, (float2*)out, out_stride, nouts, noutput, T, nrows,
                                                 h, wM, (int)iters, (const float2*)halo, inv);
    GRB_LAUNCH_CHECK();
    return 0;
}

template <int M>
static int launch_stream_os2(int T, const void* x, void* out, long long out_stride, int nouts, long long noutput,
                             long long nrows, const float* h, const float2* wM, int nsm, cudaStream_t st,
                             const int* inv)
{
    if (T <= 8)
        return launch_stream<M, 8, 256, 2>(x, out, out_stride, nouts, noutput, T, nrows, h, wM, nsm, st, nullptr, inv);
    if (T <= 16)
        return launch_stream<M, 16, 256, 2>(x, out, out_stride, nouts, noutput, T, nrows, h, wM, nsm, st, nullptr, inv);
    return launch_stream<M, 32, 256, 2>(x, out, out_stride, nouts, noutput, T, nrows, h, wM, nsm, st, nullptr, inv);
}

template <int M, int NT>
static int launch_stream_tp(int T, const void* x, void* out, long long out_stride, int nouts, long long noutput,
                            long long nrows, const float* h, const float2* wM, int nsm, cudaStream_t st, const void* halo,
                            const int* inv, bool accum = false)
{
    if (T <= 8)
        return launch_stream<M, 8, NT>(x, out, out_stride, nouts, noutput, T, nrows, h, wM, nsm, st, halo, inv, accum);
    if (T <= 16)
        return launch_stream<M, 16, NT>(x, out, out_stride, nouts, noutput, T, nrows, h, wM, nsm, st, halo, inv, accum);
    return launch_stream<M, 32, NT>(x, out, out_stride, nouts, noutput, T, nrows, h, wM, nsm, st, halo, inv, accum);
}

// oversample_rate 2 (R = M/2): 8 .. 128 channels (a CTA needs groups of both parities)
bool pfb_stream_os2_supported(int M, int T, int R, const void* x)
{
    return 2 * R == M && T >= 1 && T <= 32 && M >= 8 && M <= 128 && (M & (M - 1)) == 0 && ((uintptr_t)x & 15) == 0;
}

int pfb_stream_os2_launch(int M, int T, const void* x, void* out, long long out_stride, int nouts, long long noutput,
                          const float* h, const float2* wM, int nsm, cudaStream_t st, const int* inv)
{
    if (noutput <= 0)
        return 0;
    const long long nrows = T + (noutput + 1) / 2;
    switch (M) {
    case 8:
        return launch_stream_os2<8>(T, x, out, out_stride, nouts, noutput, nrows, h, wM, nsm, st, inv);
    case 16:
        return launch_stream_os2<16>(T, x, out, out_stride, nouts, noutput, nrows, h, wM, nsm, st, inv);
    case 32:
        return launch_stream_os2<32>(T, x, out, out_stride, nouts, noutput, nrows, h, wM, nsm, st, inv);
    case 64:
        return launch_stream_os2<64>(T, x, out, out_stride, nouts, noutput, nrows, h, wM, nsm, st, inv);
    }
    return launch_stream_os2<128>(T, x, out, out_stride, nouts, noutput, nrows, h, wM, nsm, st, inv);
}

bool pfb_stream_supported(int M, int T, int R, const void* x)
{
    return R == M && T >= 1 && T <= 32 && M >= 8 && M <= 256 && (M & (M - 1)) == 0 && ((uintptr_t)x & 15) == 0;
}

// x: interleaved input with T*M history items in front (16-byte aligned); identity channel map.
static int stream_launch_any(int M, int T, const void* x, void* out, long long out_stride, int nouts, long long noutput,
                             const float* h, const float2* wM, int nsm, cudaStream_t st, const void* halo, const int* inv,
                             bool accum)
{
    if (noutput <= 0)
        return 0;
    const long long nrows = T + noutput;
    if (M == 16)
        return launch_stream_tp<16, 256>(T, x, out, out_stride, nouts, noutput, nrows, h, wM, nsm, st, halo, inv, accum);
    if (M == 64)
        return launch_stream_tp<64, 256>(T, x, out, out_stride, nouts, noutput, nrows, h, wM, nsm, st, halo, inv, accum);
    switch (M) {
    case 8:
        return launch_stream_tp<8, 256>(T, x, out, out_stride, nouts, noutput, nrows, h, wM, nsm, st, halo, inv, accum);
    case 32:
        return launch_stream_tp<32, 256>(T, x, out, out_stride, nouts, noutput, nrows, h, wM, nsm, st, halo, inv, accum);
    case 128:
        return launch_stream_tp<128, 256>(T, x, out, out_stride, nouts, noutput, nrows, h, wM, nsm, st, halo, inv, accum);
    }
    return launch_stream_tp<256, 256>(T, x, out, out_stride, nouts, noutput, nrows, h, wM, nsm, st, halo, inv, accum);
}

int pfb_stream_launch(int M, int T, const void* x, void* out, long long out_stride, int nouts, long long noutput,
                      const float* h, const float2* wM, int nsm, cudaStream_t st, const void* halo, const int* inv)
{
    return stream_launch_any(M, T, x, out, out_stride, nouts, noutput, h, wM, nsm, st, halo, inv, false);
}

// 33 .. 64 taps per branch (a register ring holds 32): two channelizers on the same stream, the newest 32 taps first, then
// the older T - 32 on the rows 32 earlier, added to the first one's output (the DFT is linear).  Output time o of the kernel
// reads rows o + T - u for tap u: the first launch's stream starts T - 32 rows in, the second one's where the call's does.
// Identity channel map, history in place (no halo source).
bool pfb_stream_split_supported(int M, int T, int R, const void* x)
{
    return R == M && T > 32 && T <= 64 && M >= 8 && M <= 256 && (M & (M - 1)) == 0 && ((uintptr_t)x & 15) == 0;
}

int pfb_stream_split_launch(int M, int T, const void* x, void* out, long long out_stride, int nouts, long long noutput,
                            const float* h, const float2* wM, int nsm, cudaStream_t st)
{
    const float2* xa = (const float2*)x + (long long)(T - 32) * M;
    int rc = stream_launch_any(M, 32, xa, out, out_stride, nouts, noutput, h, wM, nsm, st, nullptr, nullptr, false);
    if (rc)
        return rc;
    return stream_launch_any(M, T - 32, x, out, out_stride, nouts, noutput, h + (size_t)32 * M, wM, nsm, st, nullptr, nullptr,
                             true);
}

} // namespace grb
