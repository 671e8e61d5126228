// filterbank.cu -- filterbank_vcvcf: one FIR per vector element.
//
// Replaces filterbank_vcvcf_impl::general_work (reference gr-filter/lib/filterbank_vcvcf_impl.cc:54-96) over
// kernel::filterbank (gr-filter/lib/filterbank.cc:26-79): items are vectors of nfilts complex64, history ntaps+1,
//     out[j][i] = sum_t rev_i[t] * in[j+t][i]  (t < ntaps; the newest row j+ntaps is not read -- the block's one-item delay),
// and a filter whose taps are all zero writes zeros without reading its input (:83-88).
//
// It is the branch-FIR half of the channelizer without the DFT: 8 B in + 8 B out per item, a handful of FMAs -- HBM-bound.
// A thread owns one filter and a run of consecutive output rows and walks down the run with the last TP rows of its column
// in a register ring: every input item is loaded ONCE (lanes across the filters, so a row is a coalesced read), each output
// is ntaps packed FFMA2 and one coalesced store.  Runs overlap by ntaps-1 rows (re-read from L2, < 3 % at the default run).
#include <type_traits>
#include "common.cuh"
#include "tma.cuh"

namespace grb {

template <int TP>
__global__ void __launch_bounds__(256)
filterbank_kernel(const float2* __restrict__ x, float2* __restrict__ out, long long nout, int nf, int T,
                  const float* __restrict__ hrev /* [T][nf]: hrev[t*nf + i] = rev_i[t] */,
                  const unsigned char* __restrict__ active, int run_len, long long nruns)
{
    const long long tg = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int i = (int)(tg % nf);
    const long long run = tg / nf;
    if (run >= nruns)
        return;
    const long long j0 = run * run_len;
    long long j1 = j0 + run_len;
    j1 = j1 > nout ? nout : j1;
    if (!active[i]) {
        for (long long j = j0; j < j1; j++)
            out[j * nf + i] = make_float2(0.f, 0.f);
        return;
    }
    // the TP-row window ends at the newest row: slot v <-> row j + T - TP + v; only the last T slots carry taps
    const int skip = TP - T;
    float taps[TP];
#pragma unroll
    for (int v = 0; v < TP; v++)
        taps[v] = v >= skip ? __ldg(hrev + (size_t)(v - skip) * nf + i) : 0.f;
    float2 ring[TP];
#pragma unroll
    for (int v = 0; v < TP - 1; v++) {
        const long long r = j0 + v - skip;
        ring[v] = (v >= skip) ? x[r * nf + i] : make_float2(0.f, 0.f);
    }
    ring[TP - 1] = make_float2(0.f, 0.f);
    // the TP new rows of a block are loaded together, one block ahead of the arithmetic: TP independent loads in flight per
    // thread whatever the compiler's own scheduling would have done
    float2 nx[TP];
#pragma unroll
    for (int k = 0; k < TP; k++)
        nx[k] = j0 + k < j1 ? x[(j0 + k + T - 1) * nf + i] : make_float2(0.f, 0.f);
    for (long long j = j0; j < j1; j += TP) {
        float2 cur[TP];
#pragma unroll
        for (int k = 0; k < TP; k++)
            cur[k] = nx[k];
        const long long jn = j + TP;
#pragma unroll
        for (int k = 0; k < TP; k++)
            nx[k] = jn + k < j1 ? x[(jn + k + T - 1) * nf + i] : make_float2(0.f, 0.f);
#pragma unroll
        for (int k = 0; k < TP; k++) {
            const long long jj = j + k;
            ring[(k + TP - 1) % TP] = cur[k];
            float2 a0 = make_float2(0.f, 0.f), a1 = make_float2(0.f, 0.f);
#pragma unroll
            for (int v = 0; v < TP; v += 2) {
                if (v >= skip)
                    a0 = __ffma2_rn(make_float2(taps[v], taps[v]), ring[(k + v) % TP], a0);
                if (v + 1 >= skip)
                    a1 = __ffma2_rn(make_float2(taps[v + 1], taps[v + 1]), ring[(k + v + 1) % TP], a1);
            }
            if (jj < j1)
                out[jj * nf + i] = __fadd2_rn(a0, a1);
        }
    }
}

// any tap count: one output per thread, window through L1
__global__ void __launch_bounds__(256)
filterbank_generic_kernel(const float2* __restrict__ x, float2* __restrict__ out, long long nout, int nf, int T,
                          const float* __restrict__ hrev, const unsigned char* __restrict__ active)
{
    const long long total = nout * nf;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += stride) {
        const int i = (int)(e % nf);
        const long long j = e / nf;
        float2 a = make_float2(0.f, 0.f);
        if (active[i])
            for (int t = 0; t < T; t++) {
                const float h = __ldg(hrev + (size_t)t * nf + i);
                a = __ffma2_rn(make_float2(h, h), __ldg(x + (j + t) * nf + i), a);
            }
        out[e] = a;
    }
}

template <int TP>
static int launch_fb(const void* x, void* out, long long nout, int nf, int T, const float* hrev, const unsigned char* active,
                     int nsm, cudaStream_t st)
{
    // runs long enough that the ntaps-1 warm-up rows are a few per cent, short enough to fill the machine
    long long run_len = 32LL * T;
    if (run_len < 256)
        run_len = 256;
    const long long want = (long long)nsm * 2048 * 2;
    while (run_len > 64 && ((nout + run_len - 1) / run_len) * nf < want)
        run_len /= 2;
    const long long nruns = (nout + run_len - 1) / run_len;
    const long long nthreads = nruns * nf;
    const long long nblk = (nthreads + 255) / 256;
    if (nblk > 0x7fffffffLL)
        return fail(GRB200_ERANGE, "filterbank_vcvcf: work() call too long for one launch");
    filterbank_kernel<TP><<<(unsigned)nblk, 256, 0, st>>>((const float2*)x, (float2*)out, nout, nf, T, hrev, active,
                                                          (int)run_len, nruns);
    GRB_LAUNCH_CHECK();
    return 0;
}

// ---- streaming form (nf a power of two 8 .. 256, <= 32 taps, 16-byte aligned input): pfb_stream_kernel without its DFT --------
// The register-ring kernel above is capped by its own occupancy (164 registers with the block-ahead loads: 0.66 of HBM).  Here
// the loads are not the threads' business at all: a persistent CTA owns NG contiguous runs of output rows; per iteration and
// run the KT new rows are ONE contiguous block of the input -- one TMA bulk copy into a ring of NS shared-memory stages, issued
// NS-1 iterations ahead by an elected thread -- and thread (filter i, run g) walks down its run with the last TP rows of its
// column in a register ring: one 8-byte shared load, TP packed FMAs and one coalesced 8-byte store per output.
template <int M, int NT, int NS>
struct FbStreamGeom {
    static constexpr int NG = NT / M, KT = 16, OT = NG * KT;
    static constexpr int STAGE_BYTES = NG * KT * M * 8;
    static constexpr int SMEM = NS * STAGE_BYTES + NS * 8 + 16;
};

template <int M, int TP, int NT, int NS>
__global__ void __launch_bounds__(NT)
filterbank_stream_kernel(const float2* __restrict__ x, float2* __restrict__ out, long long nout, int T,
                         const float* __restrict__ hrev, const unsigned char* __restrict__ active, int iters)
{
    using Geo = FbStreamGeom<M, NT, NS>;
    constexpr int NG = Geo::NG, KT = Geo::KT;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float2* xs = reinterpret_cast<float2*>(smem_raw); // [NS][NG][KT][M]
    unsigned long long* bar = reinterpret_cast<unsigned long long*>(smem_raw + NS * Geo::STAGE_BYTES);
    const int tid = threadIdx.x, i = tid % M, bg = tid / M;
    const long long region = (long long)iters * KT; // output rows per run
    const long long cta0 = (long long)blockIdx.x * NG * region;
    const long long gs = cta0 + (long long)bg * region; // this thread's first output row
    const long long nrows = nout + T - 1;               // input rows that are read (the newest one never is)
    if (tid == 0)
        for (int s = 0; s < NS; s++)
            mbar_init(&bar[s], 1);
    const bool live = active[i] != 0;
    // tap of the row u behind the newest one of an output's window (row j + T - 1 - u <-> rev_i[T - 1 - u])
    float taps[TP];
#pragma unroll
    for (int u = 0; u < TP; u++)
        taps[u] = (u < T && live) ? __ldg(hrev + (size_t)(T - 1 - u) * M + i) : 0.f;
    // register ring: once the newest row of output j = gs + n is in, ring[(n - u) mod TP] = row j + T - 1 - u
    float2 ring[TP];
#pragma unroll
    for (int u = 1; u < TP; u++) {
        const long long r = gs + T - 1 - u;
        ring[(TP - u) % TP] = (u < T && r < nrows) ? x[r * M + i] : make_float2(0.f, 0.f);
    }
    ring[0] = make_float2(0.f, 0.f);
    __syncthreads(); // barriers initialised

    auto issue = [&](int it) { // iteration it's NG blocks of KT new rows into stage it mod NS
        unsigned long long* b = &bar[it % NS];
        float2* dst = xs + (size_t)(it % NS) * (NG * KT * M);
        auto rows_of = [&](int g, long long& r0) {
            r0 = cta0 + (long long)g * region + (long long)it * KT + T - 1;
            const long long nv = nrows - r0;
            return (unsigned)((nv < 0 ? 0 : (nv > KT ? KT : nv)) * M * 8);
        };
        unsigned total = 0;
        long long r0;
        constexpr int UNR = NG <= 16 ? NG : 1;
#pragma unroll UNR
        for (int g = 0; g < NG; g++)
            total += rows_of(g, r0);
        mbar_expect_tx(b, total);
#pragma unroll UNR
        for (int g = 0; g < NG; g++) {
            const unsigned bytes = rows_of(g, r0);
            if (bytes)
                bulk_g2s(dst + g * KT * M, x + r0 * M, bytes, b);
        }
    };
    if (tid == 0)
        for (int it = 0; it < NS - 1 && it < iters; it++)
            issue(it);

    auto iteration = [&](int it, auto phc) {
        constexpr int PH = decltype(phc)::value; // ring position of this iteration's first row
        // stage (it - 1) mod NS was drained by every thread before the barrier that ended iteration it - 1
        if (tid == 0 && it + NS - 1 < iters)
            issue(it + NS - 1);
        mbar_wait(&bar[it % NS], (unsigned)((it / NS) & 1));
        const float2* src = xs + (size_t)(it % NS) * (NG * KT * M) + (bg * KT) * M + i;
        float2* o = out + (gs + (long long)it * KT) * M + i;
        const long long left = nout - (gs + (long long)it * KT); // rows of this run's block that exist
#pragma unroll
        for (int k = 0; k < KT; k++) {
            ring[(PH + k) % TP] = src[k * M];
            float2 a0 = make_float2(0.f, 0.f), a1 = make_float2(0.f, 0.f);
#pragma unroll
            for (int u = 0; u < TP; u += 2) {
                a0 = __ffma2_rn(make_float2(taps[u], taps[u]), ring[(PH + k - u + 2 * TP) % TP], a0);
                a1 = __ffma2_rn(make_float2(taps[u + 1], taps[u + 1]), ring[(PH + k - u - 1 + 2 * TP) % TP], a1);
            }
            // a filter whose taps are all zero writes zeros whatever its input holds (filterbank.cc:83-88: it is never run)
            const float2 y = live ? __fadd2_rn(a0, a1) : make_float2(0.f, 0.f);
            if (k < left)
                o[(long long)k * M] = y;
        }
        __syncthreads(); // the stage is drained: the next iteration's first act may refill it
    };

    if constexpr (TP <= KT) {
        for (int it = 0; it < iters; it++)
            iteration(it, std::integral_constant<int, 0>{});
    } else { // TP == 2 KT: the ring position alternates between 0 and KT
        for (int it = 0; it < iters; it += 2) {
            iteration(it, std::integral_constant<int, 0>{});
            if (it + 1 < iters)
                iteration(it + 1, std::integral_constant<int, KT>{});
        }
    }
}

template <int M, int TP>
static int launch_fb_stream(const void* x, void* out, long long nout, int T, const float* hrev, const unsigned char* active,
                            int nsm, cudaStream_t st)
{
    constexpr int NT = 256, NS = 3;
    using Geo = FbStreamGeom<M, NT, NS>;
    auto kern = filterbank_stream_kernel<M, TP, NT, NS>;
    GRB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Geo::SMEM));
    int occ = 1;
    GRB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, NT, Geo::SMEM));
    if (occ < 1)
        occ = 1;
    long long grid = (long long)nsm * occ;
    const long long need = (nout + Geo::OT - 1) / Geo::OT;
    if (grid > need)
        grid = need;
    long long iters = (nout + grid * Geo::OT - 1) / (grid * Geo::OT);
    if (TP > Geo::KT && (iters & 1))
        iters++;
    grid = (nout + iters * Geo::OT - 1) / (iters * Geo::OT); // drop CTAs the rounding left without work
    if (iters > 0x7fffffffLL / Geo::KT)
        return fail(GRB200_ERANGE, "filterbank_vcvcf: work() call too long for one launch");
    kern<<<(unsigned)grid, NT, Geo::SMEM, st>>>((const float2*)x, (float2*)out, nout, T, hrev, active, (int)iters);
    GRB_LAUNCH_CHECK();
    return 0;
}

template <int M>
static int launch_fb_stream_tp(const void* x, void* out, long long nout, int T, const float* hrev, const unsigned char* active,
                               int nsm, cudaStream_t st)
{
    if (T <= 8)
        return launch_fb_stream<M, 8>(x, out, nout, T, hrev, active, nsm, st);
    if (T <= 16)
        return launch_fb_stream<M, 16>(x, out, nout, T, hrev, active, nsm, st);
    return launch_fb_stream<M, 32>(x, out, nout, T, hrev, active, nsm, st);
}

// x: (nout + T) vectors of nf items, history first; out: nout vectors.
int filterbank_launch(const void* x, void* out, long long nout, int nf, int T, const float* hrev, const unsigned char* active,
                      int nsm, cudaStream_t st)
{
    if (nout <= 0)
        return 0;
    static const bool use_stream = [] {
        const char* e = GRB_GETENV_ONCE("GRB200_FBANK_STREAM"); // 0 = the register-ring kernels only
        return !(e && e[0] == '0');
    }();
    if (use_stream && T <= 32 && nout >= 4096 && (((uintptr_t)x) & 15) == 0) { // (TMA bulk copies: 16-byte aligned rows)
        if (nf == 8)
            return launch_fb_stream_tp<8>(x, out, nout, T, hrev, active, nsm, st);
        if (nf == 16)
            return launch_fb_stream_tp<16>(x, out, nout, T, hrev, active, nsm, st);
        if (nf == 32)
            return launch_fb_stream_tp<32>(x, out, nout, T, hrev, active, nsm, st);
        if (nf == 128)
            return launch_fb_stream_tp<128>(x, out, nout, T, hrev, active, nsm, st);
        if (nf == 64)
            return launch_fb_stream_tp<64>(x, out, nout, T, hrev, active, nsm, st);
        if (nf == 256)
            return launch_fb_stream_tp<256>(x, out, nout, T, hrev, active, nsm, st);
    }
    if (T <= 8)
        return launch_fb<8>(x, out, nout, nf, T, hrev, active, nsm, st);
    if (T <= 16)
        return launch_fb<16>(x, out, nout, nf, T, hrev, active, nsm, st);
    if (T <= 32)
        return launch_fb<32>(x, out, nout, nf, T, hrev, active, nsm, st);
    long long need = (nout * nf + 255) / 256;
    long long grid = (long long)nsm * 16;
    if (grid > need)
        grid = need;
    filterbank_generic_kernel<<<(unsigned)grid, 256, 0, st>>>((const float2*)x, (float2*)out, nout, nf, T, hrev, active);
    GRB_LAUNCH_CHECK();
    return 0;
}

} // namespace grb
