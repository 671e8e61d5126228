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
#include "common.cuh"

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

// x: (nout + T) vectors of nf items, history first; out: nout vectors.
int filterbank_launch(const void* x, void* out, long long nout, int nf, int T, const float* hrev, const unsigned char* active,
                      int nsm, cudaStream_t st)
{
    if (nout <= 0)
        return 0;
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
