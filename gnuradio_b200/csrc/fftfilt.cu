// fftfilt.cu -- fft_filter_ccc fast convolution (K3): load -> FFT -> x H -> inverse FFT -> store, one kernel.
//
// Replaces kernel::fft_filter_ccc::filter (reference gr-filter/lib/fft_filter.cc:241-285), which is
// overlap-ADD with a saved tail (d_tail).  The device uses overlap-SAVE: block b takes the L input items
// that end at its last output (ntaps-1 of them overlap the previous block), circularly convolves them with
// the taps through an L-point FFT, and keeps the V = L-ntaps+1 alias-free results.  Both forms compute
// the same linear convolution y = x (*) h; overlap-save has no carried state, so blocks are independent CTAs.
// H = FFT_L(taps/L) is computed once per set_taps with the same device FFT (reference set_taps :171-201
// folds the 1/fftsize scale into the taps the same way).  Decimation keeps every D-th y (:269-274).
//
// The forward transform runs the radix schedule in descending order and the inverse in ascending order,
// so the forward's output register pattern is exactly the inverse's input pattern: the spectrum multiply
// happens in registers and the data touches HBM once in, once out (16 B per sample for D = 1).
#include <cstdint>
#include <cstdlib>
#include "fft.cuh"
#include "tma.cuh"

namespace grb {

template <int L>
__global__ void __launch_bounds__(L / 16)
fftfilt_ols_kernel(const float2* __restrict__ in, float2* __restrict__ out, long long nout, long long nin,
                   int ntaps, int D, const float2* __restrict__ H, const float2* __restrict__ tw)
{
    constexpr int E = 16, TPV = L / E;
    constexpr int R0 = fft_first_radix(L, false);
    constexpr int RM = fft_last_radix(L, false); // == first radix of the reversed schedule
    constexpr int RL = fft_last_radix(L, true);
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float2* S = reinterpret_cast<float2*>(smem_raw);
    const int t = threadIdx.x;
    const int V = L - (ntaps - 1);
    const long long ny = nout * D; // full-rate convolution outputs needed
    const long long nblocks = (ny + V - 1) / V;

    for (long long blk = blockIdx.x; blk < nblocks; blk += gridDim.x) {
        const long long base = blk * V; // buffer index of the block's first input item
        float2 v[E];
#pragma unroll
        for (int b = 0; b < E / R0; b++)
#pragma unroll
            for (int q = 0; q < R0; q++) {
                const long long g = base + (t + b * TPV) + q * (L / R0);
                v[b * R0 + q] = (g < nin) ? in[g] : make_float2(0.f, 0.f);
            }
        cta_fft<L, true, false>(v, S, t, tw);
#pragma unroll
        for (int b = 0; b < E / RM; b++)
#pragma unroll
            for (int q = 0; q < RM; q++) {
                const int k = (t + b * TPV) + q * (L / RM);
                v[b * RM + q] = cmul(v[b * RM + q], __ldg(H + k));
            }
        __syncthreads(); // forward transform's last gather is done before the inverse reuses S
        cta_fft<L, false, true>(v, S, t, tw);
#pragma unroll
        for (int b = 0; b < E / RL; b++)
#pragma unroll
            for (int q = 0; q < RL; q++) {
                const int m = (t + b * TPV) + q * (L / RL);
                if (m >= ntaps - 1) {
                    const long long n = base + m - (ntaps - 1);
                    if (n < ny) {
                        if (D == 1)
                            out[n] = v[b * RL + q];
                        else if (n % D == 0)
                            out[n / D] = v[b * RL + q];
                    }
                }
            }
        __syncthreads();
    }
}

template <int L>
static int launch_ols(const void* in, void* out, long long nout, long long nin, int ntaps, int D, const float2* H,
                      const float2* tw, int nsm, cudaStream_t st)
{
    auto kern = fftfilt_ols_kernel<L>;
    constexpr int smem = fft_smem_elems(L) * 8;
    if (smem > 48 * 1024)
        GRB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    int occ = 1;
    GRB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, L / 16, smem));
    if (occ < 1)
        occ = 1;
    const int V = L - (ntaps - 1);
    long long nblocks = (nout * D + V - 1) / V;
    long long grid = (long long)nsm * occ;
    if (grid > nblocks)
        grid = nblocks;
    kern<<<(unsigned)grid, L / 16, smem, st>>>((const float2*)in, (float2*)out, nout, nin, ntaps, D, H, tw);
    GRB_LAUNCH_CHECK();
    return 0;
}

// ---- streaming variant (L = 4096, 8192) --------------------------------------------------------------
// ncu on the kernel above (profiles/r1_fftfilt_ncu.txt): one 512-thread CTA per SM, every block's loads
// exposed, 5.7 GB of L1 twiddle/H sectors per 2 GiB of input, MIO-throttle + barrier stalls.  Same recipe
// as fft_vcc_stream_kernel: the next block's L input items arrive by ONE TMA bulk copy while the current
// block goes through FFT -> x H -> inverse FFT; loop-invariant twiddles of both schedules sit in shared
// memory.  Block bases advance by V = L-ntaps+1 items, which may be odd, and TMA needs 16-byte aligned
// sources: the copy starts at the even item at or below the base (L+2 items) and the first pass reads
// with a 0/1 offset.
template <int L>
struct OlsStreamGeom {
    static constexpr int TPV = L / 16;
    static constexpr int XIN_BYTES = (L + 2) * 8;
    static constexpr int S_BYTES = ((fft_smem_elems(L) * 8 + 15) / 16) * 16;
    static constexpr int TWF_BYTES = fft_tws_slots(L, false) * TPV * 8;
    static constexpr int TWR_BYTES = fft_tws_slots(L, true) * TPV * 8;
    static constexpr int SMEM = XIN_BYTES + S_BYTES + TWF_BYTES + TWR_BYTES + 16;
};

template <int L>
__global__ void __launch_bounds__(L / 16)
fftfilt_ols_stream_kernel(const float2* __restrict__ in, float2* __restrict__ out, long long nout, long long nin,
                          int ntaps, int D, const float2* __restrict__ H, const float2* __restrict__ tw)
{
    using SG = OlsStreamGeom<L>;
    constexpr int E = 16, TPV = L / E;
    constexpr int R0 = fft_first_radix(L, false);
    constexpr int RM = fft_last_radix(L, false); // == first radix of the reversed schedule
    constexpr int RL = fft_last_radix(L, true);
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float2* XIN = reinterpret_cast<float2*>(smem_raw);
    float2* S = reinterpret_cast<float2*>(smem_raw + SG::XIN_BYTES);
    float2* TWF = reinterpret_cast<float2*>(smem_raw + SG::XIN_BYTES + SG::S_BYTES);
    float2* TWR = reinterpret_cast<float2*>(smem_raw + SG::XIN_BYTES + SG::S_BYTES + SG::TWF_BYTES);
    unsigned long long* bar = reinterpret_cast<unsigned long long*>(smem_raw + SG::XIN_BYTES + SG::S_BYTES +
                                                                    SG::TWF_BYTES + SG::TWR_BYTES);
    const int t = threadIdx.x;
    const int V = L - (ntaps - 1);
    const long long ny = nout * D;
    const long long nblocks = (ny + V - 1) / V;
    const long long nin_even = nin & ~1LL; // items the bulk copies may touch (the odd last one is read directly)

    auto issue = [&](long long blk) {
        const long long b0 = (blk * V) & ~1LL;
        long long n = nin_even - b0;
        n = n < 0 ? 0 : (n > L + 2 ? L + 2 : n);
        mbar_expect_tx(bar, (unsigned)(n * 8));
        if (n)
            bulk_g2s(XIN, in + b0, (unsigned)(n * 8), bar);
    };

    if (t == 0)
        mbar_init(bar, 1);
    cta_fft_tws_init<L, true, false>(TWF, t, tw);
    cta_fft_tws_init<L, false, true>(TWR, t, tw);
    __syncthreads();
    long long blk = blockIdx.x;
    if (t == 0 && blk < nblocks)
        issue(blk);

    unsigned phase = 0;
    for (; blk < nblocks; blk += gridDim.x, phase ^= 1) {
        const long long base = blk * V;
        const int off = (int)(base & 1);
        const long long b0 = base - off;
        mbar_wait(bar, phase);
        float2 v[E];
#pragma unroll
        for (int b = 0; b < E / R0; b++)
#pragma unroll
            for (int q = 0; q < R0; q++) {
                const int m = (t + b * TPV) + q * (L / R0);
                const long long g = base + m;
                float2 x = make_float2(0.f, 0.f);
                if (g < nin_even)
                    x = XIN[m + off];
                else if (g < nin)
                    x = in[g];
                v[b * R0 + q] = x;
            }
        (void)b0;
        const long long nblk = blk + gridDim.x;
        cta_fft_tws<L, true, false>(v, S, t, TWF, tw, [&]() {
            if (t == 0 && nblk < nblocks)
                issue(nblk);
        });
#pragma unroll
        for (int b = 0; b < E / RM; b++)
#pragma unroll
            for (int q = 0; q < RM; q++) {
                const int k = (t + b * TPV) + q * (L / RM);
                v[b * RM + q] = cmul(v[b * RM + q], __ldg(H + k));
            }
        __syncthreads(); // forward transform's last gather is done before the inverse reuses S
        cta_fft_tws<L, false, true>(v, S, t, TWR, tw, []() {});
#pragma unroll
        for (int b = 0; b < E / RL; b++)
#pragma unroll
            for (int q = 0; q < RL; q++) {
                const int m = (t + b * TPV) + q * (L / RL);
                if (m >= ntaps - 1) {
                    const long long n = base + m - (ntaps - 1);
                    if (n < ny) {
                        if (D == 1)
                            out[n] = v[b * RL + q];
                        else if (n % D == 0)
                            out[n / D] = v[b * RL + q];
                    }
                }
            }
        __syncthreads();
    }
}

template <int L>
static int launch_ols_stream(const void* in, void* out, long long nout, long long nin, int ntaps, int D,
                             const float2* H, const float2* tw, int nsm, cudaStream_t st)
{
    using SG = OlsStreamGeom<L>;
    auto kern = fftfilt_ols_stream_kernel<L>;
    GRB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SG::SMEM));
    int occ = 1;
    GRB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, L / 16, SG::SMEM));
    if (occ < 1)
        occ = 1;
    const int V = L - (ntaps - 1);
    long long nblocks = (nout * D + V - 1) / V;
    long long grid = (long long)nsm * occ;
    if (grid > nblocks)
        grid = nblocks;
    kern<<<(unsigned)grid, L / 16, SG::SMEM, st>>>((const float2*)in, (float2*)out, nout, nin, ntaps, D, H, tw);
    GRB_LAUNCH_CHECK();
    return 0;
}

// Device FFT length for a tap count: at least the reference's 2*2^ceil(log2 ntaps), never below 4096.
int fftfilt_device_fftsize(int ntaps)
{
    int L = 4096;
    while (L < 2 * ntaps && L < 16384)
        L *= 2;
    if (L == 8192 && ntaps - 1 > L / 4)
        L = 16384; // > 25 % of every block would be overlap: the longer transform wins (cfg2: 70 -> 78 Gsps)
    if (const char* e = getenv("GRB200_FFTFILT_L")) { // tuning knob: a longer transform wastes less on the overlap
        const int v = atoi(e);
        if ((v == 4096 || v == 8192 || v == 16384) && v >= L)
            L = v;
    }
    if (L - (ntaps - 1) < L / 4)
        return -1; // too many taps for a single-CTA transform
    return L;
}

int fftfilt_launch(int L, const void* in, void* out, long long nout, int ntaps, int D, const float2* H,
                   const float2* tw, int nsm, cudaStream_t st)
{
    if (nout <= 0)
        return 0;
    const long long nin = nout * D + ntaps - 1;
    static const bool use_stream = [] {
        const char* e = getenv("GRB200_FFTFILT_STREAM");
        return !(e && e[0] == '0');
    }();
    if (use_stream && (((uintptr_t)in) & 15) == 0) { // TMA bulk copies need a 16-byte aligned stream
        if (L == 4096)
            return launch_ols_stream<4096>(in, out, nout, nin, ntaps, D, H, tw, nsm, st);
        if (L == 8192)
            return launch_ols_stream<8192>(in, out, nout, nin, ntaps, D, H, tw, nsm, st);
    }
    switch (L) {
    case 4096:
        return launch_ols<4096>(in, out, nout, nin, ntaps, D, H, tw, nsm, st);
    case 8192:
        return launch_ols<8192>(in, out, nout, nin, ntaps, D, H, tw, nsm, st);
    case 16384:
        return launch_ols<16384>(in, out, nout, nin, ntaps, D, H, tw, nsm, st);
    }
    return fail(GRB200_ERANGE, "fft_filter: unsupported device fft length %d", L);
}

} // namespace grb
