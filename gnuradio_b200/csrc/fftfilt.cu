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
#include "fft.cuh"

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

// Device FFT length for a tap count: at least the reference's 2*2^ceil(log2 ntaps), never below 4096.
int fftfilt_device_fftsize(int ntaps)
{
    int L = 4096;
    while (L < 2 * ntaps && L < 16384)
        L *= 2;
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
