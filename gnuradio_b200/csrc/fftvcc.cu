// fftvcc.cu -- batched windowed FFT (K5), fused |.|^2 epilogue (K6) and standalone mag-squared.
//
// Replaces fft_vcc_fftw::work (reference gr-fft/lib/fft_vcc_fftw.cc:63-122): per vector
//   window multiply (:78-89; inverse+shift swaps the halves on the way in), or plain / half-swapped copy (:90-100),
//   FFT (:103), forward+shift rotates the output by ceil(N/2) (:106-115);
// and complex_to_mag_squared_impl::work (gr-blocks/lib/complex_to_mag_squared_impl.cc:37-48).
//
// One kernel per launch does load (+window, +input swap) -> cta_fft -> (+output rotate, +|.|^2) -> store,
// so a vector crosses HBM exactly once in each direction: 16 B per complex sample (12 B with the fused
// magnitude-squared), the figure the HBM roofline is taken against.  Persistent CTAs loop over vectors.
#include "fftvcc.cuh"

namespace grb {

template <int N, bool FWD, bool REAL>
__global__ void __launch_bounds__(VccGeom<N>::NT)
fft_vcc_kernel(const VccArgs a)
{
    using G = VccGeom<N>;
    const long long nvec = a.nvec;
    const float* __restrict__ window = a.window;
    const float2* __restrict__ tw = a.tw;
    const int shift = a.shift;
    constexpr int E = G::E, TPV = G::TPV, VPC = G::VPC;
    constexpr int R0 = fft_first_radix(N, false);
    constexpr int RL = fft_last_radix(N, false);
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int vl = threadIdx.x / TPV;
    const int t = threadIdx.x % TPV;
    float2* S = reinterpret_cast<float2*>(smem_raw) + vl * fft_smem_elems(N);
    const bool in_swap = (!FWD) && shift;
    const bool out_rot = FWD && shift;

    for (long long vbase = (long long)blockIdx.x * VPC; vbase < nvec; vbase += (long long)gridDim.x * VPC) {
        const long long vec = vbase + vl;
        const bool valid = vec < nvec;
        float2 v[E];
        const long long sv = (vec * a.vstride + a.voff) * N;
#pragma unroll
        for (int b = 0; b < E / R0; b++) {
#pragma unroll
            for (int q = 0; q < R0; q++) {
                int d = (t + b * TPV) + q * (N / R0);
                if (in_swap)
                    d = (d + N / 2) & (N - 1);
                float2 x = make_float2(0.f, 0.f);
                if (valid) {
                    x = REAL ? make_float2(reinterpret_cast<const float*>(a.in)[sv + d], 0.f)
                             : reinterpret_cast<const float2*>(a.in)[sv + d];
                    if (window != nullptr) {
                        const float w = __ldg(window + d);
                        x.x *= w;
                        x.y *= w;
                    }
                }
                v[b * R0 + q] = x;
            }
        }
        cta_fft<N, FWD>(v, S, t, tw);
        if (valid) {
#pragma unroll
            for (int b = 0; b < E / RL; b++) {
#pragma unroll
                for (int q = 0; q < RL; q++) {
                    int k = (t + b * TPV) + q * (N / RL);
                    if (out_rot)
                        k = (k + N / 2) & (N - 1);
                    vcc_store(a, vec, N, k, v[b * RL + q]);
                }
            }
        }
        if (N > E)
            __syncthreads(); // the next iteration's first pass overwrites S
    }
}

// Any other size: direct DFT from shared memory, one CTA per vector (O(N^2), correctness path for the
// sizes FFTW would also accept; the hot configurations are powers of two).
__global__ void __launch_bounds__(256)
dft_direct_kernel(const VccArgs a, int N, int forward, int real_in)
{
    const long long nvec = a.nvec;
    const float* __restrict__ window = a.window;
    const float2* __restrict__ tw = a.tw;
    const int shift = a.shift;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float2* x = reinterpret_cast<float2*>(smem_raw);
    float2* w = x + N;
    const bool in_swap = (!forward) && shift;
    const bool out_rot = forward && shift;
    const int len_in = N / 2;        // floor(N/2), fft_vcc_fftw.cc:79,93
    const int len_out = (N + 1) / 2; // ceil(N/2),  fft_vcc_fftw.cc:107
    for (int i = threadIdx.x; i < N; i += blockDim.x) {
        float2 c = tw[i];
        w[i] = forward ? c : make_float2(c.x, -c.y);
    }
    for (long long vec = blockIdx.x; vec < nvec; vec += gridDim.x) {
        __syncthreads();
        for (int d = threadIdx.x; d < N; d += blockDim.x) {
            int s = in_swap ? (d + len_in) % N : d;
            const long long sv = (vec * a.vstride + a.voff) * N;
            float2 c = real_in ? make_float2(reinterpret_cast<const float*>(a.in)[sv + s], 0.f)
                               : reinterpret_cast<const float2*>(a.in)[sv + s];
            if (window != nullptr) {
                const float ww = window[s];
                c.x *= ww;
                c.y *= ww;
            }
            x[d] = c;
        }
        __syncthreads();
        for (int k = threadIdx.x; k < N; k += blockDim.x) {
            float2 acc = make_float2(0.f, 0.f);
            int idx = 0;
            for (int n = 0; n < N; n++) {
                const float2 a = x[n], c = w[idx];
                acc.x = fmaf(a.x, c.x, acc.x);
                acc.x = fmaf(-a.y, c.y, acc.x);
                acc.y = fmaf(a.x, c.y, acc.y);
                acc.y = fmaf(a.y, c.x, acc.y);
                idx += k;
                if (idx >= N)
                    idx -= N;
            }
            const int o = out_rot ? (k + N - len_out) % N : k;
            vcc_store(a, vec, N, o, acc);
        }
    }
}

// complex_to_mag_squared standalone: 8 B in, 4 B out per item, pure streaming.
__global__ void __launch_bounds__(256)
magsq_kernel(const float2* __restrict__ in, float* __restrict__ out, long long n)
{
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const float2 a = in[i];
        out[i] = fmaf(a.x, a.x, a.y * a.y);
    }
}

// ---- host dispatch --------------------------------------------------------------------------------
template <int N, bool FWD, bool REAL>
static int launch_vcc(const VccArgs& a, int nsm, cudaStream_t st)
{
    using G = VccGeom<N>;
    auto kern = fft_vcc_kernel<N, FWD, REAL>;
    if (G::SMEM > 48 * 1024)
        GRB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, G::SMEM));
    int occ = 1;
    GRB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, G::NT, G::SMEM));
    if (occ < 1)
        occ = 1;
    long long need = (a.nvec + G::VPC - 1) / G::VPC;
    long long grid = (long long)nsm * occ;
    if (grid > need)
        grid = need;
    kern<<<(unsigned)grid, G::NT, G::SMEM, st>>>(a);
    GRB_LAUNCH_CHECK();
    return 0;
}

template <int N, bool REAL>
static int launch_vcc_dir(bool fwd, const VccArgs& a, int nsm, cudaStream_t st)
{
    if constexpr (N >= 32 && N <= 8192) { // (two passes at least: the prefetch is issued after the first exchange)
        static const bool use_stream = [] {
            const char* e = getenv("GRB200_FFT_STREAM");
            return !(e && e[0] == '0');
        }();
        // TMA sources must be 16-byte aligned (every vector start is, once the base is); a few vectors are not worth a
        // persistent pipeline
        if (use_stream && (((uintptr_t)a.in) & 15) == 0 && a.nvec >= 64)
            return fft_vcc_stream_launch(N, fwd, REAL, a, nsm, st);
    }
    return fwd ? launch_vcc<N, true, REAL>(a, nsm, st) : launch_vcc<N, false, REAL>(a, nsm, st);
}

bool fft_size_is_fast(int N) { return N >= 2 && N <= 16384 && (N & (N - 1)) == 0; }
int fft_direct_max() { return 8192; }

int fft_vcc_launch_ex(int N, bool fwd, bool real_in, const VccArgs& a, int nsm, cudaStream_t st)
{
    if (a.nvec <= 0)
        return 0;
#define GRB_VCC_CASE(n)                                                                   \
    case n:                                                                               \
        return real_in ? launch_vcc_dir<n, true>(fwd, a, nsm, st) : launch_vcc_dir<n, false>(fwd, a, nsm, st);
    if (fft_size_is_fast(N)) {
        switch (N) {
            GRB_VCC_CASE(2)
            GRB_VCC_CASE(4)
            GRB_VCC_CASE(8)
            GRB_VCC_CASE(16)
            GRB_VCC_CASE(32)
            GRB_VCC_CASE(64)
            GRB_VCC_CASE(128)
            GRB_VCC_CASE(256)
            GRB_VCC_CASE(512)
            GRB_VCC_CASE(1024)
            GRB_VCC_CASE(2048)
            GRB_VCC_CASE(4096)
            GRB_VCC_CASE(8192)
            GRB_VCC_CASE(16384)
        }
    }
#undef GRB_VCC_CASE
    if (fft_size_is_big(N))
        return big_launch(N, fwd, real_in, a, nsm, st);
    if (N > fft_direct_max())
        return fail(GRB200_ERANGE, "fft_vcc: fft_size %d is not supported (powers of two up to 262144, any size up to %d)",
                    N, fft_direct_max());
    static const bool use_bluestein = [] {
        const char* e = GRB_GETENV_ONCE("GRB200_FFT_BLUESTEIN"); // 0 = the O(N^2) kernel for every size
        return !(e && e[0] == '0');
    }();
    if (use_bluestein && N >= 128) // (below that the direct sum is cheaper than two 1024-point transforms)
        return bluestein_launch(N, fwd, real_in, a, nsm, st);
    const size_t smem = (size_t)N * 16;
    if (smem > 48 * 1024)
        GRB_CUDA(cudaFuncSetAttribute(dft_direct_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    long long grid = a.nvec < (long long)nsm * 4 ? a.nvec : (long long)nsm * 4;
    dft_direct_kernel<<<(unsigned)grid, 256, smem, st>>>(a, N, fwd ? 1 : 0, real_in ? 1 : 0);
    GRB_LAUNCH_CHECK();
    return 0;
}

int fft_vcc_launch(int N, bool fwd, const void* in, void* out, long long nvec, const float* window, int shift,
                   int magsq, const float2* tw, int nsm, cudaStream_t st)
{
    VccArgs a{ in, out, nvec, window, tw, shift, magsq ? 1 : 0, 1, 0, 0.f, 0.f, nullptr };
    return fft_vcc_launch_ex(N, fwd, false, a, nsm, st);
}

int magsq_launch(const void* in, float* out, long long n, int nsm, cudaStream_t st)
{
    if (n <= 0)
        return 0;
    long long need = (n + 255) / 256;
    long long grid = (long long)nsm * 16;
    if (grid > need)
        grid = need;
    magsq_kernel<<<(unsigned)grid, 256, 0, st>>>((const float2*)in, out, n);
    GRB_LAUNCH_CHECK();
    return 0;
}

} // namespace grb

