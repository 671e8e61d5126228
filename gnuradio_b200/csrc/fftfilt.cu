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

// Decimation on store: output n = nb + e is kept when D divides it.  The 64-bit division is done once per block (nb = qb D +
// rb); an element costs a 32-bit one.
__device__ __forceinline__ void store_decimated(float2* __restrict__ out, long long qb, unsigned rb, unsigned D, unsigned e,
                                                float2 v)
{
    const unsigned s = rb + e, q = s / D;
    if (s == q * D)
        out[qb + q] = v;
}

// RP ("real pair", fir_filter_fff with a long filter): `in` / `out` are FLOAT streams.  A real filter acts on the real and the
// imaginary part of a complex stream independently, so the two halves of the call -- outputs [0, half) and [half, nout) --
// ride through the transform as re and im of one complex block: two coalesced float loads per item, two float stores.
template <int L, bool RP>
__global__ void __launch_bounds__(L / 16)
fftfilt_ols_kernel(const float2* __restrict__ in, float2* __restrict__ out, long long nout, long long nin,
                   int ntaps, int D, const float2* __restrict__ H, const float2* __restrict__ tw, long long half)
{
    constexpr int E = 16, TPV = L / E;
    constexpr int R0 = fft_first_radix(L, false);
    constexpr int RM = fft_last_radix(L, false); // == first radix of the reversed schedule
    constexpr int RL = fft_last_radix(L, true);
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float2* S = reinterpret_cast<float2*>(smem_raw);
    const int t = threadIdx.x;
    const int V = L - (ntaps - 1);
    const long long ny = (RP ? half : nout) * D; // full-rate convolution outputs needed (RP: of the longer, first half)
    const long long nblocks = (ny + V - 1) / V;
    const float* __restrict__ inf = reinterpret_cast<const float*>(in);
    float* __restrict__ outf = reinterpret_cast<float*>(out);
    const long long nyB = RP ? (nout - half) * D : 0; // RP: outputs of the second half, whose stream starts half D items in

    for (long long blk = blockIdx.x; blk < nblocks; blk += gridDim.x) {
        const long long base = blk * V; // buffer index of the block's first input item
        float2 v[E];
#pragma unroll
        for (int b = 0; b < E / R0; b++)
#pragma unroll
            for (int q = 0; q < R0; q++) {
                const long long g = base + (t + b * TPV) + q * (L / R0);
                if (RP) {
                    const long long gb = g + half * D;
                    v[b * R0 + q] = make_float2(g < nin ? inf[g] : 0.f, gb < nin ? inf[gb] : 0.f);
                } else {
                    v[b * R0 + q] = (g < nin) ? in[g] : make_float2(0.f, 0.f);
                }
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
        if (RP) {
            const long long dec_q = base / D;
            const unsigned dec_r = (unsigned)(base - dec_q * D);
#pragma unroll
            for (int b = 0; b < E / RL; b++)
#pragma unroll
                for (int q = 0; q < RL; q++) {
                    const int m = (t + b * TPV) + q * (L / RL);
                    const long long n = base + m - (ntaps - 1);
                    if (m >= ntaps - 1 && n < ny) {
                        const unsigned sdec = dec_r + (unsigned)(m - (ntaps - 1)), qd = sdec / (unsigned)D;
                        if (sdec == qd * (unsigned)D) {
                            outf[dec_q + qd] = v[b * RL + q].x;
                            if (n < nyB)
                                outf[half + dec_q + qd] = v[b * RL + q].y;
                        }
                    }
                }
        } else if (D == 1) {
#pragma unroll
            for (int b = 0; b < E / RL; b++)
#pragma unroll
                for (int q = 0; q < RL; q++) {
                    const int m = (t + b * TPV) + q * (L / RL);
                    const long long n = base + m - (ntaps - 1);
                    if (m >= ntaps - 1 && n < ny)
                        out[n] = v[b * RL + q];
                }
        } else {
            const long long dec_q = base / D;
            const unsigned dec_r = (unsigned)(base - dec_q * D);
#pragma unroll
            for (int b = 0; b < E / RL; b++)
#pragma unroll
                for (int q = 0; q < RL; q++) {
                    const int m = (t + b * TPV) + q * (L / RL);
                    if (m >= ntaps - 1 && base + m - (ntaps - 1) < ny)
                        store_decimated(out, dec_q, dec_r, (unsigned)D, (unsigned)(m - (ntaps - 1)), v[b * RL + q]);
                }
        }
        __syncthreads();
    }
}

template <int L>
static int launch_ols(const void* in, void* out, long long nout, long long nin, int ntaps, int D, const float2* H,
                      const float2* tw, int nsm, cudaStream_t st)
{
    auto kern = fftfilt_ols_kernel<L, false>;
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
    kern<<<(unsigned)grid, L / 16, smem, st>>>((const float2*)in, (float2*)out, nout, nin, ntaps, D, H, tw, 0);
    GRB_LAUNCH_CHECK();
    return 0;
}

// fir_filter_fff through the overlap-save kernel: two halves of a real call as one complex stream (see RP above)
template <int L>
static int launch_ols_realpair(const void* in, void* out, long long nout, int ntaps, int D, const float2* H, const float2* tw,
                               int nsm, cudaStream_t st)
{
    auto kern = fftfilt_ols_kernel<L, true>;
    constexpr int smem = fft_smem_elems(L) * 8;
    if (smem > 48 * 1024)
        GRB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    int occ = 1;
    GRB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, L / 16, smem));
    if (occ < 1)
        occ = 1;
    const long long half = (nout + 1) / 2, nin = nout * D + ntaps - 1; // (floats)
    const int V = L - (ntaps - 1);
    long long nblocks = (half * D + V - 1) / V;
    long long grid = (long long)nsm * occ;
    if (grid > nblocks)
        grid = nblocks;
    kern<<<(unsigned)grid, L / 16, smem, st>>>((const float2*)in, (float2*)out, nout, nin, ntaps, D, H, tw, half);
    GRB_LAUNCH_CHECK();
    return 0;
}

// L for a real-pair plan (one partition, at least half of every block is output), or -1
int fftfilt_realpair_L(int ntaps)
{
    for (int L = 4096; L <= 16384; L *= 2)
        if (ntaps - 1 <= L / 2)
            return L;
    return -1;
}

static int ols3_realpair(const void* in, void* out, long long nout, int ntaps, int D, const float2* H, const float2* tw,
                         int nsm, cudaStream_t st, bool* done); // (three-group kernel, below)

int fftfilt_realpair_launch(int L, const void* in, void* out, long long nout, int ntaps, int D, const float2* H,
                            const float2* tw, int nsm, cudaStream_t st)
{
    if (nout <= 0)
        return 0;
    if (L == 4096) {
        bool done = false;
        const int rc = ols3_realpair(in, out, nout, ntaps, D, H, tw, nsm, st, &done);
        if (rc || done)
            return rc;
    }
    switch (L) {
    case 4096:
        return launch_ols_realpair<4096>(in, out, nout, ntaps, D, H, tw, nsm, st);
    case 8192:
        return launch_ols_realpair<8192>(in, out, nout, ntaps, D, H, tw, nsm, st);
    case 16384:
        return launch_ols_realpair<16384>(in, out, nout, ntaps, D, H, tw, nsm, st);
    }
    return fail(GRB200_ERANGE, "fir_filter_fff: no overlap-save kernel for fft length %d", L);
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
        if (D == 1) {
#pragma unroll
            for (int b = 0; b < E / RL; b++)
#pragma unroll
                for (int q = 0; q < RL; q++) {
                    const int m = (t + b * TPV) + q * (L / RL);
                    const long long n = base + m - (ntaps - 1);
                    if (m >= ntaps - 1 && n < ny)
                        out[n] = v[b * RL + q];
                }
        } else {
            const long long dec_q = base / D;
            const unsigned dec_r = (unsigned)(base - dec_q * D);
#pragma unroll
            for (int b = 0; b < E / RL; b++)
#pragma unroll
                for (int q = 0; q < RL; q++) {
                    const int m = (t + b * TPV) + q * (L / RL);
                    if (m >= ntaps - 1 && base + m - (ntaps - 1) < ny)
                        store_decimated(out, dec_q, dec_r, (unsigned)D, (unsigned)(m - (ntaps - 1)), v[b * RL + q]);
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

// ---- two-partition variant (long filters; L = 4096 or 8192, hop V = L/2, ntaps <= L) -------------------------------
// Plain overlap-save wastes ntaps-1 of every L transformed items; at 4096 taps that forces a 16384-point transform
// (1024 threads, 64-register cap, spills, no room for a prefetch buffer: 78 Gsps, profiles/r1_fftfilt_ncu.txt).
// Uniformly partitioned overlap-save keeps the efficient 4096-point machinery instead: the taps are split into
// h0 = h[0, V) and h1 = h[V, ntaps), H0/H1 are their L-point spectra, and
//     y_b = IFFT( X_b . H0 + X_{b-1} . H1 ),   X_b = FFT of the L items that end at block b's last output,
// whose last V items are alias-free.  One forward and one inverse transform per V outputs, like plain overlap-save at
// 50 % overlap -- but X_{b-1} has to survive from one block to the next.  Registers (128/thread at 2 CTAs/SM) and shared
// memory (100 KB/CTA) are both full, so the previous spectrum lives in TENSOR MEMORY: each thread parks its 16 complex
// values in 32 columns of its own TMEM lane (tcgen05.st.32x32b) and reads them back one block later (tcgen05.ld) -- 64
// columns per 256-thread CTA of an SM-local memory this kernel otherwise leaves idle.  Every CTA owns a contiguous run of
// blocks (one extra forward transform warms the delay line up), so consecutive blocks also share their overlap in
// shared memory: only the V NEW items of a block come from HBM (one TMA bulk copy per block into a two-slice ring).
template <int L>
struct UpolsGeom {
    static constexpr int TPV = L / 16, V = L / 2;
    static constexpr int SLICE = V + 2; // items per ring slice: the copy starts at an even item (16-byte TMA rule)
    static constexpr int XIN_BYTES = 2 * SLICE * 8;
    static constexpr int S_BYTES = ((fft_smem_elems(L) * 8 + 15) / 16) * 16;
    static constexpr int TWF_BYTES = fft_tws_slots(L, false) * TPV * 8;
    static constexpr int TWR_BYTES = fft_tws_slots(L, true) * TPV * 8;
    static constexpr int SMEM = XIN_BYTES + S_BYTES + TWF_BYTES + TWR_BYTES + 32;
    static constexpr int SLOT_COLS = (TPV / 128) * 32; // one parked spectrum: 32 columns per thread, 128 lanes
    // K == 2: the filter's two partition spectra live in tensor memory too (a thread meets the same 16 bins in every
    // block: 2 x 16 complex = 64 more columns of its lane), so the block loop reads nothing from L2
    static __host__ __device__ constexpr int h_cols(int K) { return K == 2 ? 64 : 0; }
    static __host__ __device__ constexpr int group_cols(int K) { return 32 * (K - 1) + h_cols(K); } // per 4 warps
    static __host__ __device__ constexpr int tmem_cols(int K) // power of two >= 32 (tcgen05.alloc)
    {
        int c = 32;
        while (c < (TPV / 128) * group_cols(K))
            c *= 2;
        return c;
    }
};

template <int L, int K>
__global__ void __launch_bounds__(L / 16, K == 2 ? 0 : 8192 / L) // K > 2: hold ptxas to 128 registers (two CTAs per SM)
fftfilt_upols_kernel(const float2* __restrict__ in, float2* __restrict__ out, long long nout, long long nin,
                     int ntaps, int D, const float2* __restrict__ H, const float2* __restrict__ tw, long long run, int gs)
{
    // `in` is 16-byte aligned; the stream's first item is in[gs] (gs = 1 when the caller's pointer was only 8-byte aligned;
    // nin counts from in[0]); items in front of it read as zero like the ones in front of the stream
    using G = UpolsGeom<L>;
    constexpr int E = 16, TPV = G::TPV, V = G::V, SLICE = G::SLICE;
    constexpr int R0 = fft_first_radix(L, false);
    constexpr int RM = fft_last_radix(L, false); // == first radix of the reversed schedule
    constexpr int RL = fft_last_radix(L, true);
    static_assert(R0 >= 2 && (L / R0) % V == 0 || V % (L / R0) == 0, "first-pass stride must not straddle the slices");
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float2* XIN = reinterpret_cast<float2*>(smem_raw);
    float2* S = reinterpret_cast<float2*>(smem_raw + G::XIN_BYTES);
    float2* TWF = reinterpret_cast<float2*>(smem_raw + G::XIN_BYTES + G::S_BYTES);
    float2* TWR = reinterpret_cast<float2*>(smem_raw + G::XIN_BYTES + G::S_BYTES + G::TWF_BYTES);
    unsigned long long* bar =
        reinterpret_cast<unsigned long long*>(smem_raw + G::XIN_BYTES + G::S_BYTES + G::TWF_BYTES + G::TWR_BYTES);
    unsigned* tmem_slot = reinterpret_cast<unsigned*>(bar + 2);
    const int t = threadIdx.x, warp = t >> 5;
    const long long ny = nout * D;
    const long long nblocks = (ny + V - 1) / V;
    const long long nin_even = nin & ~1LL; // items the bulk copies may touch (the odd last one is read directly)
    const int c0 = ntaps - 1 + gs, off = c0 & 1;
    const long long b0 = blockIdx.x * run, b1 = (b0 + run < nblocks) ? b0 + run : nblocks;

    // slice j = buffer items [(j-1)V + c0, jV + c0); block b transforms slices b (old half) and b+1 (new half)
    auto issue = [&](long long j) {
        const int h = (int)(j & 1);
        const long long lo = (j - 1) * V + c0 - off; // even; negative in front of the stream (zeros)
        const long long lo_c = lo < 0 ? 0 : lo;
        long long hi_c = lo + SLICE;
        hi_c = hi_c > nin_even ? nin_even : hi_c;
        const long long n = hi_c > lo_c ? hi_c - lo_c : 0;
        mbar_expect_tx(&bar[h], (unsigned)(n * 8));
        if (n)
            bulk_g2s(XIN + h * SLICE + (lo_c - lo), in + lo_c, (unsigned)(n * 8), &bar[h]);
    };

    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                     "r"(G::tmem_cols(K)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    if (t == 0) {
        mbar_init(&bar[0], 1);
        mbar_init(&bar[1], 1);
    }
    cta_fft_tws_init<L, true, false>(TWF, t, tw);
    cta_fft_tws_init<L, false, true>(TWR, t, tw);
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const unsigned tmem = *tmem_slot;
    // this thread's parking space: its own lane (sub-partition = warp % 4); per group of four warps K-1 slots of 32
    // columns, slot (it mod K-1) holding the spectrum of the block of iteration `it`
    constexpr int nslot = K - 1;
    const unsigned taddr0 = tmem + ((unsigned)((warp & 3) * 32) << 16) + (unsigned)((warp >> 2) * G::group_cols(K));
    const unsigned thaddr = taddr0 + (unsigned)(nslot * 32); // K == 2: this thread's H0 | H1 values, chunk c at +16c
    if constexpr (K == 2) {
#pragma unroll
        for (int c = 0; c < 4; c++) {
            float2 a[4], b[4];
#pragma unroll
            for (int e = 0; e < 4; e++) {
                const int i = 4 * c + e;
                const int k = (t + (i / RM) * TPV) + (i % RM) * (L / RM);
                a[e] = __ldg(H + k);
                b[e] = __ldg(H + L + k);
            }
            asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, "
                         "%13, %14, %15, %16};" ::"r"(thaddr + c * 16),
                         "f"(a[0].x), "f"(a[0].y), "f"(b[0].x), "f"(b[0].y), "f"(a[1].x), "f"(a[1].y), "f"(b[1].x),
                         "f"(b[1].y), "f"(a[2].x), "f"(a[2].y), "f"(b[2].x), "f"(b[2].y), "f"(a[3].x), "f"(a[3].y),
                         "f"(b[3].x), "f"(b[3].y)
                         : "memory");
        }
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    }

    if (b0 < b1) {
        const long long bw = b0 - nslot; // first warm-up block: K-1 forward transforms fill the delay line
        if (t == 0) {
            issue(bw);
            issue(bw + 1);
        }
        unsigned ph = 0; // bit h = parity the next wait on bar[h] expects
        int it = 0;
        for (long long blk = bw; blk < b1; blk++, it++) {
            const int hnew = (int)((blk + 1) & 1);
            const unsigned tnew = taddr0 + (unsigned)((it % nslot) * 32);
            if (blk == bw) {
                mbar_wait(&bar[hnew ^ 1], (ph >> (hnew ^ 1)) & 1);
                ph ^= 1u << (hnew ^ 1);
            }
            mbar_wait(&bar[hnew], (ph >> hnew) & 1);
            ph ^= 1u << hnew;
            const long long s = (blk - 1) * V + c0; // buffer index of the block's first item
            float2 v[E];
#pragma unroll
            for (int b = 0; b < E / R0; b++)
#pragma unroll
                for (int q = 0; q < R0; q++) {
                    const int m = (t + b * TPV) + q * (L / R0);
                    const long long g = s + m;
                    const int half = m / V, idx = m % V;
                    float2 x = make_float2(0.f, 0.f);
                    if (g >= gs && g < nin_even)
                        x = XIN[((int)((blk + half) & 1)) * SLICE + idx + off];
                    else if (g >= gs && g < nin)
                        x = in[g];
                    v[b * R0 + q] = x;
                }
            cta_fft_tws<L, true, false>(v, S, t, TWF, tw, [&]() {
                if (t == 0 && blk + 1 < b1)
                    issue(blk + 2); // into the half that held slice blk: every thread has read it by now
            });
            if (blk < b0) { // warm-up: only park the spectrum
#pragma unroll
                for (int c = 0; c < 4; c++)
                    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(
                                     tnew + c * 8),
                                 "f"(v[4 * c].x), "f"(v[4 * c].y), "f"(v[4 * c + 1].x), "f"(v[4 * c + 1].y),
                                 "f"(v[4 * c + 2].x), "f"(v[4 * c + 2].y), "f"(v[4 * c + 3].x), "f"(v[4 * c + 3].y)
                                 : "memory");
                asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
                __syncthreads(); // the transform's last gather is done before the next block reuses S
                continue;
            }
            if constexpr (K == 2) {
                // v = X_b . H0 + X_{b-1} . H1; X_b takes X_{b-1}'s place in tensor memory.  Four chunks of four values,
                // software-pipelined: chunk c+1's TMEM read and its eight H values (L2) are in flight while chunk c is
                // multiplied.  (The tcgen05 statements are volatile asm, so they keep their order among themselves; the
                // wait lists the loaded registers as in/out operands so that no use of them can move above it.)
                float2 xp[2][4], hh[2][8];
                auto fetch = [&](int c) {
                    float2* x = xp[c & 1];
                    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                                 : "=f"(x[0].x), "=f"(x[0].y), "=f"(x[1].x), "=f"(x[1].y), "=f"(x[2].x), "=f"(x[2].y),
                                   "=f"(x[3].x), "=f"(x[3].y)
                                 : "r"(tnew + c * 8));
                    float2* h = hh[c & 1];
                    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, "
                                 "%12, %13, %14, %15}, [%16];"
                                 : "=f"(h[0].x), "=f"(h[0].y), "=f"(h[1].x), "=f"(h[1].y), "=f"(h[2].x), "=f"(h[2].y),
                                   "=f"(h[3].x), "=f"(h[3].y), "=f"(h[4].x), "=f"(h[4].y), "=f"(h[5].x), "=f"(h[5].y),
                                   "=f"(h[6].x), "=f"(h[6].y), "=f"(h[7].x), "=f"(h[7].y)
                                 : "r"(thaddr + c * 16));
                };
                fetch(0);
    #pragma unroll
                for (int c = 0; c < 4; c++) {
                    if (c < 3)
                        fetch(c + 1);
                    float2* x = xp[c & 1];
                    float2* y = xp[(c + 1) & 1]; // listed too: the wait also completes the read just issued
                    asm volatile("tcgen05.wait::ld.sync.aligned;"
                                 : "+f"(x[0].x), "+f"(x[0].y), "+f"(x[1].x), "+f"(x[1].y), "+f"(x[2].x), "+f"(x[2].y),
                                   "+f"(x[3].x), "+f"(x[3].y), "+f"(y[0].x), "+f"(y[0].y), "+f"(y[1].x), "+f"(y[1].y),
                                   "+f"(y[2].x), "+f"(y[2].y), "+f"(y[3].x), "+f"(y[3].y));
                    {
                        // same wait, restated for the H registers of both buffers (a compiler fence for their uses)
                        float2* h = hh[c & 1];
                        float2* g = hh[(c + 1) & 1];
                        asm volatile("tcgen05.wait::ld.sync.aligned;"
                                     : "+f"(h[0].x), "+f"(h[0].y), "+f"(h[1].x), "+f"(h[1].y), "+f"(h[2].x), "+f"(h[2].y),
                                       "+f"(h[3].x), "+f"(h[3].y), "+f"(h[4].x), "+f"(h[4].y), "+f"(h[5].x), "+f"(h[5].y),
                                       "+f"(h[6].x), "+f"(h[6].y), "+f"(h[7].x), "+f"(h[7].y));
                        asm volatile("tcgen05.wait::ld.sync.aligned;"
                                     : "+f"(g[0].x), "+f"(g[0].y), "+f"(g[1].x), "+f"(g[1].y), "+f"(g[2].x), "+f"(g[2].y),
                                       "+f"(g[3].x), "+f"(g[3].y), "+f"(g[4].x), "+f"(g[4].y), "+f"(g[5].x), "+f"(g[5].y),
                                       "+f"(g[6].x), "+f"(g[6].y), "+f"(g[7].x), "+f"(g[7].y));
                    }
                    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(
                                     tnew + c * 8),
                                 "f"(v[4 * c].x), "f"(v[4 * c].y), "f"(v[4 * c + 1].x), "f"(v[4 * c + 1].y),
                                 "f"(v[4 * c + 2].x), "f"(v[4 * c + 2].y), "f"(v[4 * c + 3].x), "f"(v[4 * c + 3].y));
    #pragma unroll
                    for (int e = 0; e < 4; e++) {
                        const int i = 4 * c + e;
                        const float2 h0 = hh[c & 1][2 * e], h1 = hh[c & 1][2 * e + 1];
                        const float2 a = v[i], p = x[e];
                        v[i] = make_float2(fmaf(a.x, h0.x, fmaf(-a.y, h0.y, fmaf(p.x, h1.x, -p.y * h1.y))),
                                           fmaf(a.x, h0.y, fmaf(a.y, h0.x, fmaf(p.x, h1.y, p.y * h1.x))));
                    }
                }
            } else {
                // v = X_b . H_0 + sum_p X_{b-p} . H_p, p = 1..K-1; X_b then takes the oldest spectrum's slot.  Four chunks
                // of four values x (K-1) parked spectra = 4(K-1) steps, software-pipelined: step q+1's TMEM read and its H
                // values (L2) are in flight while step q is multiplied.  (The tcgen05 statements are volatile asm and keep
                // their order among themselves; the wait lists the loaded registers as in/out operands so that no use of
                // them can move above it.)
                constexpr int NQ = 4 * nslot;
                const int sbase = it % nslot;
                float2 xp[2][4], hp[2][4], h0[2][4];
                auto fetch = [&](int q) {
                    const int c = q / nslot, p = q % nslot + 1, bf = q & 1;
                    const int sl = (sbase + nslot - (p % nslot)) % nslot; // slot of iteration it - p
                    float2* x = xp[bf];
                    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                                 : "=f"(x[0].x), "=f"(x[0].y), "=f"(x[1].x), "=f"(x[1].y), "=f"(x[2].x), "=f"(x[2].y),
                                   "=f"(x[3].x), "=f"(x[3].y)
                                 : "r"(taddr0 + (unsigned)(sl * 32 + c * 8)));
    #pragma unroll
                    for (int e = 0; e < 4; e++) {
                        const int i = 4 * c + e;
                        const int k = (t + (i / RM) * TPV) + (i % RM) * (L / RM);
                        hp[bf][e] = __ldg(H + (long long)p * L + k);
                        if (p == 1)
                            h0[bf][e] = __ldg(H + k);
                    }
                };
                fetch(0);
                float2 acc[4];
    #pragma unroll
                for (int q = 0; q < NQ; q++) {
                    const int c = q / nslot, p = q % nslot + 1, bf = q & 1;
                    if (q + 1 < NQ)
                        fetch(q + 1);
                    float2* x = xp[bf];
                    float2* y = xp[bf ^ 1]; // listed too: the wait also completes the read just issued
                    asm volatile("tcgen05.wait::ld.sync.aligned;"
                                 : "+f"(x[0].x), "+f"(x[0].y), "+f"(x[1].x), "+f"(x[1].y), "+f"(x[2].x), "+f"(x[2].y),
                                   "+f"(x[3].x), "+f"(x[3].y), "+f"(y[0].x), "+f"(y[0].y), "+f"(y[1].x), "+f"(y[1].y),
                                   "+f"(y[2].x), "+f"(y[2].y), "+f"(y[3].x), "+f"(y[3].y));
    #pragma unroll
                    for (int e = 0; e < 4; e++) {
                        if (p == 1)
                            acc[e] = cmul(v[4 * c + e], h0[bf][e]);
                        acc[e].x = fmaf(x[e].x, hp[bf][e].x, fmaf(-x[e].y, hp[bf][e].y, acc[e].x));
                        acc[e].y = fmaf(x[e].x, hp[bf][e].y, fmaf(x[e].y, hp[bf][e].x, acc[e].y));
                    }
                    if (p == nslot) {
                        // slot it mod K-1 held X_{b-(K-1)}, read just above
                        asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(
                                         tnew + c * 8),
                                     "f"(v[4 * c].x), "f"(v[4 * c].y), "f"(v[4 * c + 1].x), "f"(v[4 * c + 1].y),
                                     "f"(v[4 * c + 2].x), "f"(v[4 * c + 2].y), "f"(v[4 * c + 3].x), "f"(v[4 * c + 3].y));
    #pragma unroll
                        for (int e = 0; e < 4; e++)
                            v[4 * c + e] = acc[e];
                    }
                }
            }
            asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
            __syncthreads(); // forward transform's last gather is done before the inverse reuses S
            cta_fft_tws<L, false, true>(v, S, t, TWR, tw, []() {});
            if (D == 1) {
#pragma unroll
                for (int b = 0; b < E / RL; b++)
#pragma unroll
                    for (int q = 0; q < RL; q++) {
                        const int m = (t + b * TPV) + q * (L / RL);
                        const long long n = blk * V + (m - V);
                        if (m >= V && n < ny)
                            out[n] = v[b * RL + q];
                    }
            } else {
                const long long dec_q = (blk * V) / D;
                const unsigned dec_r = (unsigned)(blk * V - dec_q * D);
#pragma unroll
                for (int b = 0; b < E / RL; b++)
#pragma unroll
                    for (int q = 0; q < RL; q++) {
                        const int m = (t + b * TPV) + q * (L / RL);
                        if (m >= V && blk * V + (m - V) < ny)
                            store_decimated(out, dec_q, dec_r, (unsigned)D, (unsigned)(m - V), v[b * RL + q]);
                    }
            }
            __syncthreads();
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (warp == 0)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(G::tmem_cols(K)));
}

template <int L, int K>
static int launch_upols(const void* in, void* out, long long nout, long long nin, int ntaps, int D,
                        const float2* H, const float2* tw, int nsm, cudaStream_t st)
{
    using G = UpolsGeom<L>;
    auto kern = fftfilt_upols_kernel<L, K>;
    GRB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, G::SMEM));
    int occ = 1;
    GRB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, L / 16, G::SMEM));
    if (occ < 1)
        occ = 1;
    // The occupancy query answers 1 for a kernel that allocates tensor memory, whatever it asks for; the hardware
    // co-schedules as many CTAs as registers, shared memory and the 512 TMEM columns allow (checked with ncu).
    cudaFuncAttributes fa;
    GRB_CUDA(cudaFuncGetAttributes(&fa, kern));
    int by_regs = 65536 / (fa.numRegs * (L / 16)), by_smem = (227 * 1024) / (G::SMEM + 1024), by_tmem = 512 / G::tmem_cols(K);
    int fit = by_regs < by_smem ? by_regs : by_smem;
    fit = fit < by_tmem ? fit : by_tmem;
    if (const char* e = GRB_GETENV_ONCE("GRB200_FFTFILT_UPOLS_OCC"))
        fit = atoi(e);
    if (fit > occ)
        occ = fit;
    const long long nblocks = (nout * D + G::V - 1) / G::V;
    long long grid = (long long)nsm * occ;
    if (grid > nblocks)
        grid = nblocks;
    const long long run = (nblocks + grid - 1) / grid;
    grid = (nblocks + run - 1) / run;
    const int gs = (int)((((uintptr_t)in) >> 3) & 1); // 8-byte aligned only: start the stream one item into an aligned one
    kern<<<(unsigned)grid, L / 16, G::SMEM, st>>>((const float2*)in - gs, (float2*)out, nout, nin + gs, ntaps, D, H, tw,
                                                  run, gs);
    GRB_LAUNCH_CHECK();
    return 0;
}

// ---- two partitions, three blocks in flight per SM (`fftfilt_upols2_kernel`, 2050..4096 taps: cfg2) -----------------
// The generic kernel above sits at 2 CTAs x 8 warps per SM (128 registers, 100 KB of shared memory per CTA) and its FMA pipe
// idles half of the time: four warps per scheduler do not cover the exchanges between the radix passes.  This variant is
// the same algorithm cut down so that THREE blocks are in flight per SM, as three independent 256-thread groups of ONE
// 768-thread CTA (named barriers 1..3 instead of __syncthreads): one CTA owns all 512 tensor-memory columns, so H0 and H1
// (a thread meets the same 16 bins in every block, and so do its two namesakes in the other groups) are written to tensor
// memory once and shared by the three groups -- the block loop reads nothing from L2 -- next to the three parked spectra
// (128 + 3 x 64 = 320 columns).  4096 = 16 x 16 x 16 written out; the twiddles of the middle pass come from a 16 x 15
// table (no products: 11 complex multiplies fewer per transform), those of the last pass from three values per thread
// (w, w^2, w^4; w^8 = (w^4)^2); one table for both directions (the inverse conjugates in registers), shared by the groups:
// 3 x 66 + 8 KB of shared memory, 80 registers.
struct Upols2 {
    static constexpr int L = 4096, TPV = 256, V = 2048, SLICE = V + 2, NG = 3;
    static constexpr int XIN_BYTES = 2 * SLICE * 8;
    static constexpr int S_BYTES = ((fft_smem_elems(L) * 8 + 15) / 16) * 16;
    static constexpr int T1_STRIDE = 18; // float2 per table row (144 bytes: 16-byte loads of 8 rows hit 8 x 4 distinct banks)
    static constexpr int T1_BYTES = 16 * T1_STRIDE * 8;
    static constexpr int TW2_BYTES = 3 * TPV * 8;
    static constexpr int SMEM = NG * (XIN_BYTES + S_BYTES) + T1_BYTES + TW2_BYTES + NG * 16 + 16;
    static constexpr int TMEM_COLS = 512; // [0,64) H0, [64,128) H1, [128 + 64 g, +64) group g's parked spectrum
};

template <bool FWD>
__device__ __forceinline__ float2 cmul_dir(float2 a, float2 w) // a * w forward, a * conj(w) inverse
{
    if constexpr (FWD)
        return cmul(a, w);
    else {
        const float2 t = __fmul2_rn(make_float2(a.y, a.y), make_float2(w.y, w.x)); // (a.y w.y, a.y w.x)
        return __ffma2_rn(make_float2(a.x, a.x), make_float2(w.x, -w.y), t);
    }
}

__device__ __forceinline__ void group_sync(int id) { asm volatile("bar.sync %0, 256;" ::"r"(id) : "memory"); }

// The two twiddle tables of fft4096_3x16, written by one 256-thread group (thread t): T1[r][q] = W256^{q r} = W4096^{16 q r}
// (rows padded to T1_STRIDE), TW2[j][t] = W4096^{2^j t}, j = 0, 1, 2.  tw: the 4096 roots.
__device__ __forceinline__ void fft4096_3x16_tables(float2* T1, float2* TW2, const float2* __restrict__ tw, int t)
{
    const int r = t >> 4, q = t & 15;
    T1[r * Upols2::T1_STRIDE + q] = tw[(16 * q * r) & (Upols2::L - 1)];
    if (q < 2)
        T1[r * Upols2::T1_STRIDE + 16 + q] = make_float2(0.f, 0.f);
    TW2[t] = tw[t];
    TW2[Upols2::TPV + t] = tw[2 * t];
    TW2[2 * Upols2::TPV + t] = tw[4 * t];
}

// Register convention as cta_fft: in v[q] = x[t + 256 q], out v[q] = X[t + 256 q]; `bid` = the group's barrier.
template <bool FWD, class After>
__device__ __forceinline__ void fft4096_3x16(float2 (&v)[16], float2* S, int t, int bid, const float2* T1,
                                             const float2* TW2, After after_first)
{
    // one base address per access pattern, constant offsets (fft_pad(i) = i + i / 16 is linear along each of them; left to
    // itself the compiler forms 16 addresses per gather, keeps them across the block loop and spills them)
    float2* Sg = S + t + (t >> 4);                                     // gather:  fft_pad(t + 256 q) = Sg + 272 q
    float2* Sa = S + 17 * t;                                           // scatter 0: fft_pad(16 t + q) = Sa + q
    float2* Sb = S + fft_pad((t >> 4) * 256 + (t & 15));               // scatter 1: fft_pad(base + 16 q) = Sb + 17 q
    dft16<FWD>(v);
    {
        // The middle pass's twiddles, applied here: output q of this thread becomes input t >> 4 of thread (16 t + q) mod
        // 256, whose twiddle W256^{(t >> 4) ((16 t + q) mod 16)} = W256^{q (t >> 4)} -- row t >> 4 of the table, two rows
        // per warp (a 16-byte load is one wavefront; by the reading thread's t mod 16 it would be four).  (volatile asm:
        // loop invariant, and ptxas would otherwise keep the twiddle registers live across the block loop and spill them)
        const unsigned row = smem_u32(T1 + (t >> 4) * Upols2::T1_STRIDE);
#pragma unroll
        for (int p = 0; p < 8; p++) {
            float4 w;
            asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                         : "=f"(w.x), "=f"(w.y), "=f"(w.z), "=f"(w.w)
                         : "r"(row + 16 * p));
            if (p > 0)
                v[2 * p] = cmul_dir<FWD>(v[2 * p], make_float2(w.x, w.y));
            v[2 * p + 1] = cmul_dir<FWD>(v[2 * p + 1], make_float2(w.z, w.w));
        }
    }
#pragma unroll
    for (int q = 0; q < 16; q++)
        Sa[q] = v[q];
    group_sync(bid);
    after_first();
#pragma unroll
    for (int q = 0; q < 16; q++)
        v[q] = Sg[272 * q];
    dft16<FWD>(v);
    group_sync(bid);
#pragma unroll
    for (int q = 0; q < 16; q++)
        Sb[17 * q] = v[q];
    group_sync(bid);
#pragma unroll
    for (int q = 0; q < 16; q++)
        v[q] = Sg[272 * q];
    {
        float2 w1, w2, w4; // W4096^{t, 2t, 4t}
        const unsigned a = smem_u32(TW2 + t);
        asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(w1.x), "=f"(w1.y) : "r"(a));
        asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(w2.x), "=f"(w2.y) : "r"(a + Upols2::TPV * 8));
        asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(w4.x), "=f"(w4.y) : "r"(a + 2 * Upols2::TPV * 8));
        if constexpr (!FWD) {
            w1 = cconj(w1);
            w2 = cconj(w2);
            w4 = cconj(w4);
        }
        const float2 w8 = cmul(w4, w4);
        apply_twiddles_l<16>(v, [&](int k) { return k == 0 ? w1 : (k == 1 ? w2 : (k == 2 ? w4 : w8)); });
    }
    dft16<FWD>(v);
}

__global__ void __launch_bounds__(768, 1)
fftfilt_upols2_kernel(const float2* __restrict__ in, float2* __restrict__ out, long long nout, long long nin, int ntaps,
                      int D, const float2* __restrict__ H, const float2* __restrict__ tw, long long run, int gs)
{
    using G = Upols2;
    constexpr int L = G::L, TPV = G::TPV, V = G::V, SLICE = G::SLICE, NG = G::NG;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int grp = threadIdx.x >> 8, t = threadIdx.x & 255, warp = t >> 5, bid = grp + 1;
    float2* XIN = reinterpret_cast<float2*>(smem_raw + grp * (G::XIN_BYTES + G::S_BYTES));
    float2* S = reinterpret_cast<float2*>(smem_raw + grp * (G::XIN_BYTES + G::S_BYTES) + G::XIN_BYTES);
    float2* T1 = reinterpret_cast<float2*>(smem_raw + NG * (G::XIN_BYTES + G::S_BYTES));
    float2* TW2 = reinterpret_cast<float2*>(smem_raw + NG * (G::XIN_BYTES + G::S_BYTES) + G::T1_BYTES);
    unsigned long long* bar0 =
        reinterpret_cast<unsigned long long*>(smem_raw + NG * (G::XIN_BYTES + G::S_BYTES) + G::T1_BYTES + G::TW2_BYTES);
    unsigned long long* bar = bar0 + 2 * grp;
    unsigned* tmem_slot = reinterpret_cast<unsigned*>(bar0 + 2 * NG);
    const long long ny = nout * D;
    const long long nblocks = (ny + V - 1) / V;
    const long long nin_even = nin & ~1LL;
    const int c0 = ntaps - 1 + gs, off = c0 & 1;
    const long long b0 = ((long long)blockIdx.x * NG + grp) * run;
    const long long b1 = (b0 + run < nblocks) ? b0 + run : nblocks;

    auto issue = [&](long long j) { // slice j = buffer items [(j-1)V + c0, jV + c0), as in the generic kernel
        const int h = (int)(j & 1);
        const long long lo = (j - 1) * V + c0 - off;
        const long long lo_c = lo < 0 ? 0 : lo;
        long long hi_c = lo + SLICE;
        hi_c = hi_c > nin_even ? nin_even : hi_c;
        const long long n = hi_c > lo_c ? hi_c - lo_c : 0;
        mbar_expect_tx(&bar[h], (unsigned)(n * 8));
        if (n)
            bulk_g2s(XIN + h * SLICE + (lo_c - lo), in + lo_c, (unsigned)(n * 8), &bar[h]);
    };

    if (threadIdx.x < 32) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                     "r"(G::TMEM_COLS));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    if (t == 0) {
        mbar_init(&bar[0], 1);
        mbar_init(&bar[1], 1);
    }
    if (grp == 0)
        fft4096_3x16_tables(T1, TW2, tw, t);
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const unsigned tmem = *tmem_slot;
    const unsigned tlane = tmem + ((unsigned)((warp & 3) * 32) << 16) + (unsigned)((warp >> 2) * 32);
    const unsigned th0 = tlane, th1 = tlane + 64, tpark = tlane + 128 + (unsigned)(grp * 64);
    if (grp == 0) { // bins k = t + 256 i, i = 0..15, of both partitions: column 2 i of this thread's lane
#pragma unroll
        for (int p = 0; p < 2; p++)
#pragma unroll
            for (int c = 0; c < 2; c++) {
                float2 a[8];
#pragma unroll
                for (int e = 0; e < 8; e++)
                    a[e] = __ldg(H + p * L + t + 256 * (8 * c + e));
                asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, "
                             "%12, %13, %14, %15, %16};" ::"r"((p ? th1 : th0) + c * 16),
                             "f"(a[0].x), "f"(a[0].y), "f"(a[1].x), "f"(a[1].y), "f"(a[2].x), "f"(a[2].y), "f"(a[3].x),
                             "f"(a[3].y), "f"(a[4].x), "f"(a[4].y), "f"(a[5].x), "f"(a[5].y), "f"(a[6].x), "f"(a[6].y),
                             "f"(a[7].x), "f"(a[7].y)
                             : "memory");
            }
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads(); // H0 | H1 are in place for the other groups
    asm volatile("tcgen05.fence::after_thread_sync;");

    if (b0 < b1) {
        // loop state in 32 bits, relative to the run's first block (64-bit stream positions only on the edge paths):
        // iteration i = -1 (warm-up: one forward transform fills the delay line) .. nrun - 1 is block b0 + i
        const int nrun = (int)(b1 - b0);
        const long long s0 = (b0 - 1) * V + c0; // buffer index of block b0's first item; block b0 + i starts at s0 + i V
        // interior iterations: every item of the block came with the bulk copies (s >= gs and s + L <= nin_even)
        long long lo64 = s0 >= gs ? 0 : (gs - s0 + V - 1) / V, hi64 = (nin_even - L - s0) >= 0 ? (nin_even - L - s0) / V + 1 : 0;
        const int i_lo = (int)(lo64 > nrun ? nrun : lo64), i_hi = (int)(hi64 > nrun ? nrun : hi64);
        // iterations whose 2048 outputs all exist: (b0 + i + 1) V <= ny
        long long full64 = ny / V - b0;
        const int i_full = D == 1 ? (int)(full64 > nrun ? nrun : (full64 < 0 ? 0 : full64)) : 0;
        float2* outp = out + b0 * V + t; // block b0 + i's outputs: outp[i V + 256 (q - 8)]
        if (t == 0) {
            issue(b0 - 1);
            issue(b0);
        }
        unsigned ph = 0;
        for (int i = -1; i < nrun; i++) {
            const int hnew = (int)((b0 + i + 1) & 1);
            if (i == -1) {
                mbar_wait(&bar[hnew ^ 1], (ph >> (hnew ^ 1)) & 1);
                ph ^= 1u << (hnew ^ 1);
            }
            mbar_wait(&bar[hnew], (ph >> hnew) & 1);
            ph ^= 1u << hnew;
            float2 v[16];
            const float2* xo = XIN + (hnew ^ 1) * SLICE + off + t; // old half: m = t + 256 q, q < 8
            const float2* xn = XIN + hnew * SLICE + off + t;       // new half: q >= 8
            if (i >= i_lo && i < i_hi) {
#pragma unroll
                for (int q = 0; q < 8; q++) {
                    v[q] = xo[256 * q];
                    v[q + 8] = xn[256 * q];
                }
            } else {
                const long long s = s0 + (long long)i * V;
#pragma unroll
                for (int q = 0; q < 16; q++) {
                    const long long g = s + t + 256 * q;
                    float2 x = make_float2(0.f, 0.f);
                    if (g >= gs && g < nin_even)
                        x = (q < 8 ? xo : xn)[256 * (q & 7)];
                    else if (g >= gs && g < nin)
                        x = in[g];
                    v[q] = x;
                }
            }
            fft4096_3x16<true>(v, S, t, bid, T1, TW2, [&]() {
                if (t == 0 && i + 1 < nrun)
                    issue(b0 + i + 2); // into the half that held slice blk: every thread of the group has read it by now
            });
            if (i < 0) { // warm-up: only park the spectrum
#pragma unroll
                for (int c = 0; c < 4; c++)
                    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(
                                     tpark + c * 8),
                                 "f"(v[4 * c].x), "f"(v[4 * c].y), "f"(v[4 * c + 1].x), "f"(v[4 * c + 1].y),
                                 "f"(v[4 * c + 2].x), "f"(v[4 * c + 2].y), "f"(v[4 * c + 3].x), "f"(v[4 * c + 3].y)
                                 : "memory");
                asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
                group_sync(bid); // the transform's last gather is done before the next block reuses S
                continue;
            }
            // v = X_b . H0 + X_{b-1} . H1; X_b takes X_{b-1}'s place.  Four chunks of four bins, everything out of tensor
            // memory (three short reads, one wait; the other two groups of the CTA fill the gap -- a second register
            // buffer to overlap the reads spills at 80 registers).
#pragma unroll
            for (int c = 0; c < 4; c++) {
                float2 x[4], h0[4], h1[4];
                auto ld8 = [&](float2* a, unsigned addr) {
                    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                                 : "=f"(a[0].x), "=f"(a[0].y), "=f"(a[1].x), "=f"(a[1].y), "=f"(a[2].x), "=f"(a[2].y),
                                   "=f"(a[3].x), "=f"(a[3].y)
                                 : "r"(addr));
                };
                ld8(x, tpark + c * 8);
                ld8(h0, th0 + c * 8);
                ld8(h1, th1 + c * 8);
                // the wait lists every register the reads write, so no use of them can move above it
                asm volatile("tcgen05.wait::ld.sync.aligned;"
                             : "+f"(x[0].x), "+f"(x[0].y), "+f"(x[1].x), "+f"(x[1].y), "+f"(x[2].x), "+f"(x[2].y),
                               "+f"(x[3].x), "+f"(x[3].y), "+f"(h0[0].x), "+f"(h0[0].y), "+f"(h0[1].x), "+f"(h0[1].y),
                               "+f"(h0[2].x), "+f"(h0[2].y), "+f"(h0[3].x), "+f"(h0[3].y), "+f"(h1[0].x), "+f"(h1[0].y),
                               "+f"(h1[1].x), "+f"(h1[1].y), "+f"(h1[2].x), "+f"(h1[2].y), "+f"(h1[3].x), "+f"(h1[3].y));
                asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(
                                 tpark + c * 8),
                             "f"(v[4 * c].x), "f"(v[4 * c].y), "f"(v[4 * c + 1].x), "f"(v[4 * c + 1].y),
                             "f"(v[4 * c + 2].x), "f"(v[4 * c + 2].y), "f"(v[4 * c + 3].x), "f"(v[4 * c + 3].y));
#pragma unroll
                for (int e = 0; e < 4; e++) {
                    const float2 a = v[4 * c + e], p = x[e], g0 = h0[e], g1 = h1[e];
                    v[4 * c + e] = make_float2(fmaf(a.x, g0.x, fmaf(-a.y, g0.y, fmaf(p.x, g1.x, -p.y * g1.y))),
                                               fmaf(a.x, g0.y, fmaf(a.y, g0.x, fmaf(p.x, g1.y, p.y * g1.x))));
                }
            }
            asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
            group_sync(bid); // forward transform's last gather is done before the inverse reuses S
            fft4096_3x16<false>(v, S, t, bid, T1, TW2, []() {});
            if (i < i_full) { // outputs m = t + 256 q, q >= 8 (the alias-free half)
                float2* o = outp + (long long)i * V;
#pragma unroll
                for (int q = 8; q < 16; q++)
                    o[256 * (q - 8)] = v[q];
            } else {
                const long long n0 = (b0 + i) * V + t;
                const long long dec_q = D == 1 ? 0 : ((b0 + i) * V) / D;
                const unsigned dec_r = D == 1 ? 0u : (unsigned)((b0 + i) * V - dec_q * D);
#pragma unroll
                for (int q = 8; q < 16; q++) {
                    const long long n = n0 + 256 * (q - 8);
                    if (n < ny) {
                        if (D == 1)
                            out[n] = v[q];
                        else
                            store_decimated(out, dec_q, dec_r, (unsigned)D, (unsigned)(t + 256 * (q - 8)), v[q]);
                    }
                }
            }
            group_sync(bid);
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (threadIdx.x < 32)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(G::TMEM_COLS));
}

static int launch_upols2(const void* in, void* out, long long nout, long long nin, int ntaps, int D, const float2* H,
                         const float2* tw, int nsm, cudaStream_t st)
{
    using G = Upols2;
    auto kern = fftfilt_upols2_kernel;
    GRB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, G::SMEM));
    const long long nblocks = (nout * D + G::V - 1) / G::V;
    long long groups = (long long)nsm * G::NG; // one CTA per SM, NG independent block runs each
    if (groups > nblocks)
        groups = nblocks;
    const long long run = (nblocks + groups - 1) / groups;
    groups = (nblocks + run - 1) / run;
    const long long grid = (groups + G::NG - 1) / G::NG;
    const int gs = (int)((((uintptr_t)in) >> 3) & 1);
    kern<<<(unsigned)grid, G::NG * G::TPV, G::SMEM, st>>>((const float2*)in - gs, (float2*)out, nout, nin + gs, ntaps, D, H,
                                                          tw, run, gs);
    GRB_LAUNCH_CHECK();
    return 0;
}

// ---- plain overlap-save on the same three-group machinery (`fftfilt_ols3_kernel`, L = 4096, up to 2048 taps) ----------
// The streaming kernel above (256-thread CTAs, generic radix schedule, H from L2) reaches a third of the FP32 peak; the
// three-group CTA of the two-partition kernel reaches half.  This is that CTA without the delay line: every group owns a run
// of blocks, block b's L (+2) items arrive by one bulk copy while block b-1 is transformed, H sits in 64 tensor-memory
// columns shared by the groups.  RP = the real-pair form of fir_filter_fff (see fftfilt_ols_kernel): the two float streams
// arrive by two bulk copies (16-byte aligned starts: up to three floats early each) and leave as two float stores.
struct Ols3 {
    static constexpr int L = 4096, TPV = 256, NG = 3;
    static constexpr int XROW = L + 4;         // floats per real-pair row; (L + 2) complex items fit in the same bytes
    static constexpr int XIN_BYTES = XROW * 8;
    static constexpr int S_BYTES = Upols2::S_BYTES, T1_BYTES = Upols2::T1_BYTES, TW2_BYTES = Upols2::TW2_BYTES;
    static constexpr int SMEM = NG * (XIN_BYTES + S_BYTES) + T1_BYTES + TW2_BYTES + NG * 8 + 16;
    static constexpr int TMEM_COLS = 64;
};

// DEC = 2, 4 or 8 (= D, or D = 8 E with DEC = 8 and E = 2 .. 32 of the results dropped on store): only every DEC-th result of a block is wanted, item m = DEC i + r with r = (ntaps-1 - b V) mod DEC for block
// b.  Decimating in time is aliasing in frequency: with Z_r[k] = X[k] H[k] e^{+j 2 pi k r / L},
//     y[DEC i + r] = sum_{k' < L/DEC} ( sum_j Z_r[k' + j L/DEC] ) e^{+j 2 pi k' i / (L/DEC)},
// an inverse transform of L/DEC points instead of L.  The DEC rotated filter spectra H_r sit side by side in tensor memory (64
// columns each: all 512 at DEC = 8), so the rotation costs nothing per block; L/DEC = 256 Q with Q = 16 / DEC, so the fold adds
// a thread's own registers (bins t + 256 q: q, q + Q, q + 2 Q, ... alias); then a radix-Q step in registers, a twiddle, and Q
// 256-point inverse transforms done by the first 16 Q threads of the group (two radix-16 passes) while the rest wait -- the
// other two groups of the CTA use the pipes meanwhile.  Forward transform and loads as before.
template <bool RP, int DEC>
__global__ void __launch_bounds__(768, 1)
fftfilt_ols3_kernel(const float2* __restrict__ in, float2* __restrict__ out, long long nout, long long nin, int ntaps, int D,
                    const float2* __restrict__ H, const float2* __restrict__ tw, long long run, int gs, long long half)
{
    using G = Ols3;
    constexpr int NROT = DEC ? DEC : 1, TMEM_COLS = 64 * NROT;
    constexpr int L = G::L, TPV = G::TPV, NG = G::NG, XROW = G::XROW;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int grp = threadIdx.x >> 8, t = threadIdx.x & 255, warp = t >> 5, bid = grp + 1;
    float2* XIN = reinterpret_cast<float2*>(smem_raw + grp * (G::XIN_BYTES + G::S_BYTES));
    float* XF = reinterpret_cast<float*>(XIN);
    float2* S = reinterpret_cast<float2*>(smem_raw + grp * (G::XIN_BYTES + G::S_BYTES) + G::XIN_BYTES);
    float2* T1 = reinterpret_cast<float2*>(smem_raw + NG * (G::XIN_BYTES + G::S_BYTES));
    float2* TW2 = reinterpret_cast<float2*>(smem_raw + NG * (G::XIN_BYTES + G::S_BYTES) + G::T1_BYTES);
    unsigned long long* bar0 =
        reinterpret_cast<unsigned long long*>(smem_raw + NG * (G::XIN_BYTES + G::S_BYTES) + G::T1_BYTES + G::TW2_BYTES);
    unsigned long long* bar = bar0 + grp;
    unsigned* tmem_slot = reinterpret_cast<unsigned*>(bar0 + NG);
    const float* __restrict__ inf = reinterpret_cast<const float*>(in);
    float* __restrict__ outf = reinterpret_cast<float*>(out);
    const int nh = ntaps - 1, V = L - nh;
    const long long ny = (RP ? half : nout) * D;      // full-rate outputs (RP: of the first, longer half)
    const long long nyB = RP ? (nout - half) * D : 0; // RP: of the second half, whose stream starts hD floats in
    const long long hD = half * D;
    const long long nblocks = (ny + V - 1) / V;
    const long long nin_al = RP ? (nin & ~3LL) : (nin & ~1LL); // what the bulk copies may touch; the rest is read directly
    const long long b0 = ((long long)blockIdx.x * NG + grp) * run;
    const long long b1 = (b0 + run < nblocks) ? b0 + run : nblocks;

    auto issue = [&](long long blk) {
        if constexpr (RP) {
            const long long a0 = (blk * V) & ~3LL, c0 = (blk * V + hD) & ~3LL;
            long long na = nin_al - a0, nc = nin_al - c0;
            na = na < 0 ? 0 : (na > XROW ? XROW : na);
            nc = nc < 0 ? 0 : (nc > XROW ? XROW : nc);
            mbar_expect_tx(bar, (unsigned)((na + nc) * 4));
            if (na)
                bulk_g2s(XF, inf + a0, (unsigned)(na * 4), bar);
            if (nc)
                bulk_g2s(XF + XROW, inf + c0, (unsigned)(nc * 4), bar);
        } else {
            const long long e0 = (blk * V + gs) & ~1LL;
            long long n = nin_al - e0;
            n = n < 0 ? 0 : (n > L + 2 ? L + 2 : n);
            mbar_expect_tx(bar, (unsigned)(n * 8));
            if (n)
                bulk_g2s(XIN, in + e0, (unsigned)(n * 8), bar);
        }
    };

    if (threadIdx.x < 32) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                     "r"(TMEM_COLS));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    if (t == 0)
        mbar_init(bar, 1);
    if (grp == 0)
        fft4096_3x16_tables(T1, TW2, tw, t);
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const unsigned tmem = *tmem_slot;
    const unsigned th0 = tmem + ((unsigned)((warp & 3) * 32) << 16) + (unsigned)((warp >> 2) * 32);
    if (grp == 0) { // bins k = t + 256 i, i = 0..15: columns 2 i, 2 i + 1 of this thread's lane (+ 64 r: rotated by r items)
#pragma unroll 1
        for (int rc = 0; rc < 2 * NROT; rc++) {
            const int r = rc >> 1, c = rc & 1;
            float2 a[8];
#pragma unroll
            for (int e = 0; e < 8; e++) {
                const int k = t + 256 * (8 * c + e);
                a[e] = __ldg(H + k);
                if (DEC) // e^{+j 2 pi k r / L} = conj(W_L^{k r})
                    a[e] = cmul_dir<false>(a[e], __ldg(tw + ((k * r) & (L - 1))));
            }
            asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, "
                         "%12, %13, %14, %15, %16};" ::"r"(th0 + 64 * r + c * 16),
                         "f"(a[0].x), "f"(a[0].y), "f"(a[1].x), "f"(a[1].y), "f"(a[2].x), "f"(a[2].y), "f"(a[3].x),
                         "f"(a[3].y), "f"(a[4].x), "f"(a[4].y), "f"(a[5].x), "f"(a[5].y), "f"(a[6].x), "f"(a[6].y),
                         "f"(a[7].x), "f"(a[7].y)
                         : "memory");
        }
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads(); // H is in place for the other groups
    asm volatile("tcgen05.fence::after_thread_sync;");

    if (b0 < b1) {
        if (t == 0)
            issue(b0);
        unsigned ph = 0;
        for (long long blk = b0; blk < b1; blk++) {
            const long long base = blk * V + gs; // buffer index of the block's first input item
            mbar_wait(bar, ph);
            ph ^= 1;
            float2 v[16];
            if constexpr (RP) {
                const float* xa = XF + (int)(base & 3) + t;
                const float* xc = XF + XROW + (int)((base + hD) & 3) + t;
                if (base + hD + L <= nin_al) {
#pragma unroll
                    for (int q = 0; q < 16; q++)
                        v[q] = make_float2(xa[256 * q], xc[256 * q]);
                } else {
#pragma unroll
                    for (int q = 0; q < 16; q++) {
                        const long long ga = base + t + 256 * q, gc = ga + hD;
                        v[q].x = ga < nin_al ? xa[256 * q] : (ga < nin ? inf[ga] : 0.f);
                        v[q].y = gc < nin_al ? xc[256 * q] : (gc < nin ? inf[gc] : 0.f);
                    }
                }
            } else {
                const float2* x = XIN + (int)(base & 1) + t;
                if (base + L <= nin_al) {
#pragma unroll
                    for (int q = 0; q < 16; q++)
                        v[q] = x[256 * q];
                } else {
#pragma unroll
                    for (int q = 0; q < 16; q++) {
                        const long long g = base + t + 256 * q;
                        float2 c = make_float2(0.f, 0.f);
                        if (g < nin_al)
                            c = x[256 * q];
                        else if (g < nin)
                            c = in[g];
                        v[q] = c;
                    }
                }
            }
            fft4096_3x16<true>(v, S, t, bid, T1, TW2, [&]() {
                if (t == 0 && blk + 1 < b1)
                    issue(blk + 1); // every thread of the group has read the block's items by now
            });
            // DEC: this block keeps items m = DEC i + rb (output blk V + m - nh is a multiple of DEC)
            const int rb = DEC ? (int)((((long long)nh - blk * V) % NROT + NROT) % NROT) : 0;
            const unsigned thr = th0 + 64u * (unsigned)rb;
#pragma unroll
            for (int c = 0; c < 4; c++) {
                float2 h[4];
                asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                             : "=f"(h[0].x), "=f"(h[0].y), "=f"(h[1].x), "=f"(h[1].y), "=f"(h[2].x), "=f"(h[2].y),
                               "=f"(h[3].x), "=f"(h[3].y)
                             : "r"(thr + c * 8));
                asm volatile("tcgen05.wait::ld.sync.aligned;"
                             : "+f"(h[0].x), "+f"(h[0].y), "+f"(h[1].x), "+f"(h[1].y), "+f"(h[2].x), "+f"(h[2].y),
                               "+f"(h[3].x), "+f"(h[3].y));
#pragma unroll
                for (int e = 0; e < 4; e++)
                    v[4 * c + e] = cmul(v[4 * c + e], h[e]);
            }
            group_sync(bid); // forward transform's last gather is done before the inverse reuses S
            if constexpr (DEC != 0) {
                constexpr int Q = 16 / DEC;
                float2 f[Q];
#pragma unroll
                for (int qq = 0; qq < Q; qq++) {
                    f[qq] = v[qq];
#pragma unroll
                    for (int j = 1; j < DEC; j++)
                        f[qq] = __fadd2_rn(f[qq], v[qq + Q * j]);
                }
                // a_{i0}[t] = e^{+j 2 pi t i0 / (256 Q)} sum_q' f[q'] e^{+j 2 pi q' i0 / Q}; the twiddle is conj(W_L^{DEC t i0})
                float2 w4 = TW2[2 * TPV + t]; // W_L^{4 t}
                const float2 w8 = cmul(w4, w4);
                if constexpr (Q == 2) {
                    dft2<false>(f[0], f[1]);
                    f[1] = cmul_dir<false>(f[1], w8);
                } else if constexpr (Q == 4) {
                    dft4<false>(f[0], f[1], f[2], f[3]);
                    f[1] = cmul_dir<false>(f[1], w4);
                    f[2] = cmul_dir<false>(f[2], w8);
                    f[3] = cmul_dir<false>(f[3], cmul(w4, w8));
                } else { // Q == 8: conj(W_L^{2 t i0}), i0 = 1..7, from W^{2t}, W^{4t}, W^{8t}
                    dft8<false>(f);
                    const float2 w2 = TW2[TPV + t], w6 = cmul(w2, w4), w12 = cmul(w4, w8);
                    f[1] = cmul_dir<false>(f[1], w2);
                    f[2] = cmul_dir<false>(f[2], w4);
                    f[3] = cmul_dir<false>(f[3], w6);
                    f[4] = cmul_dir<false>(f[4], w8);
                    f[5] = cmul_dir<false>(f[5], cmul(w2, w8));
                    f[6] = cmul_dir<false>(f[6], w12);
                    f[7] = cmul_dir<false>(f[7], cmul(w2, w12));
                }
                constexpr int P2 = 272 + DEC; // rows of S2 start DEC float2-banks apart: pass B's half-warps read without conflicts
                float2* S1 = S;        // [Q][256]
                float2* S2 = S + 2048; // [Q][16 x 17 (+ DEC)]
#pragma unroll
                for (int i0 = 0; i0 < Q; i0++)
                    S1[i0 * 256 + t] = f[i0];
                group_sync(bid);
                const bool act = t < 16 * Q;
                const int i0 = t >> 4, c = t & 15; // pass A: thread (i0, c) sums over h, t' = c + 16 h
                const int j0 = t % Q, g = t / Q;   // pass B: thread (j0, g) sums over c: its results are i = t + 16 Q d
                float2 x[16];
                if (act) {
#pragma unroll
                    for (int h = 0; h < 16; h++)
                        x[h] = S1[i0 * 256 + c + 16 * h];
                    dft16<false>(x);
                    const float2* trow = T1 + c * Upols2::T1_STRIDE; // W256^{c g}
#pragma unroll
                    for (int gg = 1; gg < 16; gg++)
                        x[gg] = cmul_dir<false>(x[gg], trow[gg]);
#pragma unroll
                    for (int gg = 0; gg < 16; gg++)
                        S2[i0 * P2 + gg * 17 + c] = x[gg];
                }
                group_sync(bid);
                if (act) {
#pragma unroll
                    for (int cc = 0; cc < 16; cc++)
                        x[cc] = S2[j0 * P2 + g * 17 + cc];
                    dft16<false>(x);
                    // x[d] = result i = j0 + Q (g + 16 d) = t + 16 Q d of the block = item m = DEC i + rb.  D = DEC: decimated
                    // output ob + i.  D = DEC E (DEC = 8, E = 2 .. 32): every E-th of them is wanted, i = ifirst + E j -- a thread
                    // keeps all of its results or none (16 Q = 32 is a multiple of E) -- and is output ob + j.
                    // (E not a power of two -- decimation 10 = 2 x 5, 20 = 4 x 5, 24 = 8 x 3 ..: one 32-bit division per result)
                    const long long nb = (long long)nh - blk * V;
                    const long long lim = RP ? half : nout, limB = RP ? nout - half : 0;
                    auto put = [&](long long o, float2 y) {
                        if constexpr (RP) {
                            outf[o] = y.x;
                            if (o < limB)
                                outf[half + o] = y.y;
                        } else {
                            out[o] = y;
                        }
                    };
                    if (D == DEC) { // the plain case, kept apart: this warp is the tail of the block, every instruction counts
                        const long long ob = (rb - nb) / DEC; // = (blk V + rb - nh) / DEC, exact
#pragma unroll
                        for (int d = 0; d < 16; d++) {
                            const int i = t + 16 * Q * d;
                            if (DEC * i + rb >= nh && ob + i < lim)
                                put(ob + i, x[d]);
                        }
                    } else {
                        const int E = D / DEC, lgE = 31 - __clz(E);
                        const bool p2 = (E & (E - 1)) == 0;
                        const int rD = p2 ? (int)(nb & (D - 1)) : (int)((nb % D + D) % D);
                        const int ifirst = (rD - rb) / DEC;
                        const long long num = rD - nb; // = blk V + rD - nh, a multiple of D
                        const long long ob = p2 ? (num >> (lgE + (DEC == 8 ? 3 : (DEC == 4 ? 2 : 1)))) : num / D;
                        if (p2) {
                            if (((t - ifirst) & (E - 1)) == 0) {
#pragma unroll
                                for (int d = 0; d < 16; d++) {
                                    const int i = t + 16 * Q * d;
                                    const long long o = ob + ((i - ifirst) >> lgE);
                                    if (DEC * i + rb >= nh && o < lim)
                                        put(o, x[d]);
                                }
                            }
                        } else { // (i - ifirst) = jj E + mm, carried from one result to the next: one division per thread
                            const unsigned uE = (unsigned)E, qE = (16u * Q) / uE, rE = 16u * Q - qE * uE;
                            const unsigned r0 = (unsigned)(t - ifirst + E); // (ifirst < E: not negative)
                            int jj = (int)(r0 / uE) - 1;
                            unsigned mm = r0 - (r0 / uE) * uE;
#pragma unroll
                            for (int d = 0; d < 16; d++) {
                                const int i = t + 16 * Q * d;
                                const long long o = ob + jj;
                                if (mm == 0 && DEC * i + rb >= nh && o < lim)
                                    put(o, x[d]);
                                mm += rE;
                                jj += (int)qE;
                                if (mm >= uE) {
                                    mm -= uE;
                                    jj++;
                                }
                            }
                        }
                    }
                }
                group_sync(bid); // S is free for the next block's first scatter
                continue;
            }
            fft4096_3x16<false>(v, S, t, bid, T1, TW2, []() {});
            // item m = t + 256 q of the block is output n = blk V + m - nh, for m >= nh
            const long long n0 = blk * V - nh + t;
            if constexpr (RP) {
                if (D == 1 && (blk + 1) * V <= nyB) {
                    float* oa = outf + n0;
                    float* oc = oa + half;
#pragma unroll
                    for (int q = 0; q < 16; q++)
                        if (t + 256 * q >= nh) {
                            oa[256 * q] = v[q].x;
                            oc[256 * q] = v[q].y;
                        }
                } else {
                    const long long dec_q = (blk * V) / D;
                    const unsigned dec_r = (unsigned)(blk * V - dec_q * D);
                    if ((256 % D) == 0 && (blk + 1) * V <= nyB) { // D | 256: a thread's 16 items are kept or dropped together
                        const int sg = (int)dec_r + t - nh;       // item m is decimated output dec_q + (sg + 256 q) / D
                        if ((sg & (D - 1)) == 0) {
                            const int step = 256 / D;
                            float* oa = outf + dec_q + sg / D;
                            float* oc = oa + half;
#pragma unroll
                            for (int q = 0; q < 16; q++)
                                if (t + 256 * q >= nh) {
                                    oa[step * q] = v[q].x;
                                    oc[step * q] = v[q].y;
                                }
                        }
                    } else if ((blk + 1) * V <= nyB) { // any other decimation, whole block inside (see the complex form below)
                        const unsigned uD = (unsigned)D, qD = 256u / uD, rD = 256u - qD * uD;
                        const unsigned lift = ((unsigned)nh + uD - 1) / uD;
                        const unsigned r0 = dec_r + (unsigned)t + lift * uD - (unsigned)nh;
                        int jj = (int)(r0 / uD) - (int)lift;
                        unsigned mm = r0 - (r0 / uD) * uD;
                        float* oa = outf + dec_q;
                        float* oc = oa + half;
#pragma unroll
                        for (int q = 0; q < 16; q++) {
                            if (mm == 0 && t + 256 * q >= nh) {
                                oa[jj] = v[q].x;
                                oc[jj] = v[q].y;
                            }
                            mm += rD;
                            jj += (int)qD;
                            if (mm >= uD) {
                                mm -= uD;
                                jj++;
                            }
                        }
                    } else {
#pragma unroll
                        for (int q = 0; q < 16; q++) {
                            const int m = t + 256 * q;
                            const long long n = n0 + 256 * q;
                            if (m >= nh && n < ny) {
                                const unsigned sdec = dec_r + (unsigned)(m - nh), qd = sdec / (unsigned)D;
                                if (sdec == qd * (unsigned)D) {
                                    outf[dec_q + qd] = v[q].x;
                                    if (n < nyB)
                                        outf[half + dec_q + qd] = v[q].y;
                                }
                            }
                        }
                    }
                }
            } else {
                if (D == 1 && (blk + 1) * V <= ny) {
                    float2* o = out + n0;
#pragma unroll
                    for (int q = 0; q < 16; q++)
                        if (t + 256 * q >= nh)
                            o[256 * q] = v[q];
                } else if (D == 1) {
#pragma unroll
                    for (int q = 0; q < 16; q++)
                        if (t + 256 * q >= nh && n0 + 256 * q < ny)
                            out[n0 + 256 * q] = v[q];
                } else {
                    const long long dec_q = (blk * V) / D;
                    const unsigned dec_r = (unsigned)(blk * V - dec_q * D);
                    if ((256 % D) == 0 && (blk + 1) * V <= ny) { // D | 256: a thread's 16 items are kept or dropped together
                        const int sg = (int)dec_r + t - nh;      // item m is decimated output dec_q + (sg + 256 q) / D
                        if ((sg & (D - 1)) == 0) {
                            const int step = 256 / D;
                            float2* o = out + dec_q + sg / D;
#pragma unroll
                            for (int q = 0; q < 16; q++)
                                if (t + 256 * q >= nh)
                                    o[step * q] = v[q];
                        }
                    } else if ((blk + 1) * V <= ny) { // any other decimation, whole block inside: dec_r + m - nh = jj D + mm,
                                                      // carried from one item of the thread to the next (256 = qD D + rD)
                        const unsigned uD = (unsigned)D, qD = 256u / uD, rD = 256u - qD * uD;
                        const unsigned lift = ((unsigned)nh + uD - 1) / uD; // (so that the first dividend is not negative)
                        const unsigned r0 = dec_r + (unsigned)t + lift * uD - (unsigned)nh;
                        int jj = (int)(r0 / uD) - (int)lift;
                        unsigned mm = r0 - (r0 / uD) * uD;
                        float2* o = out + dec_q;
#pragma unroll
                        for (int q = 0; q < 16; q++) {
                            if (mm == 0 && t + 256 * q >= nh)
                                o[jj] = v[q];
                            mm += rD;
                            jj += (int)qD;
                            if (mm >= uD) {
                                mm -= uD;
                                jj++;
                            }
                        }
                    } else {
#pragma unroll
                        for (int q = 0; q < 16; q++) {
                            const int m = t + 256 * q;
                            if (m >= nh && n0 + 256 * q < ny)
                                store_decimated(out, dec_q, dec_r, (unsigned)D, (unsigned)(m - nh), v[q]);
                        }
                    }
                }
            }
            group_sync(bid); // the inverse transform's last gather is done before the next block's first scatter
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (threadIdx.x < 32)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(TMEM_COLS));
}

static bool ols3_enabled()
{
    static const bool on = [] {
        const char* e = GRB_GETENV_ONCE("GRB200_FFTFILT_OLS3"); // 0 = the 256-thread kernels
        return !(e && e[0] == '0');
    }();
    return on;
}

// RP: `nout` real outputs, `half` as in launch_ols_realpair; otherwise complex items.
template <bool RP, int DEC>
static int launch_ols3_d(const void* in, void* out, long long nout, long long nin, int ntaps, int D, const float2* H,
                         const float2* tw, int nsm, cudaStream_t st)
{
    using G = Ols3;
    auto kern = fftfilt_ols3_kernel<RP, DEC>;
    GRB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, G::SMEM));
    const long long half = RP ? (nout + 1) / 2 : 0;
    const int V = G::L - (ntaps - 1);
    const long long nblocks = ((RP ? half : nout) * D + V - 1) / V;
    long long groups = (long long)nsm * G::NG; // one CTA per SM, NG independent block runs each
    if (groups > nblocks)
        groups = nblocks;
    const long long run = (nblocks + groups - 1) / groups;
    groups = (nblocks + run - 1) / run;
    const long long grid = (groups + G::NG - 1) / G::NG;
    const int gs = RP ? 0 : (int)((((uintptr_t)in) >> 3) & 1);
    kern<<<(unsigned)grid, G::NG * G::TPV, G::SMEM, st>>>((const float2*)in - gs, (float2*)out, nout, nin + gs, ntaps, D, H, tw,
                                                          run, gs, half);
    GRB_LAUNCH_CHECK();
    return 0;
}

template <bool RP>
static int launch_ols3(const void* in, void* out, long long nout, long long nin, int ntaps, int D, const float2* H,
                       const float2* tw, int nsm, cudaStream_t st)
{
    static const bool fold = [] {
        const char* e = GRB_GETENV_ONCE("GRB200_FFTFILT_FOLD"); // 0 = full-length inverse transform, decimation on store
        return !(e && e[0] == '0');
    }();
    // fold by the largest of 8, 4, 2 that divides the decimation; what is left of it (powers of two up to 32: whole threads;
    // anything else: per result) is dropped on store.  Odd decimations: full-length inverse.
    if (fold && D > 1 && D <= (1 << 20)) {
        const int rest8 = D / 8;
        if (D % 8 == 0 && ((rest8 & (rest8 - 1)) != 0 || rest8 <= 32))
            return launch_ols3_d<RP, 8>(in, out, nout, nin, ntaps, D, H, tw, nsm, st);
        if (D % 8 != 0 && D % 4 == 0)
            return launch_ols3_d<RP, 4>(in, out, nout, nin, ntaps, D, H, tw, nsm, st);
        if (D % 4 != 0 && D % 2 == 0)
            return launch_ols3_d<RP, 2>(in, out, nout, nin, ntaps, D, H, tw, nsm, st);
    }
    return launch_ols3_d<RP, 0>(in, out, nout, nin, ntaps, D, H, tw, nsm, st);
}

static int ols3_realpair(const void* in, void* out, long long nout, int ntaps, int D, const float2* H, const float2* tw,
                         int nsm, cudaStream_t st, bool* done)
{
    *done = false;
    if (!ols3_enabled() || (((uintptr_t)in) & 15) != 0) // the bulk copies start on 16-byte boundaries of the float stream
        return 0;
    *done = true;
    return launch_ols3<true>(in, out, nout, nout * D + ntaps - 1, ntaps, D, H, tw, nsm, st);
}

// Device plan for a tap count: FFT length L (at least the reference's 2*2^ceil(log2 ntaps), never below 4096) and
// the number of tap partitions K (1 = plain overlap-save with hop L-ntaps+1, 2 = partitioned with hop L/2).  Runs at
// create / set_taps time only (the developer knobs below are never read per work()).
int fftfilt_device_plan(int ntaps, int* parts)
{
    int L = 4096, K = 1;
    while (L < 2 * ntaps && L < 16384)
        L *= 2;
    if (L == 8192 && ntaps - 1 > L / 4)
        L = 16384; // > 25 % of every block would be overlap: the longer transform wins (cfg2: 70 -> 78 Gsps)
    const char* ue = getenv("GRB200_FFTFILT_UPOLS");           // 0 = never, 1 = whenever the geometry allows, unset = auto
    const int upols = ue ? atoi(ue) : -1;
    // 2050..10240 taps: K = ceil(ntaps / 2048) partitions on the 4096-point kernel (K-1 <= 4 parked spectra fit in the
    // 512 TMEM columns of an SM at two CTAs) beat the 16384-point transform; below 2050 plain overlap-save is faster.
    const int kneed = (ntaps + 2047) / 2048;
    if (upols != 0 && (L == 16384 || ntaps > 2048) && kneed <= 5) { // (measured: plain overlap-save on the three-group kernel 137
        L = 4096;                                                     // Gsps at 2048 taps, the two-partition kernel 130 at any count)
        K = kneed < 2 ? 2 : kneed;
    } else if (upols == 1 && ntaps > 1025 && kneed <= 5) {
        L = 4096;
        K = kneed < 2 ? 2 : kneed;
    }
    if (const char* e = getenv("GRB200_FFTFILT_L")) { // tuning knob: a longer transform wastes less on the overlap
        const int v = atoi(e);
        if (K == 1 && (v == 4096 || v == 8192 || v == 16384) && v >= L)
            L = v;
    }
    if (K == 1 && L - (ntaps - 1) < L / 4)
        return -1; // too many taps for a single-CTA transform
    if (parts)
        *parts = K;
    return L;
}

int fftfilt_launch(int L, int parts, const void* in, void* out, long long nout, int ntaps, int D, const float2* H,
                   const float2* tw, int nsm, cudaStream_t st)
{
    if (nout <= 0)
        return 0;
    const long long nin = nout * D + ntaps - 1;
    if (parts >= 2) {
        if ((((uintptr_t)in) & 7) != 0)
            return fail(GRB200_EINVAL, "fft_filter: input stream is not aligned to its item size");
        if (L == 4096) {
            switch (parts) {
            case 2: {
                static const bool three = [] {
                    const char* e = GRB_GETENV_ONCE("GRB200_FFTFILT_UPOLS2"); // 0 = the generic kernel (2 CTAs per SM)
                    return !(e && e[0] == '0');
                }();
                if (three)
                    return launch_upols2(in, out, nout, nin, ntaps, D, H, tw, nsm, st);
                return launch_upols<4096, 2>(in, out, nout, nin, ntaps, D, H, tw, nsm, st);
            }
            case 3:
                return launch_upols<4096, 3>(in, out, nout, nin, ntaps, D, H, tw, nsm, st);
            case 4:
                return launch_upols<4096, 4>(in, out, nout, nin, ntaps, D, H, tw, nsm, st);
            case 5:
                return launch_upols<4096, 5>(in, out, nout, nin, ntaps, D, H, tw, nsm, st);
            }
        }
        return fail(GRB200_ERANGE, "fft_filter: no partitioned kernel for fft length %d", L);
    }
    static const bool use_stream = [] {
        const char* e = GRB_GETENV_ONCE("GRB200_FFTFILT_STREAM");
        return !(e && e[0] == '0');
    }();
    if (L == 4096 && ols3_enabled() && (((uintptr_t)in) & 7) == 0)
        return launch_ols3<false>(in, out, nout, nin, ntaps, D, H, tw, nsm, st);
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
