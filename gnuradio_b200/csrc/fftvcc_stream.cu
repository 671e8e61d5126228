// fftvcc_stream.cu -- the streaming (persistent, TMA-prefetch) kernels of the batched FFT; see fftvcc.cu.
#include "fftvcc.cuh"

namespace grb {

// ---- streaming variant (32 <= N <= 8192; below 1024 several vectors per CTA share one bulk copy) ----------------------------------------------------------
// ncu on the kernel above (profiles/r1_fftvcc_ncu.txt): 2 CTAs/SM at 127 registers, long-scoreboard stalls
// 2.9 per issue -- every vector's loads are exposed, and window + twiddle lookups add 2.7x the input's L1
// sectors.  Here a persistent CTA keeps ONE TMA bulk copy (its next VPC vectors, contiguous in the stream)
// in flight while it transforms the current ones: the copy is issued as soon as the first pass has drained
// the landing buffer, and completes behind passes 2..n and the stores.  The window sits in shared memory,
// the loop-invariant base twiddles too (cta_fft_tws), so the only global loads left are the bulk copies.
template <int N>
struct VccStreamGeom {
    using G = VccGeom<N>;
    static constexpr int XIN_BYTES = G::VPC * N * 8;
    static constexpr int S_BYTES = G::VPC * fft_smem_elems(N) * 8;
    static constexpr int TW_BYTES = fft_tws_slots(N) * G::TPV * 8;
    static constexpr int WIN_BYTES = N * 4;
    static constexpr int SMEM = XIN_BYTES + S_BYTES + TW_BYTES + WIN_BYTES + 16;
};

// MODE is a template parameter here (straight-line store code: the 16 stores of a thread issue back to back) -- the
// per-element mode switch of vcc_store() cost the cfg5 kernel 17 % when it was first generalised.
template <int N, bool FWD, bool REAL, int MODE>
__global__ void __launch_bounds__(VccGeom<N>::NT)
fft_vcc_stream_kernel(const VccArgs a)
{
    using G = VccGeom<N>;
    using SG = VccStreamGeom<N>;
    constexpr int ISZ = REAL ? 4 : 8; // bytes per input item
    const long long nvec = a.nvec;
    const float* __restrict__ window = a.window;
    const float2* __restrict__ tw = a.tw;
    const int shift = a.shift;
    constexpr int E = G::E, TPV = G::TPV, VPC = G::VPC;
    constexpr int R0 = fft_first_radix(N, false);
    constexpr int RL = fft_last_radix(N, false);
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float2* XIN = reinterpret_cast<float2*>(smem_raw);
    float2* Sall = reinterpret_cast<float2*>(smem_raw + SG::XIN_BYTES);
    float2* TW = reinterpret_cast<float2*>(smem_raw + SG::XIN_BYTES + SG::S_BYTES);
    float* WIN = reinterpret_cast<float*>(smem_raw + SG::XIN_BYTES + SG::S_BYTES + SG::TW_BYTES);
    unsigned long long* bar = reinterpret_cast<unsigned long long*>(smem_raw + SG::XIN_BYTES + SG::S_BYTES +
                                                                    SG::TW_BYTES + SG::WIN_BYTES);
    const int vl = threadIdx.x / TPV;
    const int t = threadIdx.x % TPV;
    float2* S = Sall + vl * fft_smem_elems(N);
    const float2* X = XIN + vl * N;
    const float* XR = reinterpret_cast<const float*>(XIN) + vl * N;
    float2* __restrict__ outc = reinterpret_cast<float2*>(a.out);
    float* __restrict__ outf = reinterpret_cast<float*>(a.out);
    const float log_mul = a.log_mul, log_scale = a.log_scale;
    const bool in_swap = (!FWD) && shift;
    const bool out_rot = FWD && shift;
    const bool has_win = window != nullptr;
    // MODE 3: CTA r walks its own run of consecutive frames (the recurrence runs along them); else CTAs interleave
    const long long vstep = MODE == 3 ? 1 : (long long)gridDim.x * VPC;
    const long long vfirst = MODE == 3 ? (long long)blockIdx.x * a.run_len : (long long)blockIdx.x * VPC;
    const long long vend = MODE == 3 ? (vfirst + a.run_len < nvec ? vfirst + a.run_len : nvec) : nvec;
    const double alpha = a.alpha, beta = 1.0 - a.alpha;
    float ys[MODE == 3 ? E : 1]; // the filter memory of this thread's bins
#pragma unroll
    for (int i = 0; i < (MODE == 3 ? E : 1); i++)
        ys[i] = 0.f;

    auto issue = [&](long long vbase) {
        long long nv = nvec - vbase;
        nv = nv > VPC ? VPC : nv;
        mbar_expect_tx(bar, (unsigned)(nv * N * ISZ));
        if (a.vstride == 1) { // the CTA's vectors are contiguous in the stream: one copy
            bulk_g2s(XIN, reinterpret_cast<const char*>(a.in) + (vbase + a.voff) * N * ISZ, (unsigned)(nv * N * ISZ), bar);
        } else {
            for (int s = 0; s < (int)nv; s++)
                bulk_g2s(reinterpret_cast<char*>(XIN) + (size_t)s * N * ISZ,
                         reinterpret_cast<const char*>(a.in) + ((vbase + s) * a.vstride + a.voff) * N * ISZ, N * ISZ, bar);
        }
    };

    if (threadIdx.x == 0)
        mbar_init(bar, 1);
    if (vl == 0)
        cta_fft_tws_init<N, FWD>(TW, t, tw);
    if (has_win)
        for (int i = threadIdx.x; i < N; i += G::NT)
            WIN[i] = window[i];
    __syncthreads();
    long long vbase = vfirst;
    if (threadIdx.x == 0 && vbase < vend)
        issue(vbase);

    unsigned phase = 0;
    for (; vbase < vend; vbase += vstep, phase ^= 1) {
        const long long vec = vbase + vl;
        const bool valid = vec < nvec;
        mbar_wait(bar, phase);
        float2 v[E];
#pragma unroll
        for (int b = 0; b < E / R0; b++) {
#pragma unroll
            for (int q = 0; q < R0; q++) {
                int d = (t + b * TPV) + q * (N / R0);
                if (in_swap)
                    d = (d + N / 2) & (N - 1);
                float2 x = make_float2(0.f, 0.f);
                if (valid)
                    x = REAL ? make_float2(XR[d], 0.f) : X[d];
                if (has_win) {
                    const float w = WIN[d];
                    x = __fmul2_rn(x, make_float2(w, w));
                }
                v[b * R0 + q] = x;
            }
        }
        const long long vnext = vbase + vstep;
        cta_fft_tws<N, FWD>(v, S, t, TW, tw, [&]() {
            if (threadIdx.x == 0 && vnext < vend)
                issue(vnext); // every thread has read XIN (barrier after the first pass)
        });
        if (valid) {
#pragma unroll
            for (int b = 0; b < E / RL; b++) {
#pragma unroll
                for (int q = 0; q < RL; q++) {
                    int k = (t + b * TPV) + q * (N / RL);
                    if (out_rot)
                        k = (k + N / 2) & (N - 1);
                    const float2 Xk = v[b * RL + q];
                    if constexpr (MODE == 0) {
                        outc[vec * N + k] = Xk;
                    } else {
                        const float pw = fmaf(Xk.x, Xk.x, Xk.y * Xk.y);
                        if constexpr (MODE == 1) {
                            outf[vec * N + k] = pw;
                        } else if constexpr (MODE == 3) { // single_pole_iir.h:82 (double arithmetic, float state), as logpwr.cu
                            const float y = (float)(alpha * (double)pw + beta * (double)ys[b * RL + q]);
                            ys[b * RL + q] = y;
                            outf[vec * N + k] = y;
                        } else {
                            outf[vec * N + k] = log2f(fmaxf(pw, 1.17549435e-38f) * log_scale) * log_mul;
                            if (vec == nvec - 1)
                                a.ystate[k] = pw;
                        }
                    }
                }
            }
        }
        __syncthreads(); // the next iteration's first pass overwrites S
    }
    if constexpr (MODE == 3) {
        if (vfirst < vend) {
#pragma unroll
            for (int b = 0; b < E / RL; b++)
#pragma unroll
                for (int q = 0; q < RL; q++) {
                    int k = (t + b * TPV) + q * (N / RL);
                    if (out_rot)
                        k = (k + N / 2) & (N - 1);
                    a.agg[(long long)blockIdx.x * N + k] = ys[b * RL + q];
                }
        }
    }
}

template <int N, bool FWD, bool REAL, int MODE>
static int launch_vcc_stream_m(const VccArgs& a, int nsm, cudaStream_t st)
{
    using G = VccGeom<N>;
    using SG = VccStreamGeom<N>;
    auto kern = fft_vcc_stream_kernel<N, FWD, REAL, MODE>;
    GRB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SG::SMEM));
    int occ = 1;
    GRB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, G::NT, SG::SMEM));
    if (occ < 1)
        occ = 1;
    long long need = (a.nvec + G::VPC - 1) / G::VPC;
    long long grid = (long long)nsm * occ;
    if (grid > need)
        grid = need;
    kern<<<(unsigned)grid, G::NT, SG::SMEM, st>>>(a);
    GRB_LAUNCH_CHECK();
    return 0;
}

template <int N, bool FWD, bool REAL>
static int launch_vcc_stream(const VccArgs& a, int nsm, cudaStream_t st)
{
    if (a.mode == 0)
        return launch_vcc_stream_m<N, FWD, REAL, 0>(a, nsm, st);
    if (a.mode == 1)
        return launch_vcc_stream_m<N, FWD, REAL, 1>(a, nsm, st);
    if constexpr (FWD) // the log epilogue belongs to logpwrfft: forward transforms only
        return launch_vcc_stream_m<N, FWD, REAL, 2>(a, nsm, st);
    else
        return fail(GRB200_EINVAL, "fft_vcc: the log epilogue needs a forward transform");
}


// ---- mode 3: the run-local single-pole IIR of logpwrfft in the FFT kernel's epilogue ------------------------------------------
// (logpwr.cu's first pass -- read |X|^2, filter, write -- folded into the kernel that produces |X|^2: 8 B per bin less HBM
// traffic.)  One run per CTA, so the carries / fix-up kernels of logpwr.cu see nruns = the grid.
template <int N, bool REAL>
static int launch_vcc_stream_iir(VccArgs a, int nsm, cudaStream_t st, long long* run_len, long long* nruns, bool* done)
{
    using G = VccGeom<N>;
    using SG = VccStreamGeom<N>;
    static_assert(G::VPC == 1, "one vector per CTA");
    auto kern = fft_vcc_stream_kernel<N, true, REAL, 3>;
    GRB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SG::SMEM));
    int occ = 1;
    GRB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, G::NT, SG::SMEM));
    if (occ < 1)
        occ = 1;
    const long long grid = (long long)nsm * (occ > 4 ? 4 : occ); // (logpwr_fused_scratch_floats: room for 4 runs per SM)
    if (a.nvec < 4 * grid) // (short calls: the three-kernel form spreads a few frames over the machine better)
        return 0;
    const long long len = (a.nvec + grid - 1) / grid, runs = (a.nvec + len - 1) / len;
    a.run_len = len;
    a.mode = 3;
    kern<<<(unsigned)runs, G::NT, SG::SMEM, st>>>(a);
    GRB_LAUNCH_CHECK();
    *run_len = len;
    *nruns = runs;
    *done = true;
    return 0;
}

long long logpwr_fused_scratch_floats(int N, int nsm) { return (long long)nsm * 4 * N; }

int fft_vcc_stream_iir_launch(int N, bool real_in, VccArgs a, double alpha, float* agg, int nsm, cudaStream_t st,
                              long long* run_len, long long* nruns, bool* done)
{
    *done = false;
    static const bool on = [] {
        const char* e = GRB_GETENV_ONCE("GRB200_LOGPWR_FUSED"); // 0 = FFT kernel, then the three kernels of logpwr.cu
        return !(e && e[0] == '0');
    }();
    if (!on || (((uintptr_t)a.in) & 15) != 0)
        return 0;
    a.alpha = alpha;
    a.agg = agg;
    switch (N) {
    case 2048:
        return real_in ? launch_vcc_stream_iir<2048, true>(a, nsm, st, run_len, nruns, done)
                       : launch_vcc_stream_iir<2048, false>(a, nsm, st, run_len, nruns, done);
    case 4096:
        return real_in ? launch_vcc_stream_iir<4096, true>(a, nsm, st, run_len, nruns, done)
                       : launch_vcc_stream_iir<4096, false>(a, nsm, st, run_len, nruns, done);
    case 8192:
        return real_in ? launch_vcc_stream_iir<8192, true>(a, nsm, st, run_len, nruns, done)
                       : launch_vcc_stream_iir<8192, false>(a, nsm, st, run_len, nruns, done);
    }
    return 0;
}

int fft_vcc_stream_launch(int N, bool fwd, bool real_in, const VccArgs& a, int nsm, cudaStream_t st)
{
#define GRB_VCC_STREAM(n)                                                                                         \
    case n:                                                                                                       \
        return fwd ? (real_in ? launch_vcc_stream<n, true, true>(a, nsm, st) : launch_vcc_stream<n, true, false>(a, nsm, st)) \
                   : (real_in ? launch_vcc_stream<n, false, true>(a, nsm, st) : launch_vcc_stream<n, false, false>(a, nsm, st));
    switch (N) {
        GRB_VCC_STREAM(32)
        GRB_VCC_STREAM(64)
        GRB_VCC_STREAM(128)
        GRB_VCC_STREAM(256)
        GRB_VCC_STREAM(512)
        GRB_VCC_STREAM(1024)
        GRB_VCC_STREAM(2048)
        GRB_VCC_STREAM(4096)
        GRB_VCC_STREAM(8192)
    }
#undef GRB_VCC_STREAM
    return fail(GRB200_ERANGE, "fft_vcc: no streaming kernel for size %d", N);
}

} // namespace grb
