// pfb.cu -- pfb_channelizer_ccf: branch FIRs -> M-point inverse DFT -> channel-major store, one kernel (K4).
//
// Replaces pfb_channelizer_ccf_impl::general_work (reference gr-filter/lib/pfb_channelizer_ccf_impl.cc:128-193)
// on top of kernel::polyphase_filterbank (gr-filter/lib/polyphase_filterbank.cc:53-81).
//
// Closed form used on the device (derived from :73-77 and :157-189, checked against the literal loop by the
// parity tests).  With the M input streams seen as ONE interleaved stream x (stream j, item n at
// x[n*M + j]; what stream_to_streams split apart, gr-blocks/lib/stream_to_streams_impl.cc:35-54),
// R = M/oversample_rate, h = prototype taps zero-padded to T*M, for output index oo:
//     P      = (oo+1)*R - 1                                   (newest interleaved sample used)
//     v_i    = sum_{u<T} h[i + u*M] * x[P - i - u*M]          (branch FIR i, i = 0..M-1)
//     fin[q] = v_i  with  q = (i - (oo+2)*R) mod M            (the idxlut rotation; identity when R == M)
//     out_n[oo] = sum_q fin[q] * e^{+j 2 pi q c / M},  c = channel_map[n]
//
// First-generation fast path (M in {16, 64, 256}, oversample_rate 1, T <= 32; pfb_stream.cu now takes every power of two 8..256): persistent CTAs of 256 threads; a tile is
// OT = 256*G/M output times.  Phase 1: thread (branch i, group g) slides a register window down its
// branch's time axis -- G outputs from G+TP-1 coalesced 8-byte loads (lanes run over i, i.e. over
// consecutive interleaved samples) with its TP taps held in registers -- and drops v into shared memory
// [time][branch].  Phase 2: thread (time o, j) runs a two-pass radix-sqrt(M) inverse FFT over the branches
// with lanes running over TIME, so the final stores are 256-byte runs per channel: the channel-major
// transpose costs nothing extra.  HBM traffic is the compulsory 8 B in + 8 B out per input sample.
#include <cstdlib>
#include "fft.cuh"

namespace grb {

bool pfb_stream_supported(int M, int T, int R, const void* x);
bool pfb_stream_split_supported(int M, int T, int R, const void* x);
int pfb_stream_split_launch(int M, int T, const void* x, void* out, long long out_stride, int nouts, long long noutput,
                            const float* h, const float2* wM, int nsm, cudaStream_t st);
int pfb_stream_launch(int M, int T, const void* x, void* out, long long out_stride, int nouts, long long noutput,
                      const float* h, const float2* wM, int nsm, cudaStream_t st, const void* halo, const int* inv);
int halo_fetch_launch(void* dst, const void* src, size_t bytes, cudaStream_t st);
bool pfb_stream_os2_supported(int M, int T, int R, const void* x);
int pfb_stream_os2_launch(int M, int T, const void* x, void* out, long long out_stride, int nouts, long long noutput,
                          const float* h, const float2* wM, int nsm, cudaStream_t st, const int* inv);

// ---- generic kernel: any M, T, rate ratio, channel map -------------------------------------------------
__global__ void __launch_bounds__(256)
pfb_generic_kernel(const float2* __restrict__ x, float2* __restrict__ out, long long out_stride, int nouts,
                   const int* __restrict__ cmap, long long noutput, int M, int T, int R,
                   const float* __restrict__ h, const float2* __restrict__ wtab, int OT, int R1)
{
    // R1 > 1: M = R1 * R2 and the DFT runs in two stages through a second buffer (M (R1 + R2) multiply-adds per output time
    // instead of M^2): y[j][k1] = sum_q v[j + R2 q] W^(R2 q k1), then X[c] = sum_j y[j][c mod R1] W^(j c) for the mapped channels
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float2* v = reinterpret_cast<float2*>(smem_raw); // [OT][M+1]
    float2* W = v + (size_t)OT * (M + 1);            // [M]
    float2* v2 = W + M;                              // [OT][M+1] when R1 > 1
    const int R2 = R1 > 1 ? M / R1 : 1;
    const int NT = blockDim.x, tid = threadIdx.x;
    for (int k = tid; k < M; k += NT)
        W[k] = wtab[k];
    const long long ntiles = (noutput + OT - 1) / OT;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const long long o0 = tile * OT;
        const int rot0 = (int)(((o0 + 2) * R) % M), Rm = R % M;
        __syncthreads();
        for (int idx = tid; idx < OT * M; idx += NT) {
            const int ol = idx / M, i = idx % M;
            const long long oo = o0 + ol;
            if (oo < noutput) {
                const long long P = (oo + 1) * R - 1;
                const float2* xp = x + (P - i + (long long)T * M);
                float2 acc = make_float2(0.f, 0.f);
                int u = 0;
                for (; u + 4 <= T; u += 4) { // four independent loads of each kind in flight, then the dependent chain
                    float hh[4];
                    float2 a[4];
#pragma unroll
                    for (int e = 0; e < 4; e++) {
                        hh[e] = __ldg(h + (u + e) * M + i);
                        a[e] = __ldg(xp - (long long)(u + e) * M);
                    }
#pragma unroll
                    for (int e = 0; e < 4; e++)
                        acc = __ffma2_rn(make_float2(hh[e], hh[e]), a[e], acc);
                }
                for (; u < T; u++) {
                    const float hh = __ldg(h + u * M + i);
                    acc = __ffma2_rn(make_float2(hh, hh), __ldg(xp - (long long)u * M), acc);
                }
                int q = (i - rot0 - ol * Rm) % M; // (i - (oo + 2) R) mod M, the 64-bit part taken once per tile
                if (q < 0)
                    q += M;
                v[ol * (M + 1) + q] = acc;
            }
        }
        __syncthreads();
        if (R1 > 1) {
            for (int idx = tid; idx < OT * M; idx += NT) { // (ol, k1, j): lanes over j, then k1
                const int ol = idx / M, r = idx % M, j = r % R2, k1 = r / R2;
                const float2* vr = v + ol * (M + 1) + j;
                float2 acc = make_float2(0.f, 0.f);
                int widx = 0;
                const int wstep = (R2 * k1) % M;
                for (int q = 0; q < R1; q++) {
                    acc = __ffma2_rn(make_float2(W[widx].x, W[widx].x), vr[q * R2], acc);
                    acc = __ffma2_rn(make_float2(-W[widx].y, W[widx].y), make_float2(vr[q * R2].y, vr[q * R2].x), acc);
                    widx += wstep;
                    if (widx >= M)
                        widx -= M;
                }
                v2[ol * (M + 1) + k1 * R2 + j] = acc;
            }
            __syncthreads();
            for (int idx = tid; idx < OT * nouts; idx += NT) {
                const int ol = idx % OT, n = idx / OT;
                const long long oo = o0 + ol;
                if (oo < noutput) {
                    const int c = cmap[n];
                    const float2* yr = v2 + ol * (M + 1) + (c % R1) * R2;
                    float2 acc = make_float2(0.f, 0.f);
                    int widx = 0;
                    for (int j = 0; j < R2; j++) {
                        acc = __ffma2_rn(make_float2(W[widx].x, W[widx].x), yr[j], acc);
                        acc = __ffma2_rn(make_float2(-W[widx].y, W[widx].y), make_float2(yr[j].y, yr[j].x), acc);
                        widx += c;
                        if (widx >= M)
                            widx -= M;
                    }
                    out[(long long)n * out_stride + oo] = acc;
                }
            }
            continue;
        }
        for (int idx = tid; idx < OT * nouts; idx += NT) {
            const int ol = idx % OT, n = idx / OT;
            const long long oo = o0 + ol;
            if (oo < noutput) {
                const int c = cmap[n];
                const float2* vr = v + ol * (M + 1);
                float2 acc = make_float2(0.f, 0.f);
                int widx = 0;
                for (int q = 0; q < M; q++) {
                    const float2 a = vr[q], w = W[widx];
                    acc.x = fmaf(a.x, w.x, acc.x);
                    acc.x = fmaf(-a.y, w.y, acc.x);
                    acc.y = fmaf(a.x, w.y, acc.y);
                    acc.y = fmaf(a.y, w.x, acc.y);
                    widx += c;
                    if (widx >= M)
                        widx -= M;
                }
                out[(long long)n * out_stride + oo] = acc;
            }
        }
    }
}

// ---- any rate ratio, any channel map, any tap count -- power-of-two channel counts 8 .. 256 --------------------------
// The generic kernel above evaluates the DFT directly, M multiply-adds per output: 11 - 17 Gsps where the streaming kernel
// does 300+ (an oversampled channelizer, more than 32 taps per branch).  Same tile structure, but the branches go through
// the two in-place passes of pfb_stream.cu (M = R1 x R2, lanes over time) and the channel map is a gather at the end:
//   FIR   thread (time ol, branch i): T loads through L1 (lanes over i: consecutive interleaved samples), taps of its branch in
//         registers (its branch never changes: 256 % M == 0); the idxlut rotation q = (i - (oo+2) R) mod M is the scatter
//   DFT   pass 1 radix R1 over q = j + R2 q', pass 2 twiddle + radix R2 over j, results back in natural order
//   out   out_n[oo] = X[channel_map[n]], lanes over time (128-byte runs per channel)
template <int M, int TP>
__global__ void __launch_bounds__(256)
pfb_fftgen_kernel(const float2* __restrict__ x, float2* __restrict__ out, long long out_stride, int nouts,
                  const int* __restrict__ cmap, long long noutput, int T, int R, const float* __restrict__ h,
                  const float2* __restrict__ wtab_g)
{
    constexpr int NT = 256, OT = 32, LD = M + 1;
    constexpr int R1 = M >= 128 ? 16 : (M >= 32 ? 8 : (M == 16 ? 4 : M)), R2 = M / R1;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float2* vbuf = reinterpret_cast<float2*>(smem_raw); // [OT][LD]
    float2* wtab = vbuf + OT * LD;                      // [M]
    int* smap = reinterpret_cast<int*>(wtab + M);       // [nouts]
    const int tid = threadIdx.x;
    for (int k = tid; k < M; k += NT)
        wtab[k] = wtab_g[k];
    for (int k = tid; k < nouts; k += NT)
        smap[k] = cmap[k];
    const int i = tid % M; // this thread's branch in every FIR item it owns
    float taps[TP > 0 ? TP : 1];
    if (TP > 0) {
#pragma unroll
        for (int u = 0; u < TP; u++)
            taps[u] = u < T ? __ldg(h + u * M + i) : 0.f;
    }
    const long long ntiles = (noutput + OT - 1) / OT;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const long long o0 = tile * OT;
        __syncthreads(); // the previous tile's gather is done (and the tables are written)
        // ---------------- branch FIRs ----------------
        for (int ol = tid / M; ol < OT; ol += NT / M) {
            const long long oo = o0 + ol;
            if (oo < noutput) {
                const long long P = (oo + 1) * R - 1;
                const float2* xp = x + (P - i + (long long)T * M);
                float2 a0 = make_float2(0.f, 0.f), a1 = make_float2(0.f, 0.f);
                if (TP > 0) {
#pragma unroll
                    for (int u = 0; u < TP; u += 2) {
                        if (u < T)
                            a0 = __ffma2_rn(make_float2(taps[u], taps[u]), __ldg(xp - (long long)u * M), a0);
                        if (u + 1 < T)
                            a1 = __ffma2_rn(make_float2(taps[u + 1], taps[u + 1]), __ldg(xp - (long long)(u + 1) * M), a1);
                    }
                } else {
                    for (int u = 0; u < T; u++) {
                        const float hh = __ldg(h + u * M + i);
                        a0 = __ffma2_rn(make_float2(hh, hh), __ldg(xp - (long long)u * M), a0);
                    }
                }
                int q = (int)((i - (oo + 2) * R) % M);
                if (q < 0)
                    q += M;
                vbuf[ol * LD + q] = __fadd2_rn(a0, a1);
            }
        }
        __syncthreads();
        // ---------------- inverse DFT pass 1 (in place): y[j][k1] -> slot j + R2*k1 ----------------
        for (int idx = tid; idx < OT * R2; idx += NT) {
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
        // ---------------- pass 2: twiddle, DFT over j; natural order back into the buffer ----------------
        constexpr int NI2 = (OT * R1 + NT - 1) / NT;
        float2 a[NI2][R2];
#pragma unroll
        for (int n = 0; n < NI2; n++) {
            const int idx = tid + n * NT;
            if (idx < OT * R1) {
                const int o = idx % OT, k1 = idx / OT;
#pragma unroll
                for (int j = 0; j < R2; j++) {
                    a[n][j] = vbuf[o * LD + k1 * R2 + j];
                    if (j > 0)
                        a[n][j] = cmul(a[n][j], wtab[j * k1]);
                }
                dft_r<R2, false>(a[n]);
            }
        }
        __syncthreads();
#pragma unroll
        for (int n = 0; n < NI2; n++) {
            const int idx = tid + n * NT;
            if (idx < OT * R1) {
                const int o = idx % OT, k1 = idx / OT;
#pragma unroll
                for (int k2 = 0; k2 < R2; k2++)
                    vbuf[o * LD + k1 + R1 * k2] = a[n][k2];
            }
        }
        __syncthreads();
        // ---------------- channel map gather, lanes over time ----------------
        for (int idx = tid; idx < OT * nouts; idx += NT) {
            const int ol = idx % OT, n = idx / OT;
            const long long oo = o0 + ol;
            if (oo < noutput)
                out[(long long)n * out_stride + oo] = vbuf[ol * LD + smap[n]];
        }
    }
}

template <int M, int TP>
static int launch_fftgen(const void* x, void* out, long long out_stride, int nouts, const int* cmap, long long noutput,
                         int T, int R, const float* h, const float2* wM, int nsm, cudaStream_t st)
{
    const size_t smem = ((size_t)32 * (M + 1) + M) * 8 + (size_t)nouts * 4;
    auto kern = pfb_fftgen_kernel<M, TP>;
    if (smem > 48 * 1024)
        GRB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int occ = 1;
    GRB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, 256, smem));
    if (occ < 1)
        occ = 1;
    const long long ntiles = (noutput + 31) / 32;
    long long grid = (long long)nsm * occ;
    if (grid > ntiles)
        grid = ntiles;
    kern<<<(unsigned)grid, 256, smem, st>>>((const float2*)x, (float2*)out, out_stride, nouts, cmap, noutput, T, R, h, wM);
    GRB_LAUNCH_CHECK();
    return 0;
}

template <int M>
static int launch_fftgen_tp(const void* x, void* out, long long out_stride, int nouts, const int* cmap, long long noutput,
                            int T, int R, const float* h, const float2* wM, int nsm, cudaStream_t st)
{
    if (T <= 16)
        return launch_fftgen<M, 16>(x, out, out_stride, nouts, cmap, noutput, T, R, h, wM, nsm, st);
    if (T <= 32)
        return launch_fftgen<M, 32>(x, out, out_stride, nouts, cmap, noutput, T, R, h, wM, nsm, st);
    return launch_fftgen<M, 0>(x, out, out_stride, nouts, cmap, noutput, T, R, h, wM, nsm, st);
}

// ---- fast kernel --------------------------------------------------------------------------------------
template <int M>
struct PfbGeom {
    static constexpr int RDX = M == 16 ? 4 : (M == 64 ? 8 : 16); // sqrt(M)
    static constexpr int NT = 256;
    static constexpr int G = 16;               // outputs per thread in the FIR phase
    static constexpr int OT = NT / M * G;      // output times per tile
    static constexpr int LD = M + 1;           // padded row of the [time][branch] buffers
    static constexpr int SMEM = 2 * OT * LD * 8;
};

template <int M, int TP>
__global__ void __launch_bounds__(256)
pfb_fast_kernel(const float2* __restrict__ x, float2* __restrict__ out, long long out_stride, int nouts,
                const int* __restrict__ cmap, int identity_map, long long noutput, int T, long long nrows,
                const float* __restrict__ h, const float2* __restrict__ wM)
{
    using Geo = PfbGeom<M>;
    constexpr int G = Geo::G, OT = Geo::OT, LD = Geo::LD, RDX = Geo::RDX, NT = Geo::NT;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float2* vbuf = reinterpret_cast<float2*>(smem_raw); // [OT][LD]
    float2* ybuf = vbuf + OT * LD;                      // [OT][LD]
    const int tid = threadIdx.x;

    // FIR-phase role: branch i, output group g
    const int bi = tid % M, bg = tid / M;
    float taps[TP];
#pragma unroll
    for (int u = 0; u < TP; u++)
        taps[u] = (u < T) ? __ldg(h + u * M + bi) : 0.f;
    const int col = M - 1 - bi; // input stream feeding branch i

    const long long ntiles = (noutput + OT - 1) / OT;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const long long o0 = tile * OT;
        // ---------------- phase 1: branch FIRs ----------------
        {
            const long long og = o0 + (long long)bg * G; // first output time of this thread's group
            // buffer row of output o's newest input row is o + T; oldest row needed: og + T - (TP-1)
            const long long row0 = og + T - (TP - 1);
            float2 acc[G];
#pragma unroll
            for (int k = 0; k < G; k++)
                acc[k] = make_float2(0.f, 0.f);
#pragma unroll
            for (int rl = 0; rl < G + TP - 1; rl++) {
                const long long row = row0 + rl;
                float2 a = make_float2(0.f, 0.f);
                if (row >= 0 && row < nrows)
                    a = x[row * M + col];
                // output k (time og+k) uses this row with tap u = (og + k + T) - row = k + TP-1 - rl
#pragma unroll
                for (int k = 0; k < G; k++) {
                    const int u = k + TP - 1 - rl;
                    if (u >= 0 && u < TP) {
                        acc[k].x = fmaf(taps[u], a.x, acc[k].x);
                        acc[k].y = fmaf(taps[u], a.y, acc[k].y);
                    }
                }
            }
#pragma unroll
            for (int k = 0; k < G; k++)
                vbuf[(bg * G + k) * LD + bi] = acc[k];
        }
        __syncthreads();
        // ---------------- phase 2: inverse FFT across branches, lanes over time ----------------
        constexpr int ITEMS = OT * (M / RDX);
        float2 X[(ITEMS + NT - 1) / NT][RDX];
#pragma unroll
        for (int it = 0; it < (ITEMS + NT - 1) / NT; it++) {
            const int idx = tid + it * NT;
            const int o = idx % OT, j = idx / OT;
            if (idx < ITEMS) {
                float2 v[RDX];
#pragma unroll
                for (int q = 0; q < RDX; q++)
                    v[q] = vbuf[o * LD + j + q * (M / RDX)];
                dft_r<RDX, false>(v);
#pragma unroll
                for (int q = 0; q < RDX; q++)
                    ybuf[o * LD + j * RDX + q] = v[q];
            }
        }
        __syncthreads();
#pragma unroll
        for (int it = 0; it < (ITEMS + NT - 1) / NT; it++) {
            const int idx = tid + it * NT;
            const int o = idx % OT, j = idx / OT;
            if (idx < ITEMS) {
#pragma unroll
                for (int q = 0; q < RDX; q++) {
                    float2 a = ybuf[o * LD + j + q * RDX];
                    if (q > 0) {
                        const float2 w = __ldg(wM + q * j); // e^{+j 2 pi q j / M}
                        a = cmul(a, w);
                    }
                    X[it][q] = a;
                }
                dft_r<RDX, false>(X[it]);
            }
        }
        if (identity_map) {
#pragma unroll
            for (int it = 0; it < (ITEMS + NT - 1) / NT; it++) {
                const int idx = tid + it * NT;
                const int o = idx % OT, j = idx / OT;
                const long long oo = o0 + o;
                if (idx < ITEMS && oo < noutput) {
#pragma unroll
                    for (int q = 0; q < RDX; q++) {
                        const int c = j + q * RDX;
                        if (c < nouts)
                            out[(long long)c * out_stride + oo] = X[it][q];
                    }
                }
            }
            __syncthreads(); // vbuf/ybuf are rewritten by the next tile
        } else {
            // general channel map: park the spectrum in vbuf[time][bin], then gather per output stream
#pragma unroll
            for (int it = 0; it < (ITEMS + NT - 1) / NT; it++) {
                const int idx = tid + it * NT;
                const int o = idx % OT, j = idx / OT;
                if (idx < ITEMS) {
#pragma unroll
                    for (int q = 0; q < RDX; q++)
                        vbuf[o * LD + j + q * RDX] = X[it][q];
                }
            }
            __syncthreads();
            for (int idx = tid; idx < OT * nouts; idx += NT) {
                const int o = idx % OT, n = idx / OT;
                const long long oo = o0 + o;
                if (oo < noutput)
                    out[(long long)n * out_stride + oo] = vbuf[o * LD + cmap[n]];
            }
            __syncthreads();
        }
    }
}

// ---- stream-major [M][stride] -> interleaved [n*M + j] (the absorbed stream_to_streams inverse) ---------
__global__ void __launch_bounds__(256)
streams_to_interleaved_kernel(const float2* __restrict__ in, long long in_stride, float2* __restrict__ out,
                              int M, long long nrows)
{
    __shared__ float2 tile[32][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5; // 32 x 8
    const long long n0 = (long long)blockIdx.x * 32;
    const int j0 = blockIdx.y * 32;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int j = j0 + ty + k * 8;
        const long long n = n0 + tx;
        if (j < M && n < nrows)
            tile[ty + k * 8][tx] = in[(long long)j * in_stride + n];
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const long long n = n0 + ty + k * 8;
        const int j = j0 + tx;
        if (j < M && n < nrows)
            out[n * M + j] = tile[tx][ty + k * 8];
    }
}

// ---- host launchers -----------------------------------------------------------------------------------
template <int M, int TP>
static int launch_fast(const void* x, void* out, long long out_stride, int nouts, const int* cmap, int identity,
                       long long noutput, int T, long long nrows, const float* h, const float2* wM, int nsm,
                       cudaStream_t st)
{
    using Geo = PfbGeom<M>;
    auto kern = pfb_fast_kernel<M, TP>;
    if (Geo::SMEM > 48 * 1024)
        GRB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Geo::SMEM));
    int occ = 1;
    GRB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, Geo::NT, Geo::SMEM));
    if (occ < 1)
        occ = 1;
    long long ntiles = (noutput + Geo::OT - 1) / Geo::OT;
    long long grid = (long long)nsm * occ;
    if (grid > ntiles)
        grid = ntiles;
    kern<<<(unsigned)grid, Geo::NT, Geo::SMEM, st>>>((const float2*)x, (float2*)out, out_stride, nouts, cmap,
                                                      identity, noutput, T, nrows, h, wM);
    GRB_LAUNCH_CHECK();
    return 0;
}

template <int M>
static int launch_fast_tp(int T, const void* x, void* out, long long out_stride, int nouts, const int* cmap,
                          int identity, long long noutput, long long nrows, const float* h, const float2* wM,
                          int nsm, cudaStream_t st)
{
    if (T <= 8)
        return launch_fast<M, 8>(x, out, out_stride, nouts, cmap, identity, noutput, T, nrows, h, wM, nsm, st);
    if (T <= 16)
        return launch_fast<M, 16>(x, out, out_stride, nouts, cmap, identity, noutput, T, nrows, h, wM, nsm, st);
    return launch_fast<M, 32>(x, out, out_stride, nouts, cmap, identity, noutput, T, nrows, h, wM, nsm, st);
}

// x: interleaved input, first T*M items are history.  h: zero-padded prototype [T*M].
// wM: e^{+j 2 pi k / M}, k < M.  cmap: device int[nouts].
int pfb_launch(int M, int T, int R, const void* x, void* out, long long out_stride, int nouts, const int* cmap,
               int identity, long long noutput, const float* h, const float2* wM, int nsm, bool allow_fast,
               cudaStream_t st, const void* halo, const int* inv)
{
    // inv: the channel map inverted (output stream of DFT bin c, -1 = none) when no two outputs share a channel, else null
    const bool mapped_ok = identity || inv != nullptr;
    const int* use_inv = identity ? nullptr : inv;
    if (noutput <= 0)
        return 0;
    const long long nrows = T + (noutput * R) / M; // history rows + consumed rows
    if (allow_fast && mapped_ok && pfb_stream_supported(M, T, R, x) && !GRB_GETENV_ONCE("GRB200_PFB_V1"))
        return pfb_stream_launch(M, T, x, out, out_stride, nouts, noutput, h, wM, nsm, st, halo, use_inv);
    if (halo) // the other kernels read their history in place: bring the neighbour's rows over first (T*M items)
    {
        int rc = halo_fetch_launch(const_cast<void*>(x), halo, (size_t)T * M * 8, st);
        if (rc)
            return rc;
    }
    if (allow_fast && identity && pfb_stream_split_supported(M, T, R, x) && !GRB_GETENV_ONCE("GRB200_PFB_V1") &&
        !GRB_GETENV_ONCE("GRB200_PFB_GENERIC"))
        return pfb_stream_split_launch(M, T, x, out, out_stride, nouts, noutput, h, wM, nsm, st);
    if (allow_fast && mapped_ok && pfb_stream_os2_supported(M, T, R, x) && !GRB_GETENV_ONCE("GRB200_PFB_V1"))
        return pfb_stream_os2_launch(M, T, x, out, out_stride, nouts, noutput, h, wM, nsm, st, use_inv);
    if (allow_fast && R == M && T <= 32 && (M == 16 || M == 64 || M == 256)) {
        if (M == 16)
            return launch_fast_tp<16>(T, x, out, out_stride, nouts, cmap, identity, noutput, nrows, h, wM, nsm, st);
        if (M == 64)
            return launch_fast_tp<64>(T, x, out, out_stride, nouts, cmap, identity, noutput, nrows, h, wM, nsm, st);
        return launch_fast_tp<256>(T, x, out, out_stride, nouts, cmap, identity, noutput, nrows, h, wM, nsm, st);
    }
    if (allow_fast && M >= 8 && M <= 256 && (M & (M - 1)) == 0 && nouts <= 4096 && !GRB_GETENV_ONCE("GRB200_PFB_GENERIC")) {
        switch (M) {
        case 8:
            return launch_fftgen_tp<8>(x, out, out_stride, nouts, cmap, noutput, T, R, h, wM, nsm, st);
        case 16:
            return launch_fftgen_tp<16>(x, out, out_stride, nouts, cmap, noutput, T, R, h, wM, nsm, st);
        case 32:
            return launch_fftgen_tp<32>(x, out, out_stride, nouts, cmap, noutput, T, R, h, wM, nsm, st);
        case 64:
            return launch_fftgen_tp<64>(x, out, out_stride, nouts, cmap, noutput, T, R, h, wM, nsm, st);
        case 128:
            return launch_fftgen_tp<128>(x, out, out_stride, nouts, cmap, noutput, T, R, h, wM, nsm, st);
        case 256:
            return launch_fftgen_tp<256>(x, out, out_stride, nouts, cmap, noutput, T, R, h, wM, nsm, st);
        }
    }
    // generic path: the DFT in two stages when the channel count has a divisor (the one nearest its square root)
    int R1 = 1;
    for (int d = 2; d * d <= M; d++)
        if (M % d == 0)
            R1 = d;
    int OT = 32;
    const int nbuf = R1 > 1 ? 2 : 1;
    size_t smem = ((size_t)nbuf * OT * (M + 1) + M) * 8;
    while (smem > 200 * 1024 && OT > 1) {
        OT /= 2;
        smem = ((size_t)nbuf * OT * (M + 1) + M) * 8;
    }
    if (smem > 200 * 1024)
        return fail(GRB200_ERANGE, "pfb_channelizer: %d channels exceed the device tile", M);
    if (smem > 48 * 1024)
        GRB_CUDA(cudaFuncSetAttribute(pfb_generic_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    long long ntiles = (noutput + OT - 1) / OT;
    long long grid = (long long)nsm * 4;
    if (grid > ntiles)
        grid = ntiles;
    pfb_generic_kernel<<<(unsigned)grid, 256, smem, st>>>((const float2*)x, (float2*)out, out_stride, nouts, cmap,
                                                          noutput, M, T, R, h, wM, OT, R1);
    GRB_LAUNCH_CHECK();
    return 0;
}

int streams_to_interleaved_launch(const void* in, long long in_stride, void* out, int M, long long nrows,
                                  cudaStream_t st)
{
    if (nrows <= 0)
        return 0;
    dim3 grid((unsigned)((nrows + 31) / 32), (unsigned)((M + 31) / 32));
    streams_to_interleaved_kernel<<<grid, 256, 0, st>>>((const float2*)in, in_stride, (float2*)out, M, nrows);
    GRB_LAUNCH_CHECK();
    return 0;
}

} // namespace grb
