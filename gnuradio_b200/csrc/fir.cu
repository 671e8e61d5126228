// fir.cu -- fir_filter_ccf / fir_filter_ccc direct-convolution kernels (K1/K2 in SURVEY.md section 2a).
//
// Replaces kernel::fir_filter<gr_complex,gr_complex,float|gr_complex>::filterN/filterNdec
// (reference gr-filter/lib/fir_filter.cc:101-141,153-162):  out[i] = sum_t rev[t] * in[i*D + t],
// rev[t] = taps[ntaps-1-t].
//
// B200 mapping.  Writing t = D*u + p splits the decimating FIR into D undecimated correlations
//   out[i] = sum_p sum_u g_p[u] * x_p[i+u],   x_p[m] = in[D*m + p],   g_p[u] = rev[D*u + p],
// so every decimation uses the same sliding-window inner loop.  Persistent CTAs walk over tiles of
// TILE = NT*R consecutive outputs.  A tile's D*(TILE+Upad) input samples (tile + ntaps-1 halo) are staged
// into shared memory with cp.async (16-byte when the stream is 16-byte aligned), de-interleaved by phase;
// with two buffers the next tile streams in while the current one is being filtered, so the FP32 pipe
// never waits on HBM.  Each thread keeps R complex accumulators and an R-deep register ring of its input
// window: per tap it issues R packed FFMA2 (ccf; 2R for ccc) against ONE new 8-byte shared load, i.e.
// 4/R bytes of shared-memory traffic per FMA, which keeps the kernel on the FP32 pipe -- the roofline that
// bounds it: 2*ntaps/D (ccf) FMA per input sample against 8+8/D HBM bytes.  Rows are padded by 2 complex
// every R so the 16-byte window loads of a quarter-warp fall in distinct banks.  Results go through a
// per-warp shared staging area (no CTA barrier) so the global stores are coalesced.
#include <cstdint>
#include <cstdlib>
#include <cmath>
#include "common.cuh"

namespace grb {

template <bool CT>
struct TapT {
    using type = float;
};
template <>
struct TapT<true> {
    using type = float2;
};

// acc += h * x on the packed pipe.  ccc: (hr + j hi)(xr + j xi) = hr*(xr, xi) + hi*(-xi, xr)
__device__ __forceinline__ void mac2(float2& acc, float h, float2 x, float2)
{
    acc = __ffma2_rn(make_float2(h, h), x, acc);
}
__device__ __forceinline__ void mac2(float2& acc, float2 h, float2 x, float2 xswap)
{
    acc = __ffma2_rn(make_float2(h.x, h.x), x, acc);
    acc = __ffma2_rn(make_float2(h.y, h.y), xswap, acc);
}

__device__ __forceinline__ void cp_async8(void* dst, const void* src, unsigned src_bytes)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"((unsigned)__cvta_generic_to_shared(dst)),
                 "l"(src), "r"(src_bytes)
                 : "memory");
}
__device__ __forceinline__ void cp_async16(void* dst, const void* src, unsigned src_bytes)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"((unsigned)__cvta_generic_to_shared(dst)),
                 "l"(src), "r"(src_bytes)
                 : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait()
{
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// Stage one tile (+halo) into `xs`, de-interleaved by phase and skewed by 2 complex every R.
// Items at or beyond `nin` read as zero (cp.async src-size 0 zero-fills).
template <int R>
__device__ __forceinline__ void fir_stage_tile(float2* xs, const float2* __restrict__ in, long long g0, long long nin,
                                               int D, int Lm, int ldrow, int tid, int NT, bool aligned16)
{
    if (D == 1) {
        if (g0 + Lm <= nin) { // interior tile: no bounds, pure pointer increments (NT*2 is a multiple of R)
            if (aligned16) {
                const float2* src = in + g0 + 2 * tid;
                float2* dst = xs + 2 * tid + 2 * ((2 * tid) / R);
                const int dstep = 2 * NT + 2 * (2 * NT / R);
                for (int m = 2 * tid; m < Lm; m += 2 * NT, src += 2 * NT, dst += dstep)
                    cp_async16(dst, src, 16u);
            } else {
                const float2* src = in + g0 + tid;
                float2* dst = xs + tid + 2 * (tid / R);
                const int dstep = NT + 2 * (NT / R);
                for (int m = tid; m < Lm; m += NT, src += NT, dst += dstep)
                    cp_async8(dst, src, 8u);
            }
        } else if (aligned16) {
            for (int m = 2 * tid; m < Lm; m += 2 * NT) {
                const long long g = g0 + m;
                long long left = nin - g;
                left = left < 0 ? 0 : (left > 2 ? 2 : left);
                cp_async16(xs + m + 2 * (m / R), in + (left ? g : 0), (unsigned)left * 8);
            }
        } else {
            for (int m = tid; m < Lm; m += NT) {
                const long long g = g0 + m;
                const bool ok = g < nin;
                cp_async8(xs + m + 2 * (m / R), in + (ok ? g : 0), ok ? 8u : 0u);
            }
        }
    } else {
        const int total = D * Lm;
        int p = tid % D, m = tid / D;
        const int dp = NT % D, dm = NT / D;
        for (int e = tid; e < total; e += NT) {
            const long long g = g0 + e;
            const bool ok = g < nin;
            cp_async8(xs + p * ldrow + m + 2 * (m / R), in + (ok ? g : 0), ok ? 8u : 0u);
            p += dp;
            m += dm;
            if (p >= D) {
                p -= D;
                m += 1;
            }
        }
    }
}

// NBUF = 2: double-buffered staging (prefetch the next tile during compute); NBUF = 1: single buffer.
template <int R, bool CT, int NBUF>
__global__ void __launch_bounds__(256)
fir_tiled_kernel(const float2* __restrict__ in, float2* __restrict__ out, long long nout, long long nin, int D,
                 int Upad, const typename TapT<CT>::type* __restrict__ gtaps, int ldrow, int aligned16)
{
    using tap_t = typename TapT<CT>::type;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int NT = blockDim.x;
    const int tid = threadIdx.x;
    const int lane = tid & 31, warp = tid >> 5;
    const int TILE = NT * R;
    const int Lm = TILE + Upad; // window positions staged per phase
    const size_t bufelems = (size_t)D * ldrow;
    float2* xbuf = reinterpret_cast<float2*>(smem_raw);                      // [NBUF][D][ldrow]
    float2* stage = xbuf + NBUF * bufelems + (size_t)warp * 32 * (R + 2);    // per-warp output staging
    tap_t* gs = reinterpret_cast<tap_t*>(xbuf + NBUF * bufelems + (size_t)(NT / 32) * 32 * (R + 2));

    const long long ntiles = (nout + TILE - 1) / TILE;
    long long tile = blockIdx.x;
    if (tile >= ntiles)
        return;

    // ---- taps (phase-major, zero padded on the host), once per CTA ----
    for (int e = tid; e < D * Upad; e += NT)
        gs[e] = gtaps[e];
    fir_stage_tile<R>(xbuf, in, tile * TILE * D, nin, D, Lm, ldrow, tid, NT, aligned16 != 0);
    cp_async_commit();

    int b = 0;
    for (; tile < ntiles; tile += gridDim.x) {
        float2* xs = xbuf + (size_t)b * bufelems;
        if (NBUF == 2) {
            const long long next = tile + gridDim.x;
            if (next < ntiles)
                fir_stage_tile<R>(xbuf + (size_t)(b ^ 1) * bufelems, in, next * TILE * D, nin, D, Lm, ldrow, tid, NT,
                                  aligned16 != 0);
            cp_async_commit();
            cp_async_wait<1>();
        } else {
            cp_async_wait<0>();
        }
        __syncthreads();

        // ---- sliding-window multiply-accumulate ----
        float2 acc[R];
#pragma unroll
        for (int r = 0; r < R; r++)
            acc[r] = make_float2(0.f, 0.f);

        for (int p = 0; p < D; p++) {
            const float2* row = xs + p * ldrow + tid * (R + 2);
            const tap_t* gp = gs + p * Upad;
            float2 ring[R], rsw[CT ? R : 1];
#pragma unroll
            for (int k = 0; k < R; k += 2) {
                float4 v = *reinterpret_cast<const float4*>(row + k);
                ring[k] = make_float2(v.x, v.y);
                ring[k + 1] = make_float2(v.z, v.w);
                if (CT) {
                    rsw[k] = make_float2(-v.y, v.x);
                    rsw[k + 1] = make_float2(-v.w, v.z);
                }
            }
            for (int u0 = 0; u0 < Upad; u0 += R) {
                const float2* nx = row + (u0 / R + 1) * (R + 2);
                float2 xn[R];
#pragma unroll
                for (int k = 0; k < R; k += 2) {
                    float4 v = *reinterpret_cast<const float4*>(nx + k);
                    xn[k] = make_float2(v.x, v.y);
                    xn[k + 1] = make_float2(v.z, v.w);
                }
                tap_t hh[R];
                if (CT) {
#pragma unroll
                    for (int k = 0; k < R; k += 2) {
                        float4 v = *reinterpret_cast<const float4*>(gp + u0 + k);
                        reinterpret_cast<float2*>(hh)[k] = make_float2(v.x, v.y);
                        reinterpret_cast<float2*>(hh)[k + 1] = make_float2(v.z, v.w);
                    }
                } else {
#pragma unroll
                    for (int k = 0; k < R; k += 4) {
                        float4 v = *reinterpret_cast<const float4*>(gp + u0 + k);
                        reinterpret_cast<float*>(hh)[k] = v.x;
                        reinterpret_cast<float*>(hh)[k + 1] = v.y;
                        reinterpret_cast<float*>(hh)[k + 2] = v.z;
                        reinterpret_cast<float*>(hh)[k + 3] = v.w;
                    }
                }
#pragma unroll
                for (int uu = 0; uu < R; uu++) {
#pragma unroll
                    for (int r = 0; r < R; r++)
                        mac2(acc[r], hh[uu], ring[(uu + r) % R], rsw[CT ? (uu + r) % R : 0]);
                    ring[uu] = xn[uu];
                    if (CT)
                        rsw[uu] = make_float2(-xn[uu].y, xn[uu].x);
                }
            }
        }

        // ---- per-warp transpose through shared staging, coalesced stores ----
        {
            float2* os = stage + lane * (R + 2);
#pragma unroll
            for (int r = 0; r < R; r += 2)
                *reinterpret_cast<float4*>(os + r) = make_float4(acc[r].x, acc[r].y, acc[r + 1].x, acc[r + 1].y);
            __syncwarp();
            const long long w0 = tile * TILE + (long long)warp * 32 * R;
            if (w0 + 32 * R <= nout) {
                float2* o = out + w0 + lane;
                const float2* sp = stage + lane + 2 * (lane / R);
#pragma unroll
                for (int j = 0; j < R; j++) // 32*j advances the skewed index by 32 + 2*(32/R)
                    o[32 * j] = sp[j * (32 + 2 * (32 / R))];
            } else {
#pragma unroll
                for (int j = 0; j < R; j++) {
                    const int m = lane + 32 * j;
                    const long long i = w0 + m;
                    if (i < nout)
                        out[i] = stage[m + 2 * (m / R)];
                }
            }
            __syncwarp();
        }
        __syncthreads(); // every warp is done with this buffer before it is refilled
        if (NBUF == 2) {
            b ^= 1;
        } else {
            const long long next = tile + gridDim.x;
            if (next < ntiles)
                fir_stage_tile<R>(xbuf, in, next * TILE * D, nin, D, Lm, ldrow, tid, NT, aligned16 != 0);
            cp_async_commit();
        }
    }
    cp_async_wait<0>();
}

// Warp-private pipelines: every warp owns its own double-buffered window (32*R outputs + halo) and output
// staging (the consumed window buffer), so after the one-time tap load there is NO CTA barrier: warps drift freely and the scheduler
// always finds FFMA2 work.  Costs Upad/(32*R) extra staged samples per tile; chosen when that is small.
template <int R, bool CT>
__global__ void __launch_bounds__(256)
fir_warptile_kernel(const float2* __restrict__ in, float2* __restrict__ out, long long nout, long long nin, int D,
                    int Upad, const typename TapT<CT>::type* __restrict__ gtaps, int ldrow, int aligned16)
{
    using tap_t = typename TapT<CT>::type;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarp = blockDim.x >> 5;
    constexpr int WT = 32 * R;          // outputs per warp tile
    const int Lm = WT + Upad;           // window positions staged per phase
    const size_t bufelems = (size_t)D * ldrow;
    const size_t per_warp = 2 * bufelems; // output staging reuses the window buffer just consumed
    float2* xbuf = reinterpret_cast<float2*>(smem_raw) + (size_t)warp * per_warp; // [2][D][ldrow]
    tap_t* gs = reinterpret_cast<tap_t*>(reinterpret_cast<float2*>(smem_raw) + (size_t)nwarp * per_warp);

    for (int e = tid; e < D * Upad; e += blockDim.x)
        gs[e] = gtaps[e];
    __syncthreads();

    const long long ntiles = (nout + WT - 1) / WT;
    const long long stride = (long long)gridDim.x * nwarp;
    long long tile = (long long)blockIdx.x * nwarp + warp;
    if (tile >= ntiles)
        return;
    fir_stage_tile<R>(xbuf, in, tile * WT * D, nin, D, Lm, ldrow, lane, 32, aligned16 != 0);
    cp_async_commit();
    int b = 0;
    for (; tile < ntiles; tile += stride, b ^= 1) {
        const float2* xs = xbuf + (size_t)b * bufelems;
        const long long next = tile + stride;
        if (next < ntiles)
            fir_stage_tile<R>(xbuf + (size_t)(b ^ 1) * bufelems, in, next * WT * D, nin, D, Lm, ldrow, lane, 32,
                              aligned16 != 0);
        cp_async_commit();
        cp_async_wait<1>();
        __syncwarp();

        float2 acc[R];
#pragma unroll
        for (int r = 0; r < R; r++)
            acc[r] = make_float2(0.f, 0.f);
        for (int p = 0; p < D; p++) {
            const float2* row = xs + p * ldrow + lane * (R + 2);
            const tap_t* gp = gs + p * Upad;
            float2 ring[R], rsw[CT ? R : 1];
#pragma unroll
            for (int k = 0; k < R; k += 2) {
                float4 v = *reinterpret_cast<const float4*>(row + k);
                ring[k] = make_float2(v.x, v.y);
                ring[k + 1] = make_float2(v.z, v.w);
                if (CT) {
                    rsw[k] = make_float2(-v.y, v.x);
                    rsw[k + 1] = make_float2(-v.w, v.z);
                }
            }
            for (int u0 = 0; u0 < Upad; u0 += R) {
                const float2* nx = row + (u0 / R + 1) * (R + 2);
                float2 xn[R];
#pragma unroll
                for (int k = 0; k < R; k += 2) {
                    float4 v = *reinterpret_cast<const float4*>(nx + k);
                    xn[k] = make_float2(v.x, v.y);
                    xn[k + 1] = make_float2(v.z, v.w);
                }
                __align__(16) tap_t hh[R];
                constexpr int TPV = CT ? 2 : 4; // taps per 16-byte load
#pragma unroll
                for (int k = 0; k < R; k += TPV) {
                    float4 v = *reinterpret_cast<const float4*>(gp + u0 + k);
                    reinterpret_cast<float4*>(hh)[k / TPV] = v;
                }
#pragma unroll
                for (int uu = 0; uu < R; uu++) {
#pragma unroll
                    for (int r = 0; r < R; r++)
                        mac2(acc[r], hh[uu], ring[(uu + r) % R], rsw[CT ? (uu + r) % R : 0]);
                    ring[uu] = xn[uu];
                    if (CT)
                        rsw[uu] = make_float2(-xn[uu].y, xn[uu].x);
                }
            }
        }
        // ---- transpose through the consumed window buffer (>= 32*(R+2) items), coalesced stores ----
        __syncwarp(); // every lane is done reading the window
        float2* stage = xbuf + (size_t)b * bufelems;
        float2* os = stage + lane * (R + 2);
#pragma unroll
        for (int r = 0; r < R; r += 2)
            *reinterpret_cast<float4*>(os + r) = make_float4(acc[r].x, acc[r].y, acc[r + 1].x, acc[r + 1].y);
        __syncwarp();
        const long long w0 = tile * WT;
        if (w0 + WT <= nout) {
            float2* o = out + w0 + lane;
            const float2* sp = stage + lane + 2 * (lane / R);
#pragma unroll
            for (int j = 0; j < R; j++)
                o[32 * j] = sp[j * (32 + 2 * (32 / R))];
        } else {
#pragma unroll
            for (int j = 0; j < R; j++) {
                const int m = lane + 32 * j;
                if (w0 + m < nout)
                    out[w0 + m] = stage[m + 2 * (m / R)];
            }
        }
        __syncwarp(); // staging and this buffer are free again
    }
    cp_async_wait<0>();
}

// Long filters: split the TAPS across the warps of a CTA instead of the outputs.  Every warp works on the same
// 32*R outputs and owns a contiguous range of tap chunks (a chunk = R taps of one decimation phase); the W
// partial sums meet in shared memory.  The staged window is then only D*(32*R + Upad) items however long the
// filter is, so it can be double buffered with 2 CTAs (16 warps) per SM, where the output-tiled kernel above
// fits one single-buffered 4-warp CTA and leaves the FP32 pipe idle while it stages (ncu, profiles/
// r1_firdec_ncu.txt: 10 % occupancy, FMA pipe 63 % active, 27 % of shared wavefronts are staging conflicts).
template <int R, bool CT>
__global__ void __launch_bounds__(256)
fir_splitk_kernel(const float2* __restrict__ in, float2* __restrict__ out, long long nout, long long nin, int D,
                  int Upad, const typename TapT<CT>::type* __restrict__ gtaps, int ldrow, int aligned16)
{
    using tap_t = typename TapT<CT>::type;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int W = 8;               // warps per CTA
    constexpr int WT = 32 * R;         // outputs per tile
    constexpr int RS = 32 * (R + 2);   // one warp's partial sums (skewed like the input rows)
    const int Lm = WT + Upad;
    const size_t bufelems = (size_t)D * ldrow; // >= W*RS (host checks): the partial sums reuse the consumed buffer
    float2* xbuf = reinterpret_cast<float2*>(smem_raw); // [2][D][ldrow]
    tap_t* gs = reinterpret_cast<tap_t*>(xbuf + 2 * bufelems);

    const int UB = Upad / R;           // chunks per phase
    const int C = D * UB;
    const int per = (C + W - 1) / W;
    const int c_begin = warp * per, c_end = min(C, c_begin + per);

    for (int e = tid; e < D * Upad; e += 256)
        gs[e] = gtaps[e];
    const long long ntiles = (nout + WT - 1) / WT;
    long long tile = blockIdx.x;
    if (tile >= ntiles)
        return;
    fir_stage_tile<R>(xbuf, in, tile * WT * D, nin, D, Lm, ldrow, tid, 256, aligned16 != 0);
    cp_async_commit();

    int b = 0;
    for (; tile < ntiles; tile += gridDim.x, b ^= 1) {
        float2* xs = xbuf + (size_t)b * bufelems;
        cp_async_wait<0>();
        __syncthreads(); // this tile has landed for everyone; the previous tile's reduction is over
        const long long next = tile + gridDim.x;
        if (next < ntiles)
            fir_stage_tile<R>(xbuf + (size_t)(b ^ 1) * bufelems, in, next * WT * D, nin, D, Lm, ldrow, tid, 256,
                              aligned16 != 0);
        cp_async_commit();

        float2 acc[R];
#pragma unroll
        for (int r = 0; r < R; r++)
            acc[r] = make_float2(0.f, 0.f);
        float2 ring[R], rsw[CT ? R : 1];
        int p_cur = -1;
        for (int c = c_begin; c < c_end; c++) {
            const int p = c / UB, ub = c - p * UB;
            const float2* row = xs + p * ldrow + lane * (R + 2) + ub * (R + 2);
            if (p != p_cur) { // first chunk of a phase for this warp: load the window afresh
                p_cur = p;
#pragma unroll
                for (int k = 0; k < R; k += 2) {
                    float4 v = *reinterpret_cast<const float4*>(row + k);
                    ring[k] = make_float2(v.x, v.y);
                    ring[k + 1] = make_float2(v.z, v.w);
                    if (CT) {
                        rsw[k] = make_float2(-v.y, v.x);
                        rsw[k + 1] = make_float2(-v.w, v.z);
                    }
                }
            }
            const float2* nx = row + (R + 2);
            float2 xn[R];
#pragma unroll
            for (int k = 0; k < R; k += 2) {
                float4 v = *reinterpret_cast<const float4*>(nx + k);
                xn[k] = make_float2(v.x, v.y);
                xn[k + 1] = make_float2(v.z, v.w);
            }
            __align__(16) tap_t hh[R];
            constexpr int TPV = CT ? 2 : 4; // taps per 16-byte load
            const tap_t* gp = gs + p * Upad + ub * R;
#pragma unroll
            for (int k = 0; k < R; k += TPV)
                reinterpret_cast<float4*>(hh)[k / TPV] = *reinterpret_cast<const float4*>(gp + k);
#pragma unroll
            for (int uu = 0; uu < R; uu++) {
#pragma unroll
                for (int r = 0; r < R; r++)
                    mac2(acc[r], hh[uu], ring[(uu + r) % R], rsw[CT ? (uu + r) % R : 0]);
                ring[uu] = xn[uu];
                if (CT)
                    rsw[uu] = make_float2(-xn[uu].y, xn[uu].x);
            }
        }
        __syncthreads(); // every warp is done reading this buffer
        {
            float2* os = xs + warp * RS + lane * (R + 2);
#pragma unroll
            for (int r = 0; r < R; r += 2)
                *reinterpret_cast<float4*>(os + r) = make_float4(acc[r].x, acc[r].y, acc[r + 1].x, acc[r + 1].y);
        }
        __syncthreads();
        const long long w0 = tile * WT;
#pragma unroll
        for (int o = tid; o < WT; o += 256) {
            const float2* rp = xs + o + 2 * (o / R);
            float2 sum = rp[0];
#pragma unroll
            for (int w = 1; w < W; w++)
                sum = __fadd2_rn(sum, rp[w * RS]);
            if (w0 + o < nout)
                out[w0 + o] = sum;
        }
    }
    cp_async_wait<0>();
}

// Fallback for geometries whose tile does not fit shared memory (very large decimation or tap count):
// one warp per output, lanes stride over the taps, shuffle reduction.
template <bool CT>
__global__ void __launch_bounds__(256)
fir_warp_kernel(const float2* __restrict__ in, float2* __restrict__ out, long long nout, int D,
                int ntaps, const typename TapT<CT>::type* __restrict__ rev)
{
    const long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (w >= nout)
        return;
    const float2* x = in + w * D;
    float2 acc = make_float2(0.f, 0.f);
    for (int t = lane; t < ntaps; t += 32)
        mac2(acc, rev[t], x[t], make_float2(-x[t].y, x[t].x));
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) {
        acc.x += __shfl_xor_sync(0xffffffffu, acc.x, s);
        acc.y += __shfl_xor_sync(0xffffffffu, acc.y, s);
    }
    if (lane == 0)
        out[w] = acc;
}

// ---- host-side launch ---------------------------------------------------------------------------
struct FirPlan {
    int R = 0, NT = 0, NBUF = 0, Upad = 0, ldrow = 0;
    size_t smem = 0;
    bool tiled = false, warptile = false, splitk = false;
};

static size_t fir_smem(int D, int Upad, bool ct, int R, int NT, int NBUF, int* ldrow_out)
{
    const int Lm = NT * R + Upad;
    const int ldrow = Lm / R * (R + 2);
    if (ldrow_out)
        *ldrow_out = ldrow;
    return (size_t)NBUF * D * ldrow * 8 + (size_t)NT * (R + 2) * 8 + (size_t)D * Upad * (ct ? 8 : 4);
}

static FirPlan plan_fir(int D, int ntaps, bool ct)
{
    FirPlan best;
    const int U = (ntaps + D - 1) / D;
    best.Upad = ((U + 15) / 16) * 16;
    const size_t two_per_sm = 113 * 1024, one_per_sm = 226 * 1024;
    auto take = [&](int R, int NT, int NBUF, size_t limit) {
        if (best.tiled || (ct && R == 16))
            return;
        int ld = 0;
        const size_t s = fir_smem(D, best.Upad, ct, R, NT, NBUF, &ld);
        if (s <= limit) {
            best.R = R, best.NT = NT, best.NBUF = NBUF, best.ldrow = ld, best.smem = s, best.tiled = true;
        }
    };
    // warp-private pipelines when a warp's double-buffered window is small (short filters)
    auto take_warptile = [&](int R, int NT) {
        if (best.tiled || (ct && R == 16))
            return;
        const int Lm = 32 * R + best.Upad;
        const int ld = Lm / R * (R + 2);
        const size_t per_warp = (size_t)2 * D * ld * 8;
        if (per_warp <= 12 * 1024) {
            best.R = R, best.NT = NT, best.NBUF = 2, best.ldrow = ld, best.tiled = true, best.warptile = true;
            best.smem = per_warp * (NT / 32) + (size_t)D * best.Upad * (ct ? 8 : 4);
        }
    };
    if (const char* e = GRB_GETENV_ONCE("GRB200_FIR_PLAN")) { // tuning knob: "R,NT,NBUF" (NBUF 3 = warp-private pipelines)
        int R = 0, NT = 0, NB = 0;
        if (sscanf(e, "%d,%d,%d", &R, &NT, &NB) == 3 && (R == 8 || R == 16) && NT >= 32 && NT <= 256 && NT % 32 == 0) {
            if (NB == 1 || NB == 2)
                take(R, NT, NB, one_per_sm);
            else if (NB == 3)
                take_warptile(R, NT);
        }
    }
    if (best.tiled)
        return best;
    take_warptile(16, 128);
    take_warptile(8, 128);
    // long filters: taps split across the 8 warps of a CTA, double-buffered window, 2 CTAs per SM
    auto take_splitk = [&](int R) {
        if (best.tiled || (ct && R == 16))
            return;
        const int Lm = 32 * R + best.Upad;
        int ld = Lm / R * (R + 2);
        if (D > 1) // phase rows 16 bytes apart modulo 128: the D-strided staging stores spread over all banks
            while (ld % 16 != 2)
                ld += 2;
        const size_t buf = (size_t)D * ld * 8;
        const size_t s = 2 * buf + (size_t)D * best.Upad * (ct ? 8 : 4);
        const int chunks = D * (best.Upad / R);
        if (chunks >= 16 && buf >= (size_t)8 * 32 * (R + 2) * 8 && s <= two_per_sm) {
            best.R = R, best.NT = 256, best.NBUF = 2, best.ldrow = ld, best.smem = s, best.tiled = true, best.splitk = true;
        }
    };
    static const bool allow_splitk = [] {
        const char* e = GRB_GETENV_ONCE("GRB200_FIR_SPLITK");
        return !(e && e[0] == '0');
    }();
    if (allow_splitk) {
        take_splitk(16);
        take_splitk(8);
    }
    const int cand[4][2] = { { 16, 256 }, { 16, 128 }, { 8, 256 }, { 8, 128 } };
    for (auto& c : cand)
        take(c[0], c[1], 2, two_per_sm);
    for (auto& c : cand)
        take(c[0], c[1], 1, two_per_sm);
    take(8, 64, 2, two_per_sm);
    take(8, 64, 1, two_per_sm);
    for (auto& c : cand)
        take(c[0], c[1], 1, one_per_sm);
    take(8, 64, 1, one_per_sm);
    take(8, 32, 1, one_per_sm);
    return best;
}

template <int R, bool CT, int NBUF>
static int launch_tiled(const FirPlan& pl, const void* in, void* out, long long nout, long long nin, int D,
                        const void* gtaps, int nsm, cudaStream_t st)
{
    auto kern = fir_tiled_kernel<R, CT, NBUF>;
    static thread_local bool configured[16] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (!configured[dev & 15]) {
        GRB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(226 * 1024)));
        configured[dev & 15] = true;
    }
    int occ = 1;
    GRB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, pl.NT, pl.smem));
    if (occ < 1)
        occ = 1;
    const long long tile = (long long)pl.NT * R;
    const long long ntiles = (nout + tile - 1) / tile;
    long long grid = (long long)nsm * occ;
    if (grid > ntiles)
        grid = ntiles;
    const int aligned16 = (((uintptr_t)in) & 15) == 0;
    kern<<<(unsigned)grid, pl.NT, pl.smem, st>>>((const float2*)in, (float2*)out, nout, nin, D, pl.Upad,
                                                 (const typename TapT<CT>::type*)gtaps, pl.ldrow, aligned16);
    GRB_LAUNCH_CHECK();
    return 0;
}

template <int R, bool CT>
static int launch_warptile(const FirPlan& pl, const void* in, void* out, long long nout, long long nin, int D,
                           const void* gtaps, int nsm, cudaStream_t st)
{
    auto kern = fir_warptile_kernel<R, CT>;
    static thread_local bool configured[16] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (!configured[dev & 15]) {
        GRB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(226 * 1024)));
        configured[dev & 15] = true;
    }
    int occ = 1;
    GRB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, pl.NT, pl.smem));
    if (occ < 1)
        occ = 1;
    const long long ntiles = (nout + 32 * R - 1) / (32 * R);
    const long long wpc = pl.NT / 32;
    long long grid = (long long)nsm * occ;
    if (grid > (ntiles + wpc - 1) / wpc)
        grid = (ntiles + wpc - 1) / wpc;
    const int aligned16 = (((uintptr_t)in) & 15) == 0;
    kern<<<(unsigned)grid, pl.NT, pl.smem, st>>>((const float2*)in, (float2*)out, nout, nin, D, pl.Upad,
                                                 (const typename TapT<CT>::type*)gtaps, pl.ldrow, aligned16);
    GRB_LAUNCH_CHECK();
    return 0;
}

template <int R, bool CT>
static int launch_splitk(const FirPlan& pl, const void* in, void* out, long long nout, long long nin, int D,
                         const void* gtaps, int nsm, cudaStream_t st)
{
    auto kern = fir_splitk_kernel<R, CT>;
    static thread_local bool configured[16] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (!configured[dev & 15]) {
        GRB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(226 * 1024)));
        configured[dev & 15] = true;
    }
    int occ = 1;
    GRB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, 256, pl.smem));
    if (occ < 1)
        occ = 1;
    const long long ntiles = (nout + 32 * R - 1) / (32 * R);
    long long grid = (long long)nsm * occ;
    if (grid > ntiles)
        grid = ntiles;
    const int aligned16 = (((uintptr_t)in) & 15) == 0;
    kern<<<(unsigned)grid, 256, pl.smem, st>>>((const float2*)in, (float2*)out, nout, nin, D, pl.Upad,
                                               (const typename TapT<CT>::type*)gtaps, pl.ldrow, aligned16);
    GRB_LAUNCH_CHECK();
    return 0;
}

template <int R, bool CT>
static int launch_tiled_nb(const FirPlan& pl, const void* in, void* out, long long nout, long long nin, int D,
                           const void* gtaps, int nsm, cudaStream_t st)
{
    if (pl.splitk)
        return launch_splitk<R, CT>(pl, in, out, nout, nin, D, gtaps, nsm, st);
    if (pl.warptile)
        return launch_warptile<R, CT>(pl, in, out, nout, nin, D, gtaps, nsm, st);
    return pl.NBUF == 2 ? launch_tiled<R, CT, 2>(pl, in, out, nout, nin, D, gtaps, nsm, st)
                        : launch_tiled<R, CT, 1>(pl, in, out, nout, nin, D, gtaps, nsm, st);
}

int fir_tc_mode(bool ct, int D, int ntaps, const void* in, long long nout, int tc_pref);
int fir_tc_launch(int mode, int ntaps, const void* rev, const void* in, void* out, long long nout, long long nin, int nsm,
                  cudaStream_t st, const void* halo);
int halo_fetch_launch(void* dst, const void* src, size_t bytes, cudaStream_t st);

// gtaps: phase-major padded taps [D][Upad]; rev: reversed taps [ntaps] (both on the device).
int fir_launch(bool ct, int D, int ntaps, const void* gtaps, const void* rev, const void* in, void* out,
               long long nout, int tc_pref, cudaStream_t st, const void* halo)
{
    if (nout <= 0)
        return 0;
    const long long nin = nout * D + ntaps - 1; // readable items behind `in` (block contract)
    FirPlan pl = plan_fir(D, ntaps, ct);
    int dev = 0;
    cudaGetDevice(&dev);
    const int nsm = sm_count(dev);
    const int tc_mode = fir_tc_mode(ct, D, ntaps, in, nout, tc_pref); // tcgen05 block-Toeplitz paths (fir_tc.cu)
    if (tc_mode >= 0) // the tensor-core kernels read a history halo in place (peer memory)
        return fir_tc_launch(tc_mode, ntaps, rev, in, out, nout, nin, nsm, st, halo);
    if (halo && ntaps > 1) { // the CUDA-core kernels stage through cp.async: bring the halo over first (one-CTA kernel)
        int rc = halo_fetch_launch(const_cast<void*>(in), halo, (size_t)(ntaps - 1) * 8, st);
        if (rc)
            return rc;
    }
    if (pl.tiled) {
        if (ct)
            return launch_tiled_nb<8, true>(pl, in, out, nout, nin, D, gtaps, nsm, st);
        return pl.R == 16 ? launch_tiled_nb<16, false>(pl, in, out, nout, nin, D, gtaps, nsm, st)
                          : launch_tiled_nb<8, false>(pl, in, out, nout, nin, D, gtaps, nsm, st);
    }
    const long long nthreads = nout * 32;
    const long long nblk = (nthreads + 255) / 256;
    if (nblk > 0x7fffffffLL)
        return fail(GRB200_ERANGE, "fir: too many outputs for the fallback kernel");
    if (ct)
        fir_warp_kernel<true><<<(unsigned)nblk, 256, 0, st>>>((const float2*)in, (float2*)out, nout, D, ntaps,
                                                              (const float2*)rev);
    else
        fir_warp_kernel<false><<<(unsigned)nblk, 256, 0, st>>>((const float2*)in, (float2*)out, nout, D, ntaps,
                                                               (const float*)rev);
    GRB_LAUNCH_CHECK();
    return 0;
}

// ---- rotator (gr-blocks/include/gnuradio/blocks/rotator.h:33-53 behind freq_xlating_fir_filter_impl::work :168-178) ----
// The reference multiplies by a float phasor that it advances by complex multiplication (renormalised every 512 items):
// phase_n = incr^n up to a rounding random walk.  The device evaluates the closed form: angle in float64, reduced to
// [-pi, pi], then sincosf -- no state in the kernel, any chunking gives the same output.
__global__ void __launch_bounds__(256) rotate_kernel(float2* __restrict__ y, long long n, double theta, double delta)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n)
        return;
    const double two_pi = 6.283185307179586476925286766559;
    double a = fma((double)i, delta, theta);
    a -= two_pi * rint(a / two_pi);
    float sn, cs;
    sincosf((float)a, &sn, &cs);
    const float2 v = y[i];
    y[i] = make_float2(v.x * cs - v.y * sn, v.x * sn + v.y * cs);
}

// The history halo of a time-sharded segment, brought over from the left neighbour's memory by plain loads (NVLink peer
// access): a one-CTA kernel costs ~2 us on the stream where a peer cudaMemcpyAsync of the same 0.5 - 8 KiB costs ~13.
__global__ void __launch_bounds__(256) halo_fetch_kernel(float* __restrict__ dst, const float* __restrict__ src, int nwords)
{
    for (int i = threadIdx.x; i < nwords; i += blockDim.x)
        dst[i] = src[i];
}

int halo_fetch_launch(void* dst, const void* src, size_t bytes, cudaStream_t st)
{
    if (bytes == 0)
        return 0;
    halo_fetch_kernel<<<1, 256, 0, st>>>((float*)dst, (const float*)src, (int)(bytes / 4));
    GRB_LAUNCH_CHECK();
    return 0;
}

int rotate_launch(void* y, long long n, double theta, double delta, cudaStream_t st)
{
    if (n <= 0)
        return 0;
    const long long nblk = (n + 255) / 256;
    if (nblk > 0x7fffffffLL)
        return fail(GRB200_ERANGE, "rotator: too many items");
    rotate_kernel<<<(unsigned)nblk, 256, 0, st>>>((float2*)y, n, theta, delta);
    GRB_LAUNCH_CHECK();
    return 0;
}

double rotate_advance(double theta, double delta, long long n)
{
    const double two_pi = 6.283185307179586476925286766559;
    double a = theta + (double)n * delta;
    return a - two_pi * std::rint(a / two_pi);
}

int fir_upad(int D, int ntaps)
{
    const int U = (ntaps + D - 1) / D;
    return ((U + 15) / 16) * 16;
}

} // namespace grb
