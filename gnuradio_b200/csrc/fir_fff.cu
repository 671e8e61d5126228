// fir_fff.cu -- fir_filter_fff: real stream, real taps.
//
// Replaces kernel::fir_filter<float,float,float>::filter / filterN / filterNdec (reference gr-filter/lib/fir_filter.cc:101-130,
// volk_32f_x2_dot_prod_32f_a) under fir_filter_blk_impl<float,float,float>::work (gr-filter/lib/fir_filter_blk_impl.cc:67-90):
//     out[i] = sum_t rev[t] * in[i*D + t],   rev[t] = taps[ntaps-1-t].
//
// B200 mapping (decimation 1, the 256-tap benchmark shape of the reference's mp-sched perf data): a real FIR is FP32-bound
// almost at once (ntaps FMA per 8 HBM bytes), so the kernel is built around the packed pipe.  Two neighbouring outputs form
// one FFMA2 accumulator:
//     (out[2m], out[2m+1]) = sum_s rev[2s] * z[m+s] + rev[2s+1] * w[m+s],   z[k] = (x[2k], x[2k+1]),  w[k] = (x[2k+1], x[2k+2]),
// i.e. even taps meet the aligned pairs of the stream and odd taps the pairs shifted by one float.  A thread owns FR = 8
// consecutive output pairs and keeps BOTH pair windows in register rings (w is formed from z with two register moves per new
// item), so a step of two taps costs 16 FFMA2 against one 8-byte shared-memory load: 4 real multiply-adds per issued FMA
// instruction, the taps arrive as broadcast 16-byte loads.  Tiles of 4096 outputs (+ halo) are staged with cp.async (16-byte
// when the stream is 16-byte aligned) into a double buffer whose rows are skewed by 2 pairs every 8 -- the layout of fir.cu,
// conflict-free for the 16-byte window loads -- and results leave through a per-warp transpose so the stores are 256-byte runs.
// Algorithmic traffic 4 + 4 B per sample; 256 taps: 256 FMA per sample, FP32-bound (the HBM roofline is reported beside it).
//
// Decimation > 1 and very long filters take a plain one-output-per-thread kernel (window and taps through L1): fir_filter_fff
// with decimation is an audio-rate block in practice.
#include <cstdint>
#include "common.cuh"

namespace grb {

namespace fff {
constexpr int FR = 8;            // output pairs per thread
constexpr int NT = 256;
constexpr int TILEP = NT * FR;   // output pairs per tile
constexpr int ROW = FR + 2;      // skewed row length (pairs)
constexpr int STAGE = 32 * ROW;  // per-warp output staging (pairs)
constexpr int MAX_TAPS = 8192;
} // namespace fff

__device__ __forceinline__ void fff_cp_async4(void* dst, const void* src, unsigned src_bytes)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"((unsigned)__cvta_generic_to_shared(dst)), "l"(src),
                 "r"(src_bytes)
                 : "memory");
}
__device__ __forceinline__ void fff_cp_async16(void* dst, const void* src, unsigned src_bytes)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"((unsigned)__cvta_generic_to_shared(dst)), "l"(src),
                 "r"(src_bytes)
                 : "memory");
}

// Stage pairs [0, npairs) of the tile whose first float is g0 into the skewed buffer; floats at or beyond nin read as zero.
template <bool ALIGNED>
__device__ __forceinline__ void fff_stage(float2* zs, const float* __restrict__ in, long long g0, long long nin, int npairs,
                                          int tid)
{
    using namespace fff;
    if (ALIGNED) {
        for (int m = 2 * tid; m < npairs; m += 2 * NT) { // two pairs = 16 bytes; both fall in one 8-pair row
            const long long g = g0 + 2 * (long long)m;
            long long left = nin - g;
            left = left < 0 ? 0 : (left > 4 ? 4 : left);
            fff_cp_async16(zs + m + 2 * (m / FR), in + (left ? g : 0), (unsigned)left * 4);
        }
    } else {
        float* fs = reinterpret_cast<float*>(zs);
        for (int f = tid; f < 2 * npairs; f += NT) {
            const long long g = g0 + f;
            const bool ok = g < nin;
            const int m = f >> 1;
            fff_cp_async4(fs + 2 * (m + 2 * (m / FR)) + (f & 1), in + (ok ? g : 0), ok ? 4u : 0u);
        }
    }
}

// SP: tap pairs, a multiple of FR (taps zero padded on the host); revp: the padded reversed taps (2*SP floats).
template <bool ALIGNED>
__global__ void __launch_bounds__(fff::NT)
fir_fff_kernel(const float* __restrict__ in, float* __restrict__ out, long long nout, long long nin, int SP,
               const float* __restrict__ revp, int out_aligned)
{
    using namespace fff;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nrows = NT + SP / FR + 1;         // staged rows of FR pairs: the tile, the halo, one row for the shifted pairs
    const int npairs = nrows * FR;
    const size_t bufelems = (size_t)nrows * ROW;
    float2* zbuf = reinterpret_cast<float2*>(smem_raw);                 // [2][nrows][ROW]
    float2* stage = zbuf + 2 * bufelems + (size_t)warp * STAGE;         // per-warp output staging
    float2* gs = zbuf + 2 * bufelems + (size_t)(NT / 32) * STAGE;       // [SP] (even tap, odd tap)

    const long long npairs_out = (nout + 1) / 2;
    const long long ntiles = (npairs_out + TILEP - 1) / TILEP;
    long long tile = blockIdx.x;
    if (tile >= ntiles)
        return;
    for (int e = tid; e < SP; e += NT)
        gs[e] = make_float2(revp[2 * e], revp[2 * e + 1]);
    fff_stage<ALIGNED>(zbuf, in, tile * 2 * TILEP, nin, npairs, tid);
    asm volatile("cp.async.commit_group;" ::: "memory");

    int b = 0;
    for (; tile < ntiles; tile += gridDim.x, b ^= 1) {
        const float2* zs = zbuf + (size_t)b * bufelems;
        const long long next = tile + gridDim.x;
        if (next < ntiles)
            fff_stage<ALIGNED>(zbuf + (size_t)(b ^ 1) * bufelems, in, next * 2 * TILEP, nin, npairs, tid);
        asm volatile("cp.async.commit_group;" ::: "memory");
        asm volatile("cp.async.wait_group 1;" ::: "memory");
        __syncthreads();

        float2 acc[FR];
#pragma unroll
        for (int r = 0; r < FR; r++)
            acc[r] = make_float2(0.f, 0.f);
        const float2* row = zs + tid * ROW;
        float2 ring[FR], wring[FR];
#pragma unroll
        for (int k = 0; k < FR; k += 2) {
            const float4 v = *reinterpret_cast<const float4*>(row + k);
            ring[k] = make_float2(v.x, v.y);
            ring[k + 1] = make_float2(v.z, v.w);
        }
        {
            const float2 nx0 = row[ROW];
#pragma unroll
            for (int k = 0; k < FR - 1; k++)
                wring[k] = make_float2(ring[k].y, ring[k + 1].x);
            wring[FR - 1] = make_float2(ring[FR - 1].y, nx0.x);
        }
        for (int c = 0; c < SP / FR; c++) {
            const float2* nx = row + (c + 1) * ROW;
            float2 xn[FR + 1];
#pragma unroll
            for (int k = 0; k < FR; k += 2) {
                const float4 v = *reinterpret_cast<const float4*>(nx + k);
                xn[k] = make_float2(v.x, v.y);
                xn[k + 1] = make_float2(v.z, v.w);
            }
            xn[FR] = nx[ROW];
            float2 hh[FR];
#pragma unroll
            for (int k = 0; k < FR; k += 2) {
                const float4 v = *reinterpret_cast<const float4*>(gs + c * FR + k);
                hh[k] = make_float2(v.x, v.y);
                hh[k + 1] = make_float2(v.z, v.w);
            }
#pragma unroll
            for (int uu = 0; uu < FR; uu++) {
                const float2 he = make_float2(hh[uu].x, hh[uu].x), ho = make_float2(hh[uu].y, hh[uu].y);
#pragma unroll
                for (int r = 0; r < FR; r++) {
                    acc[r] = __ffma2_rn(he, ring[(uu + r) % FR], acc[r]);
                    acc[r] = __ffma2_rn(ho, wring[(uu + r) % FR], acc[r]);
                }
                ring[uu] = xn[uu];
                wring[uu] = make_float2(xn[uu].y, xn[uu + 1].x);
            }
        }

        // ---- per-warp transpose, coalesced stores ----
        {
            float2* os = stage + lane * ROW;
#pragma unroll
            for (int r = 0; r < FR; r += 2)
                *reinterpret_cast<float4*>(os + r) = make_float4(acc[r].x, acc[r].y, acc[r + 1].x, acc[r + 1].y);
            __syncwarp();
            const long long w0 = tile * TILEP + (long long)warp * 32 * FR; // first output pair of this warp
            if (out_aligned && 2 * (w0 + 32 * FR) <= nout) {
                float2* o = reinterpret_cast<float2*>(out) + w0 + lane;
                const float2* sp = stage + lane + 2 * (lane / FR);
#pragma unroll
                for (int j = 0; j < FR; j++)
                    o[32 * j] = sp[j * (32 + 2 * (32 / FR))];
            } else {
#pragma unroll
                for (int j = 0; j < FR; j++) {
                    const int m = lane + 32 * j;
                    const long long i = 2 * (w0 + m);
                    const float2 v = stage[m + 2 * (m / FR)];
                    if (i < nout)
                        out[i] = v.x;
                    if (i + 1 < nout)
                        out[i + 1] = v.y;
                }
            }
            __syncwarp();
        }
        __syncthreads(); // every warp is done with this buffer before it is refilled
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
}

// Everything else (decimation > 1, more than fff::MAX_TAPS taps): one output per thread, four partial sums.
__global__ void __launch_bounds__(256)
fir_fff_generic_kernel(const float* __restrict__ in, float* __restrict__ out, long long nout, int D, int ntaps,
                       const float* __restrict__ rev)
{
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nout; i += stride) {
        const float* x = in + i * D;
        float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
        int t = 0;
        for (; t + 4 <= ntaps; t += 4) {
            a0 = fmaf(__ldg(rev + t), __ldg(x + t), a0);
            a1 = fmaf(__ldg(rev + t + 1), __ldg(x + t + 1), a1);
            a2 = fmaf(__ldg(rev + t + 2), __ldg(x + t + 2), a2);
            a3 = fmaf(__ldg(rev + t + 3), __ldg(x + t + 3), a3);
        }
        for (; t < ntaps; t++)
            a0 = fmaf(__ldg(rev + t), __ldg(x + t), a0);
        out[i] = (a0 + a1) + (a2 + a3);
    }
}

// number of padded taps the fast kernel wants (a multiple of 2*FR)
int fir_fff_pad(int ntaps) { return ((ntaps + 2 * fff::FR - 1) / (2 * fff::FR)) * (2 * fff::FR); }

// revp: reversed taps zero padded to fir_fff_pad(ntaps) floats; rev: the same array (the first ntaps floats are the taps).
int fir_fff_launch(int D, int ntaps, const float* revp, const void* in, void* out, long long nout, int nsm, cudaStream_t st)
{
    using namespace fff;
    if (nout <= 0)
        return 0;
    const long long nin = nout * D + ntaps - 1;
    if (D == 1 && ntaps <= MAX_TAPS) {
        const int SP = fir_fff_pad(ntaps) / 2;
        const int nrows = NT + SP / FR + 1;
        const size_t smem = ((size_t)2 * nrows * ROW + (size_t)(NT / 32) * STAGE + SP) * sizeof(float2);
        const bool aligned = (((uintptr_t)in) & 15) == 0;
        auto kern = aligned ? fir_fff_kernel<true> : fir_fff_kernel<false>;
        GRB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        int occ = 1;
        GRB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, NT, smem));
        if (occ < 1)
            occ = 1;
        const long long ntiles = ((nout + 1) / 2 + TILEP - 1) / TILEP;
        long long grid = (long long)nsm * occ;
        if (grid > ntiles)
            grid = ntiles;
        kern<<<(unsigned)grid, NT, smem, st>>>((const float*)in, (float*)out, nout, nin, SP, revp,
                                                (((uintptr_t)out) & 7) == 0 ? 1 : 0);
        GRB_LAUNCH_CHECK();
        return 0;
    }
    long long need = (nout + 255) / 256;
    long long grid = (long long)nsm * 16;
    if (grid > need)
        grid = need;
    fir_fff_generic_kernel<<<(unsigned)grid, 256, 0, st>>>((const float*)in, (float*)out, nout, D, ntaps, revp);
    GRB_LAUNCH_CHECK();
    return 0;
}

} // namespace grb
