// fir.cu -- fir_filter_ccf / fir_filter_ccc direct-convolution kernels (K1/K2 in SURVEY.md section 2a).
//
// Replaces kernel::fir_filter<gr_complex,gr_complex,float|gr_complex>::filterN/filterNdec
// (reference gr-filter/lib/fir_filter.cc:101-141,153-162):  out[i] = sum_t rev[t] * in[i*D + t],
// rev[t] = taps[ntaps-1-t].
//
// B200 mapping.  Writing t = D*u + p splits the decimating FIR into D undecimated correlations
//   out[i] = sum_p sum_u g_p[u] * x_p[i+u],   x_p[m] = in[D*m + p],   g_p[u] = rev[D*u + p],
// so every decimation uses the same sliding-window inner loop.  A CTA owns TILE = NT*R consecutive
// outputs; it stages the D*(TILE+Upad) input samples it needs (tile + halo) into shared memory with
// coalesced 8-byte loads, de-interleaved by phase, and the phase-major taps next to them.  Each thread
// keeps R complex accumulators and an R-deep register ring of the input window: per tap it issues
// 2R (ccf) or 4R (ccc) FFMA against ONE new 8-byte shared load, i.e. 4/R bytes of shared-memory traffic
// per FFMA, which keeps the kernel on the FP32 pipe (the roofline that bounds it: 2*ntaps/D FFMA per
// input sample vs 8+8/D HBM bytes).  Rows are padded by 2 complex every R so the 16-byte window loads of
// a quarter-warp fall in distinct banks.  Results are transposed through shared memory so the global
// stores are coalesced.
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

__device__ __forceinline__ void mac(float2& acc, float h, float2 x)
{
    acc.x = fmaf(h, x.x, acc.x);
    acc.y = fmaf(h, x.y, acc.y);
}
__device__ __forceinline__ void mac(float2& acc, float2 h, float2 x)
{
    acc.x = fmaf(h.x, x.x, acc.x);
    acc.x = fmaf(-h.y, x.y, acc.x);
    acc.y = fmaf(h.x, x.y, acc.y);
    acc.y = fmaf(h.y, x.x, acc.y);
}

template <int R, bool CT>
__global__ void __launch_bounds__(256)
fir_tiled_kernel(const float2* __restrict__ in, float2* __restrict__ out, long long nout,
                 long long nin, int D, int Upad, const typename TapT<CT>::type* __restrict__ gtaps,
                 int ldrow)
{
    using tap_t = typename TapT<CT>::type;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float2* xs = reinterpret_cast<float2*>(smem_raw);
    tap_t* gs = reinterpret_cast<tap_t*>(xs + (size_t)D * ldrow);

    const int NT = blockDim.x;
    const int tid = threadIdx.x;
    const int TILE = NT * R;
    const int Lm = TILE + Upad; // window positions staged per phase
    const long long i0 = (long long)blockIdx.x * TILE;
    const long long g0 = i0 * D;

    // ---- stage taps (phase-major, zero padded on the host) ----
    for (int e = tid; e < D * Upad; e += NT)
        gs[e] = gtaps[e];

    // ---- stage input tile + halo, de-interleaved by phase ----
    if (D == 1) {
        for (int m = tid; m < Lm; m += NT) {
            long long g = g0 + m;
            float2 v = (g < nin) ? in[g] : make_float2(0.f, 0.f);
            xs[m + 2 * (m / R)] = v;
        }
    } else {
        const int total = D * Lm;
        int p = tid % D, m = tid / D;
        const int dp = NT % D, dm = NT / D;
        for (int e = tid; e < total; e += NT) {
            long long g = g0 + e;
            float2 v = (g < nin) ? in[g] : make_float2(0.f, 0.f);
            xs[p * ldrow + m + 2 * (m / R)] = v;
            p += dp;
            m += dm;
            if (p >= D) {
                p -= D;
                m += 1;
            }
        }
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
        float2 ring[R];
#pragma unroll
        for (int k = 0; k < R; k += 2) {
            float4 v = *reinterpret_cast<const float4*>(row + k);
            ring[k] = make_float2(v.x, v.y);
            ring[k + 1] = make_float2(v.z, v.w);
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
            tap_t h[R];
            if (CT) {
#pragma unroll
                for (int k = 0; k < R; k += 2) {
                    float4 v = *reinterpret_cast<const float4*>(gp + u0 + k);
                    reinterpret_cast<float2*>(h)[k] = make_float2(v.x, v.y);
                    reinterpret_cast<float2*>(h)[k + 1] = make_float2(v.z, v.w);
                }
            } else {
#pragma unroll
                for (int k = 0; k < R; k += 4) {
                    float4 v = *reinterpret_cast<const float4*>(gp + u0 + k);
                    reinterpret_cast<float*>(h)[k] = v.x;
                    reinterpret_cast<float*>(h)[k + 1] = v.y;
                    reinterpret_cast<float*>(h)[k + 2] = v.z;
                    reinterpret_cast<float*>(h)[k + 3] = v.w;
                }
            }
#pragma unroll
            for (int uu = 0; uu < R; uu++) {
#pragma unroll
                for (int r = 0; r < R; r++)
                    mac(acc[r], h[uu], ring[(uu + r) % R]);
                ring[uu] = xn[uu];
            }
        }
    }
    __syncthreads();

    // ---- transpose through shared memory, coalesced store ----
    float2* os = xs + tid * (R + 2);
#pragma unroll
    for (int r = 0; r < R; r += 2)
        *reinterpret_cast<float4*>(os + r) = make_float4(acc[r].x, acc[r].y, acc[r + 1].x, acc[r + 1].y);
    __syncthreads();
    for (int m = tid; m < TILE; m += NT) {
        long long i = i0 + m;
        if (i < nout)
            out[i] = xs[m + 2 * (m / R)];
    }
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
        mac(acc, rev[t], x[t]);
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
    int R = 0, NT = 0, Upad = 0, ldrow = 0;
    size_t smem = 0;
    bool tiled = false;
};

static FirPlan plan_fir(int D, int ntaps, bool ct)
{
    FirPlan best;
    const int U = (ntaps + D - 1) / D;
    const int Upad = ((U + 15) / 16) * 16;
    const size_t tapsz = ct ? 8 : 4;
    const int cand[6][2] = { { 16, 256 }, { 16, 128 }, { 8, 256 }, { 8, 128 }, { 8, 64 }, { 8, 32 } };
    for (int pass = 0; pass < 2 && !best.tiled; pass++) {
        const size_t limit = pass == 0 ? 110 * 1024 : 220 * 1024;
        for (auto& c : cand) {
            const int R = c[0], NT = c[1];
            const int Lm = NT * R + Upad;
            const int ldrow = Lm / R * (R + 2);
            const size_t smem = (size_t)D * ldrow * 8 + (size_t)D * Upad * tapsz;
            if (smem <= limit) {
                best.R = R;
                best.NT = NT;
                best.Upad = Upad;
                best.ldrow = ldrow;
                best.smem = smem;
                best.tiled = true;
                break;
            }
        }
    }
    if (!best.tiled)
        best.Upad = Upad;
    return best;
}

template <int R, bool CT>
static int launch_tiled(const FirPlan& pl, const void* in, void* out, long long nout, long long nin,
                        int D, const void* gtaps, cudaStream_t st)
{
    auto kern = fir_tiled_kernel<R, CT>;
    static thread_local size_t configured[16] = { 0 };
    int dev = 0;
    cudaGetDevice(&dev);
    if (pl.smem > 48 * 1024 && configured[dev & 15] < pl.smem) {
        GRB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(220 * 1024)));
        configured[dev & 15] = 220 * 1024;
    }
    const long long tile = (long long)pl.NT * R;
    const long long nblk = (nout + tile - 1) / tile;
    if (nblk > 0x7fffffffLL)
        return fail(GRB200_ERANGE, "fir: too many tiles");
    kern<<<(unsigned)nblk, pl.NT, pl.smem, st>>>((const float2*)in, (float2*)out, nout, nin, D, pl.Upad,
                                                 (const typename TapT<CT>::type*)gtaps, pl.ldrow);
    GRB_LAUNCH_CHECK();
    return 0;
}

// gtaps: phase-major padded taps [D][Upad]; rev: reversed taps [ntaps] (both on the device).
int fir_launch(bool ct, int D, int ntaps, const void* gtaps, const void* rev, const void* in, void* out,
               long long nout, cudaStream_t st)
{
    if (nout <= 0)
        return 0;
    const long long nin = nout * D + ntaps - 1; // readable items behind `in` (block contract)
    FirPlan pl = plan_fir(D, ntaps, ct);
    if (pl.tiled) {
        if (pl.R == 16)
            return ct ? launch_tiled<16, true>(pl, in, out, nout, nin, D, gtaps, st)
                      : launch_tiled<16, false>(pl, in, out, nout, nin, D, gtaps, st);
        return ct ? launch_tiled<8, true>(pl, in, out, nout, nin, D, gtaps, st)
                  : launch_tiled<8, false>(pl, in, out, nout, nin, D, gtaps, st);
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

int fir_upad(int D, int ntaps)
{
    const int U = (ntaps + D - 1) / D;
    return ((U + 15) / 16) * 16;
}

} // namespace grb
