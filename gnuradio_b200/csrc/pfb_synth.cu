// pfb_synth.cu -- pfb_synthesizer_ccf: the channelizer transposed.
//
// Replaces pfb_synthesizer_ccf_impl::work (reference gr-filter/lib/pfb_synthesizer_ccf_impl.cc:221-284) with the tap
// partitions of set_taps1 / set_taps2 (:95-184).  Closed form from the filters' delay lines (fir_filter_with_buffer.cc):
//     V_q[m]        = sum_i X_i[m] e^{+j 2 pi map[i] q / NF}          (NF = twox * M point backward FFT over the channels)
//     out[n M + i]  = sum_j h[i + j M] V_q[n - j],   q = i (1x);  q = ((parity(n) ^ 1) M + i) (2x: of the filter pair of branch
//                     i, the one with the even taps reads the half of the IFFT that alternates with n)
// (checked against the reference's own classes in tests/test_oracle.py).
//
// pfb_synth_kernel (1x, M a power of two 8..256, up to 32 taps per branch, any channel map, any number of connected inputs):
// pfb_stream_kernel read right to left.  A persistent CTA owns NG contiguous runs of input times; per iteration ONE 2-D TMA
// tensor copy per run (cp.async.bulk.tensor: box = KT times x M streams out of the stream-major input, rows beyond the
// connected inputs and items beyond the end arrive as zeros) lands the next KT columns; the M-point backward DFT runs over
// the channels in shared memory (radix sqrt(M) x sqrt(M), lanes over time, the channel map is the gather of pass 1); thread
// (branch i, run g) then walks down its run with the last T transform outputs of its branch in a register ring -- T packed
// FFMA2 per output -- and the M outputs of a time step leave as one contiguous 8M-byte row of the interleaved stream.
// 8 B in + 8 B out per sample, the channelizer's roofline figure (16 B/sample, HBM).
//
// Everything else (2x oversampling, other channel counts, unaligned streams) takes three plain launches over a scratch
// buffer: gather through the channel map -> the batched FFT of fftvcc.cu -> branch FIRs.
#include <cuda.h>
#include <cstdint>
#include <type_traits>
#include "fft.cuh"
#include "tma.cuh"
#include "capi_common.cuh"

namespace grb {

// ---- tensor map (driver entry point through the runtime: libgrb200 links libcudart only) ------------------------------
typedef CUresult (*PFN_tmapEncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                        const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                        CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_tmapEncodeTiled tmap_encoder()
{
    static PFN_tmapEncodeTiled fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            p = nullptr;
        return (PFN_tmapEncodeTiled)p;
    }();
    return fn;
}

// 2-D TMA tile copy global -> shared (coordinates in elements of the map: c0 = inner, c1 = row)
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, int c0, int c1, unsigned long long* bar)
{
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
                     smem_u32(dst)),
                 "l"(reinterpret_cast<unsigned long long>(map)), "r"(c0), "r"(c1), "r"(smem_u32(bar))
                 : "memory");
}

template <int M, int NT>
struct SynthGeom {
    // M = R1 * R2: pass 1 runs radix-R1 transforms over the bins j + R2 q, pass 2 radix-R2 ones over j (R2 = 1: none)
    static constexpr int R1 = M >= 128 ? 16 : (M >= 32 ? 8 : (M == 16 ? 4 : M));
    static constexpr int R2 = M / R1;
    static constexpr int NG = NT / M;     // runs per CTA
    static constexpr int KT = 16;         // input times per run per iteration
    static constexpr int OT = NG * KT;    // times per CTA per iteration
    static constexpr int LD = M + 1;      // padded row of the [time][branch] buffer
    static constexpr int NI1 = OT * R2 / NT, NI2 = OT * R1 / NT;
    static constexpr int XS_BYTES = OT * M * 8;
    static constexpr int VB_BYTES = OT * LD * 8;
    static constexpr int SMEM = XS_BYTES + VB_BYTES + M * 8 + M * 4 + 16;
    static_assert(NT % M == 0 && (OT * R2) % NT == 0 && (OT * R1) % NT == 0 && R1 * R2 == M && R2 <= 16, "geometry");
};

// x is described by `tmap`: [nin streams][items] floats (2 per item), T-1 history items in front of every stream.
// region: output times per run; iters*KT >= region + T-1.
template <int M, int TP, int NT>
__global__ void __launch_bounds__(NT)
pfb_synth_kernel(const __grid_constant__ CUtensorMap tmap, float2* __restrict__ out, long long ntimes, int T,
                 const float* __restrict__ h /* [T][M]: h[u*M + i] = taps[i + u*M] */, const float2* __restrict__ wM,
                 const int* __restrict__ inv /* [M]: stream feeding FFT bin c, or -1 */, long long region, int iters)
{
    using Geo = SynthGeom<M, NT>;
    constexpr int NG = Geo::NG, KT = Geo::KT, OT = Geo::OT, LD = Geo::LD, R1 = Geo::R1, R2 = Geo::R2, NI1 = Geo::NI1, NI2 = Geo::NI2;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float2* xs = reinterpret_cast<float2*>(smem_raw);                    // [NG][M][KT]
    float2* vbuf = reinterpret_cast<float2*>(smem_raw + Geo::XS_BYTES);  // [OT][LD]
    float2* wtab = vbuf + OT * LD;                                       // [M] e^{+j 2 pi k / M}
    int* sinv = reinterpret_cast<int*>(wtab + M);                        // [M]
    unsigned long long* bar = reinterpret_cast<unsigned long long*>(sinv + M);

    const int tid = threadIdx.x;
    const int bi = tid % M, bg = tid / M;
    const long long run0 = (long long)blockIdx.x * NG;      // first run of this CTA
    const long long n0 = (run0 + bg) * region;               // this thread's first output time
    long long n1 = n0 + region;
    n1 = n1 > ntimes ? ntimes : n1;

    for (int k = tid; k < M; k += NT) {
        wtab[k] = wM[k];
        sinv[k] = inv[k];
    }
    if (tid == 0)
        mbar_init(bar, 1);
    float taps[TP];
#pragma unroll
    for (int u = 0; u < TP; u++)
        taps[u] = (u < T) ? __ldg(h + u * M + bi) : 0.f;
    float2 ring[TP];
#pragma unroll
    for (int u = 0; u < TP; u++)
        ring[u] = make_float2(0.f, 0.f);
    __syncthreads();

    auto issue = [&](int it) { // NG boxes of KT items x M streams; rows / items outside the tensor are zero filled
        mbar_expect_tx(bar, (unsigned)(NG * M * KT * 8));
#pragma unroll
        for (int g = 0; g < NG; g++) {
            const long long m0 = (run0 + g) * region + (long long)it * KT; // input item index (history included)
            tma_load_2d(xs + g * M * KT, &tmap, (int)(2 * m0), 0, bar);
        }
    };
    if (tid == 0)
        issue(0);

    auto iteration = [&](int it, auto phc) {
        constexpr int PH = decltype(phc)::value;
        mbar_wait(bar, it & 1);
        // ---------------- backward DFT over the channels, pass 1: gather through the map, y[j][k1] -> slot j + R2*k1 ---
#pragma unroll
        for (int n = 0; n < NI1; n++) {
            const int idx = tid + n * NT;
            const int o = idx % OT, j = idx / OT;
            const int g = o / KT, k = o % KT;
            float2 v[R1];
#pragma unroll
            for (int q = 0; q < R1; q++) {
                const int s = sinv[j + q * R2];
                v[q] = s >= 0 ? xs[(g * M + s) * KT + k] : make_float2(0.f, 0.f);
            }
            dft_r<R1, false>(v);
#pragma unroll
            for (int q = 0; q < R1; q++)
                vbuf[o * LD + j + q * R2] = v[q];
        }
        __syncthreads(); // xs drained, pass-1 results visible
        if (tid == 0 && it + 1 < iters)
            issue(it + 1);
        // ---------------- pass 2: twiddle, DFT over j; results back in natural branch order ----------------
        float2 a[NI2][R2];
#pragma unroll
        for (int n = 0; n < NI2; n++) {
            const int idx = tid + n * NT;
            const int o = idx % OT, k1 = idx / OT;
#pragma unroll
            for (int j = 0; j < R2; j++) {
                a[n][j] = vbuf[o * LD + k1 * R2 + j];
                if (j > 0)
                    a[n][j] = cmul(a[n][j], wtab[j * k1]);
            }
            dft_r<R2, false>(a[n]);
        }
        __syncthreads(); // every slot has been read before any is overwritten
#pragma unroll
        for (int n = 0; n < NI2; n++) {
            const int idx = tid + n * NT;
            const int o = idx % OT, k1 = idx / OT;
#pragma unroll
            for (int k2 = 0; k2 < R2; k2++)
                vbuf[o * LD + k1 + R1 * k2] = a[n][k2];
        }
        __syncthreads();
        // ---------------- branch FIRs along time: one new transform output per step, TP packed FMAs ----------------
#pragma unroll
        for (int k = 0; k < KT; k++) {
            ring[(PH + k) % TP] = vbuf[(bg * KT + k) * LD + bi];
            float2 a0 = make_float2(0.f, 0.f), a1 = make_float2(0.f, 0.f);
#pragma unroll
            for (int u = 0; u < TP; u += 2) {
                a0 = __ffma2_rn(make_float2(taps[u], taps[u]), ring[(PH + k - u + 2 * TP) % TP], a0);
                a1 = __ffma2_rn(make_float2(taps[u + 1], taps[u + 1]), ring[(PH + k - u - 1 + 2 * TP) % TP], a1);
            }
            const long long n = n0 + (long long)it * KT + k - (T - 1); // output time of the newest input item
            if (n >= n0 && n < n1)
                out[n * M + bi] = __fadd2_rn(a0, a1);
        }
        __syncthreads(); // the FIR reads of vbuf are done before the next pass 1 overwrites it
    };

    if constexpr (TP <= KT) {
        for (int it = 0; it < iters; it++)
            iteration(it, std::integral_constant<int, 0>{});
    } else { // TP == 2*KT: the ring position alternates between 0 and KT
        for (int it = 0; it < iters; it += 2) {
            iteration(it, std::integral_constant<int, 0>{});
            if (it + 1 < iters)
                iteration(it + 1, std::integral_constant<int, KT>{});
        }
    }
}

template <int M, int TP, int NT>
static int launch_synth(const CUtensorMap& tmap, void* out, long long ntimes, int T, const float* h, const float2* wM,
                        const int* inv, int nsm, cudaStream_t st)
{
    using Geo = SynthGeom<M, NT>;
    auto kern = pfb_synth_kernel<M, TP, NT>;
    GRB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Geo::SMEM));
    int occ = 1;
    GRB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, NT, Geo::SMEM));
    if (occ < 1)
        occ = 1;
    long long grid = (long long)nsm * occ;
    // a run is worth starting only if it is long against its T-1 warm-up steps
    const long long min_region = 8LL * Geo::KT;
    long long nruns = grid * Geo::NG;
    if (nruns * min_region > ntimes)
        nruns = (ntimes + min_region - 1) / min_region;
    nruns = ((nruns + Geo::NG - 1) / Geo::NG) * Geo::NG;
    // whole KT-column boxes per run: every TMA box then starts on a 128-byte boundary of its stream (the tensor copy wants
    // the box start 16-byte aligned; an odd item index is not)
    long long region = (ntimes + nruns - 1) / nruns;
    region = ((region + Geo::KT - 1) / Geo::KT) * Geo::KT;
    grid = ((ntimes + region - 1) / region + Geo::NG - 1) / Geo::NG; // drop CTAs the rounding left without work
    long long iters = (region + T - 1 + Geo::KT - 1) / Geo::KT;
    if (iters > 0x7fffffffLL / Geo::KT)
        return fail(GRB200_ERANGE, "pfb_synthesizer: work() call too long for one launch");
    kern<<<(unsigned)grid, NT, Geo::SMEM, st>>>(tmap, (float2*)out, ntimes, T, h, wM, inv, region, (int)iters);
    GRB_LAUNCH_CHECK();
    return 0;
}

template <int M, int NT>
static int launch_synth_tp(int T, const CUtensorMap& tmap, void* out, long long ntimes, const float* h, const float2* wM,
                           const int* inv, int nsm, cudaStream_t st)
{
    if (T <= 8)
        return launch_synth<M, 8, NT>(tmap, out, ntimes, T, h, wM, inv, nsm, st);
    if (T <= 16)
        return launch_synth<M, 16, NT>(tmap, out, ntimes, T, h, wM, inv, nsm, st);
    return launch_synth<M, 32, NT>(tmap, out, ntimes, T, h, wM, inv, nsm, st);
}

bool pfb_synth_fast_supported(int M, int T, int twox, const void* x, long long stride, int nin)
{
    return twox == 1 && T >= 1 && T <= 32 && M >= 8 && M <= 256 && (M & (M - 1)) == 0 && nin >= 1 && nin <= M &&
           ((uintptr_t)x & 15) == 0 && (stride & 1) == 0 && tmap_encoder() != nullptr;
}

// x: nin streams, stream i at x + i*stride items, each ntimes + T-1 items with the T-1 history items first.
int pfb_synth_fast_launch(int M, int T, const void* x, long long stride, int nin, void* out, long long ntimes, const float* h,
                          const float2* wM, const int* inv, int nsm, cudaStream_t st)
{
    if (ntimes <= 0)
        return 0;
    CUtensorMap tmap;
    const cuuint64_t gdim[2] = { (cuuint64_t)(2 * (ntimes + T - 1)), (cuuint64_t)nin };
    const cuuint64_t gstr[1] = { (cuuint64_t)stride * 8 };
    const cuuint32_t box[2] = { 2 * 16, (cuuint32_t)M };
    const cuuint32_t estr[2] = { 1, 1 };
    if (gdim[0] > 0xffffffffULL)
        return fail(GRB200_ERANGE, "pfb_synthesizer: work() call too long for one launch");
    CUresult r = tmap_encoder()(&tmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(x), gdim, gstr, box, estr,
                                CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                                CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
        return fail(GRB200_ECUDA, "cuTensorMapEncodeTiled failed (%d)", (int)r);
    if (M == 8)
        return launch_synth_tp<8, 256>(T, tmap, out, ntimes, h, wM, inv, nsm, st);
    if (M == 16)
        return launch_synth_tp<16, 256>(T, tmap, out, ntimes, h, wM, inv, nsm, st);
    if (M == 32)
        return launch_synth_tp<32, 256>(T, tmap, out, ntimes, h, wM, inv, nsm, st);
    if (M == 128)
        return launch_synth_tp<128, 256>(T, tmap, out, ntimes, h, wM, inv, nsm, st);
    if (M == 64)
        return launch_synth_tp<64, 256>(T, tmap, out, ntimes, h, wM, inv, nsm, st);
    return launch_synth_tp<256, 256>(T, tmap, out, ntimes, h, wM, inv, nsm, st);
}

// ---- general path -------------------------------------------------------------------------------------------------------
// A[m][q] = X_{inv[q]}[m0 + m] or 0: the FFT input buffer of pfb_synthesizer_ccf_impl.cc:236-239
__global__ void __launch_bounds__(256)
synth_gather_kernel(const float2* __restrict__ x, long long stride, const int* __restrict__ inv, int NF, long long m0,
                    long long nm, float2* __restrict__ A)
{
    const long long total = nm * NF;
    const long long step = (long long)gridDim.x * blockDim.x;
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += step) {
        const int q = (int)(e % NF);
        const long long m = e / NF;
        const int s = inv[q];
        A[e] = s >= 0 ? x[(long long)s * stride + m0 + m] : make_float2(0.f, 0.f);
    }
}

// out[(n0 + n) M + i] = sum_j h[j M + i] B[n + T-1-j][q],  B row r = input item m0 + r;  parity of output time n0+n = (par0 + n0 + n) & 1
__global__ void __launch_bounds__(256)
synth_fir_kernel(const float2* __restrict__ B, int NF, int M, int T, const float* __restrict__ h, long long n0, long long nn,
                 int twox, int par0, float2* __restrict__ out)
{
    const long long total = nn * M;
    const long long step = (long long)gridDim.x * blockDim.x;
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += step) {
        const int i = (int)(e % M);
        const long long n = e / M;
        const int q = twox == 2 ? (((int)((par0 + n0 + n) & 1) ^ 1) * M + i) : i;
        float2 a = make_float2(0.f, 0.f);
        for (int j = T - 1; j >= 0; j--) { // oldest item first, like the reversed-tap dot product of the reference
            const float hj = __ldg(h + (size_t)j * M + i);
            a = __ffma2_rn(make_float2(hj, hj), B[(n + T - 1 - j) * NF + q], a);
        }
        out[(n0 + n) * M + i] = a;
    }
}

// scratch: 2 * slab * NF complex items (slab rows = times per pass, chosen by the caller through pfb_synth_slab)
long long pfb_synth_slab(int NF, int T)
{
    long long rows = ((long long)128 << 20) / ((long long)NF * 8); // 128 MiB per buffer
    if (rows < 4LL * T)
        rows = 4LL * T;
    return rows;
}

int pfb_synth_general_launch(int M, int twox, int T, const void* x, long long stride, void* out, long long ntimes,
                             const float* h, const float2* twNF, const int* inv, int par0, void* scratch, long long slab_rows,
                             int nsm, cudaStream_t st)
{
    const int NF = twox * M;
    float2* A = (float2*)scratch;
    float2* B = A + slab_rows * NF;
    const long long per = slab_rows - (T - 1); // output times per pass
    for (long long n0 = 0; n0 < ntimes; n0 += per) {
        const long long nn = std::min(per, ntimes - n0);
        const long long nm = nn + T - 1;
        long long need = (nm * NF + 255) / 256, grid = (long long)nsm * 16;
        if (grid > need)
            grid = need;
        synth_gather_kernel<<<(unsigned)grid, 256, 0, st>>>((const float2*)x, stride, inv, NF, n0, nm, A);
        GRB_LAUNCH_CHECK();
        int rc = fft_vcc_launch(NF, false, A, B, nm, nullptr, 0, 0, twNF, nsm, st);
        if (rc)
            return rc;
        need = (nn * M + 255) / 256;
        grid = (long long)nsm * 16;
        if (grid > need)
            grid = need;
        synth_fir_kernel<<<(unsigned)grid, 256, 0, st>>>(B, NF, M, T, h, n0, nn, twox, par0, (float2*)out);
        GRB_LAUNCH_CHECK();
    }
    return 0;
}

} // namespace grb
