// fir_tc.cu -- fir_filter_ccf / fir_filter_ccc (decimation 1, <= 64 taps) on the 5th-generation tensor cores (tcgen05, sm_100a).
//
// Why: the CUDA-core kernel (fir.cu) sits on the FP32 pipe (ncu: FMA pipe 70 % busy, 47 of 66.6 TFLOP/s measured
// peak) at 45 % of the HBM roofline; 256 flop per complex sample cannot reach the 16 B/sample HBM bound on that pipe.
// The FIR is a block-Toeplitz product, which the tensor cores evaluate at fp32 accuracy with a 3-term TF32 split.
//
// Math (reference: kernel::fir_filter<gr_complex,gr_complex,float>::filterN, gr-filter/lib/fir_filter.cc:101-141):
//   out[i] = sum_t h[t] * in[i+t],  h = reversed taps zero-padded to 64.
// Cut the input into rows of 64 complex items, X_r[k] = in[64 r + k].  Then for j in [0,64)
//   out[64 r + j] = sum_k T0[j][k] X_r[k] + sum_k T1[j][k] X_{r+1}[k],
//   T0[j][k] = h[k-j] (k >= j),  T1[j][k] = h[64+k-j] (k < j)        (two triangular Toeplitz blocks).
// Real taps act on re and im separately, so the de-interleaved re / im rows are just 2 columns of the B operand:
//   D[m][n] = sum_k A[m][k] B[n][k],  A = [T0; T1] (M = 128 TMEM lanes), n = 2 r + c (N = 128 per tile), K = 64,
// and out = D[T0 lanes][n] + D[T1 lanes][n + 2] is formed in the epilogue.  The A rows are permuted so that both
// halves of an output position sit in the SAME 32-lane TMEM sub-partition (lanes 0-15: T0 rows, lanes 16-31: T1 rows
// of 16 positions): one thread reads both with two tcgen05.ld.16x256b (second one 2 columns to the right) and adds them.
//
// Complex taps (fir_filter_ccc, fir_filter.cc:153-162): two such products per tile, Dr = A(Re h) B and Di = A(Im h) B on
// the same B operand; the (re, im) columns of an input row are adjacent registers of the epilogue thread, which forms
// out = (Dr.re - Di.im, Dr.im + Di.re).  A then takes 256 TMEM columns and Dr | Di the other 256 (one accumulator stage).
//
// Precision: kind::tf32 truncates the low 13 mantissa bits of an fp32 operand (measured: tools/tc_probe.cu), so
// x_hi is the raw value and x_lo = x - trunc(x) is exact; D = A_hi B_hi + A_lo B_hi + A_hi B_lo leaves an error of
// ~2^-21 per product (fp32 accumulation in TMEM).
//
// Pipeline per CTA (persistent, 1 CTA/SM), warp-specialised:
//   1 TMA warp      : cp.async.bulk of the raw 32 KiB input tile into a 3-deep shared-memory ring (3 tiles in flight)
//   8 converter warps: raw tile -> (hi, lo) split, re/im de-interleaved -> swizzled K-major B tiles (2 stages)
//   1 MMA warp      : 24 tcgen05.mma per tile from one elected thread; the Toeplitz operand A lives in TENSOR MEMORY
//                     (hi + lo = 128 columns, written once per CTA), so only B is read from shared memory
//   4 epilogue warps : TMEM -> registers -> add the two halves -> 64-byte-run global stores (3 accumulator stages)
// mbarriers: raw_full/raw_empty, b_full/b_empty, d_full/d_empty.  A tile = 64 input rows -> 63 output rows (tiles
// overlap by one row, so no tile depends on another).
#include "common.cuh"
#include "tma.cuh"

namespace grb {

namespace tc {

constexpr int NCONV = 8; // converter warps
constexpr int NEPI = 4;  // epilogue warps (one per TMEM sub-partition)
constexpr int W_TMA = NCONV, W_MMA = NCONV + 1, W_EPI = NCONV + 2;
constexpr int NTHREADS = (NCONV + 2 + NEPI) * 32;

// Tile geometry per MODE.
// M_CCF: 64 input rows (N = 128), one accumulator of 128 columns per stage, A = [hi | lo] (128 TMEM columns).
// M_CCC: taps hr + j hi need TWO products per tile, Dr = A(hr) B and Di = A(hi) B (out = (Dr.re - Di.im, Dr.im + Di.re)), and
//        A = [hr_hi | hr_lo | hi_hi | hi_lo] takes 256 of the 512 TMEM columns; Dr | Di take the other 256, so there is ONE
//        accumulator stage (an N = 64 tile with two stages was tried: a tcgen05.mma of N = 64 takes as long as one of N = 128,
//        56 vs 71 clocks here, so halving N halves the throughput).
// M_DEC8: fir_filter_ccf with decimation 8 and up to 1024 taps (BASELINE configs[3]).  A row of 64 inputs yields 8 outputs
//        j; out[8r + j] = sum_{d=0..16} sum_k h[64 d + k - 8 j] X_{r+d}[k].  The 16 blocks d = 0..15 are STACKED in the 128
//        lanes of one A operand (lane group (d, j)), so the MMA work per input sample equals the 64-tap case; block 16
//        (only k < 8 j, taps 968..1023: 28 multiply-adds per output) does not fit the 128 lanes and is evaluated on the CUDA
//        cores by the converter warps from the samples they still hold in registers, after they have handed the B tile to
//        the MMA warp: 28 taps per thread live in registers, the 16 lanes of a row reduce-scatter the 8 phase sums with 16
//        shuffles (a second MMA for it cost 44 %; reading the samples back from shared memory was load bound).  The epilogue sums the 16 lane groups with their row shifts: four per thread in registers
//        (tcgen05.ld column offsets are free), the last four across the epilogue warps through shared memory.
//        A tile = 64 input rows -> 48 output rows x 8 (tiles overlap by 16 rows).
enum { M_CCF = 0, M_CCC = 1, M_DEC8 = 2 };

template <int MODE>
struct Cfg {
    static constexpr bool CT = MODE == M_CCC;
    static constexpr int ROWS = 64;                            // input rows per tile
    static constexpr int OROWS = MODE == M_DEC8 ? 48 : ROWS - 1; // rows of the tile that produce output
    static constexpr int OUT_PER_ROW = MODE == M_DEC8 ? 8 : 64;
    static constexpr int NCOL = 2 * ROWS;                      // MMA N: (row, re|im)
    static constexpr int RSTAGES = MODE == M_DEC8 ? 2 : 3;     // raw input ring
    static constexpr int BSTAGES = 2;                          // converted operand ring
    static constexpr int DSTAGES = CT ? 1 : 3;                 // TMEM accumulator ring
    static constexpr int RAW_BYTES = ROWS * 64 * 8;
    static constexpr int KBLK_BYTES = NCOL * 128;        // one 32-float K block of the B tile, 128-byte swizzle
    static constexpr int OPER_BYTES = 2 * KBLK_BYTES;    // K = 64 floats
    static constexpr int BSTAGE_BYTES = 2 * OPER_BYTES;  // hi + lo
    static constexpr int DSTRIDE = CT ? 256 : 128;       // TMEM columns per accumulator stage (ccc: Dr | Di)
    static constexpr unsigned TMEM_A = DSTAGES * DSTRIDE; // ccf 384, ccc/dec8 256; 64-column blocks follow
    static constexpr int CHUNKS = ROWS * 16 / (NCONV * 32); // 32-byte chunks per converter thread and tile
    static constexpr int TSTAGES = 4;                    // M_DEC8: ring of 17th-block results (tail warps -> epilogue)
    static constexpr int NBARS = 2 * RSTAGES + 2 * BSTAGES + 2 * DSTAGES + 1 + (MODE == M_DEC8 ? 2 * TSTAGES : 0);
    static constexpr int RED_BYTES = MODE == M_DEC8 ? 4 * 384 * 8 : 0;            // 4 partial sums x 384 outputs
    static constexpr int TAIL_BYTES = MODE == M_DEC8 ? TSTAGES * 384 * 8 : 0;     // 17th-block sums per tile
    static constexpr int COEF_BYTES = 0;
    static constexpr int SMEM_BYTES =
        BSTAGES * BSTAGE_BYTES + RSTAGES * RAW_BYTES + RED_BYTES + TAIL_BYTES + COEF_BYTES + 256 + 1024;
    static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");
    static_assert(NBARS * 8 + 8 <= 256, "barrier block");
};

__device__ __forceinline__ unsigned long long desc_sw128(unsigned saddr)
{
    unsigned long long d = 0;
    d |= (unsigned long long)((saddr & 0x3FFFF) >> 4);
    d |= (unsigned long long)1 << 16;           // leading byte offset: unused for swizzled K-major
    d |= (unsigned long long)(1024 >> 4) << 32; // stride byte offset: 8 rows x 128 B
    d |= (unsigned long long)1 << 46;           // descriptor version (sm_100)
    d |= (unsigned long long)2 << 61;           // SWIZZLE_128B
    return d;
}
// D[tmem] (+)= A[tmem] * B[smem]^T, A: lane = row, one 32-bit column per k
__device__ __forceinline__ void mma_tf32_ts(unsigned tmem_d, unsigned tmem_a, unsigned long long bdesc, unsigned idesc,
                                            unsigned accumulate)
{
    asm volatile(
        "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n}\n" ::"r"(tmem_d),
        "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void umma_commit(unsigned long long* bar)
{
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long* bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_init_nofence(unsigned long long* bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void sts128(unsigned saddr, float4 v)
{
    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(saddr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ float4 lds128(unsigned saddr)
{
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(saddr) : "memory");
    return v;
}
__device__ __forceinline__ float trunc_tf32(float x) { return __uint_as_float(__float_as_uint(x) & 0xffffe000u); }

#define GRB_TMEM_LD_16x256_X4(r, taddr)                                                                                   \
    asm volatile("tcgen05.ld.sync.aligned.16x256b.x4.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];" \
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),        \
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])   \
                 : "r"(taddr))
#define GRB_TMEM_LD_16x256_X1(r, taddr)                                                                                   \
    asm volatile("tcgen05.ld.sync.aligned.16x256b.x1.b32 {%0,%1,%2,%3}, [%4];"                                            \
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])                                                         \
                 : "r"(taddr))
#define GRB_TMEM_ST_32x32_X8(taddr, r)                                                                                    \
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(taddr), "r"(r[0]), "r"(r[1]), \
                 "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])                                         \
                 : "memory")

} // namespace tc

template <int MODE, bool TRACE>
__global__ void __launch_bounds__(tc::NTHREADS, 1)
    fir_tc_kernel(const float2* __restrict__ in, float2* __restrict__ out, long long nout, long long nin, int ntaps,
                  const void* __restrict__ rev_, long long ntiles, long long* __restrict__ trace,
                  const float2* __restrict__ halo, int dbg_)
{
    using namespace tc;
    using C = Cfg<MODE>;
    constexpr bool CT = C::CT;
    constexpr int ROWS = C::ROWS, OROWS = C::OROWS, NCOL = C::NCOL;
    constexpr int RSTAGES = C::RSTAGES, BSTAGES = C::BSTAGES, DSTAGES = C::DSTAGES;
    constexpr int RAW_BYTES = C::RAW_BYTES, KBLK_BYTES = C::KBLK_BYTES, OPER_BYTES = C::OPER_BYTES,
                  BSTAGE_BYTES = C::BSTAGE_BYTES;
    extern __shared__ unsigned char smem_raw[];
    unsigned char* smem = (unsigned char*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    unsigned char* sB = smem;                          // BSTAGES x [B_hi | B_lo]
    unsigned char* sR = smem + BSTAGES * BSTAGE_BYTES; // RSTAGES raw tiles
    float2* sRed = (float2*)(sR + RSTAGES * RAW_BYTES);                       // M_DEC8: cross-warp partial sums
    float2* sTail = (float2*)(sR + RSTAGES * RAW_BYTES + C::RED_BYTES);        // M_DEC8: 17th-block sums, TSTAGES x 384
    unsigned char* sCoef = sR + RSTAGES * RAW_BYTES + C::RED_BYTES + C::TAIL_BYTES;
    unsigned long long* bars = (unsigned long long*)(sCoef + C::COEF_BYTES);
    unsigned long long* raw_full = bars;
    unsigned long long* raw_empty = raw_full + RSTAGES;
    unsigned long long* b_full = raw_empty + RSTAGES;
    unsigned long long* b_empty = b_full + BSTAGES;
    unsigned long long* d_full = b_empty + BSTAGES;
    unsigned long long* d_empty = d_full + DSTAGES;
    unsigned long long* a_ready = d_empty + DSTAGES;
    unsigned long long* t_full = a_ready + 1;          // M_DEC8
    unsigned long long* t_empty = t_full + C::TSTAGES; // M_DEC8
    // trace build only (GRB200_FIR_TC_DBG; timing experiments, wrong output): 1 = converters skip their operand stores,
    // 2 = skip the 17th-block sums, 4 = skip the raw-tile reads
    const int dbg = TRACE ? dbg_ : 0;
    const bool have_tail = MODE == M_DEC8 && ntaps > 968 && !(dbg & 2);
    unsigned* tmem_ptr = (unsigned*)(bars + C::NBARS);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const long long G = gridDim.x;

    if (warp == W_MMA) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr)), "r"(512));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    if (tid == 0) {
        for (int i = 0; i < RSTAGES; ++i) {
            mbar_init_nofence(&raw_full[i], 1);
            mbar_init_nofence(&raw_empty[i], NCONV);
        }
        for (int i = 0; i < BSTAGES; ++i) {
            mbar_init_nofence(&b_full[i], NCONV);
            mbar_init_nofence(&b_empty[i], 1);
        }
        for (int i = 0; i < DSTAGES; ++i) {
            mbar_init_nofence(&d_full[i], 1);
            mbar_init_nofence(&d_empty[i], NEPI);
        }
        mbar_init_nofence(a_ready, NEPI);
        if (MODE == M_DEC8)
            for (int i = 0; i < C::TSTAGES; ++i) {
                mbar_init_nofence(&t_full[i], NCONV);
                mbar_init_nofence(&t_empty[i], NEPI);
            }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const unsigned tmem = *tmem_ptr;

    if (warp == W_TMA) {
        // ================= loader: one bulk copy per tile, up to RSTAGES tiles ahead of the converters =================
        if (lane == 0) {
            int it = 0;
#pragma unroll 1
            for (long long tile = blockIdx.x; tile < ntiles; tile += G, ++it) {
                const int rs = it % RSTAGES, use = it / RSTAGES;
                if (use > 0)
                    mbar_wait(&raw_empty[rs], (use - 1) & 1);
                const long long e0 = tile * (long long)(OROWS * 64);
                if (e0 + ROWS * 64 <= nin) {
                    mbar_expect_tx(&raw_full[rs], RAW_BYTES);
                    bulk_g2s(sR + rs * RAW_BYTES, in + e0, RAW_BYTES, &raw_full[rs]);
                } else {
                    mbar_arrive(&raw_full[rs]); // ragged last tile: the converters read global memory with bounds checks
                }
            }
        }
    } else if (warp < NCONV) {
        // ================= converters: raw tile -> (hi, lo), re/im de-interleaved, swizzled K-major B operand =================
        const int p = tid;
        const unsigned sB_u32 = smem_u32(sB), sR_u32 = smem_u32(sR);
        const int hsel = (lane >> 2) & 1; // which 16-byte half of its 32-byte chunk a lane reads first (bank-conflict free)
        // M_DEC8, partial 17th tap block: this thread's chunk is always samples k = 4 q .. 4 q + 3 of a row (q = p & 15), so its
        // 7 x 4 taps c[j][u] = h[1024 + 4 q + u - 8 j] (zero unless 4 q + u < 8 j) stay in registers for the whole kernel
        float ct[8][4];
#pragma unroll
        for (int j = 0; j < 8; ++j)
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int k = 4 * (p & 15) + u, t = 1024 + k - 8 * j;
                ct[j][u] = (have_tail && k < 8 * j && t < ntaps) ? __ldg((const float*)rev_ + t) : 0.f;
            }
        int it = 0;
#pragma unroll 1
        for (long long tile = blockIdx.x; tile < ntiles; tile += G, ++it) {
            const int rs = it % RSTAGES, ruse = it / RSTAGES;
            const int stage = it % BSTAGES, buse = it / BSTAGES;
            const bool tr = TRACE && p == 0 && blockIdx.x == 0 && it < 64;
            const long long e0 = tile * (long long)(OROWS * 64);
            if (tr) trace[it * 16 + 0] = clock64();
            mbar_wait(&raw_full[rs], ruse & 1);
            if (tr) trace[it * 16 + 1] = clock64();
            float4 v[C::CHUNKS][2];
            // a tile that reaches into the history halo of a time-sharded segment takes the plain-load path below and reads
            // those ntaps-1 items where they are: the left neighbour's memory (halo), over NVLink
            const bool in_halo = halo != nullptr && e0 < (long long)(ntaps - 1);
            if (TRACE && (dbg & 4)) {
#pragma unroll
                for (int i = 0; i < C::CHUNKS; ++i)
                    v[i][0] = v[i][1] = make_float4(1.f, 2.f, 3.f, 4.f);
            } else if (e0 + ROWS * 64 <= nin && !in_halo) {
                const unsigned rb = sR_u32 + rs * RAW_BYTES + 32 * p;
#pragma unroll
                for (int i = 0; i < C::CHUNKS; ++i) {
                    const float4 x = lds128(rb + 8192 * i + 16 * hsel);
                    const float4 y = lds128(rb + 8192 * i + 16 * (1 - hsel));
                    v[i][0] = hsel ? y : x;
                    v[i][1] = hsel ? x : y;
                }
            } else {
#pragma unroll
                for (int i = 0; i < C::CHUNKS; ++i) {
                    const long long e = e0 + 4 * (p + 256 * i);
                    float2 z[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u)
                        z[u] = (e + u < nin) ? ((in_halo && e + u < (long long)(ntaps - 1)) ? halo[e + u] : in[e + u])
                                             : make_float2(0.f, 0.f);
                    v[i][0] = make_float4(z[0].x, z[0].y, z[1].x, z[1].y);
                    v[i][1] = make_float4(z[2].x, z[2].y, z[3].x, z[3].y);
                }
            }
            if (have_tail) {
                // (done here, before waiting for a free B stage: in the MMA-bound regime that wait hides it)
                // 17th tap block on the CUDA cores: out[8 r + j] += sum_{k < 8 j} h[1024 + k - 8 j] X_{r+16}[k].  The samples are
                // still in this thread's registers (chunk i = rows (p >> 4) + 16 i: i = 1..3 are rows 16..63 = output rows 0..47);
                // the 16 lanes that hold a row reduce-scatter their 8 phase sums (xor 8, 4, 2: halve the phases kept; xor 1: add).
                const int ts = it % C::TSTAGES, tuse = it / C::TSTAGES;
                float2 tot[3];
#pragma unroll
                for (int i = 1; i < 4; ++i) {
                    const float4 a = v[i][0], b = v[i][1];
                    float2 val[8];
#pragma unroll
                    const float2 x0 = make_float2(a.x, a.y), x1 = make_float2(a.z, a.w), x2 = make_float2(b.x, b.y),
                                 x3 = make_float2(b.z, b.w);
                    val[0] = make_float2(0.f, 0.f); // phase 0 never reaches the 17th block (k < 8 j is empty)
#pragma unroll
                    for (int j = 1; j < 8; ++j) // (re, im) together: one packed FMA per tap
                        val[j] = __ffma2_rn(make_float2(ct[j][3], ct[j][3]), x3,
                                            __ffma2_rn(make_float2(ct[j][2], ct[j][2]), x2,
                                                       __ffma2_rn(make_float2(ct[j][1], ct[j][1]), x1,
                                                                  __fmul2_rn(make_float2(ct[j][0], ct[j][0]), x0))));
                    float2 k4[4], k2[2], k1;
                    const bool b3 = lane & 8, b2 = lane & 4, b1 = lane & 2;
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const float2 keep = b3 ? val[j + 4] : val[j], send = b3 ? val[j] : val[j + 4];
                        k4[j].x = keep.x + __shfl_xor_sync(0xffffffffu, send.x, 8);
                        k4[j].y = keep.y + __shfl_xor_sync(0xffffffffu, send.y, 8);
                    }
#pragma unroll
                    for (int j = 0; j < 2; ++j) {
                        const float2 keep = b2 ? k4[j + 2] : k4[j], send = b2 ? k4[j] : k4[j + 2];
                        k2[j].x = keep.x + __shfl_xor_sync(0xffffffffu, send.x, 4);
                        k2[j].y = keep.y + __shfl_xor_sync(0xffffffffu, send.y, 4);
                    }
                    {
                        const float2 keep = b1 ? k2[1] : k2[0], send = b1 ? k2[0] : k2[1];
                        k1.x = keep.x + __shfl_xor_sync(0xffffffffu, send.x, 2);
                        k1.y = keep.y + __shfl_xor_sync(0xffffffffu, send.y, 2);
                    }
                    tot[i - 1].x = k1.x + __shfl_xor_sync(0xffffffffu, k1.x, 1);
                    tot[i - 1].y = k1.y + __shfl_xor_sync(0xffffffffu, k1.y, 1);
                }
                if (tuse > 0)
                    mbar_wait(&t_empty[ts], (tuse - 1) & 1);
                if ((lane & 1) == 0) { // lane bits (3, 2, 1) = the phase j this lane ended up with
                    const int j = ((lane >> 3) & 1) * 4 + ((lane >> 2) & 1) * 2 + ((lane >> 1) & 1);
#pragma unroll
                    for (int i = 1; i < 4; ++i)
                        sTail[ts * 384 + ((p >> 4) + 16 * (i - 1)) * 8 + j] = tot[i - 1];
                }
                __syncwarp();
                if (lane == 0)
                    mbar_arrive(&t_full[ts]);
            }
            if (buse > 0)
                mbar_wait(&b_empty[stage], (buse - 1) & 1);
            if (tr) trace[it * 16 + 2] = clock64();
            const unsigned bh = sB_u32 + stage * BSTAGE_BYTES;
#pragma unroll
            for (int i = 0; i < C::CHUNKS; ++i) {
                const int c = p + 256 * i;
                const int n0 = 2 * (c >> 4), q = c & 15;
                const unsigned base = (unsigned)((q >> 3) * KBLK_BYTES + (n0 >> 3) * 1024 + (n0 & 7) * 128);
                const unsigned off_re = base + (((q & 7) ^ (n0 & 7)) << 4);
                const unsigned off_im = base + 128 + (((q & 7) ^ ((n0 & 7) + 1)) << 4);
                const float4 a = v[i][0], b = v[i][1];
                const float4 re = make_float4(a.x, a.z, b.x, b.z), im = make_float4(a.y, a.w, b.y, b.w);
                const float4 rl = make_float4(re.x - trunc_tf32(re.x), re.y - trunc_tf32(re.y), re.z - trunc_tf32(re.z),
                                              re.w - trunc_tf32(re.w));
                const float4 il = make_float4(im.x - trunc_tf32(im.x), im.y - trunc_tf32(im.y), im.z - trunc_tf32(im.z),
                                              im.w - trunc_tf32(im.w));
                if (TRACE && (dbg & 1) && rl.x != 12345.f)
                    continue;
                sts128(bh + off_re, re); // the tensor core ignores the low 13 bits: raw value = hi part
                sts128(bh + off_im, im);
                sts128(bh + OPER_BYTES + off_re, rl);
                sts128(bh + OPER_BYTES + off_im, il);
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) {
                mbar_arrive(&b_full[stage]);
                mbar_arrive(&raw_empty[rs]); // the stores above depend on every raw read of this warp
            }
            if (tr) trace[it * 16 + 3] = clock64();
        }
    } else if (warp == W_MMA) {
        // ================= MMA issuer: one thread; per product 3 split terms x 8 K steps, A from tensor memory =================
        if (lane == 0) {
            const unsigned idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((unsigned)(NCOL >> 3) << 17) | ((128u >> 4) << 24);
            mbar_wait(a_ready, 0);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            int it = 0;
#pragma unroll 1
            for (long long tile = blockIdx.x; tile < ntiles; tile += G, ++it) {
                const int stage = it % BSTAGES, buse = it / BSTAGES;
                const int acc = it % DSTAGES, duse = it / DSTAGES;
                const bool tr = TRACE && blockIdx.x == 0 && it < 64;
                if (tr) trace[it * 16 + 4] = clock64();
                mbar_wait(&b_full[stage], buse & 1);
                if (tr) trace[it * 16 + 5] = clock64();
                if (duse > 0)
                    mbar_wait(&d_empty[acc], (duse - 1) & 1);
                if (tr) trace[it * 16 + 6] = clock64();
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const unsigned b_addr = smem_u32(sB + stage * BSTAGE_BYTES);
#pragma unroll
                for (int part = 0; part < (CT ? 2 : 1); ++part) { // ccc: Dr = A(hr) B, Di = A(hi) B
                    const unsigned d = tmem + acc * C::DSTRIDE + part * NCOL;
                    const unsigned a_hi = tmem + C::TMEM_A + part * 128, a_lo = a_hi + 64;
                    unsigned n = 0;
#pragma unroll
                    for (int term = 0; term < 3; ++term) { // A_hi B_hi + A_lo B_hi + A_hi B_lo
                        const unsigned ta = term == 1 ? a_lo : a_hi;
                        const unsigned bo = b_addr + (term == 2 ? OPER_BYTES : 0);
#pragma unroll
                        for (int kk = 0; kk < 8; ++kk)
                            mma_tf32_ts(d, ta + 8 * kk, desc_sw128(bo + (kk >> 2) * KBLK_BYTES + (kk & 3) * 32), idesc, n++);
                    }
                }
                umma_commit(&b_empty[stage]); // shared-memory stage may be refilled once these MMAs retire
                umma_commit(&d_full[acc]);    // accumulator ready for the epilogue
                if (tr) trace[it * 16 + 7] = clock64();
            }
        }
    }
    else if (warp >= W_EPI) {
        const int s = warp & 3; // the TMEM sub-partition (lanes 32 s .. 32 s + 31) this warp can access
        // ---- one-time: Toeplitz operand A = [T0; T1] into tensor memory.  Lane l < 16 of sub-partition s holds the T0 row
        //      of output position j = 16 s + l, lane 16 + l the T1 row of the same position (so one thread can add them) ----
        {
            const unsigned trow = tmem + ((unsigned)(32 * s) << 16) + C::TMEM_A;
#pragma unroll 1
            for (int part = 0; part < (CT ? 2 : 1); ++part) {
#pragma unroll 1
                for (int k0 = 0; k0 < 64; k0 += 8) {
                    unsigned hi[8], lo[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const int k = k0 + i;
                        int t; // tap index this (lane, k) element multiplies, < 0: structural zero
                        if (MODE == M_DEC8) {
                            // lane = j + 8 m holds block d = s + 4 m
                            const int j = lane & 7, m = lane >> 3;
                            t = 64 * (s + 4 * m) + k - 8 * j;
                        } else {
                            const int j = 16 * s + (lane & 15);
                            t = lane < 16 ? (k >= j ? k - j : -1) : (k < j ? 64 + k - j : -1);
                        }
                        float v = 0.f;
                        if (t >= 0 && t < ntaps) {
                            if (CT) {
                                const float2 h2 = __ldg((const float2*)rev_ + t);
                                v = part ? h2.y : h2.x;
                            } else {
                                v = __ldg((const float*)rev_ + t);
                            }
                        }
                        const float h = trunc_tf32(v);
                        hi[i] = __float_as_uint(h);
                        lo[i] = __float_as_uint(v - h);
                    }
                    GRB_TMEM_ST_32x32_X8(trow + part * 128 + k0, hi);
                    GRB_TMEM_ST_32x32_X8(trow + part * 128 + 64 + k0, lo);
                }
            }
            asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0)
                mbar_arrive(a_ready);
        }
        // ================= epilogue: D[T0 lanes][n] + D[T1 lanes][n+2] (+ the complex combination) -> global =================
        int it = 0;
#pragma unroll 1
        for (long long tile = blockIdx.x; tile < ntiles; tile += G, ++it) {
            const int acc = it % DSTAGES, duse = it / DSTAGES;
            const bool tr = TRACE && blockIdx.x == 0 && it < 64 && warp == W_EPI && lane == 0;
            if (tr) trace[it * 16 + 8] = clock64();
            mbar_wait(&d_full[acc], duse & 1);
            if (tr) trace[it * 16 + 9] = clock64();
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            if constexpr (MODE == M_DEC8) {
                // thread (sub-partition s, lane): output phase j = lane / 4; lanes j + 8 m of sub-partition s hold blocks
                // d = s + 4 m.  One .16x256b load returns lanes (j, j + 8): blocks d and d + 4, whose 4-row shift is the
                // next x1 repeat (8 columns), so register index i + 1; the second load (lane base 16) holds d + 8, d + 12.
                const unsigned tA = tmem + ((unsigned)(32 * s) << 16) + acc * C::DSTRIDE + 2 * s;
                const unsigned tB = tA + (16u << 16) + 16;
                float2* red = sRed + s * 384;
                const int j = lane >> 2, q = lane & 3;
#pragma unroll
                for (int g = 0; g < 3; ++g) {
                    unsigned a[20], b[20];
                    GRB_TMEM_LD_16x256_X4(a, tA + 32 * g);
                    GRB_TMEM_LD_16x256_X1((a + 16), tA + 32 * g + 32);
                    GRB_TMEM_LD_16x256_X4(b, tB + 32 * g);
                    GRB_TMEM_LD_16x256_X1((b + 16), tB + 32 * g + 32);
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                    if (g == 2) {
                        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                        __syncwarp();
                        if (lane == 0)
                            mbar_arrive(&d_empty[acc]);
                        if (tr) trace[it * 16 + 10] = clock64();
                    }
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const float re = __uint_as_float(a[4 * i]) + __uint_as_float(a[4 * i + 6]) + __uint_as_float(b[4 * i]) +
                                         __uint_as_float(b[4 * i + 6]);
                        const float im = __uint_as_float(a[4 * i + 1]) + __uint_as_float(a[4 * i + 7]) +
                                         __uint_as_float(b[4 * i + 1]) + __uint_as_float(b[4 * i + 7]);
                        red[(16 * g + 4 * i + q) * 8 + j] = make_float2(re, im);
                    }
                }
                const int ts = it % C::TSTAGES, tuse = it / C::TSTAGES;
                if (have_tail)
                    mbar_wait(&t_full[ts], tuse & 1);
                asm volatile("bar.sync 1, 128;" ::: "memory"); // the four epilogue warps: partial sums of this tile are in place
                const int e = (warp - W_EPI) * 32 + lane;
#pragma unroll
                for (int u = 0; u < 3; ++u) {
                    const int o = e + 128 * u;
                    const float2 p0 = sRed[o], p1 = sRed[384 + o], p2 = sRed[768 + o], p3 = sRed[1152 + o];
                    float2 y = make_float2((p0.x + p1.x) + (p2.x + p3.x), (p0.y + p1.y) + (p2.y + p3.y));
                    if (have_tail && (o & 7) != 0) { // phase 0 has no 17th-block term
                        const float2 tl = sTail[ts * 384 + o];
                        y.x += tl.x;
                        y.y += tl.y;
                    }
                    const long long idx = tile * 384 + o;
                    if (idx < nout)
                        out[idx] = y;
                }
                asm volatile("bar.sync 1, 128;" ::: "memory"); // everyone has read sRed / sTail[ts]: they may be rewritten
                if (have_tail && lane == 0)
                    mbar_arrive(&t_empty[ts]);
            } else {
                const unsigned t0 = tmem + ((unsigned)(32 * s) << 16) + acc * C::DSTRIDE;
                const unsigned t1 = t0 + (16u << 16) + 2; // T1 lanes, one input row (2 columns) to the right
                const long long row0 = tile * OROWS;
                constexpr int NG = NCOL / 32;
    #pragma unroll
                for (int g = 0; g < NG; ++g) {
                    unsigned a[16], b[16], c[CT ? 16 : 1], d[CT ? 16 : 1];
                    GRB_TMEM_LD_16x256_X4(a, t0 + 32 * g);
                    GRB_TMEM_LD_16x256_X4(b, t1 + 32 * g); // the last group reads 2 columns past the tile: discarded (row OROWS)
                    if (CT) {
                        GRB_TMEM_LD_16x256_X4(c, t0 + NCOL + 32 * g);
                        GRB_TMEM_LD_16x256_X4(d, t1 + NCOL + 32 * g);
                    }
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                    if (g == NG - 1) {
                        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                        __syncwarp();
                        if (lane == 0)
                            mbar_arrive(&d_empty[acc]);
                        if (tr) trace[it * 16 + 10] = clock64();
                    }
    #pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const int r = 16 * g + 4 * i + (lane & 3);
    #pragma unroll
                        for (int hh = 0; hh < 2; ++hh) {
                            const int j = 16 * s + (lane >> 2) + 8 * hh;
                            const int e = 4 * i + 2 * hh;
                            float re = __uint_as_float(a[e]) + __uint_as_float(b[e]);
                            float im = __uint_as_float(a[e + 1]) + __uint_as_float(b[e + 1]);
                            if (CT) { // (Dr.re - Di.im, Dr.im + Di.re)
                                const float ire = __uint_as_float(c[e]) + __uint_as_float(d[e]);
                                const float iim = __uint_as_float(c[e + 1]) + __uint_as_float(d[e + 1]);
                                re -= iim;
                                im += ire;
                            }
                            const long long idx = (row0 + r) * 64 + j;
                            if (r < OROWS && idx < nout)
                                out[idx] = make_float2(re, im);
                        }
                    }
                }
            }
            if (tr) trace[it * 16 + 11] = clock64();
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == W_MMA)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512));
}

// Which tensor-core mode serves this call, or -1 (CUDA-core kernels).  `in` must be 16-byte aligned (TMA bulk copies).
//   M_CCF / M_CCC: decimation 1, ntaps <= 64.   M_DEC8: real taps, decimation 8, ntaps <= 1024.
// GRB200_FIR_TC (developer knob: 0 = CUDA-core kernels only, 1 = tensor-core path whenever the geometry allows), read when a
// handle is created -- never per work()
int fir_tc_env()
{
    const char* e = getenv("GRB200_FIR_TC");
    return e ? (atoi(e) != 0 ? 1 : 0) : -1;
}

int fir_tc_mode(bool ct, int D, int ntaps, const void* in, long long nout, int tc_pref)
{
    int mode = -1;
    if (D == 1 && ntaps <= 64)
        mode = ct ? tc::M_CCC : tc::M_CCF;
    else if (D == 8 && !ct && ntaps <= 1024)
        mode = tc::M_DEC8;
    if (mode < 0 || ((uintptr_t)in & 15) != 0)
        return -1;
    if (tc_pref >= 0)
        return tc_pref ? mode : -1;
    if (mode == tc::M_CCC && ntaps < 20) // measured: the CUDA-core ccc kernel is as fast below ~20 taps (137 vs 154 Gsps at 21)
        return -1;
    if (mode == tc::M_DEC8 && ntaps < 256) // short decimators are already HBM-bound on the CUDA cores
        return -1;
    return nout >= 64 * 63 ? mode : -1; // below a tile or two the CUDA-core kernel's launch is cheaper
}

template <int MODE>
static int fir_tc_launch_t(int ntaps, const void* rev, const void* in, void* out, long long nout, long long nin, int nsm,
                           cudaStream_t st, const void* halo)
{
    using C = tc::Cfg<MODE>;
    static std::atomic<int> configured[64]; // the attribute is per device
    int dev = 0;
    GRB_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !configured[dev].load(std::memory_order_acquire)) {
        GRB_CUDA(cudaFuncSetAttribute(fir_tc_kernel<MODE, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES));
        GRB_CUDA(cudaFuncSetAttribute(fir_tc_kernel<MODE, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES));
        if (dev >= 0 && dev < 64)
            configured[dev].store(1, std::memory_order_release);
    }
    const long long orows = (nout + C::OUT_PER_ROW - 1) / C::OUT_PER_ROW;
    const long long ntiles = (orows + C::OROWS - 1) / C::OROWS;
    const unsigned grid = (unsigned)(ntiles < nsm ? ntiles : nsm);
    static const char* const trace_path = getenv("GRB200_FIR_TC_TRACE"); // read once, not per work()
    if (const char* path = trace_path) { // developer aid: per-tile clock64() stamps of CTA 0
        long long* dtrace = nullptr;
        GRB_CUDA(cudaMalloc(&dtrace, 64 * 16 * sizeof(long long)));
        GRB_CUDA(cudaMemsetAsync(dtrace, 0, 64 * 16 * sizeof(long long), st));
        fir_tc_kernel<MODE, true><<<grid, tc::NTHREADS, C::SMEM_BYTES, st>>>((const float2*)in, (float2*)out, nout, nin, ntaps,
                                                                            rev, ntiles, dtrace, (const float2*)halo,
                                                                            getenv("GRB200_FIR_TC_DBG") ? atoi(getenv("GRB200_FIR_TC_DBG")) : 0);
        GRB_LAUNCH_CHECK();
        std::vector<long long> h(64 * 16);
        GRB_CUDA(cudaMemcpyAsync(h.data(), dtrace, h.size() * sizeof(long long), cudaMemcpyDeviceToHost, st));
        GRB_CUDA(cudaStreamSynchronize(st));
        cudaFree(dtrace);
        if (FILE* f = fopen(path, "w")) {
            for (int i = 0; i < 64; ++i) {
                for (int j = 0; j < 12; ++j) fprintf(f, "%lld ", h[i * 16 + j] ? h[i * 16 + j] - h[0] : -1LL);
                fprintf(f, "\n");
            }
            fclose(f);
        }
        return 0;
    }
    fir_tc_kernel<MODE, false><<<grid, tc::NTHREADS, C::SMEM_BYTES, st>>>((const float2*)in, (float2*)out, nout, nin, ntaps,
                                                                         rev, ntiles, nullptr, (const float2*)halo, 0);
    GRB_LAUNCH_CHECK();
    return 0;
}

int fir_tc_launch(int mode, int ntaps, const void* rev, const void* in, void* out, long long nout, long long nin, int nsm,
                  cudaStream_t st, const void* halo)
{
    switch (mode) {
    case tc::M_CCF:
        return fir_tc_launch_t<tc::M_CCF>(ntaps, rev, in, out, nout, nin, nsm, st, halo);
    case tc::M_CCC:
        return fir_tc_launch_t<tc::M_CCC>(ntaps, rev, in, out, nout, nin, nsm, st, halo);
    case tc::M_DEC8:
        return fir_tc_launch_t<tc::M_DEC8>(ntaps, rev, in, out, nout, nin, nsm, st, halo);
    }
    return fail(GRB200_ERANGE, "fir: no tensor-core mode %d", mode);
}

} // namespace grb
