// tc_probe.cu -- one-CTA probe of tcgen05.mma kind::tf32 on sm_100a (developer tool, not part of libgrb200).
// Answers, before the FIR tensor-core kernel relies on them:
//   1. are the SW128 K-major shared-memory descriptors / the instruction descriptor right (D == A * B^T)?
//   2. does the tensor core truncate or round the low 13 mantissa bits of fp32 operands fed as tf32?
//   3. which (lane, column) does each register of tcgen05.ld.16x256b hold, also with lane base 16 and an
//      unaligned column base (+2)?
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -o tools/tc_probe tools/tc_probe.cu
#include <cuda_runtime.h>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity)
{
    unsigned done = 0;
    while (!done) {
        asm volatile(
            "{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
            : "=r"(done)
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
    }
}
// K-major, 128-byte swizzle: rows of 128 B, 8-row groups 1024 B apart
__device__ __forceinline__ unsigned long long make_desc_sw128(unsigned saddr)
{
    unsigned long long d = 0;
    d |= (unsigned long long)((saddr & 0x3FFFF) >> 4);
    d |= (unsigned long long)(1) << 16;            // LBO (ignored for swizzled K-major)
    d |= (unsigned long long)(1024 >> 4) << 32;    // SBO
    d |= (unsigned long long)1 << 46;              // descriptor version (sm_100)
    d |= (unsigned long long)2 << 61;              // SWIZZLE_128B
    return d;
}
__device__ __forceinline__ void mma_tf32(unsigned tmem_d, unsigned long long adesc, unsigned long long bdesc, unsigned idesc,
                                         unsigned accumulate)
{
    asm volatile(
        "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n}\n" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}

constexpr int M = 128, N = 128, K = 32;

__global__ void __launch_bounds__(128) probe(const float* A, const float* B, float* D, float* frag)
{
    extern __shared__ unsigned char smem_raw[];
    unsigned char* smem = (unsigned char*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    float* sA = (float*)smem;
    float* sB = (float*)(smem + 16384);
    unsigned long long* bar = (unsigned long long*)(smem + 32768);
    unsigned* tmem_ptr = (unsigned*)(smem + 32768 + 8);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    for (int idx = tid; idx < 128 * K; idx += 128) {
        const int r = idx / K, k = idx % K;
        const int off = (r / 8) * 1024 + (r % 8) * 128 + (((k / 4) ^ (r % 8)) * 16) + (k % 4) * 4;
        *(float*)((unsigned char*)sA + off) = A[idx];
        *(float*)((unsigned char*)sB + off) = B[idx];
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr)), "r"(128));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    if (tid == 0) mbar_init(bar, 1);
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const unsigned tmem = *tmem_ptr;

    if (tid == 0) {
        // idesc: c=F32 (1<<4), a=TF32 (2<<7), b=TF32 (2<<10), K-major both, N>>3 at bit 17, M>>4 at bit 24
        const unsigned idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((unsigned)(N >> 3) << 17) | ((unsigned)(M >> 4) << 24);
        for (int k = 0; k < K / 8; ++k)
            mma_tf32(tmem, make_desc_sw128(smem_u32(sA) + k * 32), make_desc_sw128(smem_u32(sB) + k * 32), idesc, k > 0);
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                     : "memory");
    }
    mbar_wait(bar, 0);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

    const unsigned tw = tmem + ((unsigned)(32 * warp) << 16);
    for (int c = 0; c < 4; ++c) {
        unsigned v[32];
        asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,"
            "%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
            : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
              "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
              "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
              "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
            : "r"(tw + c * 32));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        for (int i = 0; i < 32; ++i) D[(32 * warp + lane) * N + c * 32 + i] = __uint_as_float(v[i]);
    }
    // fragment probes: variant = (lane base 0|16) x (column base 0|2), .16x256b.x2 (8 registers, 16 columns)
    for (int var = 0; var < 4; ++var) {
        const unsigned ta = tmem + ((unsigned)(32 * warp + (var & 1) * 16) << 16) + (var >> 1) * 2;
        unsigned v[8];
        asm volatile("tcgen05.ld.sync.aligned.16x256b.x2.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                     : "r"(ta));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        for (int i = 0; i < 8; ++i) frag[((warp * 4 + var) * 32 + lane) * 8 + i] = __uint_as_float(v[i]);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(128));
}

// ---- A operand from tensor memory (tcgen05.mma [d], [a_tmem], b_desc): lane = row m, one 32-bit column per k ----
__device__ __forceinline__ void mma_tf32_ts(unsigned tmem_d, unsigned tmem_a, unsigned long long bdesc, unsigned idesc,
                                            unsigned accumulate)
{
    asm volatile(
        "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n}\n" ::"r"(tmem_d),
        "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__global__ void __launch_bounds__(128) probe_ts(const float* A, const float* B, float* D, int shift)
{
    extern __shared__ unsigned char smem_raw[];
    unsigned char* smem = (unsigned char*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    float* sB = (float*)(smem + 16384);
    unsigned long long* bar = (unsigned long long*)(smem + 32768);
    unsigned* tmem_ptr = (unsigned*)(smem + 32768 + 8);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int idx = tid; idx < 128 * K; idx += 128) {
        const int r = idx / K, k = idx % K;
        const int off = (r / 8) * 1024 + (r % 8) * 128 + (((k / 4) ^ (r % 8)) * 16) + (k % 4) * 4;
        *(float*)((unsigned char*)sB + off) = B[idx];
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr)), "r"(256));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    if (tid == 0) mbar_init(bar, 1);
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const unsigned tmem = *tmem_ptr;
    // A -> TMEM columns [128, 128+K): thread (warp, lane) owns row m = 32 warp + lane
    {
        const int m = 32 * warp + lane;
        const unsigned ta = tmem + ((unsigned)(32 * warp) << 16) + 128;
        for (int c = 0; c < K; c += 8) {
            unsigned v[8];
            for (int i = 0; i < 8; ++i) v[i] = __float_as_uint(A[m * K + c + i]);
            asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(ta + c), "r"(v[0]),
                         "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7])
                         : "memory");
        }
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    if (tid == 0) {
        const unsigned idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((unsigned)(N >> 3) << 17) | ((unsigned)(M >> 4) << 24);
        for (int k = 0; k < K / 8; ++k)
            mma_tf32_ts(tmem, tmem + 128 + k * 8, make_desc_sw128(smem_u32(sB) + k * 32), idesc, k > 0);
        if (shift) // accumulate the same product again, `shift` columns to the right: D'[m][n] = D[m][n] + D[m][n-shift]
            for (int k = 0; k < K / 8; ++k)
                mma_tf32_ts(tmem + shift, tmem + 128 + k * 8, make_desc_sw128(smem_u32(sB) + k * 32), idesc, 1);
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                     : "memory");
    }
    mbar_wait(bar, 0);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const unsigned tw = tmem + ((unsigned)(32 * warp) << 16);
    for (int c = 0; c < 128; c += 8) {
        unsigned v[8];
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                     : "r"(tw + c));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        for (int i = 0; i < 8; ++i) D[(32 * warp + lane) * N + c + i] = __uint_as_float(v[i]);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(256));
}

static float trunc_tf32(float x)
{
    uint32_t u;
    memcpy(&u, &x, 4);
    u &= 0xffffe000u;
    memcpy(&x, &u, 4);
    return x;
}
static float rna_tf32(float x)
{
    uint32_t u;
    memcpy(&u, &x, 4);
    u += 0x1000u;
    u &= 0xffffe000u;
    memcpy(&x, &u, 4);
    return x;
}

int main()
{
    std::vector<float> A(M * K), B(N * K), D(M * N), frag(4 * 4 * 32 * 8);
    srand(1);
    for (auto& v : A) v = (float)rand() / RAND_MAX * 2 - 1;
    for (auto& v : B) v = (float)rand() / RAND_MAX * 2 - 1;
    float *dA, *dB, *dD, *dF;
    cudaMalloc(&dA, A.size() * 4);
    cudaMalloc(&dB, B.size() * 4);
    cudaMalloc(&dD, D.size() * 4);
    cudaMalloc(&dF, frag.size() * 4);
    cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(dB, B.data(), B.size() * 4, cudaMemcpyHostToDevice);
    cudaMemset(dD, 0, D.size() * 4);
    const int smem = 32768 + 64 + 1024;
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    probe<<<1, 128, smem>>>(dA, dB, dD, dF);
    cudaError_t e = cudaDeviceSynchronize();
    printf("kernel: %s\n", cudaGetErrorString(e));
    if (e != cudaSuccess) return 1;
    cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost);
    cudaMemcpy(frag.data(), dF, frag.size() * 4, cudaMemcpyDeviceToHost);
    double e_full = 0, e_trunc = 0, e_rna = 0, ref2 = 0;
    for (int m = 0; m < M; ++m)
        for (int n = 0; n < N; ++n) {
            double sf = 0, st = 0, sr = 0;
            for (int k = 0; k < K; ++k) {
                sf += (double)A[m * K + k] * B[n * K + k];
                st += (double)trunc_tf32(A[m * K + k]) * trunc_tf32(B[n * K + k]);
                sr += (double)rna_tf32(A[m * K + k]) * rna_tf32(B[n * K + k]);
            }
            const double d = D[m * N + n];
            e_full += (d - sf) * (d - sf);
            e_trunc += (d - st) * (d - st);
            e_rna += (d - sr) * (d - sr);
            ref2 += sf * sf;
        }
    printf("nrms vs fp32-exact %.3e, vs truncated-tf32 %.3e, vs rna-tf32 %.3e\n", sqrt(e_full / ref2), sqrt(e_trunc / ref2),
           sqrt(e_rna / ref2));
    for (int var = 0; var < 4; ++var) {
        printf("variant lane_base=%d col_base=%d (warp 0):\n", (var & 1) * 16, (var >> 1) * 2);
        for (int lane = 0; lane < 32; lane += (lane < 5 ? 1 : 9)) {
            printf("  thread %2d:", lane);
            for (int i = 0; i < 8; ++i) {
                const float v = frag[((0 * 4 + var) * 32 + lane) * 8 + i];
                int fm = -1, fn = -1;
                for (int m = 0; m < M && fm < 0; ++m)
                    for (int n = 0; n < N; ++n)
                        if (D[m * N + n] == v) { fm = m; fn = n; break; }
                printf(" r%d=(%d,%d)", i, fm, fn);
            }
            printf("\n");
        }
    }
    {
        std::vector<float> D2(M * N);
        cudaMemset(dD, 0, D.size() * 4);
        cudaFuncSetAttribute(probe_ts, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        probe_ts<<<1, 128, smem>>>(dA, dB, dD, 0);
        e = cudaDeviceSynchronize();
        printf("probe_ts: %s\n", cudaGetErrorString(e));
        if (e != cudaSuccess) return 1;
        cudaMemcpy(D2.data(), dD, D2.size() * 4, cudaMemcpyDeviceToHost);
        double err = 0, ref = 0;
        for (int i = 0; i < M * N; ++i) { err += (double)(D2[i] - D[i]) * (D2[i] - D[i]); ref += (double)D[i] * D[i]; }
        printf("A-from-TMEM vs A-from-smem: nrms %.3e\n", sqrt(err / ref));
        for (int shift : {4, 2, 6}) {
            std::vector<float> D3(M * N);
            probe_ts<<<1, 128, smem>>>(dA, dB, dD, shift);
            e = cudaDeviceSynchronize();
            printf("probe_ts shift %d: %s\n", shift, cudaGetErrorString(e));
            if (e != cudaSuccess) return 1;
            cudaMemcpy(D3.data(), dD, D3.size() * 4, cudaMemcpyDeviceToHost);
            double err3 = 0, ref3 = 0;
            for (int m = 0; m < M; ++m)
                for (int n = shift; n < N; ++n) {
                    const double want = (double)D[m * N + n] + D[m * N + n - shift];
                    err3 += (D3[m * N + n] - want) * (D3[m * N + n] - want);
                    ref3 += want * want;
                }
            printf("accumulate with D column base +%d: nrms vs D[n] + D[n-%d] = %.3e\n", shift, shift, sqrt(err3 / ref3));
        }
    }
    return 0;
}
