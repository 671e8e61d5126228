// tools/ubench.cu -- B200 microbenchmarks that size the FP32 roofline used in DESIGN.md:
//   scalar FFMA vs packed fma.rn.f32x2 issue rate, shared-memory broadcast LDS + FFMA2 mix, and a plain
//   float4 copy (HBM).  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench ubench.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ unsigned long long pack(float a, float b)
{
    unsigned long long r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
    return r;
}
__device__ __forceinline__ void ffma2(unsigned long long& d, unsigned long long a, unsigned long long b)
{
    asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(d) : "l"(a), "l"(b));
}

template <int NACC>
__global__ void __launch_bounds__(256) k_ffma(float* out, float a, float b, int iters)
{
    float acc[NACC];
#pragma unroll
    for (int i = 0; i < NACC; i++)
        acc[i] = threadIdx.x * 1e-3f + i;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < NACC; i++)
            acc[i] = fmaf(acc[i], a, b);
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < NACC; i++)
        s += acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int NACC>
__global__ void __launch_bounds__(256) k_ffma2(float* out, float a, float b, int iters)
{
    unsigned long long acc[NACC];
#pragma unroll
    for (int i = 0; i < NACC; i++)
        acc[i] = pack(threadIdx.x * 1e-3f + i, 1.0f + i);
    const unsigned long long av = pack(a, a);
    unsigned long long bv = pack(b, -b);
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < NACC; i++)
            ffma2(acc[i], av, bv);
    }
    unsigned long long s = 0;
#pragma unroll
    for (int i = 0; i < NACC; i++)
        s ^= acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = __uint_as_float((unsigned)(s ^ (s >> 32)));
}

// FIR-like inner loop: per step one 8-byte shared load (x window) + one broadcast tap, R packed FMAs
template <int R>
__global__ void __launch_bounds__(256) k_firmix(float* out, int iters)
{
    __shared__ float2 xs[256 + 1024];
    __shared__ float hs[1024];
    for (int i = threadIdx.x; i < 256 + 1024; i += 256)
        xs[i] = make_float2(i * 1e-3f, -i * 1e-3f);
    for (int i = threadIdx.x; i < 1024; i += 256)
        hs[i] = 1.0f / (i + 1);
    __syncthreads();
    unsigned long long acc[R];
#pragma unroll
    for (int i = 0; i < R; i++)
        acc[i] = 0;
    unsigned long long ring[R];
#pragma unroll
    for (int i = 0; i < R; i++)
        ring[i] = *reinterpret_cast<unsigned long long*>(&xs[threadIdx.x + i]);
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int u = 0; u < 64; u++) {
            const float h = hs[(it & 15) * 64 + u];
            const unsigned long long hv = pack(h, h);
#pragma unroll
            for (int r = 0; r < R; r++)
                ffma2(acc[r], hv, ring[(u + r) % R]);
            ring[u % R] = *reinterpret_cast<unsigned long long*>(&xs[threadIdx.x + ((u + R + it) & 1023)]);
        }
    }
    unsigned long long s = 0;
#pragma unroll
    for (int i = 0; i < R; i++)
        s ^= acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = __uint_as_float((unsigned)(s ^ (s >> 32)));
}

__global__ void __launch_bounds__(256) k_copy(const float4* __restrict__ in, float4* __restrict__ out, long long n)
{
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
        out[i] = in[i];
}

template <class F>
static float time_ms(F&& f, int reps = 5)
{
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    f();
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int i = 0; i < reps; i++) {
        cudaEventRecord(a);
        f();
        cudaEventRecord(b);
        cudaEventSynchronize(b);
        float ms;
        cudaEventElapsedTime(&ms, a, b);
        if (ms < best)
            best = ms;
    }
    return best;
}

int main()
{
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    int clk = 0;
    cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    printf("{\"device\": \"%s\", \"sms\": %d, \"clock_khz\": %d", p.name, p.multiProcessorCount, clk);
    float* out;
    cudaMalloc(&out, 148 * 8 * 256 * 4 * 4);
    const int grid = p.multiProcessorCount * 8, iters = 4096;
    {
        float ms = time_ms([&] { k_ffma<16><<<grid, 256>>>(out, 1.0001f, 0.5f, iters); });
        printf(", \"ffma_tfma_s\": %.3f", (double)grid * 256 * 16 * iters / ms * 1e-9);
        ms = time_ms([&] { k_ffma<32><<<grid, 256>>>(out, 1.0001f, 0.5f, iters); });
        printf(", \"ffma32_tfma_s\": %.3f", (double)grid * 256 * 32 * iters / ms * 1e-9);
    }
    {
        float ms = time_ms([&] { k_ffma2<8><<<grid, 256>>>(out, 1.0001f, 0.5f, iters); });
        printf(", \"ffma2_tfma_s\": %.3f", (double)grid * 256 * 8 * 2 * iters / ms * 1e-9);
        ms = time_ms([&] { k_ffma2<16><<<grid, 256>>>(out, 1.0001f, 0.5f, iters); });
        printf(", \"ffma2x16_tfma_s\": %.3f", (double)grid * 256 * 16 * 2 * iters / ms * 1e-9);
    }
    {
        float ms = time_ms([&] { k_firmix<8><<<grid, 256>>>(out, 64); });
        printf(", \"firmix8_tfma_s\": %.3f", (double)grid * 256 * 8 * 2 * 64 * 64 / ms * 1e-9);
        ms = time_ms([&] { k_firmix<16><<<grid, 256>>>(out, 64); });
        printf(", \"firmix16_tfma_s\": %.3f", (double)grid * 256 * 16 * 2 * 64 * 64 / ms * 1e-9);
    }
    {
        const long long n = (1ll << 30) / 16 * 4; // 4 GiB each way
        float4 *a, *b;
        cudaMalloc(&a, n * 16);
        cudaMalloc(&b, n * 16);
        cudaMemset(a, 1, n * 16);
        float ms = time_ms([&] { k_copy<<<p.multiProcessorCount * 16, 256>>>(a, b, n); });
        printf(", \"copy_gbs\": %.1f", 2.0 * n * 16 / ms * 1e-6);
        ms = time_ms([&] { cudaMemcpyAsync(b, a, n * 16, cudaMemcpyDeviceToDevice); });
        printf(", \"memcpy_d2d_gbs\": %.1f", 2.0 * n * 16 / ms * 1e-6);
        // pinned H2D / D2H
        void* h;
        const size_t hb = 1ull << 30;
        cudaHostAlloc(&h, hb, cudaHostAllocDefault);
        ms = time_ms([&] { cudaMemcpyAsync(a, h, hb, cudaMemcpyHostToDevice); }, 3);
        printf(", \"h2d_gbs\": %.1f", hb / ms * 1e-6);
        ms = time_ms([&] { cudaMemcpyAsync(h, a, hb, cudaMemcpyDeviceToHost); }, 3);
        printf(", \"d2h_gbs\": %.1f", hb / ms * 1e-6);
    }
    printf(", \"err\": \"%s\"}\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
