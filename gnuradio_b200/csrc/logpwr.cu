// logpwr.cu -- the averaging tail of logpwrfft: single_pole_iir_filter_ff(alpha, vlen) -> nlog10_ff(10, vlen, k).
//
// Replaces single_pole_iir_filter_ff_impl::work (reference gr-filter/lib/single_pole_iir_filter_ff_impl.cc:47-67 over
// gr-filter/include/gnuradio/filter/single_pole_iir.h:78-86:  y = alpha*x + (1-alpha)*y_prev, double arithmetic, float state)
// and nlog10_ff_impl::work (gr-blocks/lib/nlog10_ff_impl.cc:38-60:  out = n/log2(10) * log2(max(y, FLT_MIN) * 10^(k/n)))
// behind the fused window -> FFT -> |.|^2 kernel (fftvcc.cu, mode 1), in place on the [frames][N] float array.
//
// The recurrence runs along the frame axis, one independent filter per bin.  The frames are cut into runs; three small
// kernels make it parallel in both axes (a linear recurrence: y_j = z_j + beta^(j-j0+1) * carry, z = the run filtered from a
// zero state):
//   logpwr_local_kernel  thread (run, bin): filters its run from zero in place, leaves the run's last value in agg[run][bin]
//   logpwr_carry_kernel  thread (bin): carry[r+1] = agg[r] + beta^len_r * carry[r], carry[0] = the block's filter memory;
//                        stores the new filter memory
//   logpwr_fixup_kernel  thread (run, bin): adds beta^(j-j0+1) * carry[run], takes the logarithm
// A short call (one run) is a single kernel that starts from the filter memory and writes the logarithm directly.
// Without averaging (alpha = 1) none of this runs: the FFT kernel's mode 2 writes the logarithm itself.
// Extra HBM traffic: 16 B per output bin on top of the FFT kernel's 4 B write -- per KEPT frame; logpwrfft normally keeps one
// frame in tens (frame_rate << sample_rate / fft_size), so the block's traffic is the frames it reads.
#include "common.cuh"

namespace grb {

__device__ __forceinline__ float logpwr_log(float y, float log_mul, float log_scale)
{
    return log2f(fmaxf(y, 1.17549435e-38f) * log_scale) * log_mul;
}

// p: [nframes][N] in place.  Thread (run, bin) -- bins across lanes, so every access is a coalesced row segment.
// direct != 0: a single run that starts from ystate, writes the logarithm and the new ystate.
__global__ void __launch_bounds__(256)
logpwr_local_kernel(float* __restrict__ p, long long nframes, int N, int run_len, double alpha, double beta,
                    float* __restrict__ agg, float* __restrict__ ystate, int direct, float log_mul, float log_scale)
{
    const int bin = blockIdx.x * blockDim.x + threadIdx.x;
    const long long run = blockIdx.y;
    if (bin >= N)
        return;
    const long long j0 = run * run_len;
    long long j1 = j0 + run_len;
    j1 = j1 > nframes ? nframes : j1;
    float y = direct ? ystate[bin] : 0.f;
    for (long long j = j0; j < j1; j += 8) { // eight independent loads in flight per thread, then the dependent chain
        float x[8];
#pragma unroll
        for (int u = 0; u < 8; u++)
            x[u] = j + u < j1 ? p[(j + u) * N + bin] : 0.f;
#pragma unroll
        for (int u = 0; u < 8; u++) {
            if (j + u < j1) {
                y = (float)(alpha * (double)x[u] + beta * (double)y); // single_pole_iir.h:82: double arithmetic, float state
                p[(j + u) * N + bin] = direct ? logpwr_log(y, log_mul, log_scale) : y;
            }
        }
    }
    if (direct)
        ystate[bin] = y;
    else
        agg[run * N + bin] = y;
}

// carry[r][bin] = filter memory at the start of run r.  agg is overwritten by carry in place (carry[r] replaces agg[r]).
__global__ void __launch_bounds__(256)
logpwr_carry_kernel(float* __restrict__ agg, long long nruns, long long nframes, int N, int run_len, double beta,
                    float* __restrict__ ystate)
{
    const int bin = blockIdx.x * blockDim.x + threadIdx.x;
    if (bin >= N)
        return;
    const double bfull = pow(beta, (double)run_len);
    double c = (double)ystate[bin];
    for (long long r0 = 0; r0 < nruns; r0 += 8) { // eight loads in flight, then the dependent chain (as the other two kernels)
        float a[8];
#pragma unroll
        for (int u = 0; u < 8; u++)
            a[u] = r0 + u < nruns ? agg[(r0 + u) * N + bin] : 0.f;
#pragma unroll
        for (int u = 0; u < 8; u++) {
            const long long r = r0 + u;
            if (r < nruns) {
                agg[r * N + bin] = (float)c;
                const long long len = (r + 1) * run_len <= nframes ? run_len : nframes - r * run_len;
                const double b = len == run_len ? bfull : pow(beta, (double)len);
                c = (double)a[u] + b * c;
            }
        }
    }
    ystate[bin] = (float)c;
}

// VEC = 4: a thread takes four neighbouring bins (16-byte accesses; N a multiple of 4 and p 16-byte aligned), same arithmetic
// per bin.
template <int VEC>
__global__ void __launch_bounds__(256)
logpwr_fixup_kernel(float* __restrict__ p, long long nframes, int N, int run_len, double beta,
                    const float* __restrict__ carry, float log_mul, float log_scale)
{
    const int bin = (blockIdx.x * blockDim.x + threadIdx.x) * VEC;
    const long long run = blockIdx.y;
    if (bin >= N)
        return;
    const long long j0 = run * run_len;
    long long j1 = j0 + run_len;
    j1 = j1 > nframes ? nframes : j1;
    double c[VEC];
#pragma unroll
    for (int e = 0; e < VEC; e++)
        c[e] = (double)carry[run * N + bin + e];
    double m = beta;
    constexpr int U = VEC == 4 ? 4 : 8; // frames in flight per thread
    for (long long j = j0; j < j1; j += U) {
        float z[U][VEC];
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (j + u < j1) {
                if constexpr (VEC == 4) {
                    const float4 q = *reinterpret_cast<const float4*>(p + (j + u) * N + bin);
                    z[u][0] = q.x, z[u][1] = q.y, z[u][2] = q.z, z[u][3] = q.w;
                } else {
                    z[u][0] = p[(j + u) * N + bin];
                }
            } else {
#pragma unroll
                for (int e = 0; e < VEC; e++)
                    z[u][e] = 0.f;
            }
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (j + u < j1) {
                float o[VEC];
#pragma unroll
                for (int e = 0; e < VEC; e++)
                    o[e] = logpwr_log(z[u][e] + (float)(m * c[e]), log_mul, log_scale);
                if constexpr (VEC == 4)
                    *reinterpret_cast<float4*>(p + (j + u) * N + bin) = make_float4(o[0], o[1], o[2], o[3]);
                else
                    p[(j + u) * N + bin] = o[0];
                m *= beta;
            }
        }
    }
}

static void logpwr_fixup_launch(float* p, long long nframes, int N, int run_len, long long nruns, double beta, const float* carry,
                                float log_mul, float log_scale, cudaStream_t st)
{
    if ((N & 3) == 0 && (((uintptr_t)p) & 15) == 0 && (((uintptr_t)carry) & 15) == 0) {
        const unsigned gx = (unsigned)((N / 4 + 255) / 256);
        logpwr_fixup_kernel<4><<<dim3(gx, (unsigned)nruns), 256, 0, st>>>(p, nframes, N, run_len, beta, carry, log_mul, log_scale);
    } else {
        const unsigned gx = (unsigned)((N + 255) / 256);
        logpwr_fixup_kernel<1><<<dim3(gx, (unsigned)nruns), 256, 0, st>>>(p, nframes, N, run_len, beta, carry, log_mul, log_scale);
    }
}

// In place on p [nframes][N]: y = alpha*x + (1-alpha)*y_prev along the frames, then the logarithm.  ystate: [N] filter
// memory, read and updated.  scratch: at least logpwr_scratch_floats(nframes, N) floats (run aggregates / carries).
static long long logpwr_run_len(long long nframes, int N, int nsm)
{
    // enough (run, bin) threads to fill the machine, runs no shorter than 8 frames
    const long long want_threads = (long long)nsm * 2048 * 2;
    long long nruns = (want_threads + N - 1) / N;
    if (nruns < 1)
        nruns = 1;
    if (nruns > 65535) // grid.y
        nruns = 65535;
    long long len = (nframes + nruns - 1) / nruns;
    return len < 8 ? 8 : len;
}

long long logpwr_scratch_floats(long long nframes, int N, int nsm)
{
    if (nframes <= 16)
        return 0;
    const long long len = logpwr_run_len(nframes, N, nsm);
    return ((nframes + len - 1) / len) * N;
}

int logpwr_iir_log_launch(float* p, long long nframes, int N, double alpha, float* ystate, float* scratch, float log_mul,
                          float log_scale, int nsm, cudaStream_t st)
{
    if (nframes <= 0)
        return 0;
    const double beta = 1.0 - alpha;
    const unsigned gx = (unsigned)((N + 255) / 256);
    if (nframes <= 16 || !scratch) {
        logpwr_local_kernel<<<dim3(gx, 1), 256, 0, st>>>(p, nframes, N, (int)nframes, alpha, beta, nullptr, ystate, 1, log_mul,
                                                         log_scale);
        GRB_LAUNCH_CHECK();
        return 0;
    }
    const long long len = logpwr_run_len(nframes, N, nsm);
    const long long nruns = (nframes + len - 1) / len;
    if (nruns > 65535 || len > 0x7fffffffLL)
        return fail(GRB200_ERANGE, "logpwrfft: too many frames in one call");
    logpwr_local_kernel<<<dim3(gx, (unsigned)nruns), 256, 0, st>>>(p, nframes, N, (int)len, alpha, beta, scratch, ystate, 0,
                                                                   log_mul, log_scale);
    GRB_LAUNCH_CHECK();
    logpwr_carry_kernel<<<gx, 256, 0, st>>>(scratch, nruns, nframes, N, (int)len, beta, ystate);
    GRB_LAUNCH_CHECK();
    logpwr_fixup_launch(p, nframes, N, (int)len, nruns, beta, scratch, log_mul, log_scale, st);
    GRB_LAUNCH_CHECK();
    return 0;
}

// The second half of the above for a caller whose FFT kernel already filtered the runs (fft_vcc_stream_kernel mode 3: run r =
// frames [r run_len, (r+1) run_len), local results in p, last values in agg): carries, then fix-up + logarithm.
int logpwr_carry_fixup_launch(float* p, long long nframes, int N, long long run_len, long long nruns, double alpha, float* agg,
                              float* ystate, float log_mul, float log_scale, cudaStream_t st)
{
    if (nframes <= 0)
        return 0;
    if (nruns > 65535 || run_len > 0x7fffffffLL)
        return fail(GRB200_ERANGE, "logpwrfft: too many frames in one call");
    const double beta = 1.0 - alpha;
    const unsigned gx = (unsigned)((N + 255) / 256);
    logpwr_carry_kernel<<<gx, 256, 0, st>>>(agg, nruns, nframes, N, (int)run_len, beta, ystate);
    GRB_LAUNCH_CHECK();
    logpwr_fixup_launch(p, nframes, N, (int)run_len, nruns, beta, agg, log_mul, log_scale, st);
    GRB_LAUNCH_CHECK();
    return 0;
}

// ---- the same three launches over a scratch array that is part of the library image (no allocation, ever) -------------
// For callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it):
// the run aggregates / carries go to a file-scope device array instead of the handle's lazily grown buffer.  Every GPU that
// loads the library holds these LOGPWR_STATIC_SCRATCH_FLOATS * 4 bytes until the process exits.  ONE array per device,
// shared by every handle and every call: two calls that overlap on the GPU (different streams, or different handles on
// different streams) race on it -- the caller serialises them.  The array keeps the last call's contents; nothing in it has
// to start at zero (logpwr_local_kernel writes agg[run][bin] for every run and bin before logpwr_carry_kernel reads it), so it
// is never cleared.
constexpr long long LOGPWR_STATIC_SCRATCH_FLOATS = GRB200_LOGPWR_STATIC_SCRATCH_BYTES / 4; // 1 MiB (grb200.h)
static __device__ __align__(256) float g_logpwr_static_scratch[LOGPWR_STATIC_SCRATCH_FLOATS];

long long logpwr_static_scratch_floats() { return LOGPWR_STATIC_SCRATCH_FLOATS; }

// In HOST code the array's name is a host-side stand-in, not a device address: the address the kernels are given comes from
// cudaGetSymbolAddress, asked once per device with that device current (the caller holds a DeviceScope).
static int logpwr_static_scratch_ptr(int device, float** out)
{
    constexpr int MAXDEV = 64;
    static std::mutex mu;
    static float* ptr[MAXDEV] = {};
    if (device < 0 || device >= MAXDEV)
        return fail(GRB200_EINVAL, "logpwrfft (static scratch): device %d out of range", device);
    std::lock_guard<std::mutex> lk(mu);
    if (!ptr[device]) {
        int cur = -1;
        GRB_CUDA(cudaGetDevice(&cur));
        if (cur != device)
            return fail(GRB200_EINVAL, "logpwrfft (static scratch): device %d is not current (%d is)", device, cur);
        float* q = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void**)&q, g_logpwr_static_scratch);
        if (e != cudaSuccess || !q)
            return fail(GRB200_ECUDA, "cudaGetSymbolAddress(static scratch) failed on device %d: %s", device,
                        cudaGetErrorString(e));
        ptr[device] = q;
    }
    *out = ptr[device];
    return 0;
}

// 0 when a call of `nframes` kept frames fits the static array; GRB200_ERANGE (nothing launched) otherwise.
int logpwr_static_check(long long nframes, int N, int nsm)
{
    const long long need = logpwr_scratch_floats(nframes, N, nsm);
    if (need > LOGPWR_STATIC_SCRATCH_FLOATS)
        return fail(GRB200_ERANGE,
                    "logpwrfft (static scratch): %lld frames of %d bins need %lld bytes of scratch, the static array holds %lld "
                    "(at most %lld frames per call at this fft_size)",
                    nframes, N, need * 4, LOGPWR_STATIC_SCRATCH_FLOATS * 4, 8 * (LOGPWR_STATIC_SCRATCH_FLOATS / N));
    return 0;
}

// logpwr_iir_log_launch over the static array.  `device` must be current.
int logpwr_iir_log_launch_static(float* p, long long nframes, int N, double alpha, float* ystate, float log_mul,
                                 float log_scale, int nsm, int device, cudaStream_t st)
{
    if (nframes <= 0)
        return 0;
    int rc = logpwr_static_check(nframes, N, nsm);
    if (rc)
        return rc;
    float* scratch = nullptr;
    if (logpwr_scratch_floats(nframes, N, nsm) > 0 && (rc = logpwr_static_scratch_ptr(device, &scratch)))
        return rc;
    return logpwr_iir_log_launch(p, nframes, N, alpha, ystate, scratch, log_mul, log_scale, nsm, st);
}

} // namespace grb
