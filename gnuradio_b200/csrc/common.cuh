// common.cuh -- shared host/device helpers for libgrb200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/grb200.h"

namespace grb {

// ---- error plumbing ---------------------------------------------------------------------------
extern thread_local char g_err[512];
extern std::atomic<unsigned long long> g_launches;

inline int fail(int code, const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

#define GRB_CUDA(expr)                                                                      \
    do {                                                                                    \
        cudaError_t _e = (expr);                                                            \
        if (_e != cudaSuccess)                                                              \
            return grb::fail(GRB200_ECUDA, "%s failed: %s (%s:%d)", #expr,                  \
                             cudaGetErrorString(_e), __FILE__, __LINE__);                   \
    } while (0)

#define GRB_LAUNCH_CHECK()                                                                  \
    do {                                                                                    \
        grb::g_launches.fetch_add(1, std::memory_order_relaxed);                            \
        cudaError_t _e = cudaGetLastError();                                                \
        if (_e != cudaSuccess)                                                              \
            return grb::fail(GRB200_ECUDA, "kernel launch failed: %s (%s:%d)",              \
                             cudaGetErrorString(_e), __FILE__, __LINE__);                   \
    } while (0)

// tuning / developer knobs are read from the environment ONCE per process, never per work() call
#define GRB_GETENV_ONCE(name) ([]() -> const char* { static const char* const v_ = getenv(name); return v_; }())

// Select a device and check that it is a Blackwell (sm_100) part: there is no other code path.
int use_device(int device);
int sm_count(int device);
// Every C-ABI entry point runs inside one of these: selects the handle's device for the call and puts the CALLER's
// current device back on return (a torch / NCCL process keeps its own current device).
struct DeviceScope {
    int prev = -1, dev = -1, rc = 0;
    explicit DeviceScope(int device) : dev(device)
    {
        if (cudaGetDevice(&prev) != cudaSuccess) {
            cudaGetLastError();
            prev = -1;
        }
        rc = use_device(device);
    }
    ~DeviceScope()
    {
        if (prev >= 0 && prev != dev)
            cudaSetDevice(prev);
    }
    DeviceScope(const DeviceScope&) = delete;
    DeviceScope& operator=(const DeviceScope&) = delete;
};

// ---- device-side complex helpers --------------------------------------------------------------
// Packed FP32 (sm_100 FADD2/FMUL2/FFMA2): a complex add is ONE instruction, a complex multiply TWO; the
// swaps and sign flips below fold into the instructions' operand modifiers (.LO_HI, .NP), they cost nothing.
__device__ __forceinline__ float2 cmul(float2 a, float2 b)
{
    const float2 t = __fmul2_rn(make_float2(a.y, a.y), make_float2(b.y, b.x)); // (a.y b.y, a.y b.x)
    return __ffma2_rn(make_float2(a.x, a.x), b, make_float2(-t.x, t.y));
}
__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return __fadd2_rn(a, b); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return __fadd2_rn(a, make_float2(-b.x, -b.y)); }
__device__ __forceinline__ float2 cconj(float2 a) { return make_float2(a.x, -a.y); }
// multiply by -j (forward) / +j (inverse)
template <bool FWD>
__device__ __forceinline__ float2 cmulj(float2 a)
{
    return FWD ? make_float2(a.y, -a.x) : make_float2(-a.y, a.x);
}
// a + (-+j) b and a - (-+j) b without materialising the rotation
template <bool FWD>
__device__ __forceinline__ float2 cadd_mulj(float2 a, float2 b)
{
    return __fadd2_rn(a, FWD ? make_float2(b.y, -b.x) : make_float2(-b.y, b.x));
}
template <bool FWD>
__device__ __forceinline__ float2 csub_mulj(float2 a, float2 b)
{
    return __fadd2_rn(a, FWD ? make_float2(-b.y, b.x) : make_float2(b.y, -b.x));
}

// ---- staging: pinned double-buffered host<->device pipeline (K8 in SURVEY.md section 2a) --------
// A Stager owns three streams (h2d, compute, d2h), NSLOT pinned bounce slots per direction and
// device slots.  Callers describe a job as a sequence of chunks; see staging.cu.
struct Stager;

} // namespace grb
