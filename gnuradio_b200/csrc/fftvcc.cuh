// fftvcc.cuh -- what the translation units of the batched FFT share (fftvcc.cu: single-CTA kernels and dispatch;
// fftvcc_stream.cu: the streaming kernels; fftvcc_big.cu: Bluestein and the four-step transform).  Split for build time only.
#pragma once
#include <cstdint>
#include <cmath>
#include <type_traits>
#include <cstdlib>
#include <mutex>
#include <utility>
#include <vector>
#include "fft.cuh"
#include "tma.cuh"
#include "capi_common.cuh"

namespace grb {


// What one launch does besides the transform.  in: vectors of N complex64, or N float32 when REAL (fft_vfc_fftw.cc:70-80);
// output vector j transforms input vector j*vstride + voff (keep_one_in_n in front of the FFT, logpwrfft.py:49-52);
// mode 0: complex out; 1: |X|^2 (complex_to_mag_squared fused); 2: log_mul * log2(max(|X|^2, FLT_MIN) * log_scale) -- the
// nlog10_ff at the end of logpwrfft without averaging (nlog10_ff_impl.cc:47-58) -- and the last vector's |X|^2 goes to ystate
// (the single_pole_iir memory a later set_average(true) continues from).

__device__ __forceinline__ void vcc_store(const VccArgs& a, long long vec, int N, int k, float2 X)
{
    if (a.mode == 0) {
        reinterpret_cast<float2*>(a.out)[vec * N + k] = X;
    } else {
        const float p = fmaf(X.x, X.x, X.y * X.y);
        if (a.mode == 1) {
            reinterpret_cast<float*>(a.out)[vec * N + k] = p;
        } else {
            reinterpret_cast<float*>(a.out)[vec * N + k] = log2f(fmaxf(p, 1.17549435e-38f) * a.log_scale) * a.log_mul;
            if (vec == a.nvec - 1)
                a.ystate[k] = p;
        }
    }
}

template <int N>
struct VccGeom {
    static constexpr int E = fft_elems(N);
    static constexpr int TPV = N / E;                        // threads per vector
    static constexpr int VPC = TPV >= 128 ? 1 : 128 / TPV;   // vectors per CTA
    static constexpr int NT = TPV * VPC;
    static constexpr int SMEM = VPC * fft_smem_elems(N) * 8;
};


// fftvcc_stream.cu: 32 <= N <= 8192, 16-byte aligned input, >= 64 vectors (the caller checks)
int fft_vcc_stream_launch(int N, bool fwd, bool real_in, const VccArgs& a, int nsm, cudaStream_t st);
// fftvcc_big.cu
int bluestein_launch(int N, bool fwd, bool real_in, const VccArgs& a, int nsm, cudaStream_t st);
int big_launch(int N, bool fwd, bool real_in, const VccArgs& a, int nsm, cudaStream_t st);
int fft_vcc_launch_ex(int N, bool fwd, bool real_in, const VccArgs& a, int nsm, cudaStream_t st);

} // namespace grb
