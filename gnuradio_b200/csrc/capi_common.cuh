// capi_common.cuh -- declarations shared by the C-ABI translation units.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <memory>
#include "staging.cuh"

namespace grb {
// kernels' host launchers (fir.cu, fftvcc.cu, fftfilt.cu, pfb.cu)
// tc_pref: the handle's GRB200_FIR_TC setting, read when the handle is made (fir_tc_env): -1 auto, 0 never, 1 whenever possible
int fir_launch(bool ct, int D, int ntaps, const void* gtaps, const void* rev, const void* in, void* out,
               long long nout, int tc_pref, cudaStream_t st, const void* halo = nullptr);
int fir_tc_env();
int halo_fetch_launch(void* dst, const void* src, size_t bytes, cudaStream_t st); // peer -> local by plain loads (fir.cu)
// fir_filter_fff (fir_fff.cu): revp = reversed taps zero padded to fir_fff_pad(ntaps) floats
int fir_fff_pad(int ntaps);
int fir_fff_launch(int D, int ntaps, const float* revp, const void* in, void* out, long long nout, int nsm, cudaStream_t st);
int fir_upad(int D, int ntaps);
// rotator behind freq_xlating_fir_filter: y[n] *= e^{j (theta + n delta)} in place (fir.cu); angle bookkeeping on the host
int rotate_launch(void* y, long long n, double theta, double delta, cudaStream_t st);
double rotate_advance(double theta, double delta, long long n);
bool fft_size_is_fast(int N);
bool fft_size_is_big(int N); // powers of two 32768 .. 262144: four-step over the single-CTA transform
int fft_direct_max();
int fft_vcc_launch(int N, bool fwd, const void* in, void* out, long long nvec, const float* window, int shift,
                   int magsq, const float2* tw, int nsm, cudaStream_t st);
int magsq_launch(const void* in, float* out, long long n, int nsm, cudaStream_t st);
// generalised vector FFT launch (fftvcc.cu): real input (fft_vfc), strided input vectors and the log epilogue (logpwrfft)
struct VccArgs {
    const void* in;
    void* out;
    long long nvec;
    const float* window;
    const float2* tw;
    int shift, mode; // mode 0 complex out, 1 |X|^2, 2 log_mul * log2(max(|X|^2, FLT_MIN) * log_scale)
    long long vstride, voff;
    float log_mul, log_scale;
    float* ystate;
    // mode 3 (streaming kernel only, fft_vcc_stream_iir_launch): |X|^2 through the single-pole IIR along the frames, CTA r
    // filtering its run of run_len consecutive frames from a zero state and leaving the run's last value in agg[r][bin]
    double alpha = 0.0;
    float* agg = nullptr;
    long long run_len = 0;
};
int fft_vcc_launch_ex(int N, bool fwd, bool real_in, const VccArgs& a, int nsm, cudaStream_t st);
// logpwr.cu: single_pole_iir along the frames + nlog10, in place on p [nframes][N]
// capi_resamp.cu: the streaming polyphase kernel (register ring per (phase, run)); *done = false when the geometry does not fit
int resamp_stream_try(bool ct, const void* in, void* out, long long nout, unsigned ctr0, unsigned L, unsigned D, int T,
                      const float* rev_tap_major, int nsm, cudaStream_t st, bool* done);
long long logpwr_scratch_floats(long long nframes, int N, int nsm);
// the FFT kernel with the run-local IIR in its epilogue (sizes whose vectors take a CTA each: 2048, 4096, 8192); *done = false
// when the call does not fit (then nothing was launched).  agg: *nruns x N floats, at most logpwr_fused_scratch_floats().
int fft_vcc_stream_iir_launch(int N, bool real_in, VccArgs a, double alpha, float* agg, int nsm, cudaStream_t st,
                              long long* run_len, long long* nruns, bool* done);
long long logpwr_fused_scratch_floats(int N, int nsm);
int logpwr_carry_fixup_launch(float* p, long long nframes, int N, long long run_len, long long nruns, double alpha, float* agg,
                              float* ystate, float log_mul, float log_scale, cudaStream_t st);
long long logpwr_static_scratch_floats();
int logpwr_static_check(long long nframes, int N, int nsm);
int logpwr_iir_log_launch_static(float* p, long long nframes, int N, double alpha, float* ystate, float log_mul,
                                 float log_scale, int nsm, int device, cudaStream_t st);
int logpwr_iir_log_launch(float* p, long long nframes, int N, double alpha, float* ystate, float* scratch, float log_mul,
                          float log_scale, int nsm, cudaStream_t st);
int fftfilt_device_plan(int ntaps, int* parts);
int fftfilt_launch(int L, int parts, const void* in, void* out, long long nout, int ntaps, int D, const float2* H,
                   const float2* tw, int nsm, cudaStream_t st);
int pfb_launch(int M, int T, int R, const void* x, void* out, long long out_stride, int nouts, const int* cmap,
               int identity, long long noutput, const float* h, const float2* wM, int nsm, bool allow_fast,
               cudaStream_t st, const void* halo = nullptr, const int* inv = nullptr);
bool pfb_stream_supported(int M, int T, int R, const void* x);
int pfb_stream_launch(int M, int T, const void* x, void* out, long long out_stride, int nouts, long long noutput,
                      const float* h, const float2* wM, int nsm, cudaStream_t st, const void* halo = nullptr,
                      const int* inv = nullptr);
// pfb_synthesizer_ccf (pfb_synth.cu)
bool pfb_synth_fast_supported(int M, int T, int twox, const void* x, long long stride, int nin);
int pfb_synth_fast_launch(int M, int T, const void* x, long long stride, int nin, void* out, long long ntimes, const float* h,
                          const float2* wM, const int* inv, int nsm, cudaStream_t st);
long long pfb_synth_slab(int NF, int T);
int pfb_synth_general_launch(int M, int twox, int T, const void* x, long long stride, void* out, long long ntimes,
                             const float* h, const float2* twNF, const int* inv, int par0, void* scratch, long long slab_rows,
                             int nsm, cudaStream_t st);
int filterbank_launch(const void* x, void* out, long long nout, int nf, int T, const float* hrev, const unsigned char* active,
                      int nsm, cudaStream_t st);
int streams_to_interleaved_launch(const void* in, long long in_stride, void* out, int M, long long nrows,
                                  cudaStream_t st);

// staging granularity (input bytes per chunk); GRB200_CHUNK_MB overrides it (tuning knob, 1..1024)
inline size_t chunk_bytes()
{
    static const size_t v = [] {
        const char* e = getenv("GRB200_CHUNK_MB");
        long mb = e ? atol(e) : 0;
        if (mb < 1 || mb > 1024)
            mb = 64;
        return (size_t)mb << 20;
    }();
    return v;
}
#define kChunkBytes (grb::chunk_bytes())

// Chunk size for a host-staged call that moves `total` input bytes.  Big calls use the configured chunk.  A call that
// would fit in ONE chunk but is larger than 1 MiB is still cut in four, so that its H2D copy, kernel and D2H copy overlap
// (three streams, three slots) instead of running back to back -- scheduler-sized work() calls (32 KiB .. 32 MiB) are the
// normal case inside a flowgraph.  Below 1 MiB the per-chunk launch and copy set-up costs more than the overlap gains.
inline size_t chunk_bytes_for(size_t total)
{
    const size_t cb = chunk_bytes();
    if (total <= ((size_t)1 << 20) || total >= 4 * cb)
        return cb;
    size_t c = (total + 3) / 4;
    c = (c + 4095) & ~(size_t)4095;
    return c < cb ? c : cb;
}

struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    int upload(const void* host, size_t bytes)
    {
        if (bytes > cap) {
            if (p)
                cudaFree(p);
            p = nullptr;
            cudaError_t e = cudaMalloc(&p, bytes ? bytes : 16);
            if (e != cudaSuccess) {
                cap = 0;
                return fail(GRB200_ENOMEM, "cudaMalloc(%zu) failed: %s", bytes, cudaGetErrorString(e));
            }
            cap = bytes;
        }
        if (bytes)
            GRB_CUDA(cudaMemcpy(p, host, bytes, cudaMemcpyHostToDevice));
        return 0;
    }
    void release()
    {
        if (p)
            cudaFree(p);
        p = nullptr;
        cap = 0;
    }
};
// capi_fftfilt.cu: device plan + filter spectra of the overlap-save kernels for `ntaps` complex taps
int fftfilt_plan_build(const float* taps_re_im, int ntaps, int device, cudaStream_t st, DevBuf& H, DevBuf& tw, int* L_io,
                       int* K_out, int force_L = 0);
// fftfilt.cu: fir_filter_fff on the overlap-save kernel (two halves of a real call as re / im of one complex stream)
int fftfilt_realpair_L(int ntaps);
int fftfilt_realpair_launch(int L, const void* in, void* out, long long nout, int ntaps, int D, const float2* H,
                            const float2* tw, int nsm, cudaStream_t st);


// forward twiddle table e^{-j 2 pi k / N} (or e^{+...} with sign=+1), evaluated in double
inline std::vector<float2> make_roots(int N, int sign)
{
    std::vector<float2> w((size_t)N);
    const double two_pi = 6.283185307179586476925286766559;
    for (int k = 0; k < N; k++) {
        const double a = two_pi * k / N;
        w[k] = make_float2((float)cos(a), (float)(sign * sin(a)));
    }
    return w;
}

// `stream` argument of every *_work_dev: a cudaStream_t as void*.  NULL is what it is everywhere else in CUDA -- the legacy
// default stream (torch's current stream unless the caller switched) -- so producer -> work_dev -> consumer on one stream
// are ordered.  GRB200_STREAM_OWN selects the handle's private non-blocking stream.
inline cudaStream_t pick_stream(void* user, Stager& sg)
{
    return user == GRB200_STREAM_OWN ? sg.s_exec : (cudaStream_t)user;
}

// epilogue of a device-resident launch: remember the caller's stream position so that set_* can wait for it
inline int dev_done(int rc, Stager& sg, cudaStream_t st) { return rc ? rc : sg.note_user_launch(st); }

} // namespace grb

