// fft.cuh -- CTA-level shared-memory FFT building block (K7 in SURVEY.md section 2a).
//
// Replaces gr::fft::fft_complex::execute (reference gr-fft/lib/fft.cc:151-203, an FFTW3f plan):
// unnormalised DFT, forward e^{-j2pi nk/N}, backward e^{+j2pi nk/N}.
//
// Stockham autosort, decimation in time, radix-16/8/4/2 passes.  A vector of N = 2^n points is owned by
// N/E threads (E = 16 elements per thread for N >= 64).  Pass i with radix R and Ns = product of the
// earlier radices does, for butterfly j in [0, N/R):
//     v[q]   = x[j + q*N/R]                         q = 0..R-1
//     v[q]  *= W_{Ns*R}^{q*(j mod Ns)}
//     v      = DFT_R(v)
//     y[(j div Ns)*Ns*R + (j mod Ns) + q*Ns] = v[q]
// so the first pass reads global memory with unit stride across threads and the last pass writes global
// memory with unit stride across threads: no bit reversal and no separate transpose.  Between passes the
// data goes through shared memory in place (read all -> barrier -> write all -> barrier); the index is
// padded by one complex every 16 so the R-strided writes of the first pass are bank-conflict free.
// Twiddles come from a per-size table computed in double on the host (tw[m] = e^{-j2pi m/N}); per
// butterfly only the powers 1,2,4,8 are looked up, the rest are products (<= 3 roundings deep).
#pragma once
#include "common.cuh"

namespace grb {

__host__ __device__ constexpr int fft_elems(int N) { return N >= 64 ? 16 : (N == 32 ? 8 : N); }
__host__ __device__ constexpr int fft_pick_radix(int rem) { return rem == 32 ? 8 : (rem >= 16 ? 16 : rem); }
__host__ __device__ constexpr int fft_pad(int i) { return i + (i >> 4); }
__host__ __device__ constexpr int fft_smem_elems(int N) { return N + (N >> 4) + 1; }

// ---- in-register DFTs, natural-order in and out -----------------------------------------------
template <bool FWD>
__device__ __forceinline__ void dft2(float2& a, float2& b)
{
    float2 t = a;
    a = cadd(t, b);
    b = csub(t, b);
}

template <bool FWD>
__device__ __forceinline__ void dft4(float2& v0, float2& v1, float2& v2, float2& v3)
{
    const float2 a0 = cadd(v0, v2), a1 = csub(v0, v2), a2 = cadd(v1, v3), a3 = csub(v1, v3);
    v0 = cadd(a0, a2);
    v1 = cadd_mulj<FWD>(a1, a3);
    v2 = csub(a0, a2);
    v3 = csub_mulj<FWD>(a1, a3);
}

// multiply by e^{-+j pi/4 * m}, m = 1 or 3 (sign by FWD)
template <bool FWD>
__device__ __forceinline__ float2 cmul_w8_1(float2 a)
{
    const float s = 0.70710678118654752440f; // (1 -+ j)/sqrt2 * a = s * (a + (-+j) a)
    return __fmul2_rn(make_float2(s, s), cadd_mulj<FWD>(a, a));
}
template <bool FWD>
__device__ __forceinline__ float2 cmul_w8_3(float2 a)
{
    const float s = 0.70710678118654752440f; // (-1 -+ j)/sqrt2 * a = s * ((-+j) a - a)
    return __fmul2_rn(make_float2(-s, -s), csub_mulj<FWD>(a, a));
}

template <bool FWD>
__device__ __forceinline__ void dft8(float2* v)
{
    // even/odd split (DIT)
    float2 e0 = v[0], e1 = v[2], e2 = v[4], e3 = v[6];
    float2 o0 = v[1], o1 = v[3], o2 = v[5], o3 = v[7];
    dft4<FWD>(e0, e1, e2, e3);
    dft4<FWD>(o0, o1, o2, o3);
    o1 = cmul_w8_1<FWD>(o1);
    o2 = cmulj<FWD>(o2);
    o3 = cmul_w8_3<FWD>(o3);
    v[0] = cadd(e0, o0);
    v[4] = csub(e0, o0);
    v[1] = cadd(e1, o1);
    v[5] = csub(e1, o1);
    v[2] = cadd(e2, o2);
    v[6] = csub(e2, o2);
    v[3] = cadd(e3, o3);
    v[7] = csub(e3, o3);
}

template <bool FWD>
__device__ __forceinline__ float2 cmul_const(float2 a, float c, float s)
{
    // multiply by (c - j s) forward, (c + j s) inverse:  c*a + s*((-+j) a)
    const float2 t = __fmul2_rn(make_float2(s, s), cmulj<FWD>(a));
    return __ffma2_rn(make_float2(c, c), a, t);
}

template <bool FWD>
__device__ __forceinline__ void dft16(float2* v)
{
    // 4x4 Cooley-Tukey: n = 4*n1 + n2, k = k1 + 4*k2
    const float c1 = 0.92387953251128675613f, s1 = 0.38268343236508977173f; // cos,sin(pi/8)
    const float c2 = 0.70710678118654752440f;
    float2 y[4][4]; // y[n2][k1]
#pragma unroll
    for (int n2 = 0; n2 < 4; n2++) {
        y[n2][0] = v[n2];
        y[n2][1] = v[4 + n2];
        y[n2][2] = v[8 + n2];
        y[n2][3] = v[12 + n2];
        dft4<FWD>(y[n2][0], y[n2][1], y[n2][2], y[n2][3]);
    }
    // twiddles W16^{n2*k1}
    y[1][1] = cmul_const<FWD>(y[1][1], c1, s1);  // W^1
    y[1][2] = cmul_w8_1<FWD>(y[1][2]);           // W^2
    y[1][3] = cmul_const<FWD>(y[1][3], s1, c1);  // W^3
    y[2][1] = cmul_w8_1<FWD>(y[2][1]);           // W^2
    y[2][2] = cmulj<FWD>(y[2][2]);               // W^4
    y[2][3] = cmul_w8_3<FWD>(y[2][3]);           // W^6
    y[3][1] = cmul_const<FWD>(y[3][1], s1, c1);  // W^3
    y[3][2] = cmul_w8_3<FWD>(y[3][2]);           // W^6
    y[3][3] = cmul_const<FWD>(y[3][3], -c1, -s1); // W^9 = -(c1 -+ j s1)... see note
    (void)c2;
#pragma unroll
    for (int k1 = 0; k1 < 4; k1++) {
        dft4<FWD>(y[0][k1], y[1][k1], y[2][k1], y[3][k1]);
        v[k1] = y[0][k1];
        v[k1 + 4] = y[1][k1];
        v[k1 + 8] = y[2][k1];
        v[k1 + 12] = y[3][k1];
    }
}

template <int R, bool FWD>
__device__ __forceinline__ void dft_r(float2* v)
{
    if constexpr (R == 1)
        return;
    else if constexpr (R == 2)
        dft2<FWD>(v[0], v[1]);
    else if constexpr (R == 4)
        dft4<FWD>(v[0], v[1], v[2], v[3]);
    else if constexpr (R == 8)
        dft8<FWD>(v);
    else
        dft16<FWD>(v);
}

// ---- twiddles: v[q] *= W^{q}, the powers 1,2,4,8 of W come from look(0..3), the rest are products ----
template <int R, class Look>
__device__ __forceinline__ void apply_twiddles_l(float2* v, Look look)
{
    float2 w1 = look(0);
    v[1] = cmul(v[1], w1);
    if constexpr (R >= 4) {
        float2 w2 = look(1);
        float2 w3 = cmul(w2, w1);
        v[2] = cmul(v[2], w2);
        v[3] = cmul(v[3], w3);
        if constexpr (R >= 8) {
            float2 w4 = look(2);
            float2 w5 = cmul(w4, w1), w6 = cmul(w4, w2), w7 = cmul(w4, w3);
            v[4] = cmul(v[4], w4);
            v[5] = cmul(v[5], w5);
            v[6] = cmul(v[6], w6);
            v[7] = cmul(v[7], w7);
            if constexpr (R >= 16) {
                float2 w8 = look(3);
                v[8] = cmul(v[8], w8);
                v[9] = cmul(v[9], cmul(w8, w1));
                v[10] = cmul(v[10], cmul(w8, w2));
                v[11] = cmul(v[11], cmul(w8, w3));
                v[12] = cmul(v[12], cmul(w8, w4));
                v[13] = cmul(v[13], cmul(w8, w5));
                v[14] = cmul(v[14], cmul(w8, w6));
                v[15] = cmul(v[15], cmul(w8, w7));
            }
        }
    }
}

// v[q] *= W_N^{q*m}, tw = forward table of N entries in global memory
template <int R, int N, bool FWD>
__device__ __forceinline__ void apply_twiddles(float2* v, int m, const float2* __restrict__ tw)
{
    // m = (j mod Ns) * N/(Ns*R); q*m < N for every q < R
    apply_twiddles_l<R>(v, [&](int k) {
        float2 w = __ldg(tw + (m << k));
        return FWD ? w : cconj(w);
    });
}

// ---- radix schedule ------------------------------------------------------------------------------------
// Forward order: greedy 16s, a remainder of 32 split as 8*4 (descending radices).  REV = the same radices
// in ascending order; its first-pass register pattern equals the forward order's last-pass pattern, which
// lets fft_filter chain FFT -> multiply -> inverse FFT without a shared-memory exchange in between.
__host__ __device__ constexpr int fft_num_passes(int N)
{
    int ns = 1, n = 0;
    while (ns < N) {
        ns *= fft_pick_radix(N / ns);
        n++;
    }
    return n;
}
__host__ __device__ constexpr int fft_radix_fwd(int N, int pass)
{
    int ns = 1, r = 1;
    for (int i = 0; i <= pass; i++) {
        r = fft_pick_radix(N / ns);
        ns *= r;
    }
    return r;
}
__host__ __device__ constexpr int fft_radix_at(int N, bool rev, int pass)
{
    return rev ? fft_radix_fwd(N, fft_num_passes(N) - 1 - pass) : fft_radix_fwd(N, pass);
}
__host__ __device__ constexpr int fft_first_radix(int N, bool rev) { return fft_radix_at(N, rev, 0); }
__host__ __device__ constexpr int fft_last_radix(int N, bool rev = false)
{
    return fft_radix_at(N, rev, fft_num_passes(N) - 1);
}

// ---- one vector's FFT across its N/E threads -------------------------------------------------------
// Register convention on entry: v[b*R0 + q] = x[(t + b*TPV) + q*N/R0], R0 = fft_first_radix(N, REV),
// t = thread index within the vector, TPV = N/E threads per vector.
// On exit (after the last pass): v[b*RL + q] = X[(t + b*TPV) + q*N/RL], RL = fft_last_radix(N, REV).
// `S` is this vector's shared-memory scratch of fft_smem_elems(N) complex.  All threads of the CTA must
// call this together (it contains __syncthreads()).
template <int N, bool FWD, bool REV = false, int PASS = 0, int NS = 1>
__device__ __forceinline__ void cta_fft(float2 (&v)[fft_elems(N)], float2* S, int t,
                                        const float2* __restrict__ tw)
{
    constexpr int E = fft_elems(N);
    constexpr int TPV = N / E;
    constexpr int R = fft_radix_at(N, REV, PASS);
    constexpr int NB = E / R; // butterflies per thread in this pass
    if constexpr (NS > 1) {
        // gather this pass's inputs from shared memory
#pragma unroll
        for (int b = 0; b < NB; b++) {
            const int j = t + b * TPV;
#pragma unroll
            for (int q = 0; q < R; q++)
                v[b * R + q] = S[fft_pad(j + q * (N / R))];
        }
        // twiddle
#pragma unroll
        for (int b = 0; b < NB; b++) {
            const int j = t + b * TPV;
            const int m = (j % NS) * (N / (NS * R));
            apply_twiddles<R, N, FWD>(v + b * R, m, tw);
        }
    }
#pragma unroll
    for (int b = 0; b < NB; b++)
        dft_r<R, FWD>(v + b * R);
    if constexpr (NS * R < N) {
        if constexpr (NS > 1)
            __syncthreads(); // everyone has read its inputs before anyone overwrites them
#pragma unroll
        for (int b = 0; b < NB; b++) {
            const int j = t + b * TPV;
            const int base = (j / NS) * (NS * R) + (j % NS);
#pragma unroll
            for (int q = 0; q < R; q++)
                S[fft_pad(base + q * NS)] = v[b * R + q];
        }
        __syncthreads();
        cta_fft<N, FWD, REV, PASS + 1, NS * R>(v, S, t, tw);
    }
}

// ---- variant with per-thread twiddles precomputed in shared memory -----------------------------------
// In a persistent kernel thread t meets the same butterflies for every vector, so the base twiddles
// (powers 1,2,4,8 of W^m per butterfly) are loop invariant: cta_fft_tws_init() writes them once to
// TW[slot][TPV] (column t, conflict-free), cta_fft_tws() reads them back with LDS instead of going to
// the global table every pass.  A pass whose butterflies of one thread all share m (TPV % NS == 0, true
// for every pass but the small-radix tail of the big sizes) takes log2(R) slots; the others keep the
// global table.  `after_first` runs once, right after the barrier that follows the first pass's writes
// (all threads have consumed their input by then).
__host__ __device__ constexpr int fft_ilog2(int r) { return r == 16 ? 4 : (r == 8 ? 3 : (r == 4 ? 2 : (r == 2 ? 1 : 0))); }
__host__ __device__ constexpr int fft_ns_at(int N, bool rev, int pass)
{
    int ns = 1;
    for (int p = 0; p < pass; p++)
        ns *= fft_radix_at(N, rev, p);
    return ns;
}
__host__ __device__ constexpr bool fft_tws_pass_in_smem(int N, bool rev, int pass)
{
    return pass >= 1 && ((N / fft_elems(N)) % fft_ns_at(N, rev, pass)) == 0;
}
__host__ __device__ constexpr int fft_tws_offset(int N, bool rev, int pass)
{
    int off = 0;
    for (int p = 1; p < pass; p++)
        if (fft_tws_pass_in_smem(N, rev, p))
            off += fft_ilog2(fft_radix_at(N, rev, p));
    return off;
}
__host__ __device__ constexpr int fft_tws_slots(int N, bool rev = false)
{
    return fft_tws_offset(N, rev, fft_num_passes(N));
}

template <int N, bool FWD, bool REV = false, int PASS = 1>
__device__ __forceinline__ void cta_fft_tws_init(float2* TW, int t, const float2* __restrict__ tw)
{
    if constexpr (PASS < fft_num_passes(N)) {
        constexpr int E = fft_elems(N), TPV = N / E;
        constexpr int R = fft_radix_at(N, REV, PASS), LG = fft_ilog2(R);
        constexpr int NS = fft_ns_at(N, REV, PASS);
        constexpr int OFF = fft_tws_offset(N, REV, PASS);
        if constexpr (fft_tws_pass_in_smem(N, REV, PASS)) {
            const int m = (t % NS) * (N / (NS * R));
#pragma unroll
            for (int k = 0; k < LG; k++) {
                const float2 w = tw[m << k];
                TW[(OFF + k) * TPV + t] = FWD ? w : cconj(w);
            }
        }
        cta_fft_tws_init<N, FWD, REV, PASS + 1>(TW, t, tw);
    }
}

template <int N, bool FWD, bool REV = false, int PASS = 0, int NS = 1, class After>
__device__ __forceinline__ void cta_fft_tws(float2 (&v)[fft_elems(N)], float2* S, int t, const float2* TW,
                                            const float2* __restrict__ tw, After after_first)
{
    constexpr int E = fft_elems(N);
    constexpr int TPV = N / E;
    constexpr int R = fft_radix_at(N, REV, PASS);
    constexpr int NB = E / R;
    if constexpr (NS > 1) {
        constexpr int LG = fft_ilog2(R), OFF = fft_tws_offset(N, REV, PASS);
#pragma unroll
        for (int b = 0; b < NB; b++) {
            const int j = t + b * TPV;
#pragma unroll
            for (int q = 0; q < R; q++)
                v[b * R + q] = S[fft_pad(j + q * (N / R))];
        }
        if constexpr (fft_tws_pass_in_smem(N, REV, PASS)) {
            float2 base[LG];
#pragma unroll
            for (int k = 0; k < LG; k++)
                base[k] = TW[(OFF + k) * TPV + t];
#pragma unroll
            for (int b = 0; b < NB; b++)
                apply_twiddles_l<R>(v + b * R, [&](int k) { return base[k]; });
        } else {
#pragma unroll
            for (int b = 0; b < NB; b++) {
                const int j = t + b * TPV;
                apply_twiddles<R, N, FWD>(v + b * R, (j % NS) * (N / (NS * R)), tw);
            }
        }
    }
#pragma unroll
    for (int b = 0; b < NB; b++)
        dft_r<R, FWD>(v + b * R);
    if constexpr (NS * R < N) {
        if constexpr (NS > 1)
            __syncthreads();
#pragma unroll
        for (int b = 0; b < NB; b++) {
            const int j = t + b * TPV;
            const int base = (j / NS) * (NS * R) + (j % NS);
#pragma unroll
            for (int q = 0; q < R; q++)
                S[fft_pad(base + q * NS)] = v[b * R + q];
        }
        __syncthreads();
        if constexpr (PASS == 0)
            after_first();
        cta_fft_tws<N, FWD, REV, PASS + 1, NS * R>(v, S, t, TW, tw, after_first);
    }
}

} // namespace grb
