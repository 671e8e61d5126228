// fftvcc_big.cu -- sizes the single-CTA transform does not take directly: Bluestein (not a power of two) and the four-step
// transform (powers of two above 16384); see fftvcc.cu.
#include "fftvcc.cuh"

namespace grb {

// ---- sizes that are not a power of two (128 <= N <= 8192): Bluestein over the power-of-two transform ---------------------
// FFTW takes any size at N log N; the direct kernel above is O(N^2) (1.2 Gsps at N = 1000).  With c[n] = e^{-j pi n^2 / N}:
//     X[k] = c[k] * sum_n (x[n] c[n]) * conj(c)[k - n]           (n k = (n^2 + k^2 - (k - n)^2) / 2)
// -- a convolution, done circularly at a power of two L >= 2N - 1 exactly like fftfilt.cu's overlap-save block: load
// (+ window, + swap) * chirp, zero padded to L -> cta_fft -> * B (the transformed chirp, 1/L folded in) in registers -> inverse
// with the reversed radix schedule -> * chirp -> the first N values (+ rotate, + |.|^2 / log).  One vector per CTA; the
// backward transform conjugates both tables (the chirp is even in n).  Tables per (device, N) are built at the first launch.
template <int L, bool REAL>
__global__ void __launch_bounds__(L / 16, L <= 8192 ? 8192 / L : 1) // 128 registers: 512 resident threads per SM
fft_bluestein_kernel(const VccArgs a, int N, int forward, const float2* __restrict__ chirp, const float2* __restrict__ Bhat,
                     const float2* __restrict__ twL)
{
    constexpr int E = 16, TPV = L / 16;
    constexpr int R0 = fft_first_radix(L, false), RM = fft_last_radix(L, false), RL = fft_last_radix(L, true);
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float2* S = reinterpret_cast<float2*>(smem_raw);
    const int t = threadIdx.x;
    const float* __restrict__ window = a.window;
    const bool in_swap = (!forward) && a.shift, out_rot = forward && a.shift;
    const int len_in = N / 2, len_out = (N + 1) / 2; // fft_vcc_fftw.cc:79,93,107
    for (long long vec = blockIdx.x; vec < a.nvec; vec += gridDim.x) {
        const long long sv = (vec * a.vstride + a.voff) * N;
        float2 v[E];
#pragma unroll
        for (int b = 0; b < E / R0; b++)
#pragma unroll
            for (int q = 0; q < R0; q++) {
                const int m = (t + b * TPV) + q * (L / R0);
                float2 x = make_float2(0.f, 0.f);
                if (m < N) {
                    int s = m;
                    if (in_swap) {
                        s += len_in;
                        s -= s >= N ? N : 0;
                    }
                    x = REAL ? make_float2(reinterpret_cast<const float*>(a.in)[sv + s], 0.f)
                             : reinterpret_cast<const float2*>(a.in)[sv + s];
                    if (window != nullptr) {
                        const float w = __ldg(window + s);
                        x = __fmul2_rn(x, make_float2(w, w));
                    }
                    float2 c = __ldg(chirp + m);
                    if (!forward)
                        c.y = -c.y;
                    x = cmul(x, c);
                }
                v[b * R0 + q] = x;
            }
        cta_fft<L, true, false>(v, S, t, twL);
#pragma unroll
        for (int b = 0; b < E / RM; b++)
#pragma unroll
            for (int q = 0; q < RM; q++) {
                const int k = (t + b * TPV) + q * (L / RM);
                float2 h = __ldg(Bhat + k);
                if (!forward)
                    h.y = -h.y;
                v[b * RM + q] = cmul(v[b * RM + q], h);
            }
        __syncthreads(); // the forward transform's last gather is done before the inverse reuses S
        cta_fft<L, false, true>(v, S, t, twL);
#pragma unroll
        for (int b = 0; b < E / RL; b++)
#pragma unroll
            for (int q = 0; q < RL; q++) {
                const int k = (t + b * TPV) + q * (L / RL);
                if (k < N) {
                    float2 c = __ldg(chirp + k);
                    if (!forward)
                        c.y = -c.y;
                    int o = k;
                    if (out_rot) {
                        o += N - len_out;
                        o -= o >= N ? N : 0;
                    }
                    vcc_store(a, vec, N, o, cmul(v[b * RL + q], c));
                }
            }
        __syncthreads(); // the next vector's first pass overwrites S
    }
}

struct BluesteinPlan {
    int L = 0;
    float2 *chirp = nullptr, *Bhat = nullptr, *twL = nullptr;
};

int fft_vcc_launch_ex(int N, bool fwd, bool real_in, const VccArgs& a, int nsm, cudaStream_t st);

// tables of (current device, N); built on `st` at the first use (the launch that follows is ordered behind them)
static int bluestein_plan(int N, int nsm, cudaStream_t st, BluesteinPlan* out)
{
    static std::mutex mu;
    static std::vector<std::pair<std::pair<int, int>, BluesteinPlan>> plans;
    int dev = 0;
    GRB_CUDA(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lk(mu);
    for (auto& p : plans)
        if (p.first.first == dev && p.first.second == N) {
            *out = p.second;
            return 0;
        }
    BluesteinPlan pl;
    pl.L = 1024;
    while (pl.L < 2 * N - 1)
        pl.L *= 2;
    const int L = pl.L;
    const double pi = 3.14159265358979323846;
    std::vector<float2> c((size_t)N), b((size_t)L, make_float2(0.f, 0.f)), roots((size_t)L);
    for (int n = 0; n < N; n++) {
        const long long r = ((long long)n * n) % (2LL * N); // n^2 mod 2N: the angle stays small, the table exact to fp32
        const double ang = pi * (double)r / (double)N;
        c[n] = make_float2((float)cos(ang), (float)-sin(ang));
        const float2 bn = make_float2((float)(cos(ang) / L), (float)(sin(ang) / L)); // conj(c) / L
        b[n] = bn;
        if (n)
            b[L - n] = bn;
    }
    for (int i = 0; i < L; i++) {
        const double ang = -2.0 * pi * (double)i / (double)L;
        roots[i] = make_float2((float)cos(ang), (float)sin(ang));
    }
    float2* tmp = nullptr;
    GRB_CUDA(cudaMalloc(&pl.chirp, (size_t)N * 8));
    GRB_CUDA(cudaMalloc(&pl.Bhat, (size_t)L * 8));
    GRB_CUDA(cudaMalloc(&pl.twL, (size_t)L * 8));
    GRB_CUDA(cudaMalloc(&tmp, (size_t)L * 8));
    GRB_CUDA(cudaMemcpyAsync(pl.chirp, c.data(), (size_t)N * 8, cudaMemcpyHostToDevice, st));
    GRB_CUDA(cudaMemcpyAsync(pl.twL, roots.data(), (size_t)L * 8, cudaMemcpyHostToDevice, st));
    GRB_CUDA(cudaMemcpyAsync(tmp, b.data(), (size_t)L * 8, cudaMemcpyHostToDevice, st));
    VccArgs fa{ tmp, pl.Bhat, 1, nullptr, pl.twL, 0, 0, 1, 0, 0.f, 0.f, nullptr };
    int rc = fft_vcc_launch_ex(L, true, false, fa, nsm, st);
    if (rc)
        return rc;
    GRB_CUDA(cudaStreamSynchronize(st)); // the pageable staging vectors and tmp go away
    cudaFree(tmp);
    plans.push_back({ { dev, N }, pl });
    *out = pl;
    return 0;
}

template <int L>
static int launch_bluestein(const BluesteinPlan& pl, int N, bool fwd, bool real_in, const VccArgs& a, int nsm, cudaStream_t st)
{
    const size_t smem = (size_t)fft_smem_elems(L) * 8;
    auto go = [&](auto kern) -> int {
        if (smem > 48 * 1024)
            GRB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        int occ = 1;
        GRB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, L / 16, smem));
        if (occ < 1)
            occ = 1;
        long long grid = (long long)nsm * occ;
        if (grid > a.nvec)
            grid = a.nvec;
        kern<<<(unsigned)grid, L / 16, smem, st>>>(a, N, fwd ? 1 : 0, pl.chirp, pl.Bhat, pl.twL);
        GRB_LAUNCH_CHECK();
        return 0;
    };
    return real_in ? go(fft_bluestein_kernel<L, true>) : go(fft_bluestein_kernel<L, false>);
}

int bluestein_launch(int N, bool fwd, bool real_in, const VccArgs& a, int nsm, cudaStream_t st)
{
    BluesteinPlan pl;
    int rc = bluestein_plan(N, nsm, st, &pl);
    if (rc)
        return rc;
    switch (pl.L) {
    case 1024:
        return launch_bluestein<1024>(pl, N, fwd, real_in, a, nsm, st);
    case 2048:
        return launch_bluestein<2048>(pl, N, fwd, real_in, a, nsm, st);
    case 4096:
        return launch_bluestein<4096>(pl, N, fwd, real_in, a, nsm, st);
    case 8192:
        return launch_bluestein<8192>(pl, N, fwd, real_in, a, nsm, st);
    case 16384:
        return launch_bluestein<16384>(pl, N, fwd, real_in, a, nsm, st);
    }
    return fail(GRB200_ERANGE, "fft_vcc: no transform for size %d", N);
}

// ---- powers of two above 16384 (32768 .. 262144): four-step over the single-CTA transform -------------------------------
// N = R * N2 (R = 8 or 16 columns, N2 = 4096 .. 16384), n = N2 n1 + n2, k = k1 + R k2:
//     X[k1 + R k2] = sum_n2 W_N2^(n2 k2) * [ W_N^(n2 k1) * sum_n1 x[N2 n1 + n2] W_R^(n1 k1) ]
// step 1 (big_cols_kernel): thread (vector, n2) loads its R items N2 apart (coalesced over n2; window and input swap here),
//   radix-R transform in registers, twiddle from the N-entry table, Z[k1][n2] to scratch;   step 2: the batched N2-point
//   kernel over the R rows of every vector, in place;   step 3 (big_rows_kernel): thread (vector, k2) gathers its R values
//   and stores R consecutive outputs (+ rotate, + |.|^2 / log).  Three passes over HBM instead of one: ~48 B per sample.
template <int R, bool FWD, bool REAL>
__global__ void __launch_bounds__(256)
big_cols_kernel(const VccArgs a, int N, int N2, long long vbase, long long nb, float2* __restrict__ Z)
{
    const bool in_swap = (!FWD) && a.shift;
    const int per = N2 / 256; // blocks of 256 consecutive n2 per vector (N2 is a multiple of 256)
    for (long long blk = blockIdx.x; blk < nb * per; blk += gridDim.x) {
        const long long vec = blk / per;
        const int n2 = (int)(blk - vec * per) * 256 + threadIdx.x;
        const long long sv = ((vbase + vec) * a.vstride + a.voff) * N;
        float2 v[R];
#pragma unroll
        for (int n1 = 0; n1 < R; n1++) {
            int d = n2 + N2 * n1;
            if (in_swap)
                d = (d + N / 2) & (N - 1);
            float2 x = REAL ? make_float2(reinterpret_cast<const float*>(a.in)[sv + d], 0.f)
                            : reinterpret_cast<const float2*>(a.in)[sv + d];
            if (a.window != nullptr) {
                const float w = __ldg(a.window + d);
                x = __fmul2_rn(x, make_float2(w, w));
            }
            v[n1] = x;
        }
        dft_r<R, FWD>(v);
        // v[k1] *= W_N^(n2 k1): the powers 1, 2, 4, 8 from the table (coalesced over n2), the rest are products
        apply_twiddles_l<R>(v, [&](int j) {
            float2 w = __ldg(a.tw + (((long long)n2 << j) & (N - 1)));
            if (!FWD)
                w.y = -w.y;
            return w;
        });
#pragma unroll
        for (int k1 = 0; k1 < R; k1++)
            Z[(vec * R + k1) * N2 + n2] = v[k1];
    }
}

template <int R>
__global__ void __launch_bounds__(256)
big_rows_kernel(const VccArgs a, int N, int N2, int forward, long long vbase, long long nb, const float2* __restrict__ Y)
{
    // a block owns 256 consecutive k2 of one vector = 256 R consecutive outputs k = k1 + R k2: gathered (coalesced over k2 for
    // every k1) into shared memory, rows padded by one, then stored in output order (coalesced over k)
    __shared__ float2 sh[256 * (R + 1)];
    const bool out_rot = forward && a.shift;
    const int per = N2 / 256, t = threadIdx.x;
    for (long long blk = blockIdx.x; blk < nb * per; blk += gridDim.x) {
        const long long vec = blk / per;
        const int k20 = (int)(blk - vec * per) * 256;
        __syncthreads();
#pragma unroll
        for (int k1 = 0; k1 < R; k1++)
            sh[t * (R + 1) + k1] = Y[(vec * R + k1) * N2 + k20 + t];
        __syncthreads();
#pragma unroll
        for (int i = 0; i < R; i++) {
            const int e = t + 256 * i; // k - R k20
            int k = R * k20 + e;
            if (out_rot)
                k = (k + N / 2) & (N - 1);
            vcc_store(a, vbase + vec, N, k, sh[(e / R) * (R + 1) + (e % R)]);
        }
    }
}

__global__ void big_roots_kernel(const float2* __restrict__ twN, float2* __restrict__ tw2, int N2, int R)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < N2)
        tw2[i] = twN[(long long)i * R];
}

bool fft_size_is_big(int N) { return N >= 32768 && N <= 262144 && (N & (N - 1)) == 0; }

int fft_vcc_launch_ex(int N, bool fwd, bool real_in, const VccArgs& a, int nsm, cudaStream_t st);

// scratch of (current device): grow-only, shared by every big transform on the device (calls on different streams serialise
// on the mutex only while launching; the buffer itself is used in stream order -- one stream at a time per device is assumed,
// a second stream gets its own buffer)
struct BigScratch {
    int dev;
    cudaStream_t st;
    float2* p;
    size_t cap;
    float2* tw2;
    int tw2_N;
};

int big_launch(int N, bool fwd, bool real_in, const VccArgs& a, int nsm, cudaStream_t st)
{
    static std::mutex mu;
    static std::vector<BigScratch> pool;
    int dev = 0;
    GRB_CUDA(cudaGetDevice(&dev));
    const int N2 = N <= 65536 ? 4096 : N / 16, R = N / N2;
    const size_t per_vec = (size_t)N * 8;
    long long batch = (long long)(((size_t)256 << 20) / per_vec);
    batch = batch < 1 ? 1 : (batch > a.nvec ? a.nvec : batch);
    std::lock_guard<std::mutex> lk(mu);
    BigScratch* sc = nullptr;
    for (auto& q : pool)
        if (q.dev == dev && q.st == st)
            sc = &q;
    if (!sc) {
        pool.push_back({ dev, st, nullptr, 0, nullptr, 0 });
        sc = &pool.back();
    }
    if (sc->cap < (size_t)batch * per_vec) {
        GRB_CUDA(cudaStreamSynchronize(st));
        if (sc->p)
            cudaFree(sc->p);
        sc->p = nullptr;
        sc->cap = 0;
        cudaError_t e = cudaMalloc(&sc->p, (size_t)batch * per_vec);
        if (e != cudaSuccess)
            return fail(GRB200_ENOMEM, "cudaMalloc(%zu) failed: %s", (size_t)batch * per_vec, cudaGetErrorString(e));
        sc->cap = (size_t)batch * per_vec;
    }
    if (sc->tw2_N != N) { // roots of the N2-point rows: every R-th root of N
        GRB_CUDA(cudaStreamSynchronize(st));
        if (sc->tw2)
            cudaFree(sc->tw2);
        sc->tw2 = nullptr;
        GRB_CUDA(cudaMalloc(&sc->tw2, (size_t)N2 * 8));
        big_roots_kernel<<<(N2 + 255) / 256, 256, 0, st>>>(a.tw, sc->tw2, N2, R);
        GRB_LAUNCH_CHECK();
        sc->tw2_N = N;
    }
    for (long long vbase = 0; vbase < a.nvec; vbase += batch) {
        const long long nb = a.nvec - vbase < batch ? a.nvec - vbase : batch;
        long long grid = (nb * N2 + 255) / 256;
        if (grid > (long long)nsm * 16)
            grid = (long long)nsm * 16;
        auto cols = [&](auto r) -> int {
            constexpr int Rv = decltype(r)::value;
            if (fwd) {
                if (real_in)
                    big_cols_kernel<Rv, true, true><<<(unsigned)grid, 256, 0, st>>>(a, N, N2, vbase, nb, sc->p);
                else
                    big_cols_kernel<Rv, true, false><<<(unsigned)grid, 256, 0, st>>>(a, N, N2, vbase, nb, sc->p);
            } else {
                if (real_in)
                    big_cols_kernel<Rv, false, true><<<(unsigned)grid, 256, 0, st>>>(a, N, N2, vbase, nb, sc->p);
                else
                    big_cols_kernel<Rv, false, false><<<(unsigned)grid, 256, 0, st>>>(a, N, N2, vbase, nb, sc->p);
            }
            GRB_LAUNCH_CHECK();
            return 0;
        };
        int rc = R == 8 ? cols(std::integral_constant<int, 8>{}) : cols(std::integral_constant<int, 16>{});
        if (rc)
            return rc;
        VccArgs ra{ sc->p, sc->p, nb * R, nullptr, sc->tw2, 0, 0, 1, 0, 0.f, 0.f, nullptr };
        if ((rc = fft_vcc_launch_ex(N2, fwd, false, ra, nsm, st)))
            return rc;
        if (R == 8)
            big_rows_kernel<8><<<(unsigned)grid, 256, 0, st>>>(a, N, N2, fwd ? 1 : 0, vbase, nb, sc->p);
        else
            big_rows_kernel<16><<<(unsigned)grid, 256, 0, st>>>(a, N, N2, fwd ? 1 : 0, vbase, nb, sc->p);
        GRB_LAUNCH_CHECK();
    }
    return 0;
}


} // namespace grb
