/* oracle/shim/gnuradio/fft/fft.h -- TEST INFRASTRUCTURE ONLY (never on the product path).
 *
 * Stand-in for gr::fft::{fft_complex, fft_real_fwd, fft_real_rev}
 * (/root/reference/gr-fft/include/gnuradio/fft/fft.h:56-196, gr-fft/lib/fft.cc:151-323), which in
 * the reference wrap FFTW3f (external, absent here).  Same calling convention: the caller fills
 * get_inbuf(), calls execute(), reads get_outbuf(); transforms are UNNORMALISED, forward uses
 * e^{-j2pi nk/N}, backward e^{+j2pi nk/N} (FFTW_FORWARD/FFTW_BACKWARD).
 *
 * The engine below is our own: single-precision Stockham autosort radix-2 passes (twiddles
 * evaluated in double, stored as float) for power-of-two sizes, and a double-accumulated direct
 * DFT for every other size.  It is an oracle stand-in for FFTW, cross-checked against numpy.fft
 * in tests/test_oracle.py.
 */
#ifndef GRB200_ORACLE_SHIM_FFT_H
#define GRB200_ORACLE_SHIM_FFT_H

#include <gnuradio/gr_complex.h>
#include <algorithm>
#include <cmath>
#include <memory>
#include <stdexcept>
#include <vector>

namespace gr {
namespace fft {

namespace shim_detail {

struct engine {
    int n;
    bool pow2;
    int sign; /* -1 forward, +1 backward */
    std::vector<float> wr, wi;   /* pow2: n/2 twiddles e^{sign*j*2pi*k/n} */
    std::vector<double> dr, di;  /* other sizes: full table of n roots */
    std::vector<float> sr, si, tr, ti; /* split re/im work arrays (ping-pong) */

    engine(int size, bool forward) : n(size), sign(forward ? -1 : +1)
    {
        if (size <= 0)
            throw std::out_of_range("fft_impl_fftw: invalid fft_size");
        pow2 = (n & (n - 1)) == 0;
        const double two_pi = 6.283185307179586476925286766559;
        if (pow2) {
            wr.resize(std::max(1, n / 2));
            wi.resize(std::max(1, n / 2));
            for (int k = 0; k < n / 2; k++) {
                wr[k] = (float)std::cos(two_pi * k / n);
                wi[k] = (float)(sign * std::sin(two_pi * k / n));
            }
            sr.resize(n);
            si.resize(n);
            tr.resize(n);
            ti.resize(n);
        } else {
            dr.resize(n);
            di.resize(n);
            for (int k = 0; k < n; k++) {
                dr[k] = std::cos(two_pi * k / n);
                di[k] = sign * std::sin(two_pi * k / n);
            }
        }
    }

    void run(const gr_complex* in, gr_complex* out)
    {
        if (!pow2) {
            for (int k = 0; k < n; k++) {
                double ar = 0.0, ai = 0.0;
                long idx = 0;
                for (int t = 0; t < n; t++) {
                    const double xr = in[t].real(), xi = in[t].imag();
                    ar += xr * dr[idx] - xi * di[idx];
                    ai += xr * di[idx] + xi * dr[idx];
                    idx += k;
                    if (idx >= n)
                        idx -= n;
                }
                out[k] = gr_complex((float)ar, (float)ai);
            }
            return;
        }
        if (n == 1) {
            out[0] = in[0];
            return;
        }
        float *ar = sr.data(), *ai = si.data(), *br = tr.data(), *bi = ti.data();
        for (int i = 0; i < n; i++) {
            ar[i] = in[i].real();
            ai[i] = in[i].imag();
        }
        /* Stockham autosort, decimation in frequency: at each pass the current
           sub-transform length is len (halved every pass) and s = n/len interleaved
           sequences are processed side by side; x -> y, then the roles swap. */
        const int half = n / 2;
        for (int len = n, s = 1; len > 1; len >>= 1, s <<= 1) {
            const int m = len / 2;
            for (int p = 0; p < m; p++) {
                const float cr = wr[p * s], ci = wi[p * s];
                const float* x0r = ar + s * p;
                const float* x0i = ai + s * p;
                const float* x1r = x0r + half;
                const float* x1i = x0i + half;
                float* y0r = br + s * 2 * p;
                float* y0i = bi + s * 2 * p;
                float* y1r = y0r + s;
                float* y1i = y0i + s;
                for (int q = 0; q < s; q++) {
                    const float dr_ = x0r[q] - x1r[q];
                    const float di_ = x0i[q] - x1i[q];
                    y0r[q] = x0r[q] + x1r[q];
                    y0i[q] = x0i[q] + x1i[q];
                    y1r[q] = dr_ * cr - di_ * ci;
                    y1i[q] = dr_ * ci + di_ * cr;
                }
            }
            std::swap(ar, br);
            std::swap(ai, bi);
        }
        for (int i = 0; i < n; i++)
            out[i] = gr_complex(ar[i], ai[i]);
    }
};

} // namespace shim_detail

class fft_complex
{
    int d_nthreads;
    std::vector<gr_complex> d_inbuf, d_outbuf;
    shim_detail::engine d_engine;

public:
    fft_complex(int fft_size, bool forward = true, int nthreads = 1)
        : d_nthreads(nthreads),
          d_inbuf(std::max(fft_size, 0)),
          d_outbuf(std::max(fft_size, 0)),
          d_engine(fft_size, forward)
    {
    }
    fft_complex(const fft_complex&) = delete;
    fft_complex& operator=(const fft_complex&) = delete;
    virtual ~fft_complex() {}
    gr_complex* get_inbuf() { return d_inbuf.data(); }
    gr_complex* get_outbuf() { return d_outbuf.data(); }
    int inbuf_length() const { return (int)d_inbuf.size(); }
    int outbuf_length() const { return (int)d_outbuf.size(); }
    void set_nthreads(int n)
    {
        if (n <= 0)
            throw std::out_of_range("gr::fft: invalid number of threads");
        d_nthreads = n;
    }
    int nthreads() const { return d_nthreads; }
    void execute() { d_engine.run(d_inbuf.data(), d_outbuf.data()); }
};

/* real -> half-spectrum (N/2+1 bins), forward sign */
class fft_real_fwd
{
    int d_nthreads;
    int d_n;
    std::vector<float> d_inbuf;
    std::vector<gr_complex> d_outbuf, d_cin, d_cout;
    shim_detail::engine d_engine;

public:
    fft_real_fwd(int fft_size, int nthreads = 1)
        : d_nthreads(nthreads),
          d_n(fft_size),
          d_inbuf(std::max(fft_size, 0)),
          d_outbuf(std::max(fft_size, 0) / 2 + 1),
          d_cin(std::max(fft_size, 0)),
          d_cout(std::max(fft_size, 0)),
          d_engine(fft_size, true)
    {
    }
    fft_real_fwd(const fft_real_fwd&) = delete;
    fft_real_fwd& operator=(const fft_real_fwd&) = delete;
    virtual ~fft_real_fwd() {}
    float* get_inbuf() { return d_inbuf.data(); }
    gr_complex* get_outbuf() { return d_outbuf.data(); }
    int inbuf_length() const { return (int)d_inbuf.size(); }
    int outbuf_length() const { return (int)d_outbuf.size(); }
    void set_nthreads(int n) { d_nthreads = n; }
    int nthreads() const { return d_nthreads; }
    void execute()
    {
        for (int i = 0; i < d_n; i++)
            d_cin[i] = gr_complex(d_inbuf[i], 0.0f);
        d_engine.run(d_cin.data(), d_cout.data());
        for (int i = 0; i < d_n / 2 + 1; i++)
            d_outbuf[i] = d_cout[i];
    }
};

/* half-spectrum (N/2+1 bins, Hermitian) -> real, backward sign */
class fft_real_rev
{
    int d_nthreads;
    int d_n;
    std::vector<gr_complex> d_inbuf;
    std::vector<float> d_outbuf;
    std::vector<gr_complex> d_cin, d_cout;
    shim_detail::engine d_engine;

public:
    fft_real_rev(int fft_size, int nthreads = 1)
        : d_nthreads(nthreads),
          d_n(fft_size),
          d_inbuf(std::max(fft_size, 0) / 2 + 1),
          d_outbuf(std::max(fft_size, 0)),
          d_cin(std::max(fft_size, 0)),
          d_cout(std::max(fft_size, 0)),
          d_engine(fft_size, false)
    {
    }
    fft_real_rev(const fft_real_rev&) = delete;
    fft_real_rev& operator=(const fft_real_rev&) = delete;
    virtual ~fft_real_rev() {}
    gr_complex* get_inbuf() { return d_inbuf.data(); }
    float* get_outbuf() { return d_outbuf.data(); }
    int inbuf_length() const { return (int)d_inbuf.size(); }
    int outbuf_length() const { return (int)d_outbuf.size(); }
    void set_nthreads(int n) { d_nthreads = n; }
    int nthreads() const { return d_nthreads; }
    void execute()
    {
        for (int i = 0; i < d_n / 2 + 1; i++)
            d_cin[i] = d_inbuf[i];
        for (int i = d_n / 2 + 1; i < d_n; i++)
            d_cin[i] = std::conj(d_inbuf[d_n - i]);
        d_engine.run(d_cin.data(), d_cout.data());
        for (int i = 0; i < d_n; i++)
            d_outbuf[i] = d_cout[i].real();
    }
};

} // namespace fft
} // namespace gr

#endif
