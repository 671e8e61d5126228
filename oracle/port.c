/* oracle/port.c -- TEST INFRASTRUCTURE ONLY (never imported, linked or called by the product path).
 *
 * Plain-C CPU restatement of the reference's filtering hot path.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may load the library built from this file.
 * Every function cites the reference file:line (relative to /root/reference) it follows.
 *
 * Parity pinning: this restatement is checked in tests/test_oracle.py against (1) the reference's own
 * kernel sources compiled into oracle/_ref/libgrref.so (build container only), (2) fixtures generated
 * from that library and committed under tests/golden/, (3) the golden vectors of the reference's QA
 * (gr-fft/python/fft/qa_fft.py, gr-filter/python/filter/qa_fir_filter.py, ...), (4) float64 numpy.
 *
 * Third-party arithmetic (VOLK >= 2.1.0, FFTW3f >= 3.0; un-vendored, versions are minimums from
 * /root/reference/CMakeLists.txt:58 and cmake/Modules/FindFFTW3f.cmake:7): restated as scalar float
 * loops (what VOLK_GENERIC=1 runs in the reference's QA) and as a radix-2 Stockham FFT / direct DFT.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
    float re, im;
} cf32;

#define PI_D 3.14159265358979323846

/* ------------------------------------------------------------------------------------------------
 * FIR: kernel::fir_filter<gr_complex,gr_complex,float>::filter / filterN / filterNdec
 * (gr-filter/lib/fir_filter.cc:101-141): taps stored reversed (:64), out[i] = sum_t rev[t]*in[i*D+t].
 * in[] holds nout*decim + ntaps-1 items (history = ntaps, gr-filter/lib/fir_filter_blk_impl.cc:41).
 * ---------------------------------------------------------------------------------------------- */
int grbo_fir_ccf(int decim, const float* taps, int ntaps, const cf32* in, cf32* out, long nout)
{
    float* rev = (float*)malloc(sizeof(float) * (size_t)ntaps);
    if (!rev)
        return -1;
    for (int t = 0; t < ntaps; t++)
        rev[t] = taps[ntaps - 1 - t];
    for (long i = 0; i < nout; i++) {
        const cf32* x = in + i * decim;
        float re = 0.0f, im = 0.0f;
        for (int t = 0; t < ntaps; t++) {
            re += x[t].re * rev[t];
            im += x[t].im * rev[t];
        }
        out[i].re = re;
        out[i].im = im;
    }
    free(rev);
    return 0;
}

/* kernel::fir_filter<gr_complex,gr_complex,gr_complex>::filter (gr-filter/lib/fir_filter.cc:153-162) */
int grbo_fir_ccc(int decim, const cf32* taps, int ntaps, const cf32* in, cf32* out, long nout)
{
    cf32* rev = (cf32*)malloc(sizeof(cf32) * (size_t)ntaps);
    if (!rev)
        return -1;
    for (int t = 0; t < ntaps; t++)
        rev[t] = taps[ntaps - 1 - t];
    for (long i = 0; i < nout; i++) {
        const cf32* x = in + i * decim;
        float re = 0.0f, im = 0.0f;
        for (int t = 0; t < ntaps; t++) {
            re += x[t].re * rev[t].re - x[t].im * rev[t].im;
            im += x[t].re * rev[t].im + x[t].im * rev[t].re;
        }
        out[i].re = re;
        out[i].im = im;
    }
    free(rev);
    return 0;
}

/* ------------------------------------------------------------------------------------------------
 * freq_xlating_fir_filter_impl<gr_complex,gr_complex,float|gr_complex>
 * (gr-filter/lib/freq_xlating_fir_filter_impl.cc): build_composite_fir :74-93 -- float fwT0, composite taps
 * taps[i] * e^{j i fwT0}, rotator increment e^{-j fwT0 decim} normalised (rotator.h:31); work :143-180 -- composite
 * FIR then rotator.rotateN, i.e. VOLK's generic rotator: z = in * phase; phase *= incr; phase /= |phase| after every
 * 512 items and at the end of each call.  work() is driven `chunk_out` outputs at a time like the scheduler does.
 * phase_io = rotator phase in/out ({1,0} for a fresh block).
 * ---------------------------------------------------------------------------------------------- */
static void rot_norm(cf32* p)
{
    const float m = hypotf(p->re, p->im);
    p->re /= m;
    p->im /= m;
}

int grbo_freq_xlating_fir(int cplx_taps, int decim, const float* taps, int ntaps, double center_freq,
                          double sampling_freq, const cf32* in, cf32* out, long nout, long chunk_out, float* phase_io)
{
    cf32* ct = (cf32*)malloc(sizeof(cf32) * (size_t)ntaps);
    if (!ct)
        return -1;
    const float fwT0 = (float)(2 * PI_D * center_freq / sampling_freq);
    for (unsigned int i = 0; i < (unsigned int)ntaps; i++) {
        const float a = (float)i * fwT0, c = cosf(a), sn = sinf(a);
        if (cplx_taps) {
            const float tr = taps[2 * i], ti = taps[2 * i + 1];
            ct[i].re = tr * c - ti * sn;
            ct[i].im = tr * sn + ti * c;
        } else {
            ct[i].re = taps[i] * c;
            ct[i].im = taps[i] * sn;
        }
    }
    cf32 incr, ph = { 1.0f, 0.0f };
    {
        const float a = -fwT0 * (float)decim;
        incr.re = cosf(a);
        incr.im = sinf(a);
        const float m = hypotf(incr.re, incr.im); /* std::abs(complex<float>) */
        incr.re /= m;
        incr.im /= m;
    }
    if (phase_io) {
        ph.re = phase_io[0];
        ph.im = phase_io[1];
        rot_norm(&ph); /* rotator::set_phase */
    }
    if (chunk_out < 1)
        chunk_out = nout > 0 ? nout : 1;
    for (long o0 = 0; o0 < nout; o0 += chunk_out) {
        const long n = (nout - o0 < chunk_out) ? nout - o0 : chunk_out;
        grbo_fir_ccc(decim, ct, ntaps, in + o0 * decim, out + o0, n);
        long i = 0;
        for (; i < n; i++) {
            const cf32 z = out[o0 + i];
            out[o0 + i].re = z.re * ph.re - z.im * ph.im;
            out[o0 + i].im = z.re * ph.im + z.im * ph.re;
            const cf32 q = ph;
            ph.re = q.re * incr.re - q.im * incr.im;
            ph.im = q.re * incr.im + q.im * incr.re;
            if (((i + 1) & 511) == 0)
                rot_norm(&ph);
        }
        if (n & 511)
            rot_norm(&ph);
    }
    if (phase_io) {
        phase_io[0] = ph.re;
        phase_io[1] = ph.im;
    }
    free(ct);
    return 0;
}

/* ------------------------------------------------------------------------------------------------
 * FFT: gr::fft::fft_complex (gr-fft/lib/fft.cc:151-203): unnormalised, forward = e^{-j...},
 * backward = e^{+j...}.  Power-of-two sizes: Stockham radix-2 in float; otherwise a direct DFT
 * accumulated in double.
 * ---------------------------------------------------------------------------------------------- */
typedef struct {
    int n, pow2, sign;
    float *wr, *wi;   /* n/2 twiddles */
    double *dr, *di;  /* n roots for the direct DFT */
    cf32 *a, *b;      /* ping-pong */
} fftplan;

static int plan_init(fftplan* p, int n, int forward)
{
    memset(p, 0, sizeof(*p));
    if (n <= 0)
        return -1;
    p->n = n;
    p->sign = forward ? -1 : 1;
    p->pow2 = (n & (n - 1)) == 0;
    if (p->pow2) {
        int h = n / 2 > 0 ? n / 2 : 1;
        p->wr = (float*)malloc(sizeof(float) * (size_t)h);
        p->wi = (float*)malloc(sizeof(float) * (size_t)h);
        p->a = (cf32*)malloc(sizeof(cf32) * (size_t)n);
        p->b = (cf32*)malloc(sizeof(cf32) * (size_t)n);
        for (int k = 0; k < n / 2; k++) {
            p->wr[k] = (float)cos(2.0 * PI_D * k / n);
            p->wi[k] = (float)(p->sign * sin(2.0 * PI_D * k / n));
        }
    } else {
        p->dr = (double*)malloc(sizeof(double) * (size_t)n);
        p->di = (double*)malloc(sizeof(double) * (size_t)n);
        for (int k = 0; k < n; k++) {
            p->dr[k] = cos(2.0 * PI_D * k / n);
            p->di[k] = p->sign * sin(2.0 * PI_D * k / n);
        }
    }
    return 0;
}

static void plan_free(fftplan* p)
{
    free(p->wr);
    free(p->wi);
    free(p->dr);
    free(p->di);
    free(p->a);
    free(p->b);
}

static void plan_exec(fftplan* p, const cf32* in, cf32* out)
{
    const int n = p->n;
    if (!p->pow2) {
        for (int k = 0; k < n; k++) {
            double ar = 0.0, ai = 0.0;
            long idx = 0;
            for (int t = 0; t < n; t++) {
                ar += in[t].re * p->dr[idx] - in[t].im * p->di[idx];
                ai += in[t].re * p->di[idx] + in[t].im * p->dr[idx];
                idx += k;
                if (idx >= n)
                    idx -= n;
            }
            out[k].re = (float)ar;
            out[k].im = (float)ai;
        }
        return;
    }
    cf32 *x = p->a, *y = p->b;
    memcpy(x, in, sizeof(cf32) * (size_t)n);
    const int half = n / 2;
    for (int len = n, s = 1; len > 1; len >>= 1, s <<= 1) {
        const int m = len / 2;
        for (int q0 = 0; q0 < m; q0++) {
            const float cr = p->wr[q0 * s], ci = p->wi[q0 * s];
            const cf32* x0 = x + s * q0;
            const cf32* x1 = x0 + half;
            cf32* y0 = y + 2 * s * q0;
            cf32* y1 = y0 + s;
            for (int q = 0; q < s; q++) {
                const float dr = x0[q].re - x1[q].re, di = x0[q].im - x1[q].im;
                y0[q].re = x0[q].re + x1[q].re;
                y0[q].im = x0[q].im + x1[q].im;
                y1[q].re = dr * cr - di * ci;
                y1[q].im = dr * ci + di * cr;
            }
        }
        cf32* t = x;
        x = y;
        y = t;
    }
    memcpy(out, x, sizeof(cf32) * (size_t)n);
}

/* ------------------------------------------------------------------------------------------------
 * fft_filter_ccc: compute_sizes (gr-filter/lib/fft_filter.cc:204-224), set_taps (:171-201),
 * filter (:241-285): overlap-ADD with a saved tail; a fresh dec_ctr per filter() call.
 * Driven in chunks of chunk_out items like fft_filter_ccc_impl::work
 * (gr-filter/lib/fft_filter_ccc_impl.cc:61-78).  Returns nsamples, <0 on error.
 * ---------------------------------------------------------------------------------------------- */
int grbo_fft_filter_nsamples(int ntaps)
{
    int fftsize = (int)(2 * pow(2.0, ceil(log((double)ntaps) / log(2.0))));
    return fftsize - ntaps + 1;
}

int grbo_fft_filter_ccc(int decim, const cf32* taps, int ntaps, const cf32* in, cf32* out,
                        long nitems, long chunk_out)
{
    const int fftsize = (int)(2 * pow(2.0, ceil(log((double)ntaps) / log(2.0))));
    const int nsamples = fftsize - ntaps + 1;
    const int tailsize = ntaps - 1;
    if (nitems % nsamples != 0)
        return -1;
    fftplan fwd, inv;
    if (plan_init(&fwd, fftsize, 1) || plan_init(&inv, fftsize, 0))
        return -2;
    cf32* H = (cf32*)malloc(sizeof(cf32) * (size_t)fftsize);
    cf32* buf = (cf32*)calloc((size_t)fftsize, sizeof(cf32));
    cf32* spec = (cf32*)malloc(sizeof(cf32) * (size_t)fftsize);
    cf32* tail = (cf32*)calloc((size_t)(tailsize > 0 ? tailsize : 1), sizeof(cf32));
    const float scale = (float)(1.0 / fftsize);
    for (int i = 0; i < ntaps; i++) {
        buf[i].re = taps[i].re * scale;
        buf[i].im = taps[i].im * scale;
    }
    plan_exec(&fwd, buf, H);

    long chunk = (chunk_out / nsamples) * nsamples;
    if (chunk <= 0)
        chunk = nsamples;
    for (long done = 0; done < nitems;) {
        long n = nitems - done < chunk ? nitems - done : chunk;
        /* one filter(n, in, out) call */
        int dec_ctr = 0;
        long ninput = n * decim;
        cf32* o = out + done;
        const cf32* ip = in + done * decim;
        for (long i = 0; i < ninput; i += nsamples) {
            memcpy(buf, ip + i, sizeof(cf32) * (size_t)nsamples);
            memset(buf + nsamples, 0, sizeof(cf32) * (size_t)(fftsize - nsamples));
            plan_exec(&fwd, buf, spec);
            for (int k = 0; k < fftsize; k++) {
                const float ar = spec[k].re, ai = spec[k].im;
                spec[k].re = ar * H[k].re - ai * H[k].im;
                spec[k].im = ar * H[k].im + ai * H[k].re;
            }
            plan_exec(&inv, spec, buf);
            for (int j = 0; j < tailsize; j++) {
                buf[j].re += tail[j].re;
                buf[j].im += tail[j].im;
            }
            int j = dec_ctr;
            while (j < nsamples) {
                *o++ = buf[j];
                j += decim;
            }
            dec_ctr = j - nsamples;
            if (tailsize > 0)
                memcpy(tail, buf + nsamples, sizeof(cf32) * (size_t)tailsize);
        }
        done += n;
    }
    free(H);
    free(buf);
    free(spec);
    free(tail);
    plan_free(&fwd);
    plan_free(&inv);
    return nsamples;
}

/* fft_filter_ccf (gr-filter/lib/fft_filter.cc:291-419): the taps enter the transform as (t, 0) (:317-320); everything
 * else is the ccc kernel line for line. */
int grbo_fft_filter_ccf(int decim, const float* taps, int ntaps, const cf32* in, cf32* out, long nitems, long chunk_out)
{
    cf32* t = (cf32*)malloc(sizeof(cf32) * (size_t)(ntaps > 0 ? ntaps : 1));
    if (!t)
        return -2;
    for (int i = 0; i < ntaps; i++) {
        t[i].re = taps[i];
        t[i].im = 0.0f;
    }
    int rc = grbo_fft_filter_ccc(decim, t, ntaps, in, out, nitems, chunk_out);
    free(t);
    return rc;
}

/* ------------------------------------------------------------------------------------------------
 * pfb_channelizer_ccf: polyphase_filterbank::set_taps (gr-filter/lib/polyphase_filterbank.cc:53-81),
 * ctor LUT / output multiple (gr-filter/lib/pfb_channelizer_ccf_impl.cc:49-84), general_work
 * (:128-193).  One call == one general_work().  in_streams[j] holds consumed + T items, the first T
 * being history (history = T+1, :101).  Returns items consumed per stream, <0 on error.
 * ---------------------------------------------------------------------------------------------- */
int grbo_pfb_taps_per_filter(unsigned int nfilts, int ntaps)
{
    return (int)ceil((double)ntaps / (double)nfilts);
}

int grbo_pfb_output_multiple(unsigned int nfilts, float osr)
{
    int rate_ratio = (int)rintf(nfilts / osr);
    int om = 1;
    while ((om * rate_ratio) % (int)nfilts != 0)
        om++;
    return om;
}

int grbo_pfb_channelizer_ccf(unsigned int nfilts, const float* taps, int ntaps, float osr,
                             const int* channel_map, int nmap, const cf32* const* in_streams,
                             cf32* const* out_streams, int nouts, int noutput)
{
    const int M = (int)nfilts;
    const double srate = nfilts / osr;
    if (fabs(srate - round(srate)) > 0.00001)
        return -1; /* std::invalid_argument in the reference (:51-53) */
    const int T = grbo_pfb_taps_per_filter(nfilts, ntaps);
    const int rate_ratio = (int)rintf(nfilts / osr);
    if (noutput % grbo_pfb_output_multiple(nfilts, osr) != 0)
        return -3;
    int* map = (int*)malloc(sizeof(int) * (size_t)(M > nmap ? M : nmap));
    for (int i = 0; i < M; i++)
        map[i] = i;
    if (nmap > 0) {
        for (int i = 0; i < nmap; i++) {
            if ((unsigned int)channel_map[i] >= nfilts) {
                free(map);
                return -2; /* std::invalid_argument (:117-121) */
            }
            map[i] = channel_map[i];
        }
    }
    /* branch taps d_taps[i][j] = padded[i + j*M], each branch stored reversed like fir_filter */
    float* rev = (float*)calloc((size_t)M * (size_t)T, sizeof(float));
    for (int i = 0; i < M; i++)
        for (int j = 0; j < T; j++) {
            int src = i + j * M;
            rev[(size_t)i * T + (T - 1 - j)] = src < ntaps ? taps[src] : 0.0f;
        }
    int* idxlut = (int*)malloc(sizeof(int) * (size_t)M);
    for (int i = 0; i < M; i++)
        idxlut[i] = M - ((i + rate_ratio) % M) - 1;
    fftplan plan;
    plan_init(&plan, M, 0);
    cf32* fin = (cf32*)calloc((size_t)M, sizeof(cf32));
    cf32* fout = (cf32*)calloc((size_t)M, sizeof(cf32));

    int n = 1, i = -1, j = 0, oo = 0, last;
    int toconsume = (int)rintf(noutput / osr);
    while (n <= toconsume) {
        j = 0;
        i = (i + rate_ratio) % M;
        last = i;
        while (i >= 0) {
            const cf32* x = in_streams[j] + n;
            const float* h = rev + (size_t)i * T;
            float re = 0.0f, im = 0.0f;
            for (int t = 0; t < T; t++) {
                re += x[t].re * h[t];
                im += x[t].im * h[t];
            }
            fin[idxlut[j]].re = re;
            fin[idxlut[j]].im = im;
            j++;
            i--;
        }
        i = M - 1;
        while (i > last) {
            const cf32* x = in_streams[j] + n - 1;
            const float* h = rev + (size_t)i * T;
            float re = 0.0f, im = 0.0f;
            for (int t = 0; t < T; t++) {
                re += x[t].re * h[t];
                im += x[t].im * h[t];
            }
            fin[idxlut[j]].re = re;
            fin[idxlut[j]].im = im;
            j++;
            i--;
        }
        n += (i + rate_ratio) >= M;
        plan_exec(&plan, fin, fout);
        for (int nn = 0; nn < nouts; nn++)
            out_streams[nn][oo] = fout[map[nn]];
        oo++;
    }
    free(map);
    free(rev);
    free(idxlut);
    free(fin);
    free(fout);
    plan_free(&plan);
    return toconsume;
}

/* ------------------------------------------------------------------------------------------------
 * interp_fir_filter_impl<gr_complex,gr_complex,float|gr_complex> (gr-filter/lib/interp_fir_filter_impl.cc):
 * set_taps :72-90 pads the taps with zeros to a multiple of the interpolation L; install_taps :97-117 deals them
 * round-robin, xtaps[n][k] = taps[n + k*L], history = taps per phase T; work :128-152 writes
 * out[i*L + n] = firs[n].filter(&in[i]) = sum_t rev_n[t] * in[i + t] with rev_n[t] = xtaps[n][T-1-t].
 * in: ni + T-1 items, history first.  Returns T.
 * ---------------------------------------------------------------------------------------------- */
int grbo_interp_fir(int cplx_taps, unsigned int interpolation, const float* taps, int ntaps, const cf32* in, cf32* out,
                    long ni)
{
    if (interpolation == 0)
        return -1;
    if (ntaps < 1)
        return -2;
    const int L = (int)interpolation, T = (ntaps + L - 1) / L;
    for (long i = 0; i < ni; i++) {
        for (int n = 0; n < L; n++) {
            float re = 0.0f, im = 0.0f;
            for (int t = 0; t < T; t++) {
                const int k = n + (T - 1 - t) * L; /* index into the padded prototype */
                if (k >= ntaps)
                    continue;
                const cf32 x = in[i + t];
                if (cplx_taps) {
                    const float hr = taps[2 * k], hi = taps[2 * k + 1];
                    re += x.re * hr - x.im * hi;
                    im += x.re * hi + x.im * hr;
                } else {
                    re += x.re * taps[k];
                    im += x.im * taps[k];
                }
            }
            out[i * L + n].re = re;
            out[i * L + n].im = im;
        }
    }
    return T;
}

/* ------------------------------------------------------------------------------------------------
 * rational_resampler_base_impl<gr_complex,gr_complex,float|gr_complex>
 * (gr-filter/lib/rational_resampler_base_impl.cc): set_taps :77-93 pads the taps with zeros to a multiple of the
 * interpolation L; install_taps :97-117 deals them round-robin, xtaps[n][k] = taps[n + k*L], T taps per phase;
 * general_work :141-173 walks the phase counter: out[i] = firs[ctr].filter(in), ctr += D, the input advances once per
 * wrap of ctr past L.  The loop tests count < ninput only and filter() reads T items from there: in holds
 * nin + T-1 readable items.  *ctr is d_ctr (in/out), *consumed what consume_each() gets.  Returns items produced.
 * ---------------------------------------------------------------------------------------------- */
long grbo_rational_resampler(int cplx_taps, unsigned int interpolation, unsigned int decimation, const float* taps,
                             int ntaps, const cf32* in, long nin, cf32* out, long noutput, unsigned int* ctr_io,
                             long* consumed)
{
    if (interpolation == 0 || decimation == 0)
        return -1;
    if (ntaps < 1)
        return -2;
    const int L = (int)interpolation, T = (ntaps + L - 1) / L;
    unsigned int ctr = *ctr_io;
    long count = 0, i = 0;
    while (i < noutput && count < nin) {
        float re = 0.0f, im = 0.0f;
        for (int t = 0; t < T; t++) {
            const int k = (int)ctr + (T - 1 - t) * L; /* index into the padded prototype */
            if (k >= ntaps)
                continue;
            const cf32 x = in[count + t];
            if (cplx_taps) {
                const float hr = taps[2 * k], hi = taps[2 * k + 1];
                re += x.re * hr - x.im * hi;
                im += x.re * hi + x.im * hr;
            } else {
                re += x.re * taps[k];
                im += x.im * taps[k];
            }
        }
        out[i].re = re;
        out[i].im = im;
        i++;
        ctr += decimation;
        while (ctr >= interpolation) {
            ctr -= interpolation;
            count++;
        }
    }
    *ctr_io = ctr;
    *consumed = count;
    return i;
}

/* ------------------------------------------------------------------------------------------------
 * pfb_decimator_ccf_impl::work_fir_exp (gr-filter/lib/pfb_decimator_ccf_impl.cc:135-154) over
 * polyphase_filterbank::set_taps (gr-filter/lib/polyphase_filterbank.cc:53-81):
 *   out[i] = sum_{j=M-1..0} fir_j(&in_{M-1-j}[i]) * rot[j],  taps_j[t] = padded[j + t*M],
 *   rot[j] = gr_expj(j*chan*2*pi/M) (:49-51, float phase).
 * The other three variants (:156-257) compute the same sum with an M-point backward FFT in place of the
 * rotator table and/or FFT branch filters; they differ in rounding only.
 * in_streams[s]: noutput + T-1 items, history first (set_history(T), :61).
 * ---------------------------------------------------------------------------------------------- */
int grbo_pfb_decimator_ccf(unsigned int decim, const float* taps, int ntaps, unsigned int channel,
                           const cf32* const* in_streams, cf32* out, long noutput)
{
    const int M = (int)decim, T = (ntaps + M - 1) / M;
    float* br = (float*)calloc((size_t)M * T, sizeof(float)); /* br[j*T + t] = reversed taps of branch j */
    cf32* rot = (cf32*)malloc(sizeof(cf32) * (size_t)M);
    if (!br || !rot) {
        free(br);
        free(rot);
        return -1;
    }
    for (int j = 0; j < M; j++) {
        for (int t = 0; t < T; t++) {
            const int k = j + (T - 1 - t) * M;
            br[j * T + t] = k < ntaps ? taps[k] : 0.0f;
        }
        const float ph = (float)((unsigned int)j * channel * 2 * PI_D / decim);
        rot[j].re = cosf(ph);
        rot[j].im = sinf(ph);
    }
    for (long i = 0; i < noutput; i++) {
        float ore = 0.0f, oim = 0.0f;
        for (int j = M - 1; j >= 0; j--) {
            const cf32* x = in_streams[M - 1 - j] + i;
            float re = 0.0f, im = 0.0f;
            for (int t = 0; t < T; t++) {
                re += x[t].re * br[j * T + t];
                im += x[t].im * br[j * T + t];
            }
            ore += re * rot[j].re - im * rot[j].im;
            oim += re * rot[j].im + im * rot[j].re;
        }
        out[i].re = ore;
        out[i].im = oim;
    }
    free(br);
    free(rot);
    return (int)noutput;
}

/* ------------------------------------------------------------------------------------------------
 * fft_vcc_fftw::work (gr-fft/lib/fft_vcc_fftw.cc:63-122) and set_window (:54-61).
 * ---------------------------------------------------------------------------------------------- */
int grbo_fft_vcc(int fft_size, int forward, const float* window, int nwindow, int shift,
                 const cf32* in, cf32* out, long nvec)
{
    if (!(nwindow == 0 || nwindow == fft_size))
        return -1;
    fftplan plan;
    if (plan_init(&plan, fft_size, forward))
        return -2;
    const int N = fft_size;
    cf32* dst = (cf32*)malloc(sizeof(cf32) * (size_t)N);
    cf32* res = (cf32*)malloc(sizeof(cf32) * (size_t)N);
    for (long v = 0; v < nvec; v++, in += N, out += N) {
        if (nwindow) {
            if (!forward && shift) {
                int offset = N / 2;
                int m_off = N - offset;
                for (int i = 0; i < offset; i++) {
                    dst[m_off + i].re = in[i].re * window[i];
                    dst[m_off + i].im = in[i].im * window[i];
                }
                for (int i = 0; i < N - offset; i++) {
                    dst[i].re = in[offset + i].re * window[offset + i];
                    dst[i].im = in[offset + i].im * window[offset + i];
                }
            } else {
                for (int i = 0; i < N; i++) {
                    dst[i].re = in[i].re * window[i];
                    dst[i].im = in[i].im * window[i];
                }
            }
        } else {
            if (!forward && shift) {
                int len = (int)floor(N / 2.0);
                memcpy(dst, in + len, sizeof(cf32) * (size_t)(N - len));
                memcpy(dst + (N - len), in, sizeof(cf32) * (size_t)len);
            } else {
                memcpy(dst, in, sizeof(cf32) * (size_t)N);
            }
        }
        plan_exec(&plan, dst, res);
        if (forward && shift) {
            int len = (int)ceil(N / 2.0);
            memcpy(out, res + len, sizeof(cf32) * (size_t)(N - len));
            memcpy(out + (N - len), res, sizeof(cf32) * (size_t)len);
        } else {
            memcpy(out, res, sizeof(cf32) * (size_t)N);
        }
    }
    free(dst);
    free(res);
    plan_free(&plan);
    return 0;
}

/* complex_to_mag_squared_impl::work (gr-blocks/lib/complex_to_mag_squared_impl.cc:37-48) */
int grbo_mag_squared(const cf32* in, float* out, long n)
{
    for (long i = 0; i < n; i++)
        out[i] = in[i].re * in[i].re + in[i].im * in[i].im;
    return 0;
}

/* ------------------------------------------------------------------------------------------------
 * Windows: gr::fft::window::coswindow (gr-fft/lib/window.cc:83-104; float arithmetic, cosf),
 * hamming/hann/blackman/rectangular/blackman_harris(92) and window::build's type numbering
 * (gr-fft/include/gnuradio/fft/window.h:25-35).  Returns ntaps, <0 for a type not restated here.
 * ---------------------------------------------------------------------------------------------- */
#define PI_F ((float)PI_D)

static void coswin(float* w, int ntaps, float c0, float c1, float c2, float c3)
{
    float M = (float)(ntaps - 1);
    for (int n = 0; n < ntaps; n++)
        w[n] = c0 - c1 * cosf((2.0f * PI_F * n) / M) + c2 * cosf((4.0f * PI_F * n) / M) -
               c3 * cosf((6.0f * PI_F * n) / M);
}

int grbo_window_build(int type, int ntaps, float* w)
{
    float M = (float)(ntaps - 1);
    switch (type) {
    case 0: /* WIN_HAMMING (window.cc hamming) */
        for (int n = 0; n < ntaps; n++)
            w[n] = 0.54f - 0.46f * (float)cos((2 * PI_D * n) / M);
        return ntaps;
    case 1: /* WIN_HANN */
        for (int n = 0; n < ntaps; n++)
            w[n] = 0.5f - 0.5f * (float)cos((2 * PI_D * n) / M);
        return ntaps;
    case 2: /* WIN_BLACKMAN = coswindow(0.42, 0.5, 0.08) */
        for (int n = 0; n < ntaps; n++)
            w[n] = 0.42f - 0.5f * cosf((2.0f * PI_F * n) / M) + 0.08f * cosf((4.0f * PI_F * n) / M);
        return ntaps;
    case 3: /* WIN_RECTANGULAR */
        for (int n = 0; n < ntaps; n++)
            w[n] = 1.0f;
        return ntaps;
    case 5: /* WIN_BLACKMAN_hARRIS, atten 92 (window.cc:179-180) */
        coswin(w, ntaps, 0.35875f, 0.48829f, 0.14128f, 0.01168f);
        return ntaps;
    default:
        return -1;
    }
}

/* firdes::compute_ntaps_windes (gr-filter/lib/firdes.cc:597-609) + low_pass_2 (:33-75).
 * Returns the tap count; writes at most maxn taps. */
int grbo_firdes_low_pass_2(double gain, double fs, double cutoff, double tw, double atten_db,
                           int win, float* out, int maxn)
{
    if (fs <= 0.0 || cutoff <= 0.0 || cutoff > fs / 2 || tw <= 0.0)
        return -1;
    int ntaps = (int)(atten_db * fs / (22.0 * tw));
    if ((ntaps & 1) == 0)
        ntaps++;
    float* w = (float*)malloc(sizeof(float) * (size_t)ntaps);
    float* taps = (float*)malloc(sizeof(float) * (size_t)ntaps);
    if (grbo_window_build(win, ntaps, w) < 0) {
        free(w);
        free(taps);
        return -2;
    }
    int M = (ntaps - 1) / 2;
    double fwT0 = 2 * PI_D * cutoff / fs;
    for (int n = -M; n <= M; n++) {
        if (n == 0)
            taps[n + M] = (float)(fwT0 / PI_D * w[n + M]);
        else
            taps[n + M] = (float)(sin(n * fwT0) / (n * PI_D) * w[n + M]);
    }
    double fmax = taps[0 + M];
    for (int n = 1; n <= M; n++)
        fmax += 2 * taps[n + M];
    gain /= fmax;
    for (int i = 0; i < ntaps; i++)
        taps[i] = (float)(taps[i] * gain);
    for (int i = 0; i < ntaps && i < maxn; i++)
        out[i] = taps[i];
    free(w);
    free(taps);
    return ntaps;
}

/* ------------------------------------------------------------------------------------------------
 * fir_filter_fff: kernel::fir_filter<float,float,float>::filter (gr-filter/lib/fir_filter.cc:122-130) under
 * fir_filter_blk_impl<float,float,float>::work (gr-filter/lib/fir_filter_blk_impl.cc:67-90).
 * ---------------------------------------------------------------------------------------------- */
int grbo_fir_fff(int decim, const float* taps, int ntaps, const float* in, float* out, long nout)
{
    float* rev = (float*)malloc(sizeof(float) * (size_t)ntaps);
    if (!rev)
        return -1;
    for (int t = 0; t < ntaps; t++)
        rev[t] = taps[ntaps - 1 - t];
    for (long i = 0; i < nout; i++) {
        const float* x = in + i * decim;
        float acc = 0.0f;
        for (int t = 0; t < ntaps; t++)
            acc += x[t] * rev[t];
        out[i] = acc;
    }
    free(rev);
    return 0;
}

/* ------------------------------------------------------------------------------------------------
 * fft_vfc_fftw::work (gr-fft/lib/fft_vfc_fftw.cc:57-92): float vector -> (window) -> complex FFT, all N bins.
 * ---------------------------------------------------------------------------------------------- */
int grbo_fft_vfc(int fft_size, int forward, const float* window, int nwindow, const float* in, cf32* out, long nvec)
{
    if (!(nwindow == 0 || nwindow == fft_size))
        return -1;
    fftplan plan;
    if (plan_init(&plan, fft_size, forward))
        return -2;
    const int N = fft_size;
    cf32* dst = (cf32*)malloc(sizeof(cf32) * (size_t)N);
    for (long v = 0; v < nvec; v++, in += N, out += N) {
        for (int i = 0; i < N; i++) {
            dst[i].re = nwindow ? in[i] * window[i] : in[i];
            dst[i].im = 0.0f;
        }
        plan_exec(&plan, dst, out);
    }
    free(dst);
    plan_free(&plan);
    return 0;
}

/* ------------------------------------------------------------------------------------------------
 * logpwrfft_c / logpwrfft_f (gr-fft/python/fft/logpwrfft.py:29-64), the whole hier block as one loop:
 * stream_to_vector + keep_one_in_n (gr-blocks/lib/keep_one_in_n_impl.cc:56-77: d_count counts down from n, the LAST vector
 * of every group of n is kept) -> window, forward FFT, no shift -> |.|^2 -> single_pole_iir<float,float,double>
 * (gr-filter/include/gnuradio/filter/single_pole_iir.h:78-86: y = alpha*x + (1-alpha)*y_prev in double, stored as float)
 * -> nlog10_ff (gr-blocks/lib/nlog10_ff_impl.cc:28-60: n/log2f(10) * log2f(max(x, FLT_MIN) * 10^(k/n))).
 * count_state / iir_state carry keep_one_in_n's counter and the filter memory across calls.
 * ---------------------------------------------------------------------------------------------- */
long grbo_logpwrfft(int complex_input, int fft_size, double ref_scale, int decim, double avg_alpha, int average,
                    const float* window, const void* in_, long nitems, float* out, long max_out, int* count_state,
                    float* iir_state)
{
    if (decim < 1 || fft_size < 1)
        return -1;
    const int N = fft_size;
    fftplan plan;
    if (plan_init(&plan, N, 1))
        return -2;
    double window_power = 0.0;
    for (int i = 0; i < N; i++)
        window_power += (double)window[i] * (double)window[i];
    const double kd = -20 * log10((double)N) - 10 * log10(window_power / N) - 20 * log10(ref_scale / 2);
    const float n = 10.0f, k = (float)kd;
    const float n_log2_10 = n / log2f(10.0f);
    const float ten_k_n = powf(10.0f, k / n);
    const double alpha = average ? avg_alpha : 1.0, one_minus_alpha = 1.0 - alpha;
    cf32* dst = (cf32*)malloc(sizeof(cf32) * (size_t)N);
    cf32* res = (cf32*)malloc(sizeof(cf32) * (size_t)N);
    const long nvec_in = nitems / N;
    long no = 0;
    for (long ni = 0; ni < nvec_in && no < max_out; ni++) {
        (*count_state)--;
        if (*count_state > 0)
            continue;
        *count_state = decim;
        if (complex_input) {
            const cf32* x = (const cf32*)in_ + ni * N;
            for (int i = 0; i < N; i++) {
                dst[i].re = x[i].re * window[i];
                dst[i].im = x[i].im * window[i];
            }
        } else {
            const float* x = (const float*)in_ + ni * N;
            for (int i = 0; i < N; i++) {
                dst[i].re = x[i] * window[i];
                dst[i].im = 0.0f;
            }
        }
        plan_exec(&plan, dst, res);
        float* o = out + no * N;
        for (int i = 0; i < N; i++) {
            const float p = res[i].re * res[i].re + res[i].im * res[i].im;
            const float y = (float)(alpha * (double)p + one_minus_alpha * (double)iir_state[i]);
            iir_state[i] = y;
            float v = y > 1.17549435e-38f ? y : 1.17549435e-38f;
            if (ten_k_n != 1.0f)
                v = v * ten_k_n;
            o[i] = log2f(v) * n_log2_10;
        }
        no++;
    }
    free(dst);
    free(res);
    plan_free(&plan);
    return no;
}

/* ------------------------------------------------------------------------------------------------
 * pfb_synthesizer_ccf (gr-filter/lib/pfb_synthesizer_ccf_impl.cc): set_taps1 :95-129, set_taps2 :131-184, work :221-284
 * over fir_filter_with_buffer_ccf::filter (gr-filter/lib/fir_filter_with_buffer.cc: a delay line, y = sum_j taps[j] x[n-j]).
 * A fresh block; the delay lines carry on over `chunk`-sized work() calls (chunking cannot change the result).
 * in_streams[i] holds noutput/numchans items.  Returns noutput, <0 on error.
 * ---------------------------------------------------------------------------------------------- */
int grbo_pfb_synthesizer_ccf(unsigned int numchans, const float* taps, int ntaps, int twox, const int* channel_map, int nmap,
                             const cf32* const* in_streams, int nin, cf32* out, int noutput, int chunk)
{
    (void)chunk;
    const int M = (int)numchans, tw = twox ? 2 : 1, NF = tw * M;
    if (M % tw != 0)
        return -1;
    if (noutput % M != 0)
        return -3;
    const int T = (ntaps + M - 1) / M;
    /* d_taps[f][j], f < NF */
    float* ft = (float*)calloc((size_t)NF * T, sizeof(float));
    for (int i = 0; i < M; i++) {
        int state = 0;
        for (int j = 0; j < T; j++) {
            const int idx = i + j * M;
            const float v = idx < ntaps ? taps[idx] : 0.0f;
            if (tw == 1) {
                ft[i * T + j] = v;
            } else if (state == 0) {
                ft[(M + i) * T + j] = v;
                state = 1;
            } else {
                ft[i * T + j] = v;
                state = 0;
            }
        }
    }
    int* map = (int*)malloc(sizeof(int) * (size_t)NF);
    for (int i = 0; i < NF; i++)
        map[i] = i;
    if (nmap > 0) {
        for (int i = 0; i < nmap; i++)
            if (channel_map[i] < 0 || channel_map[i] >= NF) {
                free(ft);
                free(map);
                return -2;
            }
        for (int i = 0; i < nmap && i < NF; i++)
            map[i] = channel_map[i];
    }
    fftplan plan;
    plan_init(&plan, NF, 0);
    cf32* fin = (cf32*)calloc((size_t)NF, sizeof(cf32));
    cf32* fout = (cf32*)malloc(sizeof(cf32) * (size_t)NF);
    cf32* dl = (cf32*)calloc((size_t)NF * T, sizeof(cf32)); /* delay lines, dl[f][0] = newest */
    int state = 0;
    for (int n = 0; n < noutput / M; n++) {
        for (int i = 0; i < nin; i++)
            fin[map[i]] = in_streams[i][n];
        plan_exec(&plan, fin, fout);
        for (int f = 0; f < NF; f++) {
            const int src = tw == 1 ? f : (f < M ? state * M + f : (state ^ 1) * M + (f - M));
            memmove(dl + (size_t)f * T + 1, dl + (size_t)f * T, sizeof(cf32) * (size_t)(T - 1));
            dl[(size_t)f * T] = fout[src];
        }
        for (int i = 0; i < M; i++) {
            float re = 0.0f, im = 0.0f;
            for (int j = T - 1; j >= 0; j--) { /* oldest first, like the reversed-tap dot product */
                re += dl[(size_t)i * T + j].re * ft[i * T + j];
                im += dl[(size_t)i * T + j].im * ft[i * T + j];
            }
            if (tw == 2) {
                float re2 = 0.0f, im2 = 0.0f;
                for (int j = T - 1; j >= 0; j--) {
                    re2 += dl[(size_t)(M + i) * T + j].re * ft[(M + i) * T + j];
                    im2 += dl[(size_t)(M + i) * T + j].im * ft[(M + i) * T + j];
                }
                re += re2;
                im += im2;
            }
            out[(size_t)n * M + i].re = re;
            out[(size_t)n * M + i].im = im;
        }
        state ^= 1;
    }
    free(ft);
    free(map);
    free(fin);
    free(fout);
    free(dl);
    plan_free(&plan);
    return noutput;
}

/* ------------------------------------------------------------------------------------------------
 * filterbank_vcvcf_impl::general_work (gr-filter/lib/filterbank_vcvcf_impl.cc:54-96) over kernel::filterbank
 * (gr-filter/lib/filterbank.cc:50-79): item = vector of nfilts complex; history ntaps+1;
 * out[j][i] = sum_t rev_i[t] in[j+t][i]; a filter whose taps are all zero writes zeros without looking at its input.
 * ---------------------------------------------------------------------------------------------- */
int grbo_filterbank_vcvcf(int nfilts, int ntaps, const float* taps, const cf32* in, cf32* out, int nout)
{
    if (nfilts < 1 || ntaps < 1)
        return -1;
    for (int i = 0; i < nfilts; i++) {
        const float* h = taps + (size_t)i * ntaps;
        int active = 0;
        for (int t = 0; t < ntaps; t++)
            if (h[t] != 0)
                active = 1;
        for (int j = 0; j < nout; j++) {
            float re = 0.0f, im = 0.0f;
            if (active)
                for (int t = 0; t < ntaps; t++) {
                    const cf32 x = in[(size_t)(j + t) * nfilts + i];
                    re += x.re * h[ntaps - 1 - t];
                    im += x.im * h[ntaps - 1 - t];
                }
            out[(size_t)j * nfilts + i].re = re;
            out[(size_t)j * nfilts + i].im = im;
        }
    }
    return nout;
}
