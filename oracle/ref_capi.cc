/* oracle/ref_capi.cc -- TEST INFRASTRUCTURE ONLY (never on the product path).
 *
 * C entry points around the REAL reference kernel classes, which oracle/Makefile compiles
 * unmodified from /root/reference into oracle/_ref/libgrref.so:
 *   gr-filter/lib/fir_filter.cc, gr-filter/lib/fft_filter.cc, gr-filter/lib/polyphase_filterbank.cc,
 *   gr-filter/lib/firdes.cc, gr-fft/lib/window.cc, gr-filter/lib/fir_filter_with_buffer.cc, gr-filter/lib/filterbank.cc
 *   (+ the header-only gr-filter/include/gnuradio/filter/single_pole_iir.h).
 * The block shells (work()/general_work()) depend on gr::block (Boost/PMT/GMP, absent here) and are
 * restated below next to the reference file:line each loop follows.  Third-party arithmetic comes
 * from oracle/shim (generic-C VOLK stand-ins, own FFT engine in place of FFTW3f).
 */
#include <gnuradio/blocks/rotator.h>
#include <gnuradio/fft/fft.h>
#include <gnuradio/fft/window.h>
#include <gnuradio/filter/fft_filter.h>
#include <gnuradio/filter/filterbank.h>
#include <gnuradio/filter/fir_filter.h>
#include <gnuradio/filter/fir_filter_with_buffer.h>
#include <gnuradio/filter/single_pole_iir.h>
#include <gnuradio/filter/firdes.h>
#include <gnuradio/filter/polyphase_filterbank.h>
#include <gnuradio/expj.h>
#include <gnuradio/math.h>
#include <volk/volk.h>
#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>
#include <stdexcept>
#include <vector>

using gr::filter::kernel::fft_filter_ccc;
using gr::filter::kernel::fir_filter_ccc;
using gr::filter::kernel::fir_filter_ccf;

namespace {

/* Restatement of pfb_channelizer_ccf_impl (gr-filter/lib/pfb_channelizer_ccf_impl.cc:34-193) on
 * top of the real kernel::polyphase_filterbank. */
class channelizer_shell : public gr::filter::kernel::polyphase_filterbank
{
public:
    float d_osr;
    int d_rate_ratio;
    int d_output_multiple;
    std::vector<int> d_idxlut;
    std::vector<int> d_channel_map;

    channelizer_shell(unsigned int nfilts, const std::vector<float>& taps, float osr)
        : polyphase_filterbank(nfilts, taps, false), d_osr(osr)
    {
        /* ctor checks: pfb_channelizer_ccf_impl.cc:49-53 */
        const double srate = nfilts / osr;
        const double rsrate = round(srate);
        if (fabs(srate - rsrate) > 0.00001)
            throw std::invalid_argument(
                "pfb_channelizer: oversample rate must be N/i for i in [1, N]");
        d_channel_map.resize(nfilts);
        for (unsigned int i = 0; i < nfilts; i++)
            d_channel_map[i] = i;
        /* LUT: :72-77 */
        d_rate_ratio = (int)rintf(nfilts / osr);
        d_idxlut.resize(nfilts);
        for (unsigned int i = 0; i < nfilts; i++)
            d_idxlut[i] = nfilts - ((i + d_rate_ratio) % nfilts) - 1;
        /* output multiple: :81-84 */
        d_output_multiple = 1;
        while ((d_output_multiple * d_rate_ratio) % nfilts != 0)
            d_output_multiple++;
    }

    /* general_work body: :157-192.  in[j] points at the oldest history item of stream j
     * (history = taps_per_filter + 1).  Returns items consumed per input stream. */
    int work(int noutput_items,
             const gr_complex* const* input_items,
             gr_complex* const* output_items,
             size_t noutputs)
    {
        const int nfilts = (int)d_nfilts;
        int n = 1, i = -1, j = 0, oo = 0, last;
        int toconsume = (int)rintf(noutput_items / d_osr);
        while (n <= toconsume) {
            j = 0;
            i = (i + d_rate_ratio) % nfilts;
            last = i;
            while (i >= 0) {
                const gr_complex* in = input_items[j];
                d_fft->get_inbuf()[d_idxlut[j]] = d_fir_filters[i]->filter(&in[n]);
                j++;
                i--;
            }
            i = nfilts - 1;
            while (i > last) {
                const gr_complex* in = input_items[j];
                d_fft->get_inbuf()[d_idxlut[j]] = d_fir_filters[i]->filter(&in[n - 1]);
                j++;
                i--;
            }
            n += (i + d_rate_ratio) >= nfilts;
            d_fft->execute();
            for (size_t nn = 0; nn < noutputs; nn++)
                output_items[nn][oo] = d_fft->get_outbuf()[d_channel_map[nn]];
            oo++;
        }
        return toconsume;
    }
};

/* pfb_decimator_ccf_impl (gr-filter/lib/pfb_decimator_ccf_impl.cc) without the runtime: the block derives from
 * sync_block AND polyphase_filterbank; this shell keeps the second base (the reference's own class, compiled from its
 * source) and restates the four work() bodies (:135-257) over its members. */
struct decimator_shell : public gr::filter::kernel::polyphase_filterbank {
    unsigned int d_rate, d_chan;
    std::vector<gr_complex> d_rotator;
    decimator_shell(unsigned int decim, const std::vector<float>& taps, unsigned int channel)
        : polyphase_filterbank(decim, taps), d_rate(decim), d_chan(channel), d_rotator(decim)
    {
        for (unsigned int i = 0; i < d_rate; i++) /* :49-51 */
            d_rotator[i] = gr_expj(i * d_chan * 2 * GR_M_PI / d_rate);
    }
    int taps_per_filter() const { return (int)d_taps_per_filter; }
    int fft_nsamples() const { return d_fft_filters[0]->filtersize() - d_fft_filters[0]->ntaps() + 1; } /* :58 */

    /* mode bit 0 = use_fft_rotator, bit 1 = use_fft_filters.  FIR modes: in[s] has n + T-1 items, history first.
     * FFT-filter modes: in[s] has n items (history 1, :57), n a multiple of fft_nsamples(). */
    int work(int mode, int n, const gr_complex* const* input_items, gr_complex* out)
    {
        const bool rot_fft = mode & 1, filt_fft = mode & 2;
        std::vector<gr_complex> tmp;
        if (filt_fft) { /* :187-192 / :217-220 */
            tmp.resize((size_t)n * d_rate);
            for (unsigned int j = 0; j < d_rate; j++)
                d_fft_filters[j]->filter(n, input_items[d_rate - j - 1], &tmp[(size_t)j * n]);
        }
        for (int i = 0; i < n; i++) {
            if (!rot_fft) { /* work_fir_exp :143-151 / work_fft_exp :197-202 */
                out[i] = 0;
                if (filt_fft) {
                    for (unsigned int j = 0; j < d_rate; j++)
                        out[i] += tmp[(size_t)j * n + i] * d_rotator[j];
                } else {
                    for (int j = (int)d_rate - 1; j >= 0; j--)
                        out[i] += d_fir_filters[j]->filter(&input_items[d_rate - 1 - j][i]) * d_rotator[j];
                }
            } else { /* work_fir_fft :164-176 / work_fft_fft :224-235 */
                for (unsigned int j = 0; j < d_rate; j++)
                    d_fft->get_inbuf()[j] =
                        filt_fft ? tmp[(size_t)j * n + i] : d_fir_filters[j]->filter(&input_items[d_rate - 1 - j][i]);
                d_fft->execute();
                out[i] = d_fft->get_outbuf()[d_chan];
            }
        }
        return n;
    }
};

/* pfb_synthesizer_ccf_impl (gr-filter/lib/pfb_synthesizer_ccf_impl.cc) without the runtime: the members and the three
 * methods set_taps1 (:95-129), set_taps2 (:131-184) and work (:221-284) restated over the reference's own
 * kernel::fir_filter_with_buffer_ccf (compiled from its source) and the shim fft_complex. */
struct synthesizer_shell {
    unsigned int d_numchans, d_twox, d_taps_per_filter;
    int d_state;
    std::vector<gr::filter::kernel::fir_filter_with_buffer_ccf*> d_filters;
    std::vector<std::vector<float>> d_taps;
    std::vector<int> d_channel_map;
    gr::fft::fft_complex* d_fft;

    synthesizer_shell(unsigned int numchans, const std::vector<float>& taps, bool twox) : d_numchans(numchans), d_state(0)
    {
        d_twox = twox ? 2 : 1; /* :42-47 */
        if (d_numchans % d_twox != 0)
            throw std::invalid_argument("pfb_synthesizer_ccf_impl: number of channels must be even for 2x oversampling.");
        d_filters.resize(d_twox * d_numchans);
        d_channel_map.resize(d_twox * d_numchans);
        std::vector<float> vtaps(0, d_twox * d_numchans); /* :55 (sic: an empty vector) */
        for (unsigned int i = 0; i < d_twox * d_numchans; i++) {
            d_filters[i] = new gr::filter::kernel::fir_filter_with_buffer_ccf(vtaps);
            d_channel_map[i] = i;
        }
        if (d_twox == 1)
            set_taps1(taps);
        else
            set_taps2(taps);
        d_fft = new gr::fft::fft_complex(d_twox * d_numchans, false); /* :65-66 */
        std::fill_n(d_fft->get_inbuf(), d_twox * d_numchans, 0);
    }
    ~synthesizer_shell()
    {
        delete d_fft;
        for (auto f : d_filters)
            delete f;
    }
    void set_taps1(const std::vector<float>& taps)
    {
        unsigned int ntaps = taps.size();
        d_taps_per_filter = (unsigned int)ceil((double)ntaps / (double)d_numchans);
        d_taps.resize(d_numchans);
        std::vector<float> tmp_taps = taps;
        while ((float)(tmp_taps.size()) < d_numchans * d_taps_per_filter)
            tmp_taps.push_back(0.0);
        for (unsigned int i = 0; i < d_numchans; i++) {
            d_taps[i] = std::vector<float>(d_taps_per_filter, 0);
            for (unsigned int j = 0; j < d_taps_per_filter; j++)
                d_taps[i][j] = tmp_taps[i + j * d_numchans];
            d_filters[i]->set_taps(d_taps[i]);
        }
    }
    void set_taps2(const std::vector<float>& taps)
    {
        unsigned int ntaps = taps.size();
        d_taps_per_filter = (unsigned int)ceil((double)ntaps / (double)d_numchans);
        d_taps.resize(d_twox * d_numchans);
        std::vector<float> tmp_taps = taps;
        while ((float)(tmp_taps.size()) < d_numchans * d_taps_per_filter)
            tmp_taps.push_back(0.0);
        for (unsigned int i = 0; i < d_numchans; i++) {
            d_taps[i] = std::vector<float>(d_taps_per_filter, 0);
            d_taps[d_numchans + i] = std::vector<float>(d_taps_per_filter, 0);
            int state = 0;
            for (unsigned int j = 0; j < d_taps_per_filter; j++) {
                if (state == 0) {
                    d_taps[i][j] = 0;
                    d_taps[d_numchans + i][j] = tmp_taps[i + j * d_numchans];
                    state = 1;
                } else {
                    d_taps[i][j] = tmp_taps[i + j * d_numchans];
                    d_taps[d_numchans + i][j] = 0;
                    state = 0;
                }
            }
            d_filters[i]->set_taps(d_taps[i]);
            d_filters[d_numchans + i]->set_taps(d_taps[d_numchans + i]);
        }
    }
    int set_channel_map(const std::vector<int>& map)
    { /* :202-217 */
        if (!map.empty()) {
            int mx = *std::max_element(map.begin(), map.end());
            int mn = *std::min_element(map.begin(), map.end());
            if ((mx >= static_cast<int>(d_twox * d_numchans)) || (mn < 0))
                return -2;
            d_channel_map = map;
            std::fill_n(d_fft->get_inbuf(), d_twox * d_numchans, 0);
        }
        return 0;
    }
    int work(int noutput_items, const gr_complex* const* input_items, size_t ninputs, size_t in_off, gr_complex* out)
    { /* :221-284 */
        unsigned int n, i;
        if (d_twox == 1) {
            for (n = 0; n < noutput_items / d_numchans; n++) {
                for (i = 0; i < ninputs; i++)
                    d_fft->get_inbuf()[d_channel_map[i]] = input_items[i][in_off + n];
                d_fft->execute();
                for (i = 0; i < d_numchans; i++)
                    out[i] = d_filters[i]->filter(d_fft->get_outbuf()[i]);
                out += d_numchans;
            }
        } else {
            for (n = 0; n < noutput_items / d_numchans; n++) {
                for (i = 0; i < ninputs; i++)
                    d_fft->get_inbuf()[d_channel_map[i]] = input_items[i][in_off + n];
                d_fft->execute();
                for (i = 0; i < d_numchans; i++) {
                    out[i] = d_filters[i]->filter(d_fft->get_outbuf()[d_state * d_numchans + i]);
                    out[i] += d_filters[d_numchans + i]->filter(d_fft->get_outbuf()[(d_state ^ 1) * d_numchans + i]);
                }
                d_state ^= 1;
                out += d_numchans;
            }
        }
        return noutput_items;
    }
};

/* filterbank_vcvcf_impl::general_work (gr-filter/lib/filterbank_vcvcf_impl.cc:54-96) over the reference's own
 * kernel::filterbank (gr-filter/lib/filterbank.cc, compiled from its source). */
struct filterbank_shell : public gr::filter::kernel::filterbank {
    explicit filterbank_shell(const std::vector<std::vector<float>>& taps) : filterbank(taps) {}
    unsigned int ntaps() const { return d_ntaps; }
    int work(int noutput_items, const gr_complex* in, gr_complex* out)
    {
        std::vector<gr_complex> working_v(noutput_items + d_ntaps + 16);
        gr_complex* working = working_v.data() + 8; /* headroom: filter() rounds its pointer down (fir_filter.cc:133-141) */
        for (unsigned int i = 0; i < d_nfilts; i++) {
            if (d_active[i]) {
                for (unsigned int j = 0; j < noutput_items + d_ntaps - 1; j++)
                    working[j] = in[i + j * d_nfilts];
                for (unsigned int j = 0; j < (unsigned int)(noutput_items); j++)
                    out[i + j * d_nfilts] = d_fir_filters[i]->filter(working + j);
            } else {
                for (unsigned int j = 0; j < (unsigned int)(noutput_items); j++)
                    out[i + j * d_nfilts] = 0;
            }
        }
        return noutput_items;
    }
};

} // namespace

extern "C" {

/* fir_filter_blk_impl<float,float,float>::work over kernel::fir_filter<float,float,float> (gr-filter/lib/fir_filter.cc:122-130,
 * volk_32f_x2_dot_prod_32f_a): fir_filter_fff.  in[] holds nout*decim + ntaps-1 floats. */
int grref_fir_fff(int decim, const float* taps, int ntaps, const float* in, float* out, long nout)
{
    std::vector<float> t(taps, taps + ntaps);
    gr::filter::kernel::fir_filter_fff fir(decim, t);
    if (decim == 1)
        fir.filterN(out, in, (unsigned long)nout);
    else
        fir.filterNdec(out, in, (unsigned long)nout, decim);
    return 0;
}

/* fft_vfc_fftw::work (gr-fft/lib/fft_vfc_fftw.cc:57-92) restated over the shim fft_complex. */
int grref_fft_vfc(int fft_size, int forward, const float* window, int nwindow, const float* in, void* out_, long nvec)
{
    if (!(nwindow == 0 || nwindow == fft_size))
        return -1; /* ctor throws: fft_vfc_fftw.cc:40-41 */
    gr::fft::fft_complex fft(fft_size, forward != 0, 1);
    gr_complex* out = (gr_complex*)out_;
    const unsigned int N = (unsigned int)fft_size;
    for (long count = 0; count < nvec; count++) {
        gr_complex* dst = fft.get_inbuf();
        if (nwindow) {
            for (unsigned int i = 0; i < N; i++)
                dst[i] = in[i] * window[i];
        } else {
            for (unsigned int i = 0; i < N; i++)
                dst[i] = in[i];
        }
        fft.execute();
        memcpy(out, fft.get_outbuf(), sizeof(gr_complex) * N);
        in += N;
        out += N;
    }
    return 0;
}

/* The logpwrfft_c / logpwrfft_f hier block (gr-fft/python/fft/logpwrfft.py:29-64) as one function: the chain
 *   stream_to_vector (gr-blocks/lib/stream_to_vector_impl.cc:46-58) -> keep_one_in_n (gr-blocks/lib/keep_one_in_n_impl.cc:56-77)
 *   -> fft_vcc / fft_vfc (forward, window, no shift) -> complex_to_mag_squared -> single_pole_iir_filter_ff
 *   (gr-filter/lib/single_pole_iir_filter_ff_impl.cc:47-67 over the reference's own single_pole_iir.h) -> nlog10_ff
 *   (gr-blocks/lib/nlog10_ff_impl.cc:28-60).
 * in: nitems stream items (float or gr_complex); whole vectors only are consumed.  count_state: keep_one_in_n's d_count
 * (in/out; a fresh block starts at n).  iir_state: fft_size floats, d_prev_output of every bin (in/out; zeros at start).
 * Returns the number of output vectors written (at most max_out), <0 on error. */
long grref_logpwrfft(int complex_input, int fft_size, double ref_scale, int decim, double avg_alpha, int average,
                     const float* window, const void* in_, long nitems, float* out, long max_out, int* count_state,
                     float* iir_state)
{
    if (decim < 1 || fft_size < 1)
        return -1;
    const unsigned int N = (unsigned int)fft_size;
    gr::fft::fft_complex fft(fft_size, true, 1);
    /* logpwrfft.py:47-60 (Python float arithmetic = double) */
    double window_power = 0;
    for (unsigned int i = 0; i < N; i++)
        window_power += (double)window[i] * (double)window[i];
    const double kd = -20 * log10((double)fft_size) - 10 * log10(window_power / fft_size) - 20 * log10(ref_scale / 2);
    /* nlog10_ff_impl ctor, :28-36: float n, float k */
    const float n = 10.0f, k = (float)kd;
    const float d_n_log2_10 = n / log2f(10.0f);
    const float d_10_k_n = std::pow(10.0f, k / n);
    /* single_pole_iir<float,float,double> (single_pole_iir.h): d_prev_output is private with no setter, so to carry the state
     * across calls every bin's filter is primed by one filter() call with alpha = 1 (which stores its input, :82-83) and
     * then switched to the working alpha (set_average, logpwrfft.py:101-113). */
    std::vector<gr::filter::single_pole_iir<float, float, double>> iir(N);
    for (unsigned int i = 0; i < N; i++) {
        iir[i].set_taps(1.0);
        iir[i].filter(iir_state[i]);
        iir[i].set_taps(average ? avg_alpha : 1.0);
    }
    const long nvec_in = nitems / N;
    long no = 0;
    std::vector<float> p(N);
    constexpr float float_min = std::numeric_limits<float>::min();
    for (long ni = 0; ni < nvec_in && no < max_out; ni++) {
        (*count_state)--;
        if (*count_state > 0)
            continue;
        *count_state = decim;
        gr_complex* dst = fft.get_inbuf();
        if (complex_input) { /* fft_vcc_fftw.cc:78-100, forward, no shift */
            const gr_complex* x = (const gr_complex*)in_ + ni * N;
            volk_32fc_32f_multiply_32fc(dst, x, window, N);
        } else { /* fft_vfc_fftw.cc:70-74 */
            const float* x = (const float*)in_ + ni * N;
            for (unsigned int i = 0; i < N; i++)
                dst[i] = x[i] * window[i];
        }
        fft.execute();
        volk_32fc_magnitude_squared_32f(p.data(), fft.get_outbuf(), N);
        float* o = out + no * N;
        for (unsigned int i = 0; i < N; i++) {
            float y = iir[i].filter(p[i]); /* the reference's own class: double arithmetic, float state (single_pole_iir.h:78-86) */
            /* nlog10_ff_impl::work :47-58 (volk_32f_log2_32f generic = log2f) */
            float v = std::max(y, float_min);
            if (d_10_k_n != 1.0f)
                v = v * d_10_k_n;
            o[i] = log2f(v) * d_n_log2_10;
        }
        no++;
    }
    for (unsigned int i = 0; i < N; i++)
        iir_state[i] = iir[i].prev_output();
    return no;
}

/* pfb_synthesizer_ccf: a fresh block, optional channel map, work() over `chunk` output items per call (0 = one call; the
 * filters' internal buffers carry the state between calls).  in_streams[i]: noutput/numchans items each. */
int grref_pfb_synthesizer_ccf(unsigned int numchans, const float* taps, int ntaps, int twox, const int* channel_map, int nmap,
                              const void* const* in_streams, int nin, void* out_, int noutput, int chunk)
{
    try {
        std::vector<float> t(taps, taps + ntaps);
        synthesizer_shell sh(numchans, t, twox != 0);
        if (nmap > 0) {
            int rc = sh.set_channel_map(std::vector<int>(channel_map, channel_map + nmap));
            if (rc)
                return rc;
        }
        if (noutput % (int)numchans != 0)
            return -3; /* set_output_multiple(d_numchans), :68 */
        if (chunk <= 0)
            chunk = noutput;
        chunk = std::max<int>(numchans, (chunk / (int)numchans) * (int)numchans);
        gr_complex* out = (gr_complex*)out_;
        for (int o = 0; o < noutput; o += chunk) {
            const int n = std::min(chunk, noutput - o);
            sh.work(n, (const gr_complex* const*)in_streams, (size_t)nin, (size_t)(o / numchans), out + o);
        }
        return noutput;
    } catch (const std::invalid_argument&) {
        return -1;
    }
}
int grref_pfb_synthesizer_taps(unsigned int numchans, const float* taps, int ntaps, int twox, float* out /*[twox*numchans][T]*/)
{
    try {
        std::vector<float> t(taps, taps + ntaps);
        synthesizer_shell sh(numchans, t, twox != 0);
        for (size_t i = 0; i < sh.d_taps.size(); i++)
            memcpy(out + i * sh.d_taps_per_filter, sh.d_taps[i].data(), sizeof(float) * sh.d_taps_per_filter);
        return (int)sh.d_taps_per_filter;
    } catch (const std::invalid_argument&) {
        return -1;
    }
}

/* filterbank_vcvcf: one general_work() call.  taps: [nfilts][ntaps] row-major; in: (nout + ntaps) vectors of nfilts items
 * (history = ntaps + 1, filterbank_vcvcf_impl.cc:34); out: nout vectors. */
int grref_filterbank_vcvcf(int nfilts, int ntaps, const float* taps, const void* in, void* out, int nout)
{
    try {
        std::vector<std::vector<float>> t((size_t)nfilts);
        for (int i = 0; i < nfilts; i++)
            t[i].assign(taps + (size_t)i * ntaps, taps + (size_t)(i + 1) * ntaps);
        filterbank_shell fb(t);
        return fb.work(nout, (const gr_complex*)in, (gr_complex*)out);
    } catch (const std::exception&) {
        return -1;
    }
}

/* pfb_decimator_ccf.  state: pass a pointer to a null void* for the first call and the same pointer afterwards to keep
 * the FFT filters' overlap across calls (mode bit 1); free it with grref_pfb_decimator_free.  Returns n, <0 on error. */
int grref_pfb_decimator_ccf(unsigned int decim, const float* taps, int ntaps, unsigned int channel, int mode,
                            const void* const* in_streams, void* out, int noutput, void** state)
{
    try {
        decimator_shell* d = state && *state ? (decimator_shell*)*state : nullptr;
        if (!d) {
            std::vector<float> t(taps, taps + ntaps);
            d = new decimator_shell(decim, t, channel);
        }
        int r = noutput;
        if ((mode & 2) && noutput % d->fft_nsamples() != 0)
            r = -3;
        else
            d->work(mode, noutput, (const gr_complex* const*)in_streams, (gr_complex*)out);
        if (state)
            *state = d;
        else
            delete d;
        return r;
    } catch (const std::invalid_argument&) {
        return -1;
    }
}
void grref_pfb_decimator_free(void* state) { delete (decimator_shell*)state; }
int grref_pfb_decimator_nsamples(unsigned int decim, const float* taps, int ntaps)
{
    std::vector<float> t(taps, taps + ntaps);
    decimator_shell d(decim, t, 0);
    return d.fft_nsamples();
}

/* fir_filter_blk_impl<gr_complex,gr_complex,float>::work (gr-filter/lib/fir_filter_blk_impl.cc:67-90):
 * in[] holds nout*decim + ntaps-1 items, in[0] = oldest history item. */
int grref_fir_ccf(int decim, const float* taps, int ntaps, const void* in, void* out, long nout)
{
    std::vector<float> t(taps, taps + ntaps);
    fir_filter_ccf fir(decim, t);
    if (decim == 1)
        fir.filterN((gr_complex*)out, (const gr_complex*)in, (unsigned long)nout);
    else
        fir.filterNdec((gr_complex*)out, (const gr_complex*)in, (unsigned long)nout, decim);
    return 0;
}

int grref_fir_ccc(int decim, const void* taps, int ntaps, const void* in, void* out, long nout)
{
    const gr_complex* tp = (const gr_complex*)taps;
    std::vector<gr_complex> t(tp, tp + ntaps);
    fir_filter_ccc fir(decim, t);
    if (decim == 1)
        fir.filterN((gr_complex*)out, (const gr_complex*)in, (unsigned long)nout);
    else
        fir.filterNdec((gr_complex*)out, (const gr_complex*)in, (unsigned long)nout, decim);
    return 0;
}

/* freq_xlating_fir_filter_impl<gr_complex,gr_complex,float|gr_complex> (gr-filter/lib/freq_xlating_fir_filter_impl.cc):
 * the block class needs the runtime, so build_composite_fir() (:74-93) and work() (:143-180) are restated here over
 * the reference's OWN fir_filter_ccc kernel and its OWN blocks::rotator (header-only, rotator.h).  work() is driven
 * like the scheduler does, `chunk_out` output items per call (the rotator renormalises at the end of each call).
 * phase_io (re, im) carries the rotator phase in and out; pass {1, 0} for a fresh block. */
int grref_freq_xlating_fir(int cplx_taps, int decim, const float* taps, int ntaps, double center_freq,
                           double sampling_freq, const void* in, void* out, long nout, long chunk_out, float* phase_io)
{
    std::vector<gr_complex> ctaps(ntaps);
    float fwT0 = 2 * GR_M_PI * center_freq / sampling_freq;
    for (unsigned int i = 0; i < (unsigned int)ntaps; i++) {
        if (cplx_taps)
            ctaps[i] = gr_complex(taps[2 * i], taps[2 * i + 1]) * exp(gr_complex(0, i * fwT0));
        else
            ctaps[i] = taps[i] * exp(gr_complex(0, i * fwT0));
    }
    fir_filter_ccc fir(decim, ctaps);
    gr::blocks::rotator r;
    if (phase_io)
        r.set_phase(gr_complex(phase_io[0], phase_io[1]));
    r.set_phase_incr(exp(gr_complex(0, -fwT0 * decim)));
    const gr_complex* x = (const gr_complex*)in;
    gr_complex* y = (gr_complex*)out;
    if (chunk_out < 1)
        chunk_out = nout > 0 ? nout : 1;
    for (long o0 = 0; o0 < nout; o0 += chunk_out) {
        const long n = std::min(chunk_out, nout - o0);
        if (decim == 1)
            fir.filterN(y + o0, x + o0, (unsigned long)n);
        else
            fir.filterNdec(y + o0, x + o0 * decim, (unsigned long)n, decim);
        r.rotateN(y + o0, y + o0, (int)n);
    }
    if (phase_io) {
        /* rotator keeps d_phase private; one zero-free probe item reads it back without advancing past it */
        gr_complex one(1.f, 0.f), ph;
        gr::blocks::rotator probe = r;
        ph = probe.rotate(one);
        phase_io[0] = ph.real();
        phase_io[1] = ph.imag();
    }
    return 0;
}

/* interp_fir_filter_impl<gr_complex,gr_complex,float|gr_complex> (gr-filter/lib/interp_fir_filter_impl.cc): the block
 * class needs the runtime (sync_interpolator), so set_taps (:72-90), install_taps (:97-117) and work (:128-152) are
 * restated over the reference's OWN kernel::fir_filter objects.  in: ni + taps_per_filter-1 items, history first;
 * out: ni*interpolation items.  Returns taps per filter, <0 on the reference's exceptions. */
int grref_interp_fir(int cplx_taps, unsigned interpolation, const float* taps, int ntaps, const void* in, void* out, long ni)
{
    if (interpolation == 0)
        return -1; /* std::out_of_range, :46-48 */
    if (ntaps < 1)
        return -2; /* std::runtime_error, :50-52 */
    const int padded = (int)((ntaps + interpolation - 1) / interpolation * interpolation);
    const int nt = padded / (int)interpolation;
    const gr_complex* x = (const gr_complex*)in;
    gr_complex* y = (gr_complex*)out;
    if (cplx_taps) {
        std::vector<std::vector<gr_complex>> xt(interpolation, std::vector<gr_complex>(nt, gr_complex(0, 0)));
        for (int i = 0; i < ntaps; i++)
            xt[i % interpolation][i / interpolation] = gr_complex(taps[2 * i], taps[2 * i + 1]);
        std::vector<fir_filter_ccc*> firs(interpolation);
        for (unsigned n = 0; n < interpolation; n++)
            firs[n] = new fir_filter_ccc(1, xt[n]);
        for (long i = 0; i < ni; i++)
            for (unsigned nf = 0; nf < interpolation; nf++)
                y[i * interpolation + nf] = firs[nf]->filter(&x[i]);
        for (auto f : firs)
            delete f;
    } else {
        std::vector<std::vector<float>> xt(interpolation, std::vector<float>(nt, 0.f));
        for (int i = 0; i < ntaps; i++)
            xt[i % interpolation][i / interpolation] = taps[i];
        std::vector<fir_filter_ccf*> firs(interpolation);
        for (unsigned n = 0; n < interpolation; n++)
            firs[n] = new fir_filter_ccf(1, xt[n]);
        for (long i = 0; i < ni; i++)
            for (unsigned nf = 0; nf < interpolation; nf++)
                y[i * interpolation + nf] = firs[nf]->filter(&x[i]);
        for (auto f : firs)
            delete f;
    }
    return nt;
}

/* rational_resampler_base_impl<gr_complex,gr_complex,float|gr_complex> (gr-filter/lib/rational_resampler_base_impl.cc):
 * the block class needs the runtime (gr::block), so set_taps (:77-93), install_taps (:97-117) and the general_work loop
 * (:156-172) are restated over the reference's OWN kernel::fir_filter objects.  in: nin + taps_per_filter-1 readable
 * items; *ctr_io is d_ctr, *consumed the consume_each() argument.  Returns items produced, <0 on the reference's
 * exceptions. */
} /* extern "C": the helper below is a template */
template <class FIR, class TAP>
static long resamp_run(unsigned interpolation, unsigned decimation, const std::vector<TAP>& taps, const gr_complex* in,
                       long nin, gr_complex* out, long noutput, unsigned* ctr_io, long* consumed)
{
    std::vector<TAP> padded(taps);
    int n = padded.size() % interpolation;
    if (n > 0) {
        n = interpolation - n;
        while (n-- > 0)
            padded.insert(padded.end(), TAP(0));
    }
    const int nt = (int)(padded.size() / interpolation);
    std::vector<std::vector<TAP>> xt(interpolation, std::vector<TAP>(nt));
    for (int i = 0; i < (int)padded.size(); i++)
        xt[i % interpolation][i / interpolation] = padded[i];
    std::vector<FIR*> firs(interpolation);
    for (unsigned k = 0; k < interpolation; k++)
        firs[k] = new FIR(1, xt[k]);
    unsigned ctr = *ctr_io;
    long count = 0, i = 0;
    while (i < noutput && count < nin) {
        out[i++] = firs[ctr]->filter(in);
        ctr += decimation;
        while (ctr >= interpolation) {
            ctr -= interpolation;
            in++;
            count++;
        }
    }
    for (auto f : firs)
        delete f;
    *ctr_io = ctr;
    *consumed = count;
    return i;
}
extern "C" {

long grref_rational_resampler(int cplx_taps, unsigned interpolation, unsigned decimation, const float* taps, int ntaps,
                              const void* in, long nin, void* out, long noutput, unsigned* ctr_io, long* consumed)
{
    if (interpolation == 0 || decimation == 0)
        return -1; /* std::out_of_range, :46-51 */
    if (ntaps < 1)
        return -2;
    if (cplx_taps) {
        std::vector<gr_complex> t(ntaps);
        for (int i = 0; i < ntaps; i++)
            t[i] = gr_complex(taps[2 * i], taps[2 * i + 1]);
        return resamp_run<fir_filter_ccc, gr_complex>(interpolation, decimation, t, (const gr_complex*)in, nin,
                                                      (gr_complex*)out, noutput, ctr_io, consumed);
    }
    std::vector<float> t(taps, taps + ntaps);
    return resamp_run<fir_filter_ccf, float>(interpolation, decimation, t, (const gr_complex*)in, nin, (gr_complex*)out,
                                             noutput, ctr_io, consumed);
}

/* fft_filter_ccc_impl::work driven like the scheduler does (gr-filter/lib/fft_filter_ccc_impl.cc:61-78):
 * one filter() call per chunk of `chunk_out` output items (rounded down to a multiple of nsamples,
 * at least one block).  nitems must be a multiple of nsamples; consumes nitems*decim inputs.
 * Returns nsamples, or -1 on a bad nitems. */
int grref_fft_filter_ccc(int decim,
                         const void* taps,
                         int ntaps,
                         const void* in,
                         void* out,
                         long nitems,
                         long chunk_out)
{
    const gr_complex* tp = (const gr_complex*)taps;
    std::vector<gr_complex> t(tp, tp + ntaps);
    fft_filter_ccc filt(decim, t, 1);
    const int nsamples = filt.set_taps(t);
    if (nitems % nsamples != 0)
        return -1;
    long chunk = (chunk_out / nsamples) * nsamples;
    if (chunk <= 0)
        chunk = nsamples;
    const gr_complex* ip = (const gr_complex*)in;
    gr_complex* op = (gr_complex*)out;
    for (long done = 0; done < nitems;) {
        long n = std::min(chunk, nitems - done);
        filt.filter((int)n, ip, op);
        ip += n * decim;
        op += n;
        done += n;
    }
    return nsamples;
}

/* same driver for kernel::fft_filter_ccf (gr-filter/lib/fft_filter.cc:291-419; shell fft_filter_ccf_impl.cc:61-78) */
int grref_fft_filter_ccf(int decim, const float* taps, int ntaps, const void* in, void* out, long nitems, long chunk_out)
{
    std::vector<float> t(taps, taps + ntaps);
    gr::filter::kernel::fft_filter_ccf filt(decim, t, 1);
    const int nsamples = filt.set_taps(t);
    if (nitems % nsamples != 0)
        return -1;
    long chunk = (chunk_out / nsamples) * nsamples;
    if (chunk <= 0)
        chunk = nsamples;
    const gr_complex* ip = (const gr_complex*)in;
    gr_complex* op = (gr_complex*)out;
    for (long done = 0; done < nitems;) {
        long n = std::min(chunk, nitems - done);
        filt.filter((int)n, ip, op);
        ip += n * decim;
        op += n;
        done += n;
    }
    return nsamples;
}

int grref_fft_filter_nsamples(int ntaps)
{
    std::vector<gr_complex> t(ntaps, gr_complex(1.0f, 0.0f));
    fft_filter_ccc filt(1, t, 1);
    return filt.set_taps(t);
}

/* One pfb_channelizer_ccf general_work() call.  in_streams[j] -> consumed + taps_per_filter items
 * (the first taps_per_filter being history).  Returns items consumed per stream, <0 on ctor error. */
int grref_pfb_channelizer_ccf(unsigned int nfilts,
                              const float* taps,
                              int ntaps,
                              float osr,
                              const int* channel_map,
                              int nmap,
                              const void* const* in_streams,
                              void* const* out_streams,
                              int nouts,
                              int noutput)
{
    try {
        std::vector<float> t(taps, taps + ntaps);
        channelizer_shell ch(nfilts, t, osr);
        if (nmap > 0) {
            /* set_channel_map: pfb_channelizer_ccf_impl.cc:112-124 */
            std::vector<int> map(channel_map, channel_map + nmap);
            unsigned int mx = 0;
            for (int v : map)
                mx = std::max(mx, (unsigned int)v);
            if (mx >= nfilts)
                return -2;
            ch.d_channel_map = map;
        }
        if (noutput % ch.d_output_multiple != 0)
            return -3;
        return ch.work(noutput,
                       (const gr_complex* const*)in_streams,
                       (gr_complex* const*)out_streams,
                       (size_t)nouts);
    } catch (const std::invalid_argument&) {
        return -1;
    }
}

int grref_pfb_taps_per_filter(unsigned int nfilts, int ntaps)
{
    return (int)ceil((double)ntaps / (double)nfilts);
}

/* fft_vcc_fftw::work (gr-fft/lib/fft_vcc_fftw.cc:63-122) restated over the shim fft_complex. */
int grref_fft_vcc(int fft_size,
                  int forward,
                  const float* window,
                  int nwindow,
                  int shift,
                  const void* in_,
                  void* out_,
                  long nvec)
{
    if (!(nwindow == 0 || nwindow == fft_size))
        return -1; /* ctor throws std::runtime_error: fft_vcc_fftw.cc:46-47 */
    gr::fft::fft_complex fft(fft_size, forward != 0, 1);
    const gr_complex* in = (const gr_complex*)in_;
    gr_complex* out = (gr_complex*)out_;
    const unsigned int N = (unsigned int)fft_size;
    for (long count = 0; count < nvec; count++) {
        gr_complex* dst = fft.get_inbuf();
        if (nwindow) {
            if (!forward && shift) {
                unsigned int offset = N / 2;
                int fft_m_offset = N - offset;
                volk_32fc_32f_multiply_32fc(&dst[fft_m_offset], &in[0], &window[0], offset);
                volk_32fc_32f_multiply_32fc(&dst[0], &in[offset], &window[offset], N - offset);
            } else {
                volk_32fc_32f_multiply_32fc(&dst[0], in, &window[0], N);
            }
        } else {
            if (!forward && shift) {
                unsigned int len = (unsigned int)(floor(N / 2.0));
                memcpy(&dst[0], &in[len], sizeof(gr_complex) * (N - len));
                memcpy(&dst[N - len], &in[0], sizeof(gr_complex) * len);
            } else {
                memcpy(dst, in, sizeof(gr_complex) * N);
            }
        }
        fft.execute();
        if (forward && shift) {
            unsigned int len = (unsigned int)(ceil(N / 2.0));
            memcpy(&out[0], &fft.get_outbuf()[len], sizeof(gr_complex) * (N - len));
            memcpy(&out[N - len], &fft.get_outbuf()[0], sizeof(gr_complex) * len);
        } else {
            memcpy(out, fft.get_outbuf(), sizeof(gr_complex) * N);
        }
        in += N;
        out += N;
    }
    return 0;
}

/* complex_to_mag_squared_impl::work (gr-blocks/lib/complex_to_mag_squared_impl.cc:37-48) */
int grref_mag_squared(const void* in, float* out, long n)
{
    volk_32fc_magnitude_squared_32f(out, (const gr_complex*)in, (unsigned int)n);
    return 0;
}

/* gr::fft::window::build (gr-fft/lib/window.cc) -- real reference code. */
int grref_window_build(int type, int ntaps, double beta, float* out)
{
    try {
        std::vector<float> w =
            gr::fft::window::build((gr::fft::window::win_type)type, ntaps, beta);
        memcpy(out, w.data(), sizeof(float) * w.size());
        return (int)w.size();
    } catch (...) {
        return -1;
    }
}

/* gr::filter::firdes::low_pass / low_pass_2 (gr-filter/lib/firdes.cc:33-119) -- real reference code.
 * Returns the tap count (writes at most maxn taps), <0 on error. */
int grref_firdes_low_pass(double gain, double fs, double cutoff, double tw, int win, double beta,
                          float* out, int maxn)
{
    try {
        std::vector<float> t = gr::filter::firdes::low_pass(
            gain, fs, cutoff, tw, (gr::filter::firdes::win_type)win, beta);
        for (int i = 0; i < (int)t.size() && i < maxn; i++)
            out[i] = t[i];
        return (int)t.size();
    } catch (...) {
        return -1;
    }
}

int grref_firdes_low_pass_2(double gain, double fs, double cutoff, double tw, double atten_db,
                            int win, double beta, float* out, int maxn)
{
    try {
        std::vector<float> t = gr::filter::firdes::low_pass_2(
            gain, fs, cutoff, tw, atten_db, (gr::filter::firdes::win_type)win, beta);
        for (int i = 0; i < (int)t.size() && i < maxn; i++)
            out[i] = t[i];
        return (int)t.size();
    } catch (...) {
        return -1;
    }
}

} /* extern "C" */
