// blocks_capi.cc -- drives the C++ block wrappers of include/grb200/blocks.hpp the way block_executor would
// (general_work in chunks, the d_updated "return 0 once" protocol, consume/produce bookkeeping), behind a
// few extern "C" entry points so tests/ can reach the C++ host side through ctypes.
// Built into gnuradio_b200/cxx/libgrb200_blocks.so by the Makefile next to this file.
#include <cstring>
#include <grb200/blocks.hpp>

static thread_local std::string g_what;

template <class F>
static long long guarded(F&& f)
{
    try {
        return f();
    } catch (const std::invalid_argument& e) {
        g_what = e.what();
        return -1;
    } catch (const std::out_of_range& e) {
        g_what = e.what();
        return -2;
    } catch (const std::runtime_error& e) {
        g_what = e.what();
        return -3;
    } catch (const std::exception& e) {
        g_what = e.what();
        return -4;
    }
}

// sync blocks: input holds (history-1) + consumed items; chunked like the scheduler (max `chunk` outputs per call,
// rounded down to the output multiple)
template <class BLK>
static long long run_sync(BLK& blk, const char* in, char* out, long long nout, int chunk, size_t in_item, size_t out_item)
{
    long long produced = 0;
    int zero_calls = 0;
    const int D = (int)blk.decimation();
    while (produced < nout) {
        long long n = std::min<long long>(chunk, nout - produced);
        n = n / blk.output_multiple() * blk.output_multiple();
        if (n == 0)
            break;
        gr_vector_const_void_star ii(1, in + (size_t)produced * D * in_item);
        gr_vector_void_star oo(1, out + (size_t)produced * out_item);
        gr_vector_int ni(1, (int)(n * D + blk.history() - 1));
        int r = blk.general_work((int)n, ni, ii, oo);
        if (r < 0)
            break;
        if (r == 0 && ++zero_calls > 4)
            break;
        produced += r;
    }
    return produced;
}

extern "C" {

const char* grbx_what(void) { return g_what.c_str(); }

long long grbx_fir_run(int complex_taps, int decim, const float* taps, int ntaps, const void* in, void* out,
                       long long nout, int chunk, int retap_after_first)
{
    return guarded([&]() -> long long {
        if (complex_taps) {
            std::vector<gr_complex> t((const gr_complex*)taps, (const gr_complex*)taps + ntaps);
            auto b = gr::filter::fir_filter_ccc::make(decim, t);
            if (b->history() != (unsigned)ntaps || b->taps() != t)
                throw std::logic_error("fir_filter_ccc: history/taps accessor mismatch");
            return run_sync(*b, (const char*)in, (char*)out, nout, chunk, 8, 8);
        }
        std::vector<float> t(taps, taps + ntaps);
        auto b = gr::filter::fir_filter_ccf::make(decim, t);
        if (b->history() != (unsigned)ntaps || b->taps() != t)
            throw std::logic_error("fir_filter_ccf: history/taps accessor mismatch");
        if (retap_after_first) { // set_taps mid-stream: the next work() returns 0 once, then carries on
            b->set_taps(t);
            gr_vector_const_void_star ii(1, in);
            gr_vector_void_star oo(1, out);
            if (b->work(1, ii, oo) != 0)
                throw std::logic_error("fir_filter_ccf: work() after set_taps() must return 0 once");
        }
        return run_sync(*b, (const char*)in, (char*)out, nout, chunk, 8, 8);
    });
}

// freq_xlating_fir_filter_ccf/ccc: starts at center_freq 0 and retunes through set_center_freq() before the first item,
// so the deferred rebuild (work() returns 0 once) is on the tested path
long long grbx_freq_xlating_run(int complex_taps, int decim, const float* taps, int ntaps, double center_freq,
                                double sampling_freq, const void* in, void* out, long long nout, int chunk)
{
    return guarded([&]() -> long long {
        gr_vector_const_void_star ii(1, in);
        gr_vector_void_star oo(1, out);
        if (complex_taps) {
            std::vector<gr_complex> t((const gr_complex*)taps, (const gr_complex*)taps + ntaps);
            auto b = gr::filter::freq_xlating_fir_filter_ccc::make(decim, t, 0.0, sampling_freq);
            b->set_center_freq(center_freq);
            if (b->history() != (unsigned)ntaps || b->taps() != t || b->center_freq() != center_freq)
                throw std::logic_error("freq_xlating_fir_filter_ccc: accessor mismatch");
            if (b->work(1, ii, oo) != 0)
                throw std::logic_error("freq_xlating_fir_filter_ccc: work() after set_center_freq() must return 0 once");
            return run_sync(*b, (const char*)in, (char*)out, nout, chunk, 8, 8);
        }
        std::vector<float> t(taps, taps + ntaps);
        auto b = gr::filter::freq_xlating_fir_filter_ccf::make(decim, t, 0.0, sampling_freq);
        b->set_center_freq(center_freq);
        if (b->history() != (unsigned)ntaps || b->taps() != t || b->center_freq() != center_freq)
            throw std::logic_error("freq_xlating_fir_filter_ccf: accessor mismatch");
        if (b->work(1, ii, oo) != 0)
            throw std::logic_error("freq_xlating_fir_filter_ccf: work() after set_center_freq() must return 0 once");
        return run_sync(*b, (const char*)in, (char*)out, nout, chunk, 8, 8);
    });
}

long long grbx_fft_filter_run(int decim, const float* taps_re_im, int ntaps, const void* in, void* out, long long nout,
                              int chunk, int* nsamples)
{
    return guarded([&]() -> long long {
        std::vector<gr_complex> t((const gr_complex*)taps_re_im, (const gr_complex*)taps_re_im + ntaps);
        auto b = gr::filter::fft_filter_ccc::make(decim, t, 1);
        if (nsamples)
            *nsamples = b->output_multiple();
        return run_sync(*b, (const char*)in, (char*)out, nout, chunk, 8, 8);
    });
}

long long grbx_fft_vcc_run(int fft_size, int forward, const float* window, int nwindow, int shift, const void* in,
                           void* out, long long nvec, int chunk)
{
    return guarded([&]() -> long long {
        std::vector<float> w(window, window + nwindow);
        auto b = gr::fft::fft_vcc::make(fft_size, forward != 0, w, shift != 0, 1);
        return run_sync(*b, (const char*)in, (char*)out, nvec, chunk, (size_t)fft_size * 8, (size_t)fft_size * 8);
    });
}

long long grbx_magsq_run(int vlen, const void* in, float* out, long long nitems, int chunk)
{
    return guarded([&]() -> long long {
        auto b = gr::blocks::complex_to_mag_squared::make((size_t)vlen);
        return run_sync(*b, (const char*)in, (char*)out, nitems, chunk, (size_t)vlen * 8, (size_t)vlen * 4);
    });
}

// channelizer: in_streams[j] holds T history items + consumed items; outputs written to out_streams[k][0..nout)
long long grbx_pfb_run(unsigned nfilts, const float* taps, int ntaps, float osr, const int* cmap, int ncmap,
                       const void* const* in_streams, void* const* out_streams, int nouts, long long nout, int chunk,
                       long long* consumed_total)
{
    return guarded([&]() -> long long {
        std::vector<float> t(taps, taps + ntaps);
        auto b = gr::filter::pfb_channelizer_ccf::make(nfilts, t, osr);
        if (ncmap > 0)
            b->set_channel_map(std::vector<int>(cmap, cmap + ncmap));
        long long produced = 0, consumed = 0;
        int zero_calls = 0;
        while (produced < nout) {
            long long n = std::min<long long>(chunk, nout - produced);
            n = n / b->output_multiple() * b->output_multiple();
            if (n == 0)
                break;
            gr_vector_const_void_star ii(nfilts);
            gr_vector_void_star oo(nouts);
            for (unsigned j = 0; j < nfilts; j++)
                ii[j] = (const gr_complex*)in_streams[j] + consumed;
            for (int k = 0; k < nouts; k++)
                oo[k] = (gr_complex*)out_streams[k] + produced;
            gr_vector_int ni(nfilts, 0);
            int r = b->general_work((int)n, ni, ii, oo);
            if (r < 0)
                break;
            if (r == 0) {
                if (++zero_calls > 4)
                    break;
                continue;
            }
            produced += r;
            consumed += b->nitems_consumed_last();
        }
        if (consumed_total)
            *consumed_total = consumed;
        return produced;
    });
}

// interp_fir_filter_ccf/ccc over a fresh stream of nin items (no history in front): returns items produced
long long grbx_interp_fir_run(int complex_taps, unsigned interpolation, const float* taps, int ntaps, const void* in,
                              void* out, long long nin, int chunk_in)
{
    return guarded([&]() -> long long {
        auto drive = [&](auto& b) -> long long {
            const long long hist = b->history();
            std::vector<gr_complex> buf((size_t)(hist - 1), gr_complex(0, 0));
            buf.insert(buf.end(), (const gr_complex*)in, (const gr_complex*)in + nin);
            long long done = 0;
            while (done < nin) {
                const long long n = std::min<long long>(chunk_in, nin - done);
                gr_vector_const_void_star ii(1, buf.data() + done);
                gr_vector_void_star oo(1, (gr_complex*)out + done * interpolation);
                gr_vector_int ni(1, 0);
                const int r = b->general_work((int)(n * interpolation), ni, ii, oo);
                if (r <= 0)
                    break;
                done += b->nitems_consumed_last();
            }
            return done * interpolation;
        };
        if (complex_taps) {
            std::vector<gr_complex> t((const gr_complex*)taps, (const gr_complex*)taps + ntaps);
            auto b = gr::filter::interp_fir_filter_ccc::make(interpolation, t);
            return drive(b);
        }
        std::vector<float> t(taps, taps + ntaps);
        auto b = gr::filter::interp_fir_filter_ccf::make(interpolation, t);
        if ((int)b->taps().size() != (ntaps + (int)interpolation - 1) / (int)interpolation * (int)interpolation)
            throw std::logic_error("interp_fir_filter_ccf: taps() must return the padded taps");
        return drive(b);
    });
}

// rational_resampler_base_ccf/ccc over a stream of nin items (the scheduler sees history 1: no zeros in front), driven in
// calls of at most max_noutput outputs stepped down until forecast() fits what is left: returns items produced
long long grbx_rational_resampler_run(int complex_taps, unsigned interpolation, unsigned decimation, const float* taps,
                                      int ntaps, const void* in, void* out, long long nin, int max_noutput)
{
    return guarded([&]() -> long long {
        auto drive = [&](auto& b) -> long long {
            long long consumed = 0, produced = 0;
            while (true) {
                const long long avail = nin - consumed;
                gr_vector_int need(1, 0);
                int nout = max_noutput;
                for (; nout > 0; nout--) {
                    b->forecast(nout, need);
                    if (need[0] <= avail)
                        break;
                }
                if (nout <= 0 || avail < (long long)b->history())
                    break;
                gr_vector_const_void_star ii(1, (const gr_complex*)in + consumed);
                gr_vector_void_star oo(1, (gr_complex*)out + produced);
                gr_vector_int ni(1, (int)(avail - (b->history() - 1))); // the array ends: no reads past it
                const int r = b->general_work(nout, ni, ii, oo);
                if (r <= 0)
                    break;
                produced += r;
                consumed += b->nitems_consumed_last();
            }
            return produced;
        };
        if (complex_taps) {
            std::vector<gr_complex> t((const gr_complex*)taps, (const gr_complex*)taps + ntaps);
            auto b = gr::filter::rational_resampler_base_ccc::make(interpolation, decimation, t);
            return drive(b);
        }
        std::vector<float> t(taps, taps + ntaps);
        auto b = gr::filter::rational_resampler_base_ccf::make(interpolation, decimation, t);
        if ((int)b->taps().size() != (ntaps + (int)interpolation - 1) / (int)interpolation * (int)interpolation)
            throw std::logic_error("rational_resampler_base_ccf: taps() must return the padded taps");
        return drive(b);
    });
}

// pfb_decimator_ccf over in_streams[s] (no history in front: a fresh stream), scheduler-sized calls
long long grbx_pfb_dec_run(unsigned decim, const float* taps, int ntaps, unsigned channel, int use_fft_rotator,
                           int use_fft_filters, const void* const* in_streams, void* out, long long nitems, int chunk)
{
    return guarded([&]() -> long long {
        std::vector<float> t(taps, taps + ntaps);
        auto b = gr::filter::pfb_decimator_ccf::make(decim, t, channel, use_fft_rotator != 0, use_fft_filters != 0);
        const long long hist = b->history();
        std::vector<std::vector<gr_complex>> buf(decim);
        for (unsigned s = 0; s < decim; s++) {
            buf[s].assign((size_t)(hist - 1), gr_complex(0, 0));
            buf[s].insert(buf[s].end(), (const gr_complex*)in_streams[s], (const gr_complex*)in_streams[s] + nitems);
        }
        long long produced = 0;
        while (produced < nitems) {
            long long n = std::min<long long>(chunk, nitems - produced);
            n = n / b->output_multiple() * b->output_multiple();
            if (n == 0)
                break;
            gr_vector_const_void_star ii(decim);
            gr_vector_void_star oo(1, (gr_complex*)out + produced);
            for (unsigned s = 0; s < decim; s++)
                ii[s] = buf[s].data() + produced;
            const int r = b->work((int)n, ii, oo);
            if (r <= 0)
                break;
            produced += r;
        }
        return produced;
    });
}


// fir_filter_fff: float stream
long long grbx_fir_fff_run(int decim, const float* taps, int ntaps, const float* in, float* out, long long nout, int chunk)
{
    return guarded([&]() -> long long {
        std::vector<float> t(taps, taps + ntaps);
        auto b = gr::filter::fir_filter_fff::make(decim, t);
        if (b->history() != (unsigned)ntaps || b->taps() != t || b->input_signature().item_size != 4)
            throw std::logic_error("fir_filter_fff: history/taps/io_signature mismatch");
        return run_sync(*b, (const char*)in, (char*)out, nout, chunk, 4, 4);
    });
}

long long grbx_fft_vfc_run(int fft_size, int forward, const float* window, int nwindow, const float* in, void* out,
                           long long nvec, int chunk)
{
    return guarded([&]() -> long long {
        std::vector<float> w(window, window + nwindow);
        auto b = gr::fft::fft_vfc::make(fft_size, forward != 0, w, 1);
        return run_sync(*b, (const char*)in, (char*)out, nvec, chunk, (size_t)fft_size * 4, (size_t)fft_size * 8);
    });
}

// pfb_synthesizer_ccf: in_streams[i] holds nout/numchans items; work() in chunks of `chunk` output items
long long grbx_pfb_synth_run(unsigned numchans, const float* taps, int ntaps, int twox, const int* cmap, int ncmap,
                             const void* const* in_streams, int nin, void* out, long long nout, int chunk)
{
    return guarded([&]() -> long long {
        std::vector<float> t(taps, taps + ntaps);
        auto b = gr::filter::pfb_synthesizer_ccf::make(numchans, t, twox != 0);
        if (ncmap > 0) {
            b->set_channel_map(std::vector<int>(cmap, cmap + ncmap));
            if (b->channel_map() != std::vector<int>(cmap, cmap + ncmap))
                throw std::logic_error("pfb_synthesizer_ccf: channel_map accessor mismatch");
        }
        if (b->output_multiple() != (int)numchans || b->history() != 1 || b->interpolation() != numchans)
            throw std::logic_error("pfb_synthesizer_ccf: rate contract mismatch");
        long long produced = 0;
        int zero_calls = 0;
        while (produced < nout) {
            long long n = std::min<long long>(chunk, nout - produced);
            n = n / numchans * numchans;
            if (n == 0)
                break;
            gr_vector_const_void_star ii((size_t)nin);
            for (int i = 0; i < nin; i++)
                ii[i] = (const char*)in_streams[i] + (size_t)(produced / numchans) * 8;
            gr_vector_void_star oo(1, (char*)out + (size_t)produced * 8);
            gr_vector_int ni((size_t)nin, (int)(n / numchans));
            int r = b->general_work((int)n, ni, ii, oo);
            if (r < 0)
                break;
            if (r == 0 && ++zero_calls > 4)
                break;
            if (r > 0 && b->nitems_consumed_last() != r / (int)numchans)
                throw std::logic_error("pfb_synthesizer_ccf: consume_each mismatch");
            produced += r;
        }
        return produced;
    });
}

// filterbank_vcvcf: in holds ntaps history vectors + nout vectors; set_taps(taps2) after `retap_at` outputs when taps2 != NULL
long long grbx_filterbank_run(int nfilts, int ntaps, const float* taps, const float* taps2, long long retap_at, const void* in,
                              void* out, long long nout, int chunk)
{
    return guarded([&]() -> long long {
        auto rows = [&](const float* p) {
            std::vector<std::vector<float>> t((size_t)nfilts);
            for (int i = 0; i < nfilts; i++)
                t[i].assign(p + (size_t)i * ntaps, p + (size_t)(i + 1) * ntaps);
            return t;
        };
        auto b = gr::filter::filterbank_vcvcf::make(rows(taps));
        if (b->history() != (unsigned)ntaps + 1 || b->taps() != rows(taps))
            throw std::logic_error("filterbank_vcvcf: history/taps accessor mismatch");
        const size_t vsz = (size_t)nfilts * 8;
        long long produced = 0;
        int zero_calls = 0;
        bool retapped = false;
        while (produced < nout) {
            if (taps2 && !retapped && produced >= retap_at) {
                b->set_taps(rows(taps2));
                retapped = true;
            }
            long long n = std::min<long long>(chunk, nout - produced);
            if (taps2 && !retapped)
                n = std::min<long long>(n, retap_at - produced);
            gr_vector_const_void_star ii(1, (const char*)in + (size_t)produced * vsz);
            gr_vector_void_star oo(1, (char*)out + (size_t)produced * vsz);
            gr_vector_int ni(1, (int)(n + ntaps));
            int r = b->general_work((int)n, ni, ii, oo);
            if (r < 0)
                break;
            if (r == 0 && ++zero_calls > 4)
                break;
            produced += r;
        }
        return produced;
    });
}

// fir_filter_ccf in lag-by-one mode, driven like the scheduler: the output pointer advances by what general_work returns,
// the input pointer by what it consumes; `room` = output items offered per call.  Returns the items produced before stop().
long long grbx_fir_lag_run(int decim, const float* taps, int ntaps, const void* in, void* out, long long nout, int lag_items,
                           int room)
{
    return guarded([&]() -> long long {
        std::vector<float> t(taps, taps + ntaps);
        auto b = gr::filter::fir_filter_ccf::make(decim, t);
        b->set_lag(lag_items);
        if (b->lag() != lag_items || b->min_output_buffer() != 4L * lag_items)
            throw std::logic_error("fir_filter_ccf: lag bookkeeping mismatch");
        long long produced = 0, consumed = 0;
        int idle = 0;
        while (consumed < nout * decim) {
            // block_executor: noutput = min(output space, what the available input covers) for a fixed-rate block
            long long offer = std::min<long long>(room, (nout * decim - consumed) / decim);
            if (offer <= 0)
                break;
            gr_vector_const_void_star ii(1, (const char*)in + (size_t)consumed * 8);
            gr_vector_void_star oo(1, (char*)out + (size_t)produced * 8);
            gr_vector_int ni(1, (int)(offer * decim + b->history() - 1));
            int r = b->general_work((int)offer, ni, ii, oo);
            if (r < 0)
                break;
            if (r == 0 && b->nitems_consumed_last() == 0 && ++idle > 2)
                break;
            produced += r;
            consumed += b->nitems_consumed_last();
        }
        b->stop();
        return produced;
    });
}

} // extern "C"
