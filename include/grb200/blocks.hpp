// grb200/blocks.hpp -- the reference's block classes re-hosted on the C-ABI (include/grb200.h).
//
// Host side of the drop-in boundary, C++ like the reference.  Each class keeps the reference's name,
// make()/set_taps()/taps()/work() signatures, io_signature, history / output_multiple / relative_rate
// declarations, locking and exception types; only the body of work() changed: it hands the scheduler's
// borrowed buffers to libgrb200.
//
//   gr::filter::fir_filter_blk<IN,OUT,TAP>  gr-filter/include/gnuradio/filter/fir_filter_blk.h:46-68
//                                           gr-filter/lib/fir_filter_blk_impl.{h,cc}
//   gr::filter::fft_filter_ccc              gr-filter/include/gnuradio/filter/fft_filter_ccc.h:43-70
//                                           gr-filter/lib/fft_filter_ccc_impl.{h,cc}
//   gr::filter::pfb_channelizer_ccf         gr-filter/include/gnuradio/filter/pfb_channelizer_ccf.h:100-192
//                                           gr-filter/lib/pfb_channelizer_ccf_impl.{h,cc}
//   gr::fft::fft_vcc                        gr-fft/include/gnuradio/fft/fft_vcc.h:47-70, gr-fft/lib/fft_vcc_fftw.{h,cc}
//   gr::fft::fft_vfc                        gr-fft/include/gnuradio/fft/fft_vfc.h, gr-fft/lib/fft_vfc_fftw.{h,cc}
//   gr::filter::pfb_synthesizer_ccf         gr-filter/include/gnuradio/filter/pfb_synthesizer_ccf.h, lib/pfb_synthesizer_ccf_impl.{h,cc}
//   gr::filter::filterbank_vcvcf            gr-filter/include/gnuradio/filter/filterbank_vcvcf.h, lib/filterbank_vcvcf_impl.{h,cc}
//   gr::blocks::complex_to_mag_squared      gr-blocks/lib/complex_to_mag_squared_impl.{h,cc}
//
// Build modes
//   -DGRB200_WITH_GNURADIO : include the real <gnuradio/sync_decimator.h> etc. and derive from gr::block
//                            (what a maintainer compiles inside gr-filter / gr-fft).
//   default                : a ~60-line stand-in of the gr::block surface these classes touch
//                            (grb200::rt below), so the wrappers compile and are unit-tested where GNU Radio
//                            cannot be built (no Boost/VOLK/FFTW in this image).
#ifndef GRB200_BLOCKS_HPP
#define GRB200_BLOCKS_HPP

#include <cmath>
#include <complex>
#include <cstdio>
#include <cstdlib>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <vector>

#include "../grb200.h"

#ifdef GRB200_WITH_GNURADIO
#include <gnuradio/block.h>
#include <gnuradio/io_signature.h>
#include <gnuradio/sptr_magic.h>
#include <gnuradio/sync_block.h>
#include <gnuradio/sync_decimator.h>
namespace grb200 {
namespace rt = ::gr;
}
#else
typedef std::complex<float> gr_complex;
typedef std::vector<int> gr_vector_int;
typedef std::vector<void*> gr_vector_void_star;
typedef std::vector<const void*> gr_vector_const_void_star;
namespace grb200 {
namespace rt {
struct io_signature {
    int min_streams, max_streams, item_size;
    static io_signature make(int mn, int mx, int sz) { return io_signature{ mn, mx, sz }; }
};
class block
{
public:
    enum { WORK_CALLED_PRODUCE = -2, WORK_DONE = -1 };
    enum tag_propagation_policy_t { TPP_DONT = 0, TPP_ALL_TO_ALL = 1, TPP_ONE_TO_ONE = 2 };
    block(const std::string& name, io_signature in, io_signature out) : d_name(name), d_in(in), d_out(out) {}
    virtual ~block() {}
    const std::string& name() const { return d_name; }
    const io_signature& input_signature() const { return d_in; }
    const io_signature& output_signature() const { return d_out; }
    unsigned history() const { return d_history; }
    void set_history(unsigned h) { d_history = h; }
    int output_multiple() const { return d_output_multiple; }
    void set_output_multiple(int m)
    {
        if (m < 1)
            throw std::invalid_argument("block::set_output_multiple");
        d_output_multiple = m;
    }
    double relative_rate() const { return d_relative_rate; }
    void set_relative_rate(double r)
    {
        if (r <= 0.0)
            throw std::invalid_argument("block::set_relative_rate: relative rate must be > 0.0");
        d_relative_rate = r;
    }
    void set_tag_propagation_policy(tag_propagation_policy_t p) { d_tpp = p; }
    void consume_each(int n) { d_consumed = n; }
    int nitems_consumed_last() const { return d_consumed; }
    void set_min_output_buffer(long n) { d_min_output_buffer = n; } // block.h: grows the downstream buffer allocation
    long min_output_buffer() const { return d_min_output_buffer; }
    virtual bool start() { return true; }
    virtual bool stop() { return true; }
    virtual int general_work(int noutput_items, gr_vector_int& ninput_items, gr_vector_const_void_star& input_items,
                             gr_vector_void_star& output_items) = 0;

protected:
    std::mutex d_setlock;

private:
    std::string d_name;
    io_signature d_in, d_out;
    unsigned d_history = 1;
    int d_output_multiple = 1;
    double d_relative_rate = 1.0;
    tag_propagation_policy_t d_tpp = TPP_ALL_TO_ALL;
    int d_consumed = 0;
    long d_min_output_buffer = -1;
};
class sync_block : public block
{
public:
    using block::block;
    virtual int work(int noutput_items, gr_vector_const_void_star& input_items, gr_vector_void_star& output_items) = 0;
    virtual unsigned decimation() const { return 1; }
    int general_work(int noutput_items, gr_vector_int&, gr_vector_const_void_star& input_items,
                     gr_vector_void_star& output_items) override
    {
        int r = work(noutput_items, input_items, output_items); // sync_decimator.cc:45-54
        consume_each(r > 0 ? r * (int)decimation() : 0);
        return r;
    }
};
class sync_decimator : public sync_block
{
public:
    sync_decimator(const std::string& name, io_signature in, io_signature out, unsigned decimation)
        : sync_block(name, in, out), d_decimation(decimation)
    {
        set_relative_rate(1.0 / decimation);
    }
    unsigned decimation() const override { return d_decimation; }

private:
    unsigned d_decimation;
};
class sync_interpolator : public sync_block
{
public:
    sync_interpolator(const std::string& name, io_signature in, io_signature out, unsigned interpolation)
        : sync_block(name, in, out), d_interpolation(interpolation)
    {
        set_relative_rate((double)interpolation);
        set_output_multiple((int)interpolation); // sync_interpolator.cc:27-33
    }
    unsigned interpolation() const { return d_interpolation; }
    int general_work(int noutput_items, gr_vector_int&, gr_vector_const_void_star& input_items,
                     gr_vector_void_star& output_items) override
    {
        int r = work(noutput_items, input_items, output_items); // sync_interpolator.cc:57-66
        consume_each(r > 0 ? r / (int)d_interpolation : 0);
        return r;
    }

private:
    unsigned d_interpolation;
};
} // namespace rt
} // namespace grb200
#endif

namespace grb200 {

// status -> the exception type the reference block throws at the same site
inline void check(int rc)
{
    if (rc == GRB200_OK)
        return;
    const std::string msg = grb200_last_error();
    if (rc == GRB200_EINVAL)
        throw std::invalid_argument(msg);
    if (rc == GRB200_ERANGE)
        throw std::out_of_range(msg);
    throw std::runtime_error(msg);
}

// Device the next make() puts its block on: GRB200_DEVICE in the environment (read once), or whatever the application set
// with set_default_device() -- e.g. one flowgraph per GPU, or the upper and lower half of a channel plan on two GPUs.
inline int& default_device_ref()
{
    static int dev = [] {
        const char* e = std::getenv("GRB200_DEVICE");
        return e ? std::atoi(e) : 0;
    }();
    return dev;
}
inline int default_device() { return default_device_ref(); }
inline void set_default_device(int device) { default_device_ref() = device; }

// Lag-by-one host mode for the blocks that have it (fir_filter_blk, pfb_channelizer_ccf): GRB200_LAG=<items per call> in the
// environment switches it on for every block made afterwards, set_lag() per block.  work() then takes the input of the items
// it was offered, returns the PREVIOUS call's outputs (grb200_*_work_lag) and asks the scheduler for buffers of at least four
// calls (set_min_output_buffer) so that a call is worth a launch and the results can be delivered in place.  Trade: one
// call of latency, and the outputs of the last call are dropped at stop() -- a streaming flowgraph ends mid-stream anyway.
inline long lag_items_env()
{
    static const long v = [] {
        const char* e = std::getenv("GRB200_LAG");
        return e ? std::atol(e) : 0L;
    }();
    return v;
}

// make() hands the new block to the runtime exactly like the reference: gnuradio::get_initial_sptr (gnuradio-runtime/include/
// gnuradio/sptr_magic.h) registers it for hier-block self-connection; the stand-in build has no such registry.
#ifdef GRB200_WITH_GNURADIO
template <class T>
inline std::shared_ptr<T> make_sptr(T* p)
{
    return gnuradio::get_initial_sptr(p);
}
#else
template <class T>
inline std::shared_ptr<T> make_sptr(T* p)
{
    return std::shared_ptr<T>(p);
}
#endif

} // namespace grb200

namespace gr {
namespace filter {

// ---------------------------------------------------------------------------------------------------
// fir_filter_blk<gr_complex, gr_complex, float|gr_complex> and <float, float, float>  (fir_filter_ccf / _ccc / _fff)
// ---------------------------------------------------------------------------------------------------
template <class IN_T, class OUT_T, class TAP_T>
class fir_filter_blk : public grb200::rt::sync_decimator
{
    static constexpr bool is_fff =
        std::is_same<IN_T, float>::value && std::is_same<OUT_T, float>::value && std::is_same<TAP_T, float>::value;
    static_assert(is_fff || (std::is_same<IN_T, gr_complex>::value && std::is_same<OUT_T, gr_complex>::value &&
                             (std::is_same<TAP_T, float>::value || std::is_same<TAP_T, gr_complex>::value)),
                  "libgrb200 re-hosts the ccf, ccc and fff instantiations (SURVEY.md section 8)");

public:
    typedef std::shared_ptr<fir_filter_blk<IN_T, OUT_T, TAP_T>> sptr;

    static sptr make(int decimation, const std::vector<TAP_T>& taps)
    {
        return grb200::make_sptr(new fir_filter_blk<IN_T, OUT_T, TAP_T>(decimation, taps));
    }

    ~fir_filter_blk() override { grb200_fir_destroy(d_h); }

    void set_taps(const std::vector<TAP_T>& taps)
    {
        std::lock_guard<std::mutex> l(d_setlock); // fir_filter_blk_impl.cc:56
        grb200::check(grb200_fir_set_taps(d_h, reinterpret_cast<const float*>(taps.data()), (int)taps.size()));
        d_taps = taps;
        d_updated = true;
    }
    std::vector<TAP_T> taps() const { return d_taps; }

    int work(int noutput_items, gr_vector_const_void_star& input_items, gr_vector_void_star& output_items) override
    {
        std::lock_guard<std::mutex> l(d_setlock); // fir_filter_blk_impl.cc:72
        if (d_updated) {                           // :77-81
            set_history(d_taps.size());
            d_updated = false;
            return 0; // history requirements may have changed
        }
        grb200::check(grb200_fir_work(d_h, input_items[0], output_items[0], noutput_items));
        return noutput_items;
    }

    grb200_fir* handle() { return d_h; } // device-resident chaining (grb200_fir_work_dev)

    // ---- lag-by-one host mode (grb200.h "Host staging") ----
    void set_lag(long items_per_call)
    {
        std::lock_guard<std::mutex> l(d_setlock);
        d_lag = items_per_call;
        if (d_lag > 0)
            set_min_output_buffer(4 * d_lag); // flat_flowgraph.cc:131-145 honours it when the buffers are allocated
    }
    long lag() const { return d_lag; }
    int general_work(int noutput_items, gr_vector_int& ninput_items, gr_vector_const_void_star& input_items,
                     gr_vector_void_star& output_items) override
    {
        if (d_lag <= 0)
            return grb200::rt::sync_decimator::general_work(noutput_items, ninput_items, input_items, output_items);
        std::lock_guard<std::mutex> l(d_setlock);
        if (d_updated) {
            set_history(d_taps.size());
            d_updated = false;
            consume_each(0);
            return 0;
        }
        // take as much input as leaves room for its results behind the ones returned now (delivered in place when the
        // buffer is page-locked), at most one lag unit
        if (noutput_items < d_parked) { // no room yet for what is waiting (or the stream is ending): no progress this call
            consume_each(0);
            return 0;
        }
        long long n = (long long)noutput_items - d_parked;
        n = n > d_lag ? d_lag : n;
        long long produced = 0;
        grb200::check(grb200_fir_work_lag(d_h, input_items[0], n, output_items[0], noutput_items, &produced));
        consume_each((int)(n * (long long)decimation()));
        d_parked = n;
        return (int)produced; // gr::block: fewer items than offered is fine (block.h:160-175)
    }
    bool stop() override
    {
        d_parked = 0; // the call still in flight is dropped with the flowgraph
        return true;
    }

private:
    fir_filter_blk(int decimation, const std::vector<TAP_T>& taps)
        : sync_decimator(is_fff ? "fir_filter_fff" : std::is_same<TAP_T, float>::value ? "fir_filter_ccf" : "fir_filter_ccc",
                         grb200::rt::io_signature::make(1, 1, sizeof(IN_T)),
                         grb200::rt::io_signature::make(1, 1, sizeof(OUT_T)), decimation),
          d_taps(taps)
    {
        const float* t = reinterpret_cast<const float*>(taps.data());
        if (is_fff)
            grb200::check(grb200_fir_fff_create(decimation, t, (int)taps.size(), grb200::default_device(), &d_h));
        else if (std::is_same<TAP_T, float>::value)
            grb200::check(grb200_fir_ccf_create(decimation, t, (int)taps.size(), grb200::default_device(), &d_h));
        else
            grb200::check(grb200_fir_ccc_create(decimation, t, (int)taps.size(), grb200::default_device(), &d_h));
        set_history(taps.size()); // fir_filter_blk_impl.cc:41
        if (grb200::lag_items_env() > 0)
            set_lag(grb200::lag_items_env());
    }
    grb200_fir* d_h = nullptr;
    std::vector<TAP_T> d_taps;
    bool d_updated = false;
    long d_lag = 0;
    long long d_parked = 0; // items of the previous lag call not yet returned
};
typedef fir_filter_blk<gr_complex, gr_complex, float> fir_filter_ccf;
typedef fir_filter_blk<gr_complex, gr_complex, gr_complex> fir_filter_ccc;
typedef fir_filter_blk<float, float, float> fir_filter_fff;

// ---------------------------------------------------------------------------------------------------
// freq_xlating_fir_filter<gr_complex, gr_complex, float|gr_complex>  (freq_xlating_fir_filter_ccf / _ccc)
// gr-filter/include/gnuradio/filter/freq_xlating_fir_filter.h:56-90, lib/freq_xlating_fir_filter_impl.cc
// ---------------------------------------------------------------------------------------------------
template <class IN_T, class OUT_T, class TAP_T>
class freq_xlating_fir_filter : public grb200::rt::sync_decimator
{
    static constexpr bool is_fff =
        std::is_same<IN_T, float>::value && std::is_same<OUT_T, float>::value && std::is_same<TAP_T, float>::value;
    static_assert(is_fff || (std::is_same<IN_T, gr_complex>::value && std::is_same<OUT_T, gr_complex>::value &&
                             (std::is_same<TAP_T, float>::value || std::is_same<TAP_T, gr_complex>::value)),
                  "libgrb200 re-hosts the ccf, ccc and fff instantiations (SURVEY.md section 8)");

public:
    typedef std::shared_ptr<freq_xlating_fir_filter<IN_T, OUT_T, TAP_T>> sptr;

    static sptr make(int decimation, const std::vector<TAP_T>& taps, double center_freq, double sampling_freq)
    {
        return grb200::make_sptr(new freq_xlating_fir_filter<IN_T, OUT_T, TAP_T>(decimation, taps, center_freq, sampling_freq));
    }

    ~freq_xlating_fir_filter() override { grb200_freq_xlating_fir_destroy(d_h); }

    void set_center_freq(double center_freq) // freq_xlating_fir_filter_impl.cc:96-100: picked up by the next work()
    {
        d_center_freq = center_freq;
        d_updated = true;
    }
    double center_freq() const { return d_center_freq; }

    void set_taps(const std::vector<TAP_T>& taps) // :109-113
    {
        d_proto_taps = taps;
        d_updated = true;
    }
    std::vector<TAP_T> taps() const { return d_proto_taps; }

    int work(int noutput_items, gr_vector_const_void_star& input_items, gr_vector_void_star& output_items) override
    {
        if (d_updated) { // :155-160: rebuild composite FIR, history requirements may have changed
            set_history(d_proto_taps.size());
            grb200::check(grb200_freq_xlating_fir_set_center_freq(d_h, d_center_freq));
            grb200::check(grb200_freq_xlating_fir_set_taps(
                d_h, reinterpret_cast<const float*>(d_proto_taps.data()), (int)d_proto_taps.size()));
            d_updated = false;
            return 0;
        }
        grb200::check(grb200_freq_xlating_fir_work(d_h, input_items[0], output_items[0], noutput_items));
        return noutput_items;
    }

    grb200_xlat* handle() { return d_h; } // device-resident chaining (grb200_freq_xlating_fir_work_dev)

private:
    freq_xlating_fir_filter(int decimation, const std::vector<TAP_T>& taps, double center_freq, double sampling_freq)
        : sync_decimator(std::is_same<TAP_T, float>::value ? "freq_xlating_fir_filter_ccf" : "freq_xlating_fir_filter_ccc",
                         grb200::rt::io_signature::make(1, 1, sizeof(IN_T)),
                         grb200::rt::io_signature::make(1, 1, sizeof(OUT_T)), decimation),
          d_proto_taps(taps), d_center_freq(center_freq)
    {
        const float* t = reinterpret_cast<const float*>(taps.data());
        if (std::is_same<TAP_T, float>::value)
            grb200::check(grb200_freq_xlating_fir_ccf_create(decimation, t, (int)taps.size(), center_freq, sampling_freq,
                                                             grb200::default_device(), &d_h));
        else
            grb200::check(grb200_freq_xlating_fir_ccc_create(decimation, t, (int)taps.size(), center_freq, sampling_freq,
                                                             grb200::default_device(), &d_h));
        set_history(taps.size()); // freq_xlating_fir_filter_impl.cc:53
    }
    grb200_xlat* d_h = nullptr;
    std::vector<TAP_T> d_proto_taps;
    double d_center_freq;
    bool d_updated = false;
};
typedef freq_xlating_fir_filter<gr_complex, gr_complex, float> freq_xlating_fir_filter_ccf;
typedef freq_xlating_fir_filter<gr_complex, gr_complex, gr_complex> freq_xlating_fir_filter_ccc;

// ---------------------------------------------------------------------------------------------------
// interp_fir_filter<gr_complex, gr_complex, float|gr_complex>  (interp_fir_filter_ccf / _ccc)
// gr-filter/include/gnuradio/filter/interp_fir_filter.h:43-72, lib/interp_fir_filter_impl.cc
// ---------------------------------------------------------------------------------------------------
template <class IN_T, class OUT_T, class TAP_T>
class interp_fir_filter : public grb200::rt::sync_interpolator
{
    static constexpr bool is_fff =
        std::is_same<IN_T, float>::value && std::is_same<OUT_T, float>::value && std::is_same<TAP_T, float>::value;
    static_assert(is_fff || (std::is_same<IN_T, gr_complex>::value && std::is_same<OUT_T, gr_complex>::value &&
                             (std::is_same<TAP_T, float>::value || std::is_same<TAP_T, gr_complex>::value)),
                  "libgrb200 re-hosts the ccf, ccc and fff instantiations (SURVEY.md section 8)");

public:
    typedef std::shared_ptr<interp_fir_filter<IN_T, OUT_T, TAP_T>> sptr;

    static sptr make(unsigned interpolation, const std::vector<TAP_T>& taps)
    {
        if (interpolation == 0) // :46-48
            throw std::out_of_range("interp_fir_filter_impl: interpolation must be > 0");
        if (taps.empty()) // :50-52
            throw std::runtime_error("interp_fir_filter_impl: no filter taps provided.");
        return grb200::make_sptr(new interp_fir_filter<IN_T, OUT_T, TAP_T>(interpolation, taps));
    }

    ~interp_fir_filter() override { grb200_interp_fir_destroy(d_h); }

    void set_taps(const std::vector<TAP_T>& taps) // :72-90: padded to a multiple of the interpolation
    {
        d_new_taps = taps;
        d_updated = true;
        int n = taps.size() % this->interpolation();
        if (n > 0) {
            n = this->interpolation() - n;
            while (n-- > 0)
                d_new_taps.insert(d_new_taps.end(), 0);
        }
    }
    std::vector<TAP_T> taps() const { return d_new_taps; }

    int work(int noutput_items, gr_vector_const_void_star& input_items, gr_vector_void_star& output_items) override
    {
        if (d_updated) { // :136-139: install_taps, history requirements may have changed
            grb200::check(grb200_interp_fir_set_taps(d_h, reinterpret_cast<const float*>(d_new_taps.data()),
                                                     (int)d_new_taps.size()));
            set_history(grb200_interp_fir_taps_per_filter(d_h));
            d_updated = false;
            return 0;
        }
        grb200::check(grb200_interp_fir_work(d_h, input_items[0], output_items[0], noutput_items));
        return noutput_items;
    }

    grb200_interp* handle() { return d_h; } // device-resident chaining (grb200_interp_fir_work_dev)

private:
    interp_fir_filter(unsigned interpolation, const std::vector<TAP_T>& taps)
        : sync_interpolator(std::is_same<TAP_T, float>::value ? "interp_fir_filter_ccf" : "interp_fir_filter_ccc",
                            grb200::rt::io_signature::make(1, 1, sizeof(IN_T)),
                            grb200::rt::io_signature::make(1, 1, sizeof(OUT_T)), interpolation)
    {
        set_taps(taps);
        const float* t = reinterpret_cast<const float*>(d_new_taps.data());
        if (std::is_same<TAP_T, float>::value)
            grb200::check(grb200_interp_fir_ccf_create(interpolation, t, (int)d_new_taps.size(), grb200::default_device(), &d_h));
        else
            grb200::check(grb200_interp_fir_ccc_create(interpolation, t, (int)d_new_taps.size(), grb200::default_device(), &d_h));
        set_history(grb200_interp_fir_taps_per_filter(d_h)); // install_taps, :115
        d_updated = false;
    }
    grb200_interp* d_h = nullptr;
    std::vector<TAP_T> d_new_taps;
    bool d_updated = false;
};
typedef interp_fir_filter<gr_complex, gr_complex, float> interp_fir_filter_ccf;
typedef interp_fir_filter<gr_complex, gr_complex, gr_complex> interp_fir_filter_ccc;

// ---------------------------------------------------------------------------------------------------
// rational_resampler_base<gr_complex, gr_complex, float|gr_complex>  (rational_resampler_base_ccf / _ccc)
// gr-filter/include/gnuradio/filter/rational_resampler_base.h, lib/rational_resampler_base_impl.{h,cc}
// ---------------------------------------------------------------------------------------------------
template <class IN_T, class OUT_T, class TAP_T>
class rational_resampler_base : public grb200::rt::block
{
    static_assert(std::is_same<IN_T, gr_complex>::value && std::is_same<OUT_T, gr_complex>::value,
                  "only the ccf and ccc instantiations are built");

public:
    typedef std::shared_ptr<rational_resampler_base<IN_T, OUT_T, TAP_T>> sptr;

    static sptr make(unsigned interpolation, unsigned decimation, const std::vector<TAP_T>& taps)
    {
        if (interpolation == 0) // :46-48
            throw std::out_of_range("rational_resampler_base_impl<IN_T,OUT_T,TAP_T>: interpolation must be > 0");
        if (decimation == 0) // :49-51
            throw std::out_of_range("rational_resampler_base_impl<IN_T,OUT_T,TAP_T>: decimation must be > 0");
        if (taps.empty())
            throw std::runtime_error("rational_resampler_base_impl: no filter taps provided.");
        return grb200::make_sptr(new rational_resampler_base<IN_T, OUT_T, TAP_T>(interpolation, decimation, taps));
    }

    ~rational_resampler_base() override { grb200_rational_resampler_destroy(d_h); }

    // shadow the block's accessors like rational_resampler_base_impl.h:48-49: the scheduler keeps seeing history 1
    unsigned history() const { return d_history; }
    void set_history(unsigned history) { d_history = history; }

    unsigned interpolation() const { return d_interpolation; }
    unsigned decimation() const { return d_decimation; }

    void set_taps(const std::vector<TAP_T>& taps) // :77-93: padded to a multiple of the interpolation
    {
        d_new_taps = taps;
        d_updated = true;
        int n = taps.size() % d_interpolation;
        if (n > 0) {
            n = d_interpolation - n;
            while (n-- > 0)
                d_new_taps.insert(d_new_taps.end(), TAP_T(0));
        }
    }
    std::vector<TAP_T> taps() const { return d_new_taps; }

    void forecast(int noutput_items, gr_vector_int& ninput_items_required) // :128-138
    {
        int nreqd = std::max((unsigned)1, (int)((double)(noutput_items + 1) * d_decimation / d_interpolation) + history() - 1);
        for (auto& r : ninput_items_required)
            r = nreqd;
    }

    // input_items[0] must hold ninput_items[0] + history()-1 readable items (what forecast() asked for covers it)
    int general_work(int noutput_items, gr_vector_int& ninput_items, gr_vector_const_void_star& input_items,
                     gr_vector_void_star& output_items) override
    {
        if (d_updated) { // :151-154
            grb200::check(grb200_rational_resampler_set_taps(d_h, reinterpret_cast<const float*>(d_new_taps.data()),
                                                             (int)d_new_taps.size()));
            set_history(grb200_rational_resampler_taps_per_filter(d_h));
            d_updated = false;
            return 0;
        }
        long long consumed = 0, produced = 0;
        grb200::check(grb200_rational_resampler_work(d_h, input_items[0], ninput_items[0], output_items[0], noutput_items,
                                                     &consumed, &produced));
        consume_each((int)consumed);
        return (int)produced;
    }

    grb200_resamp* handle() { return d_h; } // device-resident chaining (grb200_rational_resampler_work_dev)

private:
    rational_resampler_base(unsigned interpolation, unsigned decimation, const std::vector<TAP_T>& taps)
        : block(std::is_same<TAP_T, float>::value ? "rational_resampler_base_ccf" : "rational_resampler_base_ccc",
                grb200::rt::io_signature::make(1, 1, sizeof(IN_T)), grb200::rt::io_signature::make(1, 1, sizeof(OUT_T))),
          d_interpolation(interpolation), d_decimation(decimation)
    {
        set_relative_rate((double)interpolation / (double)decimation); // :53
        set_output_multiple(1);
        set_taps(taps);
        const float* t = reinterpret_cast<const float*>(d_new_taps.data());
        if (std::is_same<TAP_T, float>::value)
            grb200::check(grb200_rational_resampler_ccf_create(interpolation, decimation, t, (int)d_new_taps.size(),
                                                               grb200::default_device(), &d_h));
        else
            grb200::check(grb200_rational_resampler_ccc_create(interpolation, decimation, t, (int)d_new_taps.size(),
                                                               grb200::default_device(), &d_h));
        set_history(grb200_rational_resampler_taps_per_filter(d_h)); // install_taps, :115
        d_updated = false;
    }
    unsigned d_history = 1, d_interpolation, d_decimation;
    grb200_resamp* d_h = nullptr;
    std::vector<TAP_T> d_new_taps;
    bool d_updated = false;
};
typedef rational_resampler_base<gr_complex, gr_complex, float> rational_resampler_base_ccf;
typedef rational_resampler_base<gr_complex, gr_complex, gr_complex> rational_resampler_base_ccc;

// ---------------------------------------------------------------------------------------------------
// fft_filter_ccc
// ---------------------------------------------------------------------------------------------------
class fft_filter_ccc : public grb200::rt::sync_decimator
{
public:
    typedef std::shared_ptr<fft_filter_ccc> sptr;
    static sptr make(int decimation, const std::vector<gr_complex>& taps, int nthreads = 1)
    {
        return grb200::make_sptr(new fft_filter_ccc(decimation, taps, nthreads));
    }
    ~fft_filter_ccc() override { grb200_fftfilt_destroy(d_h); }

    void set_taps(const std::vector<gr_complex>& taps) // fft_filter_ccc_impl.cc:49-53
    {
        d_new_taps = taps;
        d_updated = true;
    }
    std::vector<gr_complex> taps() const { return d_new_taps; }
    void set_nthreads(int n) { d_nthreads = n; } // FFTW knob; the device transform has none
    int nthreads() const { return d_nthreads; }

    int work(int noutput_items, gr_vector_const_void_star& input_items, gr_vector_void_star& output_items) override
    {
        if (d_updated) { // fft_filter_ccc_impl.cc:68-73
            grb200::check(grb200_fftfilt_set_taps(d_h, reinterpret_cast<const float*>(d_new_taps.data()),
                                                  (int)d_new_taps.size()));
            d_nsamples = grb200_fftfilt_nsamples(d_h);
            d_updated = false;
            set_output_multiple(d_nsamples);
            return 0; // output multiple may have changed
        }
        grb200::check(grb200_fftfilt_work(d_h, input_items[0], output_items[0], noutput_items));
        return noutput_items;
    }
    grb200_fftfilt* handle() { return d_h; }

private:
    fft_filter_ccc(int decimation, const std::vector<gr_complex>& taps, int nthreads)
        : sync_decimator("fft_filter_ccc", grb200::rt::io_signature::make(1, 1, sizeof(gr_complex)),
                         grb200::rt::io_signature::make(1, 1, sizeof(gr_complex)), decimation),
          d_new_taps(taps), d_nthreads(nthreads)
    {
        set_history(1);
        grb200::check(grb200_fftfilt_ccc_create(decimation, reinterpret_cast<const float*>(taps.data()),
                                                (int)taps.size(), grb200::default_device(), &d_h));
        d_nsamples = grb200_fftfilt_nsamples(d_h);
        set_output_multiple(d_nsamples);
    }
    grb200_fftfilt* d_h = nullptr;
    std::vector<gr_complex> d_new_taps;
    int d_nsamples = 0, d_nthreads = 1;
    bool d_updated = false;
};

// fft_filter_ccf (gr-filter/include/gnuradio/filter/fft_filter_ccf.h, lib/fft_filter_ccf_impl.cc:32-78): real taps
class fft_filter_ccf : public grb200::rt::sync_decimator
{
public:
    typedef std::shared_ptr<fft_filter_ccf> sptr;
    static sptr make(int decimation, const std::vector<float>& taps, int nthreads = 1)
    {
        return grb200::make_sptr(new fft_filter_ccf(decimation, taps, nthreads));
    }
    ~fft_filter_ccf() override { grb200_fftfilt_destroy(d_h); }

    void set_taps(const std::vector<float>& taps) // fft_filter_ccf_impl.cc:49-53
    {
        d_new_taps = taps;
        d_updated = true;
    }
    std::vector<float> taps() const { return d_new_taps; }
    void set_nthreads(int n) { d_nthreads = n; }
    int nthreads() const { return d_nthreads; }

    int work(int noutput_items, gr_vector_const_void_star& input_items, gr_vector_void_star& output_items) override
    {
        if (d_updated) { // fft_filter_ccf_impl.cc:68-73
            grb200::check(grb200_fftfilt_set_taps_real(d_h, d_new_taps.data(), (int)d_new_taps.size()));
            d_nsamples = grb200_fftfilt_nsamples(d_h);
            d_updated = false;
            set_output_multiple(d_nsamples);
            return 0;
        }
        grb200::check(grb200_fftfilt_work(d_h, input_items[0], output_items[0], noutput_items));
        return noutput_items;
    }
    grb200_fftfilt* handle() { return d_h; }

private:
    fft_filter_ccf(int decimation, const std::vector<float>& taps, int nthreads)
        : sync_decimator("fft_filter_ccf", grb200::rt::io_signature::make(1, 1, sizeof(gr_complex)),
                         grb200::rt::io_signature::make(1, 1, sizeof(gr_complex)), decimation),
          d_new_taps(taps), d_nthreads(nthreads)
    {
        set_history(1);
        grb200::check(grb200_fftfilt_ccf_create(decimation, taps.data(), (int)taps.size(), grb200::default_device(), &d_h));
        d_nsamples = grb200_fftfilt_nsamples(d_h);
        set_output_multiple(d_nsamples);
    }
    grb200_fftfilt* d_h = nullptr;
    std::vector<float> d_new_taps;
    int d_nsamples = 0, d_nthreads = 1;
    bool d_updated = false;
};

// ---------------------------------------------------------------------------------------------------
// pfb_channelizer_ccf
// ---------------------------------------------------------------------------------------------------
class pfb_channelizer_ccf : public grb200::rt::block
{
public:
    typedef std::shared_ptr<pfb_channelizer_ccf> sptr;
    static sptr make(unsigned int numchans, const std::vector<float>& taps, float oversample_rate)
    {
        return grb200::make_sptr(new pfb_channelizer_ccf(numchans, taps, oversample_rate));
    }
    ~pfb_channelizer_ccf() override { grb200_pfb_chan_destroy(d_h); }

    void set_taps(const std::vector<float>& taps) // pfb_channelizer_ccf_impl.cc:96-103
    {
        std::lock_guard<std::mutex> l(d_setlock);
        grb200::check(grb200_pfb_chan_set_taps(d_h, taps.data(), (int)taps.size()));
        set_history(grb200_pfb_chan_taps_per_filter(d_h) + 1);
        d_updated = true;
    }
    std::vector<std::vector<float>> taps() const
    {
        const int T = grb200_pfb_chan_taps_per_filter(d_h);
        std::vector<float> flat((size_t)d_nfilts * T);
        grb200::check(grb200_pfb_chan_branch_taps(d_h, flat.data()));
        std::vector<std::vector<float>> out(d_nfilts);
        for (unsigned i = 0; i < d_nfilts; i++)
            out[i].assign(flat.begin() + (size_t)i * T, flat.begin() + (size_t)(i + 1) * T);
        return out;
    }
    void print_taps()
    {
        auto t = taps();
        for (unsigned i = 0; i < d_nfilts; i++) {
            printf("filter[%u]: [", i);
            for (float v : t[i])
                printf(" %.4e", v);
            printf("]\n\n");
        }
    }
    void set_channel_map(const std::vector<int>& map) // :112-124
    {
        std::lock_guard<std::mutex> l(d_setlock);
        grb200::check(grb200_pfb_chan_set_channel_map(d_h, map.data(), (int)map.size()));
    }
    std::vector<int> channel_map() const
    {
        std::vector<int> m(d_nfilts);
        int n = grb200_pfb_chan_get_channel_map(d_h, m.data(), (int)m.size());
        m.resize(n < 0 ? 0 : n);
        return m;
    }

    int general_work(int noutput_items, gr_vector_int&, gr_vector_const_void_star& input_items,
                     gr_vector_void_star& output_items) override
    {
        std::lock_guard<std::mutex> l(d_setlock); // :133
        if (d_updated) {                           // :140-143
            d_updated = false;
            return 0; // history requirements may have changed
        }
        long long consumed = 0;
        if (d_lag > 0) { // lag-by-one host mode: see fir_filter_blk::general_work
            const int om = output_multiple();
            if (noutput_items < d_parked) {
                consume_each(0);
                return 0;
            }
            long long n = (long long)noutput_items - d_parked;
            n = n > d_lag ? d_lag : n;
            n = n / om * om;
            long long produced = 0;
            grb200::check(grb200_pfb_chan_work_lag(d_h, input_items.data(), n, output_items.data(), (int)output_items.size(),
                                                   noutput_items, &consumed, &produced));
            consume_each((int)consumed);
            d_parked = n;
            return (int)produced;
        }
        grb200::check(grb200_pfb_chan_work(d_h, input_items.data(), output_items.data(), (int)output_items.size(),
                                           noutput_items, &consumed));
        consume_each((int)consumed); // :191
        return noutput_items;
    }
    grb200_pfb* handle() { return d_h; }
    void set_lag(long items_per_call)
    {
        std::lock_guard<std::mutex> l(d_setlock);
        d_lag = items_per_call;
        if (d_lag > 0)
            set_min_output_buffer(4 * d_lag);
    }
    long lag() const { return d_lag; }
    bool stop() override
    {
        d_parked = 0;
        return true;
    }

private:
    pfb_channelizer_ccf(unsigned int nfilts, const std::vector<float>& taps, float oversample_rate)
        : block("pfb_channelizer_ccf", grb200::rt::io_signature::make(nfilts, nfilts, sizeof(gr_complex)),
                grb200::rt::io_signature::make(1, nfilts, sizeof(gr_complex))),
          d_nfilts(nfilts)
    {
        grb200::check(grb200_pfb_chan_ccf_create(nfilts, taps.data(), (int)taps.size(), oversample_rate,
                                                 grb200::default_device(), &d_h)); // throws invalid_argument (:49-53)
        set_relative_rate(oversample_rate);                               // :55
        set_output_multiple(grb200_pfb_chan_output_multiple(d_h));        // :81-84
        set_history(grb200_pfb_chan_taps_per_filter(d_h) + 1);            // :101
        d_updated = true;                                                 // ctor goes through set_taps (:87)
        set_tag_propagation_policy(TPP_ONE_TO_ONE);                       // :91
        if (grb200::lag_items_env() > 0)
            set_lag(grb200::lag_items_env());
    }
    long d_lag = 0;
    long long d_parked = 0;
    grb200_pfb* d_h = nullptr;
    unsigned d_nfilts;
    bool d_updated = false;
};

// ---------------------------------------------------------------------------------------------------
// pfb_decimator_ccf  (gr-filter/include/gnuradio/filter/pfb_decimator_ccf.h:94-138, lib/pfb_decimator_ccf_impl.cc)
// ---------------------------------------------------------------------------------------------------
class pfb_decimator_ccf : public grb200::rt::sync_block
{
public:
    typedef std::shared_ptr<pfb_decimator_ccf> sptr;
    static sptr make(unsigned int decim, const std::vector<float>& taps, unsigned int channel,
                     bool use_fft_rotator = true, bool use_fft_filters = true)
    {
        return grb200::make_sptr(new pfb_decimator_ccf(decim, taps, channel, use_fft_rotator, use_fft_filters));
    }
    ~pfb_decimator_ccf() override { grb200_pfb_dec_destroy(d_h); }

    void set_taps(const std::vector<float>& taps) // pfb_decimator_ccf_impl.cc:88-95
    {
        std::lock_guard<std::mutex> l(d_setlock);
        grb200::check(grb200_pfb_dec_set_taps(d_h, taps.data(), (int)taps.size()));
        d_proto = taps;
        set_history(grb200_pfb_dec_taps_per_filter(d_h));
        d_updated = true;
    }
    std::vector<std::vector<float>> taps() const // d_taps[i][j] = padded[i + j*M], polyphase_filterbank.cc:69-73
    {
        const int T = grb200_pfb_dec_taps_per_filter(d_h);
        std::vector<std::vector<float>> out(d_rate, std::vector<float>((size_t)T, 0.f));
        for (size_t k = 0; k < d_proto.size(); k++)
            out[k % d_rate][k / d_rate] = d_proto[k];
        return out;
    }
    void print_taps()
    {
        auto t = taps();
        for (unsigned i = 0; i < d_rate; i++) {
            printf("filter[%u]: [", i);
            for (float v : t[i])
                printf(" %.4e", v);
            printf("]\n\n");
        }
    }
    void set_channel(const unsigned int channel) // :106-110
    {
        std::lock_guard<std::mutex> l(d_setlock);
        grb200::check(grb200_pfb_dec_set_channel(d_h, channel));
    }

    int work(int noutput_items, gr_vector_const_void_star& input_items, gr_vector_void_star& output_items) override
    {
        std::lock_guard<std::mutex> l(d_setlock);
        if (d_updated) { // :114-117
            d_updated = false;
            return 0; // history requirements may have changed.
        }
        if (input_items.size() != d_rate)
            throw std::invalid_argument("pfb_decimator_ccf: one input stream per branch is required");
        // history 1 (the FFT-filter variants, :57-59): the handle carries the overlap between calls
        const int carry = history() == 1 && grb200_pfb_dec_taps_per_filter(d_h) > 1;
        grb200::check(grb200_pfb_dec_work(d_h, input_items.data(), output_items[0], noutput_items, carry));
        return noutput_items;
    }

    grb200_pfbdec* handle() { return d_h; } // device-resident chaining (grb200_pfb_dec_work_dev)

private:
    pfb_decimator_ccf(unsigned int decim, const std::vector<float>& taps, unsigned int channel, bool use_fft_rotator,
                      bool use_fft_filters)
        : sync_block("pfb_decimator_ccf", grb200::rt::io_signature::make(decim, decim, sizeof(gr_complex)),
                     grb200::rt::io_signature::make(1, 1, sizeof(gr_complex))),
          d_rate(decim), d_proto(taps), d_use_fft_rotator(use_fft_rotator), d_use_fft_filters(use_fft_filters)
    {
        grb200::check(grb200_pfb_dec_ccf_create(decim, taps.data(), (int)taps.size(), channel, grb200::default_device(),
                                                &d_h));
        const int T = grb200_pfb_dec_taps_per_filter(d_h);
        if (d_use_fft_filters) { // :56-59, fft_filter_ccf geometry for T taps (fft_filter.cc:336-341)
            const int fftsize = (int)(2 * pow(2.0, ceil(log((double)T) / log(2.0))));
            set_history(1);
            set_output_multiple(fftsize - T + 1);
        } else {
            set_history(T); // :61
        }
    }
    grb200_pfbdec* d_h = nullptr;
    unsigned int d_rate;
    std::vector<float> d_proto;
    bool d_use_fft_rotator, d_use_fft_filters; // the reference's four variants differ in rounding only
    bool d_updated = false;
};

// ---------------------------------------------------------------------------------------------------
// pfb_synthesizer_ccf  (gr-filter/include/gnuradio/filter/pfb_synthesizer_ccf.h, lib/pfb_synthesizer_ccf_impl.{h,cc})
// ---------------------------------------------------------------------------------------------------
class pfb_synthesizer_ccf : public grb200::rt::sync_interpolator
{
public:
    typedef std::shared_ptr<pfb_synthesizer_ccf> sptr;
    static sptr make(unsigned int numchans, const std::vector<float>& taps, bool twox = false)
    {
        return grb200::make_sptr(new pfb_synthesizer_ccf(numchans, taps, twox));
    }
    ~pfb_synthesizer_ccf() override { grb200_pfb_synth_destroy(d_h); }

    void set_taps(const std::vector<float>& taps) // pfb_synthesizer_ccf_impl.cc:77-93
    {
        std::lock_guard<std::mutex> l(d_setlock);
        grb200::check(grb200_pfb_synth_set_taps(d_h, taps.data(), (int)taps.size()));
        set_history(1);
        d_updated = true;
    }
    std::vector<std::vector<float>> taps() const // d_taps, :95-184
    {
        const int rows = grb200_pfb_synth_twox(d_h) * grb200_pfb_synth_numchans(d_h), T = grb200_pfb_synth_taps_per_filter(d_h);
        std::vector<float> flat((size_t)rows * T);
        grb200::check(grb200_pfb_synth_branch_taps(d_h, flat.data()));
        std::vector<std::vector<float>> out((size_t)rows);
        for (int i = 0; i < rows; i++)
            out[i].assign(flat.begin() + (size_t)i * T, flat.begin() + (size_t)(i + 1) * T);
        return out;
    }
    void print_taps()
    {
        const std::vector<std::vector<float>> t = taps();
        for (size_t i = 0; i < t.size(); i++) {
            printf("filter[%d]: [", (int)i);
            for (float v : t[i])
                printf(" %.4e", v);
            printf("]\n\n");
        }
    }
    void set_channel_map(const std::vector<int>& map) // :202-217
    {
        std::lock_guard<std::mutex> l(d_setlock);
        grb200::check(grb200_pfb_synth_set_channel_map(d_h, map.data(), (int)map.size()));
    }
    std::vector<int> channel_map() const
    {
        std::vector<int> m((size_t)grb200_pfb_synth_twox(d_h) * grb200_pfb_synth_numchans(d_h) * 2 + 16);
        const int n = grb200_pfb_synth_get_channel_map(d_h, m.data(), (int)m.size());
        grb200::check(n < 0 ? n : 0);
        m.resize((size_t)n);
        return m;
    }

    int work(int noutput_items, gr_vector_const_void_star& input_items, gr_vector_void_star& output_items) override
    {
        std::lock_guard<std::mutex> l(d_setlock); // :225
        if (d_updated) {                           // :229-232
            d_updated = false;
            return 0; // history requirements may have changed
        }
        grb200::check(grb200_pfb_synth_work(d_h, input_items.data(), (int)input_items.size(), output_items[0], noutput_items));
        return noutput_items;
    }
    grb200_pfbsynth* handle() { return d_h; } // device-resident chaining (grb200_pfb_synth_work_dev)

private:
    pfb_synthesizer_ccf(unsigned int numchans, const std::vector<float>& taps, bool twox)
        : sync_interpolator("pfb_synthesizer_ccf", grb200::rt::io_signature::make(1, numchans, sizeof(gr_complex)),
                            grb200::rt::io_signature::make(1, 1, sizeof(gr_complex)), numchans)
    {
        // an odd channel count with twox throws std::invalid_argument (:44-47)
        grb200::check(grb200_pfb_synth_ccf_create(numchans, taps.data(), (int)taps.size(), twox ? 1 : 0,
                                                  grb200::default_device(), &d_h));
        set_history(1);                 // :92
        set_output_multiple(numchans);  // :68
        d_updated = true;               // the constructor goes through set_taps (:63)
    }
    grb200_pfbsynth* d_h = nullptr;
    bool d_updated = false;
};

// ---------------------------------------------------------------------------------------------------
// filterbank_vcvcf  (gr-filter/include/gnuradio/filter/filterbank_vcvcf.h, lib/filterbank_vcvcf_impl.{h,cc}, lib/filterbank.cc)
// ---------------------------------------------------------------------------------------------------
class filterbank_vcvcf : public grb200::rt::block
{
public:
    typedef std::shared_ptr<filterbank_vcvcf> sptr;
    static sptr make(const std::vector<std::vector<float>>& taps) { return grb200::make_sptr(new filterbank_vcvcf(taps)); }
    ~filterbank_vcvcf() override { grb200_filterbank_vcvcf_destroy(d_h); }

    void set_taps(const std::vector<std::vector<float>>& taps) // filterbank_vcvcf_impl.cc:39-45 over filterbank.cc:50-79
    {
        std::lock_guard<std::mutex> l(d_setlock);
        if (taps.size() != d_taps.size())
            throw std::runtime_error("The number of filters is incorrect.");
        const std::vector<float> flat = flatten(taps);
        grb200::check(grb200_filterbank_vcvcf_set_taps(d_h, (int)taps.size(), (int)taps[0].size(), flat.data()));
        d_taps = taps;
        set_history(taps[0].size() + 1);
        d_updated = true;
    }
    void print_taps()
    {
        for (size_t i = 0; i < d_taps.size(); i++) {
            printf("filter[%d]: [", (int)i);
            for (float v : d_taps[i])
                printf(" %.4e", v);
            printf("]\n\n");
        }
    }
    std::vector<std::vector<float>> taps() const { return d_taps; }

    int general_work(int noutput_items, gr_vector_int&, gr_vector_const_void_star& input_items,
                     gr_vector_void_star& output_items) override
    {
        std::lock_guard<std::mutex> l(d_setlock); // :59
        if (d_updated) {                           // :64-67
            d_updated = false;
            return 0; // history requirements may have changed
        }
        grb200::check(grb200_filterbank_vcvcf_work(d_h, input_items[0], output_items[0], noutput_items));
        consume_each(noutput_items); // :94
        return noutput_items;
    }
    grb200_fbank* handle() { return d_h; }

private:
    static std::vector<float> flatten(const std::vector<std::vector<float>>& taps)
    {
        if (taps.empty())
            throw std::invalid_argument("The taps vector may not be empty."); // filterbank.cc:30-32
        const size_t nt = taps[0].size();
        std::vector<float> flat;
        flat.reserve(taps.size() * nt);
        for (const std::vector<float>& row : taps) {
            if (row.size() != nt)
                throw std::runtime_error("All sets of taps must be of the same length."); // :60-64
            flat.insert(flat.end(), row.begin(), row.end());
        }
        return flat;
    }
    explicit filterbank_vcvcf(const std::vector<std::vector<float>>& taps)
        : block("filterbank_vcvcf", grb200::rt::io_signature::make(1, 1, sizeof(gr_complex) * taps.size()),
                grb200::rt::io_signature::make(1, 1, sizeof(gr_complex) * taps.size())),
          d_taps(taps)
    {
        const std::vector<float> flat = flatten(taps);
        grb200::check(grb200_filterbank_vcvcf_create((int)taps.size(), (int)taps[0].size(), flat.data(),
                                                     grb200::default_device(), &d_h));
        set_history(taps[0].size() + 1); // filterbank_vcvcf_impl.cc:34
    }
    grb200_fbank* d_h = nullptr;
    std::vector<std::vector<float>> d_taps;
    bool d_updated = false;
};

} // namespace filter

namespace fft {

// ---------------------------------------------------------------------------------------------------
// fft_vcc
// ---------------------------------------------------------------------------------------------------
class fft_vcc : public grb200::rt::sync_block
{
public:
    typedef std::shared_ptr<fft_vcc> sptr;
    static sptr make(int fft_size, bool forward, const std::vector<float>& window, bool shift = false,
                     int nthreads = 1)
    {
        return grb200::make_sptr(new fft_vcc(fft_size, forward, window, shift, nthreads));
    }
    ~fft_vcc() override { grb200_fft_vcc_destroy(d_h); }
    void set_nthreads(int n) { d_nthreads = n; }
    int nthreads() const { return d_nthreads; }
    bool set_window(const std::vector<float>& window) // fft_vcc_fftw.cc:54-61
    {
        if (window.empty() || window.size() == d_fft_size) {
            grb200::check(grb200_fft_vcc_set_window(d_h, window.data(), (int)window.size()));
            return true;
        }
        return false;
    }
    int work(int noutput_items, gr_vector_const_void_star& input_items, gr_vector_void_star& output_items) override
    {
        grb200::check(grb200_fft_vcc_work(d_h, input_items[0], output_items[0], noutput_items));
        return noutput_items;
    }
    grb200_fftvcc* handle() { return d_h; }

private:
    fft_vcc(int fft_size, bool forward, const std::vector<float>& window, bool shift, int nthreads)
        : sync_block("fft_vcc_fftw", grb200::rt::io_signature::make(1, 1, fft_size * sizeof(gr_complex)),
                     grb200::rt::io_signature::make(1, 1, fft_size * sizeof(gr_complex))),
          d_fft_size(fft_size > 0 ? fft_size : 0), d_nthreads(nthreads)
    {
        // wrong window length -> std::runtime_error("fft_vcc: window not the same length as fft_size"), :46-47
        grb200::check(grb200_fft_vcc_create(fft_size, forward ? 1 : 0, window.data(), (int)window.size(), shift ? 1 : 0,
                                            grb200::default_device(), &d_h));
    }
    grb200_fftvcc* d_h = nullptr;
    size_t d_fft_size;
    int d_nthreads;
};

// ---------------------------------------------------------------------------------------------------
// fft_vfc  (gr-fft/include/gnuradio/fft/fft_vfc.h, gr-fft/lib/fft_vfc_fftw.{h,cc})
// ---------------------------------------------------------------------------------------------------
class fft_vfc : public grb200::rt::sync_block
{
public:
    typedef std::shared_ptr<fft_vfc> sptr;
    static sptr make(int fft_size, bool forward, const std::vector<float>& window, int nthreads = 1)
    {
        return grb200::make_sptr(new fft_vfc(fft_size, forward, window, nthreads));
    }
    ~fft_vfc() override { grb200_fft_vcc_destroy(d_h); }
    void set_nthreads(int n) { d_nthreads = n; }
    int nthreads() const { return d_nthreads; }
    bool set_window(const std::vector<float>& window) // fft_vfc_fftw.cc:48-55
    {
        if (window.empty() || window.size() == d_fft_size) {
            grb200::check(grb200_fft_vcc_set_window(d_h, window.data(), (int)window.size()));
            return true;
        }
        return false;
    }
    int work(int noutput_items, gr_vector_const_void_star& input_items, gr_vector_void_star& output_items) override
    {
        grb200::check(grb200_fft_vcc_work(d_h, input_items[0], output_items[0], noutput_items)); // :57-92
        return noutput_items;
    }
    grb200_fftvcc* handle() { return d_h; }

private:
    fft_vfc(int fft_size, bool forward, const std::vector<float>& window, int nthreads)
        : sync_block("fft_vfc_fftw", grb200::rt::io_signature::make(1, 1, fft_size * sizeof(float)),
                     grb200::rt::io_signature::make(1, 1, fft_size * sizeof(gr_complex))),
          d_fft_size(fft_size > 0 ? fft_size : 0), d_nthreads(nthreads)
    {
        // wrong window length -> std::runtime_error("fft_vfc: window not the same length as fft_size"), :40-41
        grb200::check(grb200_fft_vfc_create(fft_size, forward ? 1 : 0, window.data(), (int)window.size(),
                                            grb200::default_device(), &d_h));
    }
    grb200_fftvcc* d_h = nullptr;
    size_t d_fft_size;
    int d_nthreads;
};

} // namespace fft

namespace blocks {

class complex_to_mag_squared : public grb200::rt::sync_block
{
public:
    typedef std::shared_ptr<complex_to_mag_squared> sptr;
    static sptr make(size_t vlen = 1) { return grb200::make_sptr(new complex_to_mag_squared(vlen)); }
    int work(int noutput_items, gr_vector_const_void_star& input_items, gr_vector_void_star& output_items) override
    {
        grb200::check(grb200_magsq_work(grb200::default_device(), input_items[0], (float*)output_items[0],
                                        (long long)noutput_items * (long long)d_vlen));
        return noutput_items;
    }

private:
    explicit complex_to_mag_squared(size_t vlen)
        : sync_block("complex_to_mag_squared", grb200::rt::io_signature::make(1, 1, sizeof(gr_complex) * vlen),
                     grb200::rt::io_signature::make(1, 1, sizeof(float) * vlen)),
          d_vlen(vlen)
    {
    }
    size_t d_vlen;
};

} // namespace blocks
} // namespace gr

#endif // GRB200_BLOCKS_HPP
