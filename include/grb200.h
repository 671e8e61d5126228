/* grb200.h -- C-ABI of libgrb200: B200-native (sm_100a) kernels for the GNU Radio filtering hot path.
 *
 * This is the drop-in boundary: the re-hosted GNU Radio block classes (include/grb200/blocks.hpp, and
 * the bindings shown in INTEGRATION.md) call ONLY these functions.  Plain pointers and sizes, no CUDA
 * or torch types.  Every entry point names the reference interface it replaces (paths relative to the
 * reference tree, GNU Radio 3.9.0-git).
 *
 * Conventions
 *  - complex64 items are interleaved float pairs (gr_complex, gnuradio-runtime/include/gnuradio/gr_complex.h:15).
 *  - Every function returns 0 (GRB200_OK) or a negative grb200_status; grb200_last_error() gives the
 *    message of the calling thread's last failure.  The block wrappers turn GRB200_EINVAL into
 *    std::invalid_argument, GRB200_ERANGE into std::out_of_range and the rest into std::runtime_error,
 *    matching the exceptions of the reference blocks.
 *  - `*_work`     : HOST buffers borrowed for the duration of the call (what gr::block::general_work hands
 *                   over, gnuradio-runtime/include/gnuradio/block.h:176-179); blocks until the outputs are valid.
 *                   Inputs are staged through pinned double-buffered async copies on the handle's streams.
 *  - `*_work_dev` : DEVICE buffers (device-resident chaining mode); asynchronous on `stream`
 *                   (a cudaStream_t passed as void*; NULL = CUDA's legacy default stream, as everywhere else in
 *                   CUDA -- so a producer, work_dev and a consumer on one stream are ordered; GRB200_STREAM_OWN =
 *                   the handle's private non-blocking stream).  set_* waits for the handle's own work only.
 *  - Host staging.  A `*_work` call larger than 1 MiB is cut into chunks (at least four; 64 MiB for big calls,
 *    GRB200_CHUNK_MB) that run H2D | kernel | D2H concurrently on three streams.  Page-locked caller memory (grb200_host_alloc,
 *    grb200_host_register, cudaHostRegister, torch pinned tensors) is copied directly; pageable memory -- what a gr::buffer is
 *    -- goes through pinned bounce slots, filled by a small pool of copy threads (GRB200_COPY_THREADS, default 8).
 *    `*_work_lag` (fir_filter, pfb_channelizer): the call returns as soon as its INPUT has been taken and delivers the
 *    previous call's results, so nothing is waited for inside work() but the copy of its own input; results are one call
 *    late.  A gr::block may do this (general_work produces fewer items than offered and consumes on its own,
 *    gnuradio-runtime/include/gnuradio/block.h:160-175); INTEGRATION.md shows the block side.
 *  - One handle <-> one calling thread for work(); set_* may come from any thread (internally serialised,
 *    mirroring d_setlock / d_mutex in the reference blocks).
 *  - There is no CPU fallback: every call fails with GRB200_ECUDA when no sm_100 device is usable.
 */
#ifndef GRB200_H
#define GRB200_H

#ifdef __cplusplus
extern "C" {
#endif

/* `stream` sentinel: run on the handle's private stream (never a valid cudaStream_t) */
#define GRB200_STREAM_OWN ((void*)(~(unsigned long long)0))

typedef enum {
    GRB200_OK = 0,
    GRB200_EINVAL = -1, /* bad argument            -> std::invalid_argument */
    GRB200_ERANGE = -2, /* size out of range       -> std::out_of_range     */
    GRB200_ECUDA = -3,  /* CUDA runtime failure    -> std::runtime_error    */
    GRB200_ENOMEM = -4, /* allocation failure      -> std::runtime_error    */
    GRB200_EWINDOW = -5, /* window length mismatch -> std::runtime_error ("fft_vcc: window not the same length as fft_size") */
    GRB200_ERUNTIME = -6 /* any other std::runtime_error site of the reference blocks */
} grb200_status;

/* ---- library / device ------------------------------------------------------------------------- */
const char* grb200_last_error(void);
const char* grb200_version(void);
int grb200_device_count(void);
/* Number of kernels this library has launched so far in the process (all handles). */
unsigned long long grb200_launch_count(void);
/* Pinned host buffers for callers that want zero-bounce staging (optional). */
int grb200_host_alloc(void** ptr, unsigned long long bytes);
int grb200_host_free(void* ptr);
/* Page-lock memory the caller owns (a gr::buffer's circular buffer, a numpy array) for direct DMA; unregister before freeing. */
int grb200_host_register(void* ptr, unsigned long long bytes);
int grb200_host_unregister(void* ptr);
/* Device buffers for the device-resident mode (thin cudaMalloc/cudaFree/cudaMemcpy wrappers, so a C++
 * block can chain handles without linking the CUDA runtime itself). */
int grb200_dev_alloc(int device, void** ptr, unsigned long long bytes);
int grb200_dev_free(int device, void* ptr);
int grb200_memcpy_h2d(int device, void* dst_dev, const void* src_host, unsigned long long bytes);
int grb200_memcpy_d2h(int device, void* dst_host, const void* src_dev, unsigned long long bytes);
int grb200_dev_sync(int device);

/* ---- peer memory: the halo of a time-sharded stream without a host round trip ----------------
 * A stream sharded by time over several GPUs needs ONE exchange: every segment's history (ntaps-1 items; taps_per_filter *
 * numchans items for the channelizer) is the left neighbour's newest items.  grb200_*_set_halo_source(h, peer_tail) makes
 * *_work_dev fetch it itself, on its own stream, from the neighbour GPU's memory over NVLink -- the channelizer's streaming
 * kernel reads the rows in place with plain loads, the FIR family copies the (at most 8 KiB) region in front of in_dev first
 * (which it therefore overwrites).  The caller orders the fetch behind the neighbour's producer (an event, a flag, a barrier).
 * peer_tail: a device pointer valid in THIS process -- from grb200_ipc_open (one process per GPU: the exporter calls
 * grb200_ipc_export on its tail pointer and ships the 72-byte handle, e.g. through torch.distributed) or, inside one process,
 * the neighbour's own pointer after grb200_enable_peer. */
int grb200_ipc_export(int device, const void* dev_ptr, void* handle72);
int grb200_ipc_open(int device, const void* handle72, void** peer_ptr);
int grb200_ipc_close(int device, void* peer_ptr);
int grb200_enable_peer(int device, int peer_device);

/* ---- fir_filter_ccf / fir_filter_ccc -----------------------------------------------------------
 * Replaces gr::filter::kernel::fir_filter<gr_complex,gr_complex,float|gr_complex>
 *   ctor/set_taps  gr-filter/lib/fir_filter.cc:21-75
 *   filterN        gr-filter/lib/fir_filter.cc:101-108
 *   filterNdec     gr-filter/lib/fir_filter.cc:111-121
 * as called from fir_filter_blk_impl::work, gr-filter/lib/fir_filter_blk_impl.cc:67-90.
 * work(): `in` holds noutput*decim + ntaps-1 items, in[0] being the oldest history item
 * (history = ntaps, fir_filter_blk_impl.cc:41);  out[i] = sum_k taps[k] * in[i*decim + ntaps-1-k]. */
typedef struct grb200_fir grb200_fir;
int grb200_fir_ccf_create(int decim, const float* taps, int ntaps, int device, grb200_fir** out);
int grb200_fir_ccc_create(int decim, const float* taps_re_im, int ntaps, int device, grb200_fir** out);
/* fir_filter_fff: fir_filter_blk<float,float,float> (gr-filter/include/gnuradio/filter/fir_filter_blk.h:66)
 * over kernel::fir_filter<float,float,float>::filter (gr-filter/lib/fir_filter.cc:122-130).  Same handle type and calls;
 * items are float32 (in: noutput*decim + ntaps-1 floats, out: noutput floats). */
int grb200_fir_fff_create(int decim, const float* taps, int ntaps, int device, grb200_fir** out);
int grb200_fir_set_taps(grb200_fir* h, const float* taps, int ntaps); /* re/im pairs for ccc */
int grb200_fir_set_halo_source(grb200_fir* h, const void* peer_tail); /* NULL clears; see "peer memory" above */
int grb200_fir_ntaps(const grb200_fir* h);
int grb200_fir_decimation(const grb200_fir* h);
int grb200_fir_work(grb200_fir* h, const void* in_host, void* out_host, long long noutput);
int grb200_fir_work_dev(grb200_fir* h, const void* in_dev, void* out_dev, long long noutput, void* stream);
/* Lag-by-one host mode (see "Host staging" above): queues this call's `noutput` items and delivers the PREVIOUS call's results
 * at out_host (room for out_capacity items; GRB200_EINVAL when more are waiting); *produced = how many.  The first call
 * produces 0.  flush() hands out the call still in flight (end of stream; *produced = 0 when nothing is parked).
 * Pointer contract: the out_host of call k+1 must be the out_host of call k advanced by what call k returned -- how a
 * gr::buffer's write pointer moves.  When out_host is page-locked and out_capacity leaves room behind the items being
 * returned, the NEW results are DMA'd straight to where the next call will look (zero CPU copies); otherwise they are parked
 * in the handle and copied out by the next call. */
int grb200_fir_work_lag(grb200_fir* h, const void* in_host, long long noutput, void* out_host, long long out_capacity,
                        long long* produced);
int grb200_fir_flush(grb200_fir* h, void* out_host, long long out_capacity, long long* produced);
void grb200_fir_destroy(grb200_fir* h);

/* ---- freq_xlating_fir_filter_ccf / freq_xlating_fir_filter_ccc ----------------------------------
 * Replaces gr::filter::freq_xlating_fir_filter_impl<gr_complex,gr_complex,float|gr_complex>
 *   build_composite_fir  gr-filter/lib/freq_xlating_fir_filter_impl.cc:74-93
 *       ctaps[i] = taps[i] * exp(j i fwT0), float fwT0 = 2 pi center_freq / sampling_freq;
 *       rotator increment exp(-j fwT0 decimation)
 *   work                 gr-filter/lib/freq_xlating_fir_filter_impl.cc:143-180
 *       out[i] = composite_fir.filter(&in[i*decimation]);  rotator.rotateN(out, out, noutput)
 *   rotator              gr-blocks/include/gnuradio/blocks/rotator.h:33-53
 * Same buffer convention as fir_filter (history = ntaps).  set_taps / set_center_freq rebuild the composite taps at
 * once; the rotator phase carries on across calls and across rebuilds as in the reference.  The block mirrors above
 * this ABI keep the reference's "return 0 from the work() that sees d_updated" behaviour. */
typedef struct grb200_xlat grb200_xlat;
int grb200_freq_xlating_fir_ccf_create(int decim, const float* taps, int ntaps, double center_freq, double sampling_freq,
                                       int device, grb200_xlat** out);
int grb200_freq_xlating_fir_ccc_create(int decim, const float* taps_re_im, int ntaps, double center_freq,
                                       double sampling_freq, int device, grb200_xlat** out);
int grb200_freq_xlating_fir_set_taps(grb200_xlat* h, const float* taps, int ntaps); /* re/im pairs for ccc */
int grb200_freq_xlating_fir_set_center_freq(grb200_xlat* h, double center_freq);
/* Moves the rotator on as if `noutput` items had been produced (no data touched): a worker that takes over a stream at
 * output item n0 -- time-segment sharding across GPUs, DESIGN.md section 6 -- calls this once with n0 (negative: back). */
int grb200_freq_xlating_fir_advance(grb200_xlat* h, long long noutput);
double grb200_freq_xlating_fir_center_freq(const grb200_xlat* h);
int grb200_freq_xlating_fir_ntaps(const grb200_xlat* h);
int grb200_freq_xlating_fir_decimation(const grb200_xlat* h);
int grb200_freq_xlating_fir_composite_taps(const grb200_xlat* h, float* taps_re_im_out, int capacity);
int grb200_freq_xlating_fir_work(grb200_xlat* h, const void* in_host, void* out_host, long long noutput);
int grb200_freq_xlating_fir_work_dev(grb200_xlat* h, const void* in_dev, void* out_dev, long long noutput, void* stream);
void grb200_freq_xlating_fir_destroy(grb200_xlat* h);

/* ---- fft_filter_ccc ----------------------------------------------------------------------------
 * Replaces gr::filter::kernel::fft_filter_ccc
 *   compute_sizes  gr-filter/lib/fft_filter.cc:204-224   (fftsize = 2*2^ceil(log2 ntaps), nsamples = fftsize-ntaps+1)
 *   set_taps       gr-filter/lib/fft_filter.cc:171-201   (zeroes the overlap state)
 *   filter         gr-filter/lib/fft_filter.cc:241-285
 * as called from fft_filter_ccc_impl::work, gr-filter/lib/fft_filter_ccc_impl.cc:61-78.
 * work(): consumes noutput*decim input items (history 1) and carries the ntaps-1 item overlap between
 * calls like the reference's d_tail; noutput must be a multiple of nsamples() (the block's output_multiple).
 * The device computes the same linear convolution by overlap-save; the FFT length used on the device may be
 * larger than the reference's (parity is on outputs, not on block structure). */
typedef struct grb200_fftfilt grb200_fftfilt;
int grb200_fftfilt_ccc_create(int decim, const float* taps_re_im, int ntaps, int device, grb200_fftfilt** out);
int grb200_fftfilt_set_taps(grb200_fftfilt* h, const float* taps_re_im, int ntaps);
/* fft_filter_ccf: kernel::fft_filter_ccf (gr-filter/lib/fft_filter.cc:291-419) and the block shell
 * gr-filter/lib/fft_filter_ccf_impl.cc -- real taps, complex stream.  The reference transforms the taps as (t, 0) complex
 * numbers (fft_filter.cc:317-320) and is otherwise the ccc kernel, so both constructors give the same handle type; every
 * other grb200_fftfilt_* call applies unchanged. */
int grb200_fftfilt_ccf_create(int decim, const float* taps, int ntaps, int device, grb200_fftfilt** out);
int grb200_fftfilt_set_taps_real(grb200_fftfilt* h, const float* taps, int ntaps);
int grb200_fftfilt_nsamples(const grb200_fftfilt* h);
int grb200_fftfilt_ntaps(const grb200_fftfilt* h);
int grb200_fftfilt_work(grb200_fftfilt* h, const void* in_host, void* out_host, long long noutput);
/* Stateless device form: in_dev[0] is the oldest of ntaps-1 history items (caller supplies the halo),
 * noutput*decim + ntaps-1 items readable. */
int grb200_fftfilt_work_dev(grb200_fftfilt* h, const void* in_dev, void* out_dev, long long noutput, void* stream);
void grb200_fftfilt_destroy(grb200_fftfilt* h);

/* ---- pfb_channelizer_ccf -----------------------------------------------------------------------
 * Replaces gr::filter::kernel::polyphase_filterbank (ctor/set_taps, gr-filter/lib/polyphase_filterbank.cc:22-81)
 * and pfb_channelizer_ccf_impl: ctor checks/LUT/output multiple gr-filter/lib/pfb_channelizer_ccf_impl.cc:49-84,
 * set_channel_map :112-124, general_work :128-193.
 * work(): in_streams[j] (j < numchans) points at consumed + taps_per_filter items of input stream j, the first
 * taps_per_filter being history (history = taps_per_filter+1, :101); out_streams[n] (n < nout_streams <= numchans)
 * receives noutput items of FFT bin channel_map[n].  noutput must be a multiple of output_multiple();
 * consumed = noutput / oversample_rate is returned through *consumed.
 * work_interleaved_dev(): the stream_to_streams step (gr-blocks/lib/stream_to_streams_impl.cc:35-54) absorbed:
 * `in_dev` is ONE interleaved stream (sample n of stream j at in_dev[n*numchans + j]) with
 * taps_per_filter*numchans history items in front; `out_dev` is channel-major [nout_streams][out_stride]. */
typedef struct grb200_pfb grb200_pfb;
int grb200_pfb_chan_ccf_create(unsigned int numchans, const float* taps, int ntaps, float oversample_rate,
                               int device, grb200_pfb** out);
int grb200_pfb_chan_set_taps(grb200_pfb* h, const float* taps, int ntaps);
int grb200_pfb_chan_set_channel_map(grb200_pfb* h, const int* map, int n);
int grb200_pfb_chan_get_channel_map(const grb200_pfb* h, int* map, int n);
int grb200_pfb_chan_set_halo_source(grb200_pfb* h, const void* peer_tail); /* for work_interleaved_dev; NULL clears */
int grb200_pfb_chan_taps_per_filter(const grb200_pfb* h);
int grb200_pfb_chan_output_multiple(const grb200_pfb* h);
/* branch taps d_taps[i][j] = padded[i + j*numchans]; `out` holds numchans*taps_per_filter floats */
int grb200_pfb_chan_branch_taps(const grb200_pfb* h, float* out);
int grb200_pfb_chan_work(grb200_pfb* h, const void* const* in_streams, void* const* out_streams,
                         int nout_streams, long long noutput, long long* consumed);
/* Host form of the same absorbed stream_to_streams step -- what the hier block pfb.channelizer_ccf
 * (gr-filter/python/filter/pfb.py:43-50) computes from its single input: `in_host` holds
 * (taps_per_filter + consumed) * numchans interleaved items, history rows first. */
int grb200_pfb_chan_work_interleaved(grb200_pfb* h, const void* in_host, void* const* out_streams,
                                     int nout_streams, long long noutput, long long* consumed);
/* Lag-by-one host forms of the two calls above (see grb200_fir_work_lag): this call's inputs are queued, the previous call's
 * outputs delivered; out_capacity = room per output stream. */
int grb200_pfb_chan_work_lag(grb200_pfb* h, const void* const* in_streams, long long noutput, void* const* out_streams,
                             int nout_streams, long long out_capacity, long long* consumed, long long* produced);
int grb200_pfb_chan_work_interleaved_lag(grb200_pfb* h, const void* in_host, long long noutput, void* const* out_streams,
                                         int nout_streams, long long out_capacity, long long* consumed, long long* produced);
int grb200_pfb_chan_flush(grb200_pfb* h, void* const* out_streams, int nout_streams, long long out_capacity,
                          long long* produced);
int grb200_pfb_chan_work_interleaved_dev(grb200_pfb* h, const void* in_dev, void* out_dev,
                                         long long out_stride, int nout_streams, long long noutput,
                                         void* stream);
void grb200_pfb_chan_destroy(grb200_pfb* h);

/* ---- pfb_decimator_ccf ---------------------------------------------------------------------------
 * Replaces gr::filter::pfb_decimator_ccf_impl::work and its four variants work_fir_exp / work_fir_fft /
 * work_fft_exp / work_fft_fft (gr-filter/lib/pfb_decimator_ccf_impl.cc:107-257) over
 * kernel::polyphase_filterbank (gr-filter/lib/polyphase_filterbank.cc:22-81):
 *     out[i] = sum_j rot[j] * (branch filter j over input stream M-1-j)[i],  rot[j] = e^{+j 2 pi j channel / M}.
 * The variants differ only in rounding (table vs M-point FFT, FIR vs FFT branch filters); one device kernel serves all.
 * in_streams[s] (s < decim): the block's input stream s.  carry_state = 0: noutput + taps_per_filter-1 items each,
 * history first (set_history(taps_per_filter), :61).  carry_state = 1: noutput items each, the overlap is kept in the
 * handle like the fft_filter kernels keep theirs (set_history(1), :57-59).
 * work_dev: x_dev holds the streams stream-major, stream s at x_dev + s*stride items, history first. */
typedef struct grb200_pfbdec grb200_pfbdec;
int grb200_pfb_dec_ccf_create(unsigned int decim, const float* taps, int ntaps, unsigned int channel, int device,
                              grb200_pfbdec** out);
int grb200_pfb_dec_set_taps(grb200_pfbdec* h, const float* taps, int ntaps);
int grb200_pfb_dec_set_channel(grb200_pfbdec* h, unsigned int channel);
int grb200_pfb_dec_taps_per_filter(const grb200_pfbdec* h);
int grb200_pfb_dec_decimation(const grb200_pfbdec* h);
int grb200_pfb_dec_work(grb200_pfbdec* h, const void* const* in_streams, void* out_host, long long noutput,
                        int carry_state);
int grb200_pfb_dec_work_dev(grb200_pfbdec* h, const void* x_dev, long long stride, void* out_dev, long long noutput,
                            void* stream);
void grb200_pfb_dec_destroy(grb200_pfbdec* h);

/* ---- pfb_synthesizer_ccf ---------------------------------------------
 * Replaces pfb_synthesizer_ccf_impl: ctor gr-filter/lib/pfb_synthesizer_ccf_impl.cc:31-69, set_taps/set_taps1/set_taps2
 * :77-184, set_channel_map :202-217, work :221-284 (over kernel::fir_filter_with_buffer_ccf delay lines and a
 * twox*numchans-point backward fft_complex).  sync_interpolator by numchans: every input item of the (up to twox*numchans)
 * input streams makes numchans output items;  out[n*numchans + i] = sum_j taps[i + j*numchans] V_q[n-j],
 * V = IFFT over the channel-mapped inputs, q = i (1x) or the half of the 2*numchans-point IFFT that alternates with n (2x).
 * work(): in_streams[i] holds noutput/numchans NEW items (history 1, :92); the filters' delay lines are kept in the handle
 * (as the last taps_per_filter-1 items of every stream) and zeroed by set_taps like the reference's.  noutput must be a
 * multiple of numchans (set_output_multiple, :68).  work_dev() is stateless: stream i at x_dev + i*stride items with
 * taps_per_filter-1 history items in front; `parity` = d_state at the first output (2x only).
 * The block mirrors keep the reference's "return 0 from the work() that sees d_updated". */
typedef struct grb200_pfbsynth grb200_pfbsynth;
int grb200_pfb_synth_ccf_create(unsigned int numchans, const float* taps, int ntaps, int twox, int device,
                                grb200_pfbsynth** out);
int grb200_pfb_synth_set_taps(grb200_pfbsynth* h, const float* taps, int ntaps);
int grb200_pfb_synth_set_channel_map(grb200_pfbsynth* h, const int* map, int n);
int grb200_pfb_synth_get_channel_map(const grb200_pfbsynth* h, int* map, int n); /* returns the map length */
int grb200_pfb_synth_taps_per_filter(const grb200_pfbsynth* h);
int grb200_pfb_synth_numchans(const grb200_pfbsynth* h);
int grb200_pfb_synth_twox(const grb200_pfbsynth* h);
/* d_taps as taps() returns them: twox*numchans rows of taps_per_filter floats */
int grb200_pfb_synth_branch_taps(const grb200_pfbsynth* h, float* out);
int grb200_pfb_synth_work(grb200_pfbsynth* h, const void* const* in_streams, int nin, void* out_host, long long noutput);
int grb200_pfb_synth_work_dev(grb200_pfbsynth* h, const void* x_dev, long long stride, int nin, void* out_dev,
                              long long noutput, int parity, void* stream);
void grb200_pfb_synth_destroy(grb200_pfbsynth* h);

/* ---- filterbank_vcvcf ------------------------------------------------
 * Replaces gr::filter::kernel::filterbank (ctor/set_taps, gr-filter/lib/filterbank.cc:26-79) and
 * filterbank_vcvcf_impl::general_work (gr-filter/lib/filterbank_vcvcf_impl.cc:54-96): items are vectors of nfilts complex64,
 *     out[j][i] = sum_k taps[i][k] * in[j + ntaps-1-k][i];
 * `in` holds noutput + ntaps vectors, the first ntaps being history (set_history(ntaps + 1), :34 -- the newest vector of a
 * window is never read: the block delays by one item); a filter whose taps are all zero writes zeros (:83-88).
 * taps: [nfilts][ntaps] row-major.  set_taps with a different nfilts returns GRB200_ERUNTIME (std::runtime_error,
 * filterbank.cc:54-56); rows of unequal length are refused by the block mirrors above this ABI (:60-64). */
typedef struct grb200_fbank grb200_fbank;
int grb200_filterbank_vcvcf_create(int nfilts, int ntaps, const float* taps, int device, grb200_fbank** out);
int grb200_filterbank_vcvcf_set_taps(grb200_fbank* h, int nfilts, int ntaps, const float* taps);
int grb200_filterbank_vcvcf_nfilts(const grb200_fbank* h);
int grb200_filterbank_vcvcf_ntaps(const grb200_fbank* h);
int grb200_filterbank_vcvcf_work(grb200_fbank* h, const void* in_host, void* out_host, long long noutput);
int grb200_filterbank_vcvcf_work_dev(grb200_fbank* h, const void* in_dev, void* out_dev, long long noutput, void* stream);
void grb200_filterbank_vcvcf_destroy(grb200_fbank* h);

/* ---- interp_fir_filter_ccf / interp_fir_filter_ccc -------------------------------------------------
 * Replaces gr::filter::interp_fir_filter_impl<gr_complex,gr_complex,float|gr_complex>
 *   set_taps / install_taps  gr-filter/lib/interp_fir_filter_impl.cc:72-117 (zero-pad to a multiple of the
 *                            interpolation, xtaps[n][k] = taps[n + k*interpolation], history = taps per phase)
 *   work                     gr-filter/lib/interp_fir_filter_impl.cc:128-152
 *       out[i*interpolation + n] = firs[n].filter(&in[i])
 * in: noutput/interpolation + taps_per_filter-1 items, history first; noutput must be a multiple of the interpolation
 * (sync_interpolator).  A zero interpolation returns GRB200_ERANGE (std::out_of_range in the reference, :46-48). */
typedef struct grb200_interp grb200_interp;
int grb200_interp_fir_ccf_create(unsigned int interpolation, const float* taps, int ntaps, int device, grb200_interp** out);
int grb200_interp_fir_ccc_create(unsigned int interpolation, const float* taps_re_im, int ntaps, int device,
                                 grb200_interp** out);
int grb200_interp_fir_set_taps(grb200_interp* h, const float* taps, int ntaps); /* re/im pairs for ccc */
int grb200_interp_fir_taps_per_filter(const grb200_interp* h);
int grb200_interp_fir_interpolation(const grb200_interp* h);
int grb200_interp_fir_work(grb200_interp* h, const void* in_host, void* out_host, long long noutput);
int grb200_interp_fir_work_dev(grb200_interp* h, const void* in_dev, void* out_dev, long long noutput, void* stream);
void grb200_interp_fir_destroy(grb200_interp* h);

/* ---- rational_resampler_base_ccf / rational_resampler_base_ccc ------------------------------------
 * Replaces gr::filter::rational_resampler_base_impl<gr_complex,gr_complex,float|gr_complex>
 *   set_taps / install_taps  gr-filter/lib/rational_resampler_base_impl.cc:77-117 (zero-pad to a multiple of the
 *                            interpolation, xtaps[n][k] = taps[n + k*interpolation])
 *   general_work             gr-filter/lib/rational_resampler_base_impl.cc:141-173
 *       while (i < noutput && count < ninput) { out[i++] = firs[ctr].filter(in); ctr += decimation;
 *                                               while (ctr >= interpolation) { ctr -= interpolation; in++; count++; } }
 * The reference class shadows history() (rational_resampler_base_impl.h:48-49), so the scheduler puts NO zero history in
 * front of the stream: the first output is the first full window.  The loop tests count < ninput only and filter() reads
 * taps_per_filter items from there, so `in` must hold nin + taps_per_filter-1 readable items (forecast :128-138 asks the
 * scheduler for that much).  *consumed is what general_work passes to consume_each(), *produced what it returns; the
 * phase counter d_ctr is carried in the handle (ctr / set_ctr expose it so a time-sharded stream can hand it over).
 * Zero interpolation or decimation returns GRB200_ERANGE (std::out_of_range in the reference, :46-51). */
typedef struct grb200_resamp grb200_resamp;
int grb200_rational_resampler_ccf_create(unsigned int interpolation, unsigned int decimation, const float* taps, int ntaps,
                                         int device, grb200_resamp** out);
int grb200_rational_resampler_ccc_create(unsigned int interpolation, unsigned int decimation, const float* taps_re_im,
                                         int ntaps, int device, grb200_resamp** out);
int grb200_rational_resampler_set_taps(grb200_resamp* h, const float* taps, int ntaps); /* re/im pairs for ccc */
int grb200_rational_resampler_taps_per_filter(const grb200_resamp* h);
unsigned int grb200_rational_resampler_interpolation(const grb200_resamp* h);
unsigned int grb200_rational_resampler_decimation(const grb200_resamp* h);
unsigned int grb200_rational_resampler_ctr(const grb200_resamp* h);
int grb200_rational_resampler_set_ctr(grb200_resamp* h, unsigned int ctr);
int grb200_rational_resampler_work(grb200_resamp* h, const void* in_host, long long nin, void* out_host, long long noutput,
                                   long long* consumed, long long* produced);
int grb200_rational_resampler_work_dev(grb200_resamp* h, const void* in_dev, long long nin, void* out_dev,
                                       long long noutput, long long* consumed, long long* produced, void* stream);
void grb200_rational_resampler_destroy(grb200_resamp* h);

/* ---- fft_vcc (+ complex_to_mag_squared) --------------------------------------------------------
 * Replaces gr::fft::fft_complex (gr-fft/lib/fft.cc:151-203) and fft_vcc_fftw: ctor/set_window
 * gr-fft/lib/fft_vcc_fftw.cc:33-61, work :63-122; and complex_to_mag_squared_impl::work
 * (gr-blocks/lib/complex_to_mag_squared_impl.cc:37-48), standalone or fused behind the FFT.
 * Items are vectors of fft_size complex64.  Unnormalised; forward e^{-j..}, reverse e^{+j..}. */
typedef struct grb200_fftvcc grb200_fftvcc;
int grb200_fft_vcc_create(int fft_size, int forward, const float* window, int nwindow, int shift,
                          int device, grb200_fftvcc** out);
int grb200_fft_vcc_set_window(grb200_fftvcc* h, const float* window, int nwindow); /* GRB200_EWINDOW if bad */
int grb200_fft_vcc_fft_size(const grb200_fftvcc* h);
int grb200_fft_vcc_work(grb200_fftvcc* h, const void* in_host, void* out_host, long long nvectors);
int grb200_fft_vcc_work_dev(grb200_fftvcc* h, const void* in_dev, void* out_dev, long long nvectors, void* stream);
/* fused fft_vcc -> complex_to_mag_squared(vlen=fft_size): out is float32 */
int grb200_fft_vcc_work_magsq(grb200_fftvcc* h, const void* in_host, float* out_host, long long nvectors);
int grb200_fft_vcc_work_magsq_dev(grb200_fftvcc* h, const void* in_dev, float* out_dev, long long nvectors, void* stream);
void grb200_fft_vcc_destroy(grb200_fftvcc* h);

/* ---- fft_vfc -------------------------------------------------------
 * Replaces fft_vfc_fftw: ctor/set_window gr-fft/lib/fft_vfc_fftw.cc:30-55, work :57-92 -- float32 vectors in, the full
 * fft_size-point complex transform out (no shift option).  Same handle type as fft_vcc: set_window, fft_size, work,
 * work_dev and destroy are the grb200_fft_vcc_* calls; `in` holds fft_size floats per vector. */
int grb200_fft_vfc_create(int fft_size, int forward, const float* window, int nwindow, int device, grb200_fftvcc** out);

/* ---- logpwrfft_c / logpwrfft_f -------------------------------------
 * Replaces the hier block gr-fft/python/fft/logpwrfft.py:29-64 as one handle:
 *   stream_to_vector_decimator (gr-blocks/python/blocks/stream_to_vector_decimator.py:15-96: stream_to_vector + keep_one_in_n,
 *   gr-blocks/lib/keep_one_in_n_impl.cc:56-77 -- the LAST vector of every group of `decimation` is kept)
 *   -> fft_vcc | fft_vfc (forward, window, no shift) -> complex_to_mag_squared
 *   -> single_pole_iir_filter_ff (gr-filter/lib/single_pole_iir_filter_ff_impl.cc:47-67; alpha = avg_alpha when averaging, else 1)
 *   -> nlog10_ff(10, fft_size, -20log10(fft_size) - 10log10(sum(w^2)/fft_size) - 20log10(ref_scale/2)) (gr-blocks/lib/nlog10_ff_impl.cc).
 * `window` must hold fft_size floats (the Python block always passes one; default window.blackmanharris).
 * work: `in` holds nitems stream items (complex64 or float32); whole vectors only are consumed; at most max_out frames of
 * fft_size floats are written.  *consumed = stream items used, *produced = frames written.  The decimator's counter and the
 * filter memory live in the handle like in the reference blocks.  Setters mirror the Python methods :66-113. */
typedef struct grb200_logpwr grb200_logpwr;
int grb200_logpwrfft_create(int complex_input, double sample_rate, int fft_size, double ref_scale, double frame_rate,
                            double avg_alpha, int average, const float* window, int nwindow, int device, grb200_logpwr** out);
int grb200_logpwrfft_set_decimation(grb200_logpwr* h, double decim);
int grb200_logpwrfft_set_vec_rate(grb200_logpwr* h, double vec_rate);
int grb200_logpwrfft_set_sample_rate(grb200_logpwr* h, double sample_rate);
int grb200_logpwrfft_set_average(grb200_logpwr* h, int average);
int grb200_logpwrfft_set_avg_alpha(grb200_logpwr* h, double avg_alpha);
int grb200_logpwrfft_decimation(const grb200_logpwr* h);
double grb200_logpwrfft_sample_rate(const grb200_logpwr* h);
double grb200_logpwrfft_frame_rate(const grb200_logpwr* h);
int grb200_logpwrfft_average(const grb200_logpwr* h);
double grb200_logpwrfft_avg_alpha(const grb200_logpwr* h);
int grb200_logpwrfft_fft_size(const grb200_logpwr* h);
int grb200_logpwrfft_work(grb200_logpwr* h, const void* in_host, long long nitems, float* out_host, long long max_out,
                          long long* consumed, long long* produced);
int grb200_logpwrfft_work_dev(grb200_logpwr* h, const void* in_dev, long long nitems, float* out_dev, long long max_out,
                              long long* consumed, long long* produced, void* stream);
void grb200_logpwrfft_destroy(grb200_logpwr* h);
/* grb200_logpwrfft_work_dev that ALLOCATES NOTHING.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  grb200_logpwrfft_work_dev grows a per-handle scratch buffer (the averaging filter's run carries) with
 * cudaMalloc on demand; this variant keeps that scratch in a file-scope `static __device__` array of
 * GRB200_LOGPWR_STATIC_SCRATCH_BYTES inside the library image -- every GPU that loads the library holds it until the process
 * exits -- and calls no cudaMalloc / cudaFree / allocating routine.  Arguments, results, handle state (decimator counter,
 * filter memory) and output bits are those of grb200_logpwrfft_work_dev; the two may be mixed on one handle.
 * Limit: with averaging on, a call may keep at most 8 * (GRB200_LOGPWR_STATIC_SCRATCH_BYTES / 4 / fft_size) frames (2048
 * frames of 1024 bins; 128 of 16384); beyond that the call returns GRB200_ERANGE before anything is launched and the handle is
 * unchanged -- split the call.  Calls of <= 16 kept frames and handles without averaging use no scratch at all.
 * The array is ONE per GPU, shared by every handle and every call: calls that overlap on the GPU (different streams) race
 * on it -- serialise them.  It keeps the last call's contents and needs no clearing (every word is written before it is
 * read).  Its device address is resolved (cudaGetSymbolAddress, once per device) by the first averaging call on that
 * device: make one such call before starting a graph capture. */
#define GRB200_LOGPWR_STATIC_SCRATCH_BYTES (1 << 20)
int grb200_logpwrfft_work_dev_static(grb200_logpwr* h, const void* in_dev, long long nitems, float* out_dev,
                                     long long max_out, long long* consumed, long long* produced, void* stream);

/* ---- stream_to_vector / vector_to_stream / vector_to_streams --------
 * stream_to_vector_impl::work (gr-blocks/lib/stream_to_vector_impl.cc:46-58) is memcpy(out, in, noutput*block_size): a stream
 * of items and the stream of vectors over it are THE SAME BYTES.  In the device-resident mode the block therefore is a pointer
 * reinterpretation: hand the upstream block's device buffer to grb200_fft_vcc_work_dev (etc.) as it is -- no entry point, no
 * kernel, no copy.  vector_to_stream (vector_to_stream_impl.cc) is the same statement read backwards.
 * vector_to_streams_impl::work (gr-blocks/lib/vector_to_streams_impl.cc:34-55): out_j[n] = in[n*nstreams + j] -- a transpose;
 * streams_to_vector is its mirror image (and, with itemsize 8, the stream_to_streams step the channelizer absorbs).
 * itemsize: 4 (float) or 8 (gr_complex).  Device forms: stream j at base + j*stride items. */
int grb200_vector_to_streams_work(int device, int itemsize, int nstreams, const void* in_host, void* const* out_streams,
                                  long long noutput);
int grb200_vector_to_streams_work_dev(int device, int itemsize, int nstreams, const void* in_dev, void* out_dev,
                                      long long out_stride, long long noutput, void* stream);
int grb200_streams_to_vector_work_dev(int device, int itemsize, int nstreams, const void* in_dev, long long in_stride,
                                      void* out_dev, long long noutput, void* stream);

int grb200_magsq_work(int device, const void* in_host, float* out_host, long long nitems);
int grb200_magsq_work_dev(int device, const void* in_dev, float* out_dev, long long nitems, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* GRB200_H */
