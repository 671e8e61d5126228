// staging.cuh -- K8: the gr::buffer chunk staged through pinned, double-buffered cudaMemcpyAsync.
//
// The scheduler hands work() borrowed HOST pointers into its circular buffers (reference
// gnuradio-runtime/lib/buffer.cc:101-140, block.h:176-179).  A Stager splits one work() call into chunks and
// runs  H2D(chunk c+1) | kernel(chunk c) | D2H(chunk c-1)  concurrently on three streams with three device
// slots.  Host memory that is already page-locked (grb200_host_alloc, cudaHostRegister, torch pinned
// tensors) is copied from/to directly; pageable memory goes through pinned bounce slots owned by the Stager.
#pragma once
#include <functional>
#include "common.cuh"

namespace grb {

struct Piece {
    const void* src; // H2D: host source / D2H: device source (offset within the slot is in dst/src below)
    void* dst;
    size_t bytes;
    // D2H only, lag-by-one mode: which output stream of the call this piece belongs to and where in it (bytes).  The result
    // is parked in the handle and handed to the NEXT call's output buffers, whose addresses differ from this call's.
    int ostream = 0;
    size_t ooff = 0;
    Piece(const void* s, void* d, size_t b) : src(s), dst(d), bytes(b) {}
    Piece(const void* s, void* d, size_t b, int os, size_t oo) : src(s), dst(d), bytes(b), ostream(os), ooff(oo) {}
};

struct Stager {
    int device = 0;
    cudaStream_t s_h2d = nullptr, s_exec = nullptr, s_d2h = nullptr;
    static constexpr int NSLOT = 3; // one slot in H2D, one in D2H, one of slack so neither engine waits on the host
    cudaEvent_t ev_h2d[NSLOT] = {}, ev_exec[NSLOT] = {}, ev_d2h[NSLOT] = {};
    void* pin_in[NSLOT] = {};
    void* pin_out[NSLOT] = {};
    size_t pin_in_cap = 0, pin_out_cap = 0;
    void* dev[NSLOT][3] = {}; // [slot][in, out, scratch]
    size_t dev_cap[3] = { 0, 0, 0 };
    bool ready = false;
    // slot state survives a run(): in lag-by-one mode a call returns with its last chunks still in flight
    struct Bounce {
        void* host_dst;
        const void* pin_src;
        size_t bytes;
    };
    std::vector<Bounce> bounce[NSLOT];
    bool pending[NSLOT] = {};
    long long next_chunk = 0; // chunks issued through this stager so far (slot = next_chunk % NSLOT)

    // ---- lag-by-one mode (SURVEY.md section 7 "hard parts"): work() queues this call's chunk and hands
    // back the PREVIOUS call's results, so H2D, kernel and D2H of call k overlap the scheduler's bookkeeping and call k+1
    // instead of being waited for inside call k.  Legal for a gr::block: general_work may produce fewer items than it was
    // offered and consume independently (gnuradio-runtime/include/gnuradio/block.h:160-175).
    struct LagSeg {
        int ostream;
        size_t ooff, boff, bytes;
        bool direct; // copied straight into the caller's page-locked output, nothing to deliver later
    };
    struct LagResult {
        void* pin = nullptr;
        size_t cap = 0, used = 0, need = 0; // the parking buffer is allocated when the first piece has to be parked
        cudaEvent_t ev = nullptr;
        bool pending = false;
        long long nitems = 0;
        std::vector<LagSeg> segs;
        bool pin_ok() const { return pin != nullptr && cap >= need; }
    };
    LagResult lag[2];
    int lag_cur = 0;
    bool lag_mode = false;
    LagResult* lag_active = nullptr; // set between lag_begin() and the end of run()
    // Zero-copy delivery: when the caller's output memory is page-locked and has room BEHIND the previous call's items, this
    // call's results are copied by the DMA engine straight to where the caller will look for them next time -- its output
    // pointer then is this call's pointer advanced by what this call returns, which is how a gr::buffer write pointer moves
    // (and how any caller that appends results to one array moves).  lag_direct() arms it for the coming run().
    void* const* lag_outs = nullptr;
    int lag_nouts = 0;
    size_t lag_shift = 0; // bytes of the previous call's items in front of this call's
    void lag_direct(void* const* outs, int nouts, size_t shift_bytes)
    {
        lag_outs = outs;
        lag_nouts = nouts;
        lag_shift = shift_bytes;
    }
    int lag_begin(size_t out_bytes, long long npieces, long long nitems); // reserve the parking buffer of this call
    // hand the previous call's results to `outs` (base address of every output stream); returns its item count in *nitems
    int lag_collect(void* const* outs, int nouts, long long* nitems);
    int lag_flush(void* const* outs, int nouts, long long* nitems); // the call still in flight (end of stream)
    void lag_drop();                                                // forget parked results (set_taps, stop)

    // last *_work_dev launch on a CALLER's stream: setters wait for it (and for the handle's own streams) instead of
    // stalling the whole device with cudaDeviceSynchronize
    cudaEvent_t ev_user = nullptr;
    bool user_pending = false;
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr; // work_dev side branch (halo fetch overlapped with the main kernel)

    int init(int dev_id);
    void destroy();
    ~Stager() { destroy(); } // create-failure paths delete the handle without calling destroy()
    int note_user_launch(cudaStream_t st); // record ev_user behind a device-resident launch
    int quiesce();                         // block until nothing launched through this handle is still running
    int reserve_dev(int which, size_t bytes);  // both slots
    int reserve_pin(bool in, size_t bytes);    // both slots
    static bool is_pinned(const void* p);
    int retire(int slot);

    // One chunk's description, filled by the `plan` callback:
    //   h2d : pieces {host src, device dst (inside dev[slot][0]), bytes}
    //   d2h : pieces {device src (inside dev[slot][1]), host dst, bytes}
    // `launch` enqueues the chunk's kernels on s_exec.
    struct Chunk {
        std::vector<Piece> h2d, d2h;
    };
    using PlanFn = std::function<int(long long c, int slot, Chunk& ch)>;
    using LaunchFn = std::function<int(long long c, int slot, cudaStream_t st)>;
    int run(long long nchunks, const PlanFn& plan, const LaunchFn& launch);
};

struct CopyJob {
    void* dst;
    const void* src;
    size_t bytes;
};
void par_memcpy(void* dst, const void* src, size_t bytes); // big copies sliced over a small helper-thread pool
void par_memcpy_many(const std::vector<CopyJob>& list);    // many copies (e.g. 64 stream pieces) spread over the pool
void pin_cache_invalidate();                               // after cudaHostRegister / Unregister / free of pinned memory

// One Stager per device for the handle-less entry points (complex_to_mag_squared, vector_to_streams): created on first use,
// lives for the process.  Host-staged calls through it hold shared_stager_mutex(device) for their duration.
int shared_stager(int device, Stager** out);
std::mutex& shared_stager_mutex(int device);

} // namespace grb
