// staging.cu -- library globals, device selection and the pinned double-buffered staging pipeline (K8).
#include <algorithm>
#include <condition_variable>
#include <cstdint>
#include <thread>
#include "staging.cuh"

namespace grb {

thread_local char g_err[512] = "";
std::atomic<unsigned long long> g_launches{ 0 };

static std::mutex g_dev_mutex;
static int g_dev_checked[64];
static int g_dev_sms[64];

int use_device(int device)
{
    if (device < 0 || device >= 64)
        return fail(GRB200_EINVAL, "invalid device index %d", device);
    GRB_CUDA(cudaSetDevice(device));
    std::lock_guard<std::mutex> lk(g_dev_mutex);
    if (!g_dev_checked[device]) {
        cudaDeviceProp prop;
        GRB_CUDA(cudaGetDeviceProperties(&prop, device));
        if (prop.major != 10)
            return fail(GRB200_ECUDA,
                        "device %d is sm_%d%d; libgrb200 is built for sm_100a (B200) only and has no fallback",
                        device, prop.major, prop.minor);
        g_dev_sms[device] = prop.multiProcessorCount;
        g_dev_checked[device] = 1;
    }
    return 0;
}

int sm_count(int device) { return (device >= 0 && device < 64 && g_dev_checked[device]) ? g_dev_sms[device] : 148; }

// ---- Stager ----------------------------------------------------------------------------------------
int Stager::init(int dev_id)
{
    if (ready)
        return 0;
    device = dev_id;
    int rc = use_device(device);
    if (rc)
        return rc;
    GRB_CUDA(cudaStreamCreateWithFlags(&s_h2d, cudaStreamNonBlocking));
    GRB_CUDA(cudaStreamCreateWithFlags(&s_exec, cudaStreamNonBlocking));
    GRB_CUDA(cudaStreamCreateWithFlags(&s_d2h, cudaStreamNonBlocking));
    for (int i = 0; i < NSLOT; i++) {
        GRB_CUDA(cudaEventCreateWithFlags(&ev_h2d[i], cudaEventDisableTiming));
        GRB_CUDA(cudaEventCreateWithFlags(&ev_exec[i], cudaEventDisableTiming));
        GRB_CUDA(cudaEventCreateWithFlags(&ev_d2h[i], cudaEventDisableTiming));
    }
    GRB_CUDA(cudaEventCreateWithFlags(&ev_user, cudaEventDisableTiming));
    GRB_CUDA(cudaEventCreateWithFlags(&ev_fork, cudaEventDisableTiming));
    GRB_CUDA(cudaEventCreateWithFlags(&ev_join, cudaEventDisableTiming));
    ready = true;
    return 0;
}

int Stager::note_user_launch(cudaStream_t st)
{
    if (st == s_exec)
        return 0;
    GRB_CUDA(cudaEventRecord(ev_user, st));
    user_pending = true;
    return 0;
}

int Stager::quiesce()
{
    if (!ready)
        return 0;
    GRB_CUDA(cudaStreamSynchronize(s_h2d));
    GRB_CUDA(cudaStreamSynchronize(s_exec));
    GRB_CUDA(cudaStreamSynchronize(s_d2h));
    if (user_pending) {
        GRB_CUDA(cudaEventSynchronize(ev_user));
        user_pending = false;
    }
    return 0;
}

void Stager::destroy()
{
    if (!ready)
        return;
    DeviceScope ds(device);
    cudaStreamSynchronize(s_h2d);
    cudaStreamSynchronize(s_exec);
    cudaStreamSynchronize(s_d2h);
    for (int i = 0; i < NSLOT; i++) {
        cudaEventDestroy(ev_h2d[i]);
        cudaEventDestroy(ev_exec[i]);
        cudaEventDestroy(ev_d2h[i]);
        if (pin_in[i])
            cudaFreeHost(pin_in[i]);
        if (pin_out[i])
            cudaFreeHost(pin_out[i]);
        for (int k = 0; k < 3; k++)
            if (dev[i][k])
                cudaFree(dev[i][k]);
    }
    if (user_pending)
        cudaEventSynchronize(ev_user);
    cudaEventDestroy(ev_user);
    cudaEventDestroy(ev_fork);
    cudaEventDestroy(ev_join);
    for (auto& r : lag) {
        if (r.pin)
            cudaFreeHost(r.pin);
        if (r.ev)
            cudaEventDestroy(r.ev);
        r = LagResult();
    }
    for (int i = 0; i < NSLOT; i++) {
        pending[i] = false;
        bounce[i].clear();
    }
    cudaStreamDestroy(s_h2d);
    cudaStreamDestroy(s_exec);
    cudaStreamDestroy(s_d2h);
    for (int i = 0; i < NSLOT; i++) {
        pin_in[i] = pin_out[i] = nullptr;
        for (int k = 0; k < 3; k++)
            dev[i][k] = nullptr;
    }
    pin_in_cap = pin_out_cap = 0;
    dev_cap[0] = dev_cap[1] = dev_cap[2] = 0;
    ready = false;
}

int Stager::reserve_dev(int which, size_t bytes)
{
    if (bytes <= dev_cap[which])
        return 0;
    size_t cap = bytes + bytes / 8 + 256;
    for (int i = 0; i < NSLOT; i++) {
        if (dev[i][which]) {
            GRB_CUDA(cudaStreamSynchronize(s_exec));
            GRB_CUDA(cudaFree(dev[i][which]));
            dev[i][which] = nullptr;
        }
        cudaError_t e = cudaMalloc(&dev[i][which], cap);
        if (e != cudaSuccess) {
            dev_cap[which] = 0;
            return fail(GRB200_ENOMEM, "cudaMalloc(%zu) failed: %s", cap, cudaGetErrorString(e));
        }
    }
    dev_cap[which] = cap;
    return 0;
}

int Stager::reserve_pin(bool in, size_t bytes)
{
    size_t& cap = in ? pin_in_cap : pin_out_cap;
    void** p = in ? pin_in : pin_out;
    if (bytes <= cap)
        return 0;
    size_t ncap = bytes + bytes / 8 + 256;
    for (int i = 0; i < NSLOT; i++) {
        if (p[i]) {
            GRB_CUDA(cudaFreeHost(p[i]));
            p[i] = nullptr;
        }
        cudaError_t e = cudaHostAlloc(&p[i], ncap, cudaHostAllocDefault);
        if (e != cudaSuccess) {
            cap = 0;
            return fail(GRB200_ENOMEM, "cudaHostAlloc(%zu) failed: %s", ncap, cudaGetErrorString(e));
        }
    }
    cap = ncap;
    return 0;
}

// Is this host address page-locked (cudaHostAlloc / cudaHostRegister / torch pinned)?  cudaPointerGetAttributes costs about a
// microsecond; a scheduler-sized work() call asks 2..128 times, always about the same few buffers.  The answer is cached per
// 64 KiB granule in a small direct-mapped table.  A stale entry is harmless: "pinned" for memory that no longer is just
// makes cudaMemcpyAsync stage it itself, "pageable" for memory that now is pinned takes the bounce path.
static std::atomic<unsigned long long> g_pin_cache[1024]; // (granule << 2) | valid << 1 | pinned
static std::atomic<unsigned long long> g_pin_epoch{ 0 };

void pin_cache_invalidate()
{
    for (auto& e : g_pin_cache)
        e.store(0, std::memory_order_relaxed);
    g_pin_epoch.fetch_add(1);
}

bool Stager::is_pinned(const void* p)
{
    const unsigned long long gran = (unsigned long long)(uintptr_t)p >> 16;
    std::atomic<unsigned long long>& slot = g_pin_cache[(gran ^ (gran >> 10)) & 1023];
    const unsigned long long v = slot.load(std::memory_order_relaxed);
    if ((v >> 2) == gran && (v & 2))
        return (v & 1) != 0;
    cudaPointerAttributes attr;
    cudaError_t e = cudaPointerGetAttributes(&attr, p);
    bool pinned = false;
    if (e != cudaSuccess)
        cudaGetLastError();
    else
        pinned = attr.type == cudaMemoryTypeHost;
    slot.store((gran << 2) | 2 | (pinned ? 1 : 0), std::memory_order_relaxed);
    return pinned;
}

// ---- parallel memcpy for the pageable bounce ---------------------------------------------------------------------
// One thread copies 10-12 GB/s; PCIe 5 moves 50+.  Copies of a MiB or more are cut into slices for a small pool of helper
// threads (GRB200_COPY_THREADS, default 8, at most half the cores, 0 = inline), started on first use and shared by every handle of the process.
namespace {
struct CopyPool {
    struct Job {
        char* dst;
        const char* src;
        size_t bytes;
    };
    std::vector<std::thread> threads;
    std::mutex mu;
    std::condition_variable cv_work, cv_done;
    std::vector<Job> jobs;
    size_t next = 0, done = 0;
    bool quit = false;
    int n = 0;
    CopyPool()
    {
        const char* e = getenv("GRB200_COPY_THREADS");
        n = e ? atoi(e) : 8;
        unsigned hw = std::thread::hardware_concurrency();
        if (hw && (unsigned)n > hw / 2)
            n = (int)(hw / 2);
        if (n < 0)
            n = 0;
        for (int i = 0; i < n; i++)
            threads.emplace_back([this] { loop(); });
    }
    ~CopyPool()
    {
        {
            std::lock_guard<std::mutex> lk(mu);
            quit = true;
        }
        cv_work.notify_all();
        for (auto& t : threads)
            t.join();
    }
    void loop()
    {
        std::unique_lock<std::mutex> lk(mu);
        for (;;) {
            cv_work.wait(lk, [this] { return quit || next < jobs.size(); });
            if (quit)
                return;
            const Job j = jobs[next++];
            lk.unlock();
            memcpy(j.dst, j.src, j.bytes);
            lk.lock();
            if (++done == jobs.size())
                cv_done.notify_all();
        }
    }
    // run a list of copies: big ones are cut into slices; helpers and the caller all pull from the same list
    void run(const std::vector<Job>& list)
    {
        const size_t kSlice = (size_t)1 << 20;
        std::unique_lock<std::mutex> lk(mu);
        jobs.clear();
        next = done = 0;
        for (const Job& j : list) {
            if (j.bytes == 0)
                continue;
            if (j.bytes <= 2 * kSlice) {
                jobs.push_back(j);
                continue;
            }
            const size_t parts = std::min<size_t>((size_t)(n + 1) * 2, j.bytes / kSlice);
            const size_t per = ((j.bytes + parts - 1) / parts + 63) & ~(size_t)63;
            for (size_t o = 0; o < j.bytes; o += per)
                jobs.push_back({ j.dst + o, j.src + o, std::min(per, j.bytes - o) });
        }
        if (jobs.empty())
            return;
        lk.unlock();
        cv_work.notify_all();
        lk.lock();
        while (next < jobs.size()) { // the caller works too
            const Job j = jobs[next++];
            lk.unlock();
            memcpy(j.dst, j.src, j.bytes);
            lk.lock();
            ++done;
        }
        cv_done.wait(lk, [this] { return done == jobs.size(); });
        jobs.clear();
        next = done = 0;
    }
};
std::mutex g_pool_mu; // one big copy at a time through the pool
CopyPool& copy_pool()
{
    static CopyPool p;
    return p;
}
} // namespace

// Copies below this total are done inline: waking the helpers costs more than it saves
static const size_t kParMin = (size_t)512 << 10;

void par_memcpy_many(const std::vector<CopyJob>& list)
{
    size_t total = 0;
    for (const CopyJob& j : list)
        total += j.bytes;
    if (total < kParMin || copy_pool().n == 0) {
        for (const CopyJob& j : list)
            if (j.bytes)
                memcpy(j.dst, j.src, j.bytes);
        return;
    }
    std::vector<CopyPool::Job> jobs;
    jobs.reserve(list.size());
    for (const CopyJob& j : list)
        jobs.push_back({ (char*)j.dst, (const char*)j.src, j.bytes });
    std::lock_guard<std::mutex> lk(g_pool_mu);
    copy_pool().run(jobs);
}

void par_memcpy(void* dst, const void* src, size_t bytes)
{
    std::vector<CopyJob> one(1, CopyJob{ dst, src, bytes });
    par_memcpy_many(one);
}

static std::mutex g_shared_mu;
static std::mutex g_shared_work_mu[64];
static Stager* g_shared_sg[64];

int shared_stager(int device, Stager** out)
{
    DeviceScope ds(device);
    if (ds.rc)
        return ds.rc;
    std::lock_guard<std::mutex> lk(g_shared_mu);
    if (!g_shared_sg[device]) {
        Stager* s = new Stager;
        int rc = s->init(device);
        if (rc) {
            delete s;
            return rc;
        }
        g_shared_sg[device] = s;
    }
    *out = g_shared_sg[device];
    return 0;
}

std::mutex& shared_stager_mutex(int device) { return g_shared_work_mu[device >= 0 && device < 64 ? device : 0]; }

// Pieces that form an arithmetic progression on both sides (equal sizes, constant positive pitches) -- e.g. the M output
// streams of the channelizer when the caller's buffers are rows of one array -- go out as ONE cudaMemcpy2DAsync: the copy
// engine then streams the whole chunk instead of paying the per-copy setup cost M times (measured on B200: 64 x 1 MiB
// copies per chunk reach 32 GB/s, the 2-D copy 45+ GB/s of the 50 GB/s the link gives with both directions busy).
static bool strided_group(const std::vector<Piece>& v, const std::vector<char>& pinned, size_t* spitch, size_t* dpitch)
{
    if (v.size() < 4)
        return false;
    const size_t bytes = v[0].bytes;
    if (bytes == 0 || !pinned[0])
        return false;
    const ptrdiff_t sp = (const char*)v[1].src - (const char*)v[0].src, dp = (char*)v[1].dst - (char*)v[0].dst;
    if (sp < (ptrdiff_t)bytes || dp < (ptrdiff_t)bytes)
        return false;
    for (size_t k = 1; k < v.size(); k++) {
        if (v[k].bytes != bytes || !pinned[k] || (const char*)v[k].src - (const char*)v[k - 1].src != sp ||
            (char*)v[k].dst - (char*)v[k - 1].dst != dp)
            return false;
    }
    *spitch = (size_t)sp;
    *dpitch = (size_t)dp;
    return true;
}

int Stager::retire(int slot)
{
    if (!pending[slot])
        return 0;
    GRB_CUDA(cudaEventSynchronize(ev_d2h[slot]));
    if (!bounce[slot].empty()) {
        std::vector<CopyJob> jobs;
        jobs.reserve(bounce[slot].size());
        for (auto& b : bounce[slot])
            jobs.push_back({ b.host_dst, b.pin_src, b.bytes });
        par_memcpy_many(jobs);
    }
    bounce[slot].clear();
    pending[slot] = false;
    return 0;
}

int Stager::run(long long nchunks, const PlanFn& plan, const LaunchFn& launch)
{
    Chunk ch;
    LagResult* lr = lag_active;
    lag_active = nullptr; // whatever happens below, the next call starts clean
    void* const* louts = lag_outs;
    const int lnouts = lag_nouts;
    const size_t lshift = lag_shift;
    lag_outs = nullptr;
    int last_slot = -1;
    bool direct_in = false; // some H2D copy reads the caller's memory itself (page-locked input)
    for (long long c = 0; c < nchunks; c++, next_chunk++) {
        const int slot = (int)(next_chunk % NSLOT);
        int rc = retire(slot);
        if (rc)
            return rc;
        ch.h2d.clear();
        ch.d2h.clear();
        rc = plan(c, slot, ch);
        if (rc)
            return rc;
        // pageable pieces need bounce space
        size_t need_in = 0, need_out = 0;
        std::vector<char> pinned_in(ch.h2d.size()), pinned_out(ch.d2h.size());
        for (size_t k = 0; k < ch.h2d.size(); k++) {
            pinned_in[k] = is_pinned(ch.h2d[k].src);
            if (!pinned_in[k])
                need_in += (ch.h2d[k].bytes + 255) & ~(size_t)255;
        }
        for (size_t k = 0; k < ch.d2h.size() && !lr; k++) {
            pinned_out[k] = is_pinned(ch.d2h[k].dst);
            if (!pinned_out[k])
                need_out += (ch.d2h[k].bytes + 255) & ~(size_t)255;
        }
        if (need_in > pin_in_cap || need_out > pin_out_cap) {
            // growing the bounce slots invalidates in-flight copies: drain first
            for (int k = 0; k < NSLOT; k++)
                if ((rc = retire(k)))
                    return rc;
            if ((rc = reserve_pin(true, need_in)) || (rc = reserve_pin(false, need_out)))
                return rc;
        }
        size_t off = 0, spitch = 0, dpitch = 0;
        const bool h2d_2d = strided_group(ch.h2d, pinned_in, &spitch, &dpitch);
        if (h2d_2d) {
            GRB_CUDA(cudaMemcpy2DAsync(ch.h2d[0].dst, dpitch, ch.h2d[0].src, spitch, ch.h2d[0].bytes, ch.h2d.size(),
                                       cudaMemcpyHostToDevice, s_h2d));
            direct_in = true;
        }
        if (!h2d_2d && need_in) { // all pageable pieces of the chunk into the bounce slot at once, over the copy threads
            std::vector<CopyJob> jobs;
            size_t o2 = 0;
            for (size_t k = 0; k < ch.h2d.size(); k++) {
                if (pinned_in[k] || ch.h2d[k].bytes == 0)
                    continue;
                jobs.push_back({ (char*)pin_in[slot] + o2, ch.h2d[k].src, ch.h2d[k].bytes });
                o2 += (ch.h2d[k].bytes + 255) & ~(size_t)255;
            }
            par_memcpy_many(jobs);
        }
        bool bounce_2d = false;
        if (!h2d_2d && need_in && ch.h2d.size() >= 4) { // every piece bounced, equal sizes, constant device pitch: one copy
            std::vector<Piece> bp = ch.h2d;
            std::vector<char> allp(bp.size(), 1);
            size_t o2 = 0;
            bool all_pageable = true;
            for (size_t k = 0; k < bp.size(); k++) {
                if (pinned_in[k] || bp[k].bytes == 0)
                    all_pageable = false;
                bp[k].src = (char*)pin_in[slot] + o2;
                o2 += (bp[k].bytes + 255) & ~(size_t)255;
            }
            if (all_pageable && strided_group(bp, allp, &spitch, &dpitch)) {
                GRB_CUDA(cudaMemcpy2DAsync(bp[0].dst, dpitch, bp[0].src, spitch, bp[0].bytes, bp.size(), cudaMemcpyHostToDevice,
                                           s_h2d));
                bounce_2d = true;
            }
        }
        for (size_t k = 0; k < ch.h2d.size() && !h2d_2d && !bounce_2d; k++) {
            const Piece& p = ch.h2d[k];
            if (p.bytes == 0)
                continue;
            if (pinned_in[k]) {
                GRB_CUDA(cudaMemcpyAsync(p.dst, p.src, p.bytes, cudaMemcpyHostToDevice, s_h2d));
                direct_in = true;
            } else {
                char* b = (char*)pin_in[slot] + off;
                GRB_CUDA(cudaMemcpyAsync(p.dst, b, p.bytes, cudaMemcpyHostToDevice, s_h2d));
                off += (p.bytes + 255) & ~(size_t)255;
            }
        }
        GRB_CUDA(cudaEventRecord(ev_h2d[slot], s_h2d));
        GRB_CUDA(cudaStreamWaitEvent(s_exec, ev_h2d[slot], 0));
        rc = launch(c, slot, s_exec);
        if (rc)
            return rc;
        GRB_CUDA(cudaEventRecord(ev_exec[slot], s_exec));
        GRB_CUDA(cudaStreamWaitEvent(s_d2h, ev_exec[slot], 0));
        off = 0;
        if (lr) { // lag-by-one: delivered in place where the caller's output is page-locked, else parked in the handle
            // in place: the caller's buffers, shifted behind the previous call's items
            std::vector<char> direct(ch.d2h.size(), 0);
            std::vector<Piece> tgt = ch.d2h;
            bool all_direct = louts != nullptr && !ch.d2h.empty();
            for (size_t k = 0; k < ch.d2h.size(); k++) {
                const Piece& p = ch.d2h[k];
                if (louts && p.ostream >= 0 && p.ostream < lnouts && louts[p.ostream] && p.bytes) {
                    char* target = (char*)louts[p.ostream] + lshift + p.ooff;
                    if (is_pinned(target) && is_pinned(target + p.bytes - 1)) {
                        direct[k] = 1;
                        tgt[k].dst = target;
                    }
                }
                if (!direct[k] && p.bytes)
                    all_direct = false;
            }
            if (!all_direct && !lr->pin_ok()) { // first piece that must be parked: make room for the whole call
                if (lr->pin)
                    GRB_CUDA(cudaFreeHost(lr->pin));
                lr->pin = nullptr;
                lr->cap = 0;
                const size_t want = lr->need + lr->need / 4;
                cudaError_t e = cudaHostAlloc(&lr->pin, want, cudaHostAllocDefault);
                if (e != cudaSuccess)
                    return fail(GRB200_ENOMEM, "cudaHostAlloc(%zu) failed: %s", want, cudaGetErrorString(e));
                lr->cap = want;
            }
            if (!all_direct) // parked pieces sit back to back (64-byte aligned) in the handle's buffer
                for (size_t k = 0; k < ch.d2h.size(); k++)
                    if (!direct[k] && ch.d2h[k].bytes) {
                        if (lr->used + ch.d2h[k].bytes > lr->cap)
                            return fail(GRB200_ECUDA, "staging: lag buffer overflow (%zu + %zu > %zu)", lr->used,
                                        ch.d2h[k].bytes, lr->cap);
                        tgt[k].dst = (char*)lr->pin + lr->used;
                        lr->used += (ch.d2h[k].bytes + 63) & ~(size_t)63;
                    }
            std::vector<char> allp(tgt.size(), 1);
            if (strided_group(tgt, allp, &spitch, &dpitch)) { // e.g. the 64 streams of a channelizer chunk: ONE 2-D copy
                GRB_CUDA(cudaMemcpy2DAsync(tgt[0].dst, dpitch, tgt[0].src, spitch, tgt[0].bytes, tgt.size(),
                                           cudaMemcpyDeviceToHost, s_d2h));
            } else {
                for (const Piece& p : tgt)
                    if (p.bytes)
                        GRB_CUDA(cudaMemcpyAsync(p.dst, p.src, p.bytes, cudaMemcpyDeviceToHost, s_d2h));
            }
            for (size_t k = 0; k < tgt.size(); k++)
                if (tgt[k].bytes)
                    lr->segs.push_back({ tgt[k].ostream, tgt[k].ooff,
                                         direct[k] ? 0 : (size_t)((char*)tgt[k].dst - (char*)lr->pin), tgt[k].bytes, direct[k] != 0 });
        } else {
            const bool d2h_2d = strided_group(ch.d2h, pinned_out, &spitch, &dpitch);
            if (d2h_2d)
                GRB_CUDA(cudaMemcpy2DAsync(ch.d2h[0].dst, dpitch, ch.d2h[0].src, spitch, ch.d2h[0].bytes, ch.d2h.size(),
                                           cudaMemcpyDeviceToHost, s_d2h));
            bool bounce_out_2d = false;
            if (!d2h_2d && need_out && ch.d2h.size() >= 4) {
                std::vector<Piece> bp = ch.d2h;
                std::vector<char> allp(bp.size(), 1);
                size_t o2 = 0;
                bool all_pageable = true;
                for (size_t k = 0; k < bp.size(); k++) {
                    if (pinned_out[k] || bp[k].bytes == 0)
                        all_pageable = false;
                    bp[k].dst = (char*)pin_out[slot] + o2;
                    o2 += (bp[k].bytes + 255) & ~(size_t)255;
                }
                if (all_pageable && strided_group(bp, allp, &spitch, &dpitch)) {
                    GRB_CUDA(cudaMemcpy2DAsync(bp[0].dst, dpitch, bp[0].src, spitch, bp[0].bytes, bp.size(),
                                               cudaMemcpyDeviceToHost, s_d2h));
                    for (size_t k = 0; k < bp.size(); k++)
                        bounce[slot].push_back({ ch.d2h[k].dst, bp[k].dst, bp[k].bytes });
                    bounce_out_2d = true;
                }
            }
            for (size_t k = 0; k < ch.d2h.size() && !d2h_2d && !bounce_out_2d; k++) {
                const Piece& p = ch.d2h[k];
                if (p.bytes == 0)
                    continue;
                if (pinned_out[k]) {
                    GRB_CUDA(cudaMemcpyAsync(p.dst, p.src, p.bytes, cudaMemcpyDeviceToHost, s_d2h));
                } else {
                    char* b = (char*)pin_out[slot] + off;
                    GRB_CUDA(cudaMemcpyAsync(b, p.src, p.bytes, cudaMemcpyDeviceToHost, s_d2h));
                    bounce[slot].push_back({ p.dst, b, p.bytes });
                    off += (p.bytes + 255) & ~(size_t)255;
                }
            }
        }
        GRB_CUDA(cudaEventRecord(ev_d2h[slot], s_d2h));
        pending[slot] = true;
        last_slot = slot;
    }
    if (lr) {
        // the caller's INPUT buffers are borrowed for this call only: every host->device copy that reads them directly
        // (pinned memory; pageable went through the bounce already) must be over before we return.  Kernel and D2H run on.
        if (last_slot >= 0 && direct_in)
            GRB_CUDA(cudaEventSynchronize(ev_h2d[last_slot]));
        GRB_CUDA(cudaEventRecord(lr->ev, s_d2h));
        lr->pending = true;
        return 0;
    }
    for (int k = 0; k < NSLOT; k++) { // drain: the caller's output buffers are valid when work() returns
        int rc = retire(k);
        if (rc)
            return rc;
    }
    return 0;
}

// ---- lag-by-one mode -------------------------------------------------------------------------------------------------
int Stager::lag_begin(size_t out_bytes, long long npieces, long long nitems)
{
    LagResult& r = lag[lag_cur];
    if (r.pending) // two calls back: must have been collected; never overwrite a result nobody fetched
        return fail(GRB200_EINVAL, "staging: lag result not collected before the buffer comes round again");
    r.need = out_bytes + 64 * (size_t)npieces + 4096; // every piece starts on a 64-byte boundary
    if (!r.ev)
        GRB_CUDA(cudaEventCreateWithFlags(&r.ev, cudaEventDisableTiming));
    r.used = 0;
    r.nitems = nitems;
    r.segs.clear();
    lag_active = &r;
    lag_cur ^= 1;
    return 0;
}

static int lag_deliver(Stager::LagResult& r, void* const* outs, int nouts, long long* nitems)
{
    *nitems = 0;
    if (!r.pending)
        return 0;
    GRB_CUDA(cudaEventSynchronize(r.ev));
    std::vector<CopyJob> jobs;
    jobs.reserve(r.segs.size());
    for (const Stager::LagSeg& s : r.segs) {
        if (s.direct)
            continue; // went straight to the caller's page-locked buffer
        if (s.ostream < 0 || s.ostream >= nouts || !outs[s.ostream])
            return fail(GRB200_EINVAL, "staging: lag result for output stream %d, %d offered", s.ostream, nouts);
        jobs.push_back({ (char*)outs[s.ostream] + s.ooff, (const char*)r.pin + s.boff, s.bytes });
    }
    par_memcpy_many(jobs);
    r.pending = false;
    *nitems = r.nitems;
    return 0;
}

int Stager::lag_collect(void* const* outs, int nouts, long long* nitems)
{
    // the result that is NOT the one just begun: lag_cur was flipped by lag_begin, so the previous call's is lag[lag_cur]
    return lag_deliver(lag[lag_cur], outs, nouts, nitems);
}

int Stager::lag_flush(void* const* outs, int nouts, long long* nitems)
{
    // in submission order: the older of the two first (normally already collected), then the newest
    long long a = 0, b = 0;
    int rc = lag_deliver(lag[lag_cur], outs, nouts, &a);
    if (rc)
        return rc;
    if (a) { // a result two calls old was still parked: hand that one out; the newest stays for another flush
        *nitems = a;
        return 0;
    }
    rc = lag_deliver(lag[lag_cur ^ 1], outs, nouts, &b);
    *nitems = b;
    return rc;
}

void Stager::lag_drop()
{
    for (auto& r : lag) {
        if (r.pending && r.ev)
            cudaEventSynchronize(r.ev);
        r.pending = false;
        r.segs.clear();
    }
    lag_active = nullptr;
}

} // namespace grb
