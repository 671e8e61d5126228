// staging.cu -- library globals, device selection and the pinned double-buffered staging pipeline (K8).
#include <algorithm>
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

bool Stager::is_pinned(const void* p)
{
    cudaPointerAttributes attr;
    cudaError_t e = cudaPointerGetAttributes(&attr, p);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return attr.type == cudaMemoryTypeHost;
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

int Stager::run(long long nchunks, const PlanFn& plan, const LaunchFn& launch)
{
    struct Bounce {
        void* host_dst;
        const void* pin_src;
        size_t bytes;
    };
    std::vector<Bounce> bounce[NSLOT];
    bool pending[NSLOT] = {};
    Chunk ch;
    auto retire = [&](int slot) -> int {
        if (!pending[slot])
            return 0;
        GRB_CUDA(cudaEventSynchronize(ev_d2h[slot]));
        for (auto& b : bounce[slot])
            memcpy(b.host_dst, b.pin_src, b.bytes);
        bounce[slot].clear();
        pending[slot] = false;
        return 0;
    };
    for (long long c = 0; c < nchunks; c++) {
        const int slot = (int)(c % NSLOT);
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
        for (size_t k = 0; k < ch.d2h.size(); k++) {
            pinned_out[k] = is_pinned(ch.d2h[k].dst);
            if (!pinned_out[k])
                need_out += (ch.d2h[k].bytes + 255) & ~(size_t)255;
        }
        if (need_in > pin_in_cap || need_out > pin_out_cap) {
            // growing the bounce slots invalidates in-flight copies: drain first
            for (long long k = c - NSLOT + 1; k < c; k++)
                if (k >= 0 && (rc = retire((int)(k % NSLOT))))
                    return rc;
            if ((rc = reserve_pin(true, need_in)) || (rc = reserve_pin(false, need_out)))
                return rc;
        }
        size_t off = 0, spitch = 0, dpitch = 0;
        const bool h2d_2d = strided_group(ch.h2d, pinned_in, &spitch, &dpitch);
        if (h2d_2d)
            GRB_CUDA(cudaMemcpy2DAsync(ch.h2d[0].dst, dpitch, ch.h2d[0].src, spitch, ch.h2d[0].bytes, ch.h2d.size(),
                                       cudaMemcpyHostToDevice, s_h2d));
        for (size_t k = 0; k < ch.h2d.size() && !h2d_2d; k++) {
            const Piece& p = ch.h2d[k];
            if (p.bytes == 0)
                continue;
            if (pinned_in[k]) {
                GRB_CUDA(cudaMemcpyAsync(p.dst, p.src, p.bytes, cudaMemcpyHostToDevice, s_h2d));
            } else {
                char* b = (char*)pin_in[slot] + off;
                memcpy(b, p.src, p.bytes);
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
        const bool d2h_2d = strided_group(ch.d2h, pinned_out, &spitch, &dpitch);
        if (d2h_2d)
            GRB_CUDA(cudaMemcpy2DAsync(ch.d2h[0].dst, dpitch, ch.d2h[0].src, spitch, ch.d2h[0].bytes, ch.d2h.size(),
                                       cudaMemcpyDeviceToHost, s_d2h));
        for (size_t k = 0; k < ch.d2h.size() && !d2h_2d; k++) {
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
        GRB_CUDA(cudaEventRecord(ev_d2h[slot], s_d2h));
        pending[slot] = true;
    }
    // drain in submission order
    for (long long k = std::max<long long>(0, nchunks - NSLOT); k < nchunks; k++) {
        int rc = retire((int)(k % NSLOT));
        if (rc)
            return rc;
    }
    return 0;
}

} // namespace grb
