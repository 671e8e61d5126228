// peer.cu -- peer memory for the one exchange step the path has: the history halo of a time-sharded stream
// (SURVEY.md section 8(e)).  Rank g's first ntaps-1 (FIR) / taps_per_filter*numchans (channelizer) history items are the
// left neighbour's newest items.  Instead of a host-launched send/recv pair per step, the neighbour's tail is mapped into
// this process once (CUDA IPC between the one-process-per-GPU ranks of a torchrun job; cudaDeviceEnablePeerAccess inside one
// process) and read over NVLink by the block itself: grb200_*_set_halo_source().
#include <cuda.h>
#include <map>
#include "capi_common.cuh"

using namespace grb;

namespace {
std::mutex g_ipc_mu;
std::map<void*, void*> g_ipc_base; // pointer handed out -> base returned by cudaIpcOpenMemHandle
} // namespace

struct grb200_ipc_blob { // what grb200_ipc_export writes: 64-byte CUDA handle + the pointer's offset inside the allocation
    cudaIpcMemHandle_t handle;
    unsigned long long offset;
};
static_assert(sizeof(grb200_ipc_blob) == 72, "grb200_ipc_handle is 72 bytes");

extern "C" {

int grb200_ipc_export(int device, const void* dev_ptr, void* handle72)
{
    if (!dev_ptr || !handle72)
        return fail(GRB200_EINVAL, "ipc_export: bad arguments");
    grb::DeviceScope _ds(device);
    if (_ds.rc)
        return _ds.rc;
    // the handle names the whole allocation; remember where inside it the pointer sits (torch sub-allocates its segments)
    typedef CUresult (*PFN_range)(CUdeviceptr*, size_t*, CUdeviceptr);
    static PFN_range get_range = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuMemGetAddressRange", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            p = nullptr;
        return (PFN_range)p;
    }();
    if (!get_range)
        return fail(GRB200_ECUDA, "ipc_export: cuMemGetAddressRange is not available");
    CUdeviceptr base = 0;
    size_t size = 0;
    if (get_range(&base, &size, (CUdeviceptr)(uintptr_t)dev_ptr) != CUDA_SUCCESS)
        return fail(GRB200_ECUDA, "ipc_export: not a device allocation");
    grb200_ipc_blob* b = (grb200_ipc_blob*)handle72;
    GRB_CUDA(cudaIpcGetMemHandle(&b->handle, (void*)(uintptr_t)base));
    b->offset = (unsigned long long)((uintptr_t)dev_ptr - (uintptr_t)base);
    return 0;
}

int grb200_ipc_open(int device, const void* handle72, void** peer_ptr)
{
    if (!handle72 || !peer_ptr)
        return fail(GRB200_EINVAL, "ipc_open: bad arguments");
    grb::DeviceScope _ds(device);
    if (_ds.rc)
        return _ds.rc;
    const grb200_ipc_blob* b = (const grb200_ipc_blob*)handle72;
    void* base = nullptr;
    GRB_CUDA(cudaIpcOpenMemHandle(&base, b->handle, cudaIpcMemLazyEnablePeerAccess));
    *peer_ptr = (char*)base + b->offset;
    std::lock_guard<std::mutex> lk(g_ipc_mu);
    g_ipc_base[*peer_ptr] = base;
    return 0;
}

int grb200_ipc_close(int device, void* peer_ptr)
{
    grb::DeviceScope _ds(device);
    if (_ds.rc)
        return _ds.rc;
    void* base = nullptr;
    {
        std::lock_guard<std::mutex> lk(g_ipc_mu);
        auto it = g_ipc_base.find(peer_ptr);
        if (it == g_ipc_base.end())
            return fail(GRB200_EINVAL, "ipc_close: pointer was not opened with grb200_ipc_open");
        base = it->second;
        g_ipc_base.erase(it);
    }
    GRB_CUDA(cudaIpcCloseMemHandle(base));
    return 0;
}

// one process driving several GPUs (a C++ flowgraph with blocks on GPUs 0..7): let `device` read `peer_device`'s memory
int grb200_enable_peer(int device, int peer_device)
{
    grb::DeviceScope _ds(device);
    if (_ds.rc)
        return _ds.rc;
    int can = 0;
    GRB_CUDA(cudaDeviceCanAccessPeer(&can, device, peer_device));
    if (!can)
        return fail(GRB200_ECUDA, "device %d cannot access device %d", device, peer_device);
    cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) {
        cudaGetLastError();
        e = cudaSuccess;
    }
    GRB_CUDA(e);
    return 0;
}

} // extern "C"
