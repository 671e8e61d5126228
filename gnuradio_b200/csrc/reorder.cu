// reorder.cu -- the gr-blocks plumbing either side of the path, C-ABI included.
//
//   stream_to_vector   gr-blocks/lib/stream_to_vector_impl.cc:46-58   memcpy(out, in, noutput * block_size): on the device a
//                      stream of items IS the stream of vectors -- a pointer reinterpretation, no kernel, no copy.  The host
//                      entry point exists for flowgraphs that keep the block: one staged copy.
//   vector_to_stream   the same statement the other way round.
//   vector_to_streams  gr-blocks/lib/vector_to_streams_impl.cc:34-55  out_j[n] = in[n*nstreams + j]: a transpose
//   streams_to_vector / stream_to_streams (gr-blocks/lib/stream_to_streams_impl.cc:35-54): its mirror image.
// The transposes go through a 32 x 33 shared-memory tile so both sides are coalesced; 2 x itemsize bytes per item, HBM-bound.
#include "capi_common.cuh"

namespace grb {

// interleaved [n*M + j] -> stream-major, stream j at out + j*out_stride
template <class T>
__global__ void __launch_bounds__(256)
interleaved_to_streams_kernel(const T* __restrict__ in, T* __restrict__ out, long long out_stride, int M, long long nrows)
{
    __shared__ T tile[32][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5; // 32 x 8
    const long long n0 = (long long)blockIdx.x * 32;
    const int j0 = blockIdx.y * 32;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const long long n = n0 + ty + k * 8;
        const int j = j0 + tx;
        if (j < M && n < nrows)
            tile[ty + k * 8][tx] = in[n * M + j];
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int j = j0 + ty + k * 8;
        const long long n = n0 + tx;
        if (j < M && n < nrows)
            out[(long long)j * out_stride + n] = tile[tx][ty + k * 8];
    }
}

template <class T>
__global__ void __launch_bounds__(256)
streams_to_interleaved_kernel_t(const T* __restrict__ in, long long in_stride, T* __restrict__ out, int M, long long nrows)
{
    __shared__ T tile[32][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const long long n0 = (long long)blockIdx.x * 32;
    const int j0 = blockIdx.y * 32;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int j = j0 + ty + k * 8;
        const long long n = n0 + tx;
        if (j < M && n < nrows)
            tile[ty + k * 8][tx] = in[(long long)j * in_stride + n];
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const long long n = n0 + ty + k * 8;
        const int j = j0 + tx;
        if (j < M && n < nrows)
            out[n * M + j] = tile[tx][ty + k * 8];
    }
}

static int reorder_launch(bool to_streams, int itemsize, const void* in, void* out, long long stride, int M, long long nrows,
                          cudaStream_t st)
{
    if (nrows <= 0)
        return 0;
    const long long gx = (nrows + 31) / 32;
    if (gx > 0x7fffffffLL || (M + 31) / 32 > 65535)
        return fail(GRB200_ERANGE, "vector_to_streams: call too large for one launch");
    dim3 grid((unsigned)gx, (unsigned)((M + 31) / 32));
    if (itemsize == 8) {
        if (to_streams)
            interleaved_to_streams_kernel<float2><<<grid, 256, 0, st>>>((const float2*)in, (float2*)out, stride, M, nrows);
        else
            streams_to_interleaved_kernel_t<float2><<<grid, 256, 0, st>>>((const float2*)in, stride, (float2*)out, M, nrows);
    } else if (itemsize == 4) {
        if (to_streams)
            interleaved_to_streams_kernel<float><<<grid, 256, 0, st>>>((const float*)in, (float*)out, stride, M, nrows);
        else
            streams_to_interleaved_kernel_t<float><<<grid, 256, 0, st>>>((const float*)in, stride, (float*)out, M, nrows);
    } else {
        return fail(GRB200_EINVAL, "vector_to_streams: item size %d not supported on the device (4 = float, 8 = gr_complex)",
                    itemsize);
    }
    GRB_LAUNCH_CHECK();
    return 0;
}

} // namespace grb

using namespace grb;

extern "C" {

int grb200_vector_to_streams_work_dev(int device, int itemsize, int nstreams, const void* in_dev, void* out_dev,
                                      long long out_stride, long long noutput, void* stream)
{
    if (nstreams < 1 || noutput < 0 || out_stride < noutput || (noutput > 0 && (!in_dev || !out_dev)))
        return fail(GRB200_EINVAL, "vector_to_streams::work_dev: bad arguments");
    Stager* sg = nullptr;
    int rc = shared_stager(device, &sg);
    if (rc)
        return rc;
    grb::DeviceScope _ds(device);
    cudaStream_t st_ = pick_stream(stream, *sg);
    return dev_done(reorder_launch(true, itemsize, in_dev, out_dev, out_stride, nstreams, noutput, st_), *sg, st_);
}

int grb200_streams_to_vector_work_dev(int device, int itemsize, int nstreams, const void* in_dev, long long in_stride,
                                      void* out_dev, long long noutput, void* stream)
{
    if (nstreams < 1 || noutput < 0 || in_stride < noutput || (noutput > 0 && (!in_dev || !out_dev)))
        return fail(GRB200_EINVAL, "streams_to_vector::work_dev: bad arguments");
    Stager* sg = nullptr;
    int rc = shared_stager(device, &sg);
    if (rc)
        return rc;
    grb::DeviceScope _ds(device);
    cudaStream_t st_ = pick_stream(stream, *sg);
    return dev_done(reorder_launch(false, itemsize, in_dev, out_dev, in_stride, nstreams, noutput, st_), *sg, st_);
}

// vector_to_streams_impl::work: `in_host` holds noutput vectors of nstreams items; out_streams[j] receives noutput items
int grb200_vector_to_streams_work(int device, int itemsize, int nstreams, const void* in_host, void* const* out_streams,
                                  long long noutput)
{
    if (nstreams < 1 || noutput < 0 || !out_streams || (noutput > 0 && !in_host))
        return fail(GRB200_EINVAL, "vector_to_streams::work: bad arguments");
    if (itemsize != 4 && itemsize != 8)
        return fail(GRB200_EINVAL, "vector_to_streams: item size %d not supported on the device (4 = float, 8 = gr_complex)",
                    itemsize);
    if (noutput == 0)
        return 0;
    Stager* sgp = nullptr;
    int rc = shared_stager(device, &sgp);
    if (rc)
        return rc;
    grb::DeviceScope _ds(device);
    std::lock_guard<std::mutex> lk(shared_stager_mutex(device));
    Stager& sg = *sgp;
    const size_t vsz = (size_t)itemsize * nstreams;
    long long chunk = std::max<long long>(1, (long long)(chunk_bytes_for((size_t)noutput * vsz) / vsz));
    if (chunk > noutput)
        chunk = noutput;
    const long long nchunks = (noutput + chunk - 1) / chunk;
    const long long stride = (chunk + 3) & ~3LL;
    if ((rc = sg.reserve_dev(0, (size_t)chunk * vsz)) || (rc = sg.reserve_dev(1, (size_t)stride * vsz)))
        return rc;
    return sg.run(
        nchunks,
        [&](long long c, int slot, Stager::Chunk& ch) {
            const long long n0 = c * chunk, n = std::min(chunk, noutput - n0);
            ch.h2d.push_back({ (const char*)in_host + (size_t)n0 * vsz, sg.dev[slot][0], (size_t)n * vsz });
            for (int j = 0; j < nstreams; j++)
                ch.d2h.push_back({ (const char*)sg.dev[slot][1] + (size_t)j * stride * itemsize,
                                   (char*)out_streams[j] + (size_t)n0 * itemsize, (size_t)n * itemsize });
            return 0;
        },
        [&](long long c, int slot, cudaStream_t st) {
            const long long n0 = c * chunk, n = std::min(chunk, noutput - n0);
            return reorder_launch(true, itemsize, sg.dev[slot][0], sg.dev[slot][1], stride, nstreams, n, st);
        });
}

} // extern "C"
