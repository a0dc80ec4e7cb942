// Device selection, memory and stream plumbing behind the C ABI.
// Replaces the gmalloc/gfree/gmemset/gmemcpy macros of the reference
// (utilities/src/debug.h:311-348) and get_num_gpus (utilities/src/device.c:26-49).
#include <mutex>
#include "common.cuh"

namespace grt {

static constexpr int kMaxDevices = 64;
static cudaStream_t g_streams[kMaxDevices];
static bool g_have_stream[kMaxDevices];
static int g_sm_count[kMaxDevices];
static std::mutex g_mutex;

int stream_of(int device, cudaStream_t *stream)
{
    if (device < 0 || device >= kMaxDevices)
    {
        return fail("device id %d out of range (HOST_ONLY is not supported: there is no CPU path)",
                    device);
    }
    std::lock_guard<std::mutex> lock(g_mutex);
    GRT_CUDA(cudaSetDevice(device));
    if (!g_have_stream[device])
    {
        GRT_CUDA(cudaStreamCreateWithFlags(&g_streams[device], cudaStreamNonBlocking));
        cudaDeviceProp prop;
        GRT_CUDA(cudaGetDeviceProperties(&prop, device));
        g_sm_count[device] = prop.multiProcessorCount;
        // grt_malloc / grt_free are stream-ordered (below): keep freed blocks in the device's
        // pool instead of returning them to the driver at every synchronisation
        cudaMemPool_t pool;
        GRT_CUDA(cudaDeviceGetDefaultMemPool(&pool, device));
        unsigned long long keep = ~0ull;
        GRT_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep));
        g_have_stream[device] = true;
    }
    *stream = g_streams[device];
    return 0;
}

int sm_count(int device)
{
    cudaStream_t s;
    if (stream_of(device, &s) != 0)
    {
        return 148;
    }
    return g_sm_count[device];
}

}  // namespace grt

using grt::fail;

extern "C" {

int grt_device_count(int *count)
{
    if (count == nullptr) return fail("null count");
    cudaError_t e = cudaGetDeviceCount(count);
    if (e != cudaSuccess)
    {
        *count = 0;
        return fail("cuda: %s", cudaGetErrorString(e));
    }
    return 0;
}

int grt_device_name(int device, char *buf, int len)
{
    cudaDeviceProp prop;
    GRT_CUDA(cudaGetDeviceProperties(&prop, device));
    snprintf(buf, len, "%s", prop.name);
    return 0;
}

// Device arrays of the public objects (Optics_t, solver buffers): allocated and freed in the
// order of the device's stream, on which every kernel and copy of this library runs, from the
// device's memory pool.  add_optics allocates its result at every call and the caller destroys
// it (optics.c:84-148): with cudaMalloc / cudaFree that was two driver calls and a full
// synchronisation per 6.5 GB object of a config-2 step.
int grt_malloc(int device, void **ptr_dev, size_t bytes)
{
    if (ptr_dev == nullptr) return fail("null pointer");
    cudaStream_t s;
    GRT_TRY(grt::stream_of(device, &s));
    GRT_CUDA(cudaMallocAsync(ptr_dev, bytes > 0 ? bytes : 1, s));
    return 0;
}

int grt_free(int device, void *ptr_dev)
{
    cudaStream_t s;
    GRT_TRY(grt::stream_of(device, &s));
    GRT_CUDA(cudaFreeAsync(ptr_dev, s));
    return 0;
}

int grt_memset(int device, void *ptr_dev, int value, size_t bytes)
{
    cudaStream_t s;
    GRT_TRY(grt::stream_of(device, &s));
    GRT_CUDA(cudaMemsetAsync(ptr_dev, value, bytes, s));
    return 0;
}

int grt_copy_to_device(int device, void *dst_dev, void const *src, size_t bytes)
{
    cudaStream_t s;
    GRT_TRY(grt::stream_of(device, &s));
    GRT_CUDA(cudaMemcpyAsync(dst_dev, src, bytes, cudaMemcpyHostToDevice, s));
    GRT_CUDA(cudaStreamSynchronize(s));
    return 0;
}

int grt_copy_to_host(int device, void *dst, void const *src_dev, size_t bytes)
{
    cudaStream_t s;
    GRT_TRY(grt::stream_of(device, &s));
    GRT_CUDA(cudaMemcpyAsync(dst, src_dev, bytes, cudaMemcpyDeviceToHost, s));
    GRT_CUDA(cudaStreamSynchronize(s));
    return 0;
}

int grt_copy_on_device(int device, void *dst_dev, void const *src_dev, size_t bytes)
{
    cudaStream_t s;
    GRT_TRY(grt::stream_of(device, &s));
    GRT_CUDA(cudaMemcpyAsync(dst_dev, src_dev, bytes, cudaMemcpyDeviceToDevice, s));
    return 0;
}

int grt_synchronize(int device)
{
    cudaStream_t s;
    GRT_TRY(grt::stream_of(device, &s));
    GRT_CUDA(cudaStreamSynchronize(s));
    return 0;
}

int grt_stream_handle(int device, void **stream)
{
    if (stream == nullptr) return fail("null pointer");
    cudaStream_t s;
    GRT_TRY(grt::stream_of(device, &s));
    *stream = (void *)s;
    return 0;
}

int grt_host_alloc(void **ptr, size_t bytes)
{
    if (ptr == nullptr) return fail("null pointer");
    GRT_CUDA(cudaMallocHost(ptr, bytes > 0 ? bytes : 1));
    return 0;
}

int grt_host_free(void *ptr)
{
    GRT_CUDA(cudaFreeHost(ptr));
    return 0;
}

int grt_last_error(char *buf, int len)
{
    if (buf == nullptr || len < 1) return GRT_SM100_ERR;
    snprintf(buf, len, "%s", grt::error_buffer());
    return 0;
}

}  // extern "C"
