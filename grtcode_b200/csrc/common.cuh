// Shared plumbing of the sm_100a layer: error capture, the per-device stream.
#pragma once
#include <cuda_runtime.h>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include "grt_sm100.h"

// Working precision of the data arrays that cross the C ABI (grt_real in grt_sm100.h): double,
// or float when the whole stack is compiled with the reference's -DSINGLE_PRECISION
// (configure.ac:25-32, utilities/src/floating_point_type.h:23-29).
#ifdef SINGLE_PRECISION
#define GRT_SINGLE 1
#endif

namespace grt {

typedef grt_real real_t;

// Last error message of the calling thread (read by grt_last_error()).
inline char *error_buffer()
{
    static thread_local char buf[1024] = "";
    return buf;
}

inline int fail(char const *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(error_buffer(), 1024, fmt, ap);
    va_end(ap);
    return GRT_SM100_ERR;
}

#define GRT_CUDA(call) \
    do { \
        cudaError_t e__ = (call); \
        if (e__ != cudaSuccess) { \
            return ::grt::fail("cuda: %s (%s:%d)", cudaGetErrorString(e__), __FILE__, __LINE__); \
        } \
    } while (0)

#define GRT_TRY(call) \
    do { \
        int r__ = (call); \
        if (r__ != 0) return r__; \
    } while (0)

// One stream per device, created on first use; every kernel and copy of that
// device is ordered on it.
int stream_of(int device, cudaStream_t *stream);

// Number of SMs of the device (cached).
int sm_count(int device);

}  // namespace grt
