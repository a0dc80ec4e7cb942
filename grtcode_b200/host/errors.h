/* Internal error plumbing of the host library.
 * Behaviour mirrors the reference's return-code discipline
 * (utilities/src/debug.h:39-101): a failing function records one message, every
 * frame on the way up appends "file: line" to a process-global buffer that
 * grtcode_errstr() returns, and the numeric code travels to the caller.  The
 * library never aborts. */
#ifndef GRT_ERRORS_H_
#define GRT_ERRORS_H_
#include <math.h>
#include <stdio.h>
#include "return_codes.h"
#include "verbosity.h"

/* Math functions of the working precision (utilities/src/grtcode_config.h:25-47). */
#ifdef SINGLE_PRECISION
#define GRT_EXP expf
#define GRT_POW powf
#define GRT_SQRT sqrtf
#define GRT_ABS fabsf
#define GRT_LOG logf
#else
#define GRT_EXP exp
#define GRT_POW pow
#define GRT_SQRT sqrt
#define GRT_ABS fabs
#define GRT_LOG log
#endif

void grt_error_reset(void);
void grt_error_append(char const *text);
void grt_error_frame(char const *file, int line);
/* Converts a failure of the CUDA layer into GRTCODE_GPU_ERR with its message. */
int grt_gpu_failure(char const *file, int line);

#define FAIL(code, ...) \
    do { \
        char m__[1024]; \
        snprintf(m__, sizeof(m__), __VA_ARGS__); \
        grt_error_reset(); \
        grt_error_append("Error: "); \
        grt_error_append(m__); \
        grt_error_append("\nBacktrace:"); \
        grt_error_frame(__FILE__, __LINE__); \
        return (code); \
    } while (0)

#define TRY(call) \
    do { \
        int rc__ = (call); \
        if (rc__ != GRTCODE_SUCCESS) { \
            grt_error_frame(__FILE__, __LINE__); \
            return rc__; \
        } \
    } while (0)

/* Calls into libgrt_sm100. */
#define GPU(call) \
    do { \
        if ((call) != 0) return grt_gpu_failure(__FILE__, __LINE__); \
    } while (0)

#define REQUIRE_PTR(p) \
    do { \
        if ((p) == NULL) FAIL(GRTCODE_NULL_ERR, "null pointer: %s.", #p); \
    } while (0)

#define REQUIRE_RANGE(v, lo, hi) \
    do { \
        double v__ = (double)(v); \
        if (isnan(v__)) FAIL(GRTCODE_INVALID_ERR, "input value %s is NaN.", #v); \
        if (v__ < (double)(lo)) \
            FAIL(GRTCODE_RANGE_ERR, "value (%e) less than minimum allowed (%e).", v__, (double)(lo)); \
        if (v__ > (double)(hi)) \
            FAIL(GRTCODE_RANGE_ERR, "value (%e) greater than maximum allowed (%e).", v__, (double)(hi)); \
    } while (0)

#define REQUIRE_EQUAL(a, b) \
    do { \
        if ((a) != (b)) \
            FAIL(GRTCODE_VALUE_ERR, "values (%lld, %lld) are not equal.", (long long)(a), (long long)(b)); \
    } while (0)

#define REQUIRE_GPU(device) \
    do { \
        if ((device) < 0) \
            FAIL(GRTCODE_VALUE_ERR, "device %d: this library has no host (HOST_ONLY) code path; " \
                 "pass a GPU id.", (int)(device)); \
    } while (0)

#define LOG_AT(level, stream, tag, ...) \
    do { \
        if (grtcode_verbosity() >= (level)) { \
            char m__[1024]; \
            snprintf(m__, sizeof(m__), __VA_ARGS__); \
            fprintf(stream, "[%s:%d] %s%s\n", __FILE__, __LINE__, tag, m__); \
        } \
    } while (0)
#define LOG_INFO(...) LOG_AT(GRTCODE_INFO, stderr, "info: ", __VA_ARGS__)
#define LOG_WARN(...) LOG_AT(GRTCODE_WARN, stderr, "warning: ", __VA_ARGS__)
/* The reference prints these at every verbosity level (debug.h:64-70); here
 * they need at least GRTCODE_WARN so that batch drivers stay quiet by default. */
#define LOG_MESG(...) LOG_AT(GRTCODE_WARN, stdout, "", __VA_ARGS__)

/* Limits shared with the reference (utilities/src/grtcode_config.h:50-67). */
#define MIN_NUM_LAYERS 1
#define MAX_NUM_LAYERS 200
#define MIN_NUM_LEVELS (MIN_NUM_LAYERS + 1)
#define MAX_NUM_LEVELS (MAX_NUM_LAYERS + 1)
#define MIN_WAVENUMBER 1.
#define MAX_WAVENUMBER 50000.
#define MIN_RESOLUTION 0.001
#define MAX_RESOLUTION 10.
#endif
