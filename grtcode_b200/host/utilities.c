/* Host helpers: bit fields, string -> number conversion, piecewise resampling
 * and integration (reference: utilities/src/utilities.c:34-381).
 *
 * interpolate2() fixes the numerical recipe of every table the GPU later
 * reads: per source interval (x[j], x[j+1]] the line m*w + b with
 * m = (y1 - y0)/(x1 - x0), b = y0 - m*x0; no extrapolation unless asked for. */
#include <errno.h>
#include <float.h>
#include <limits.h>
#include <stdlib.h>
#include <string.h>
#include "errors.h"
#include "utilities.h"

int activate(uint64_t * const bit_field, int const index)
{
    REQUIRE_PTR(bit_field);
    REQUIRE_RANGE(index, 0, 63);
    *bit_field |= ((uint64_t)1 << index);
    return GRTCODE_SUCCESS;
}

int is_active(uint64_t const bit_field, int const index)
{
    if (index < 0 || index > 63)
    {
        return 0;
    }
    return (bit_field >> index) & 1 ? 1 : 0;
}

fp_t angstrom_exponent(fp_t tau1, fp_t tau2, fp_t lambda1, fp_t lambda2)
{
    return -1.*log(tau1/tau2)/log(lambda1/lambda2);
}

int angstrom_exponent_sample(fp_t const * const x, fp_t const * const y,
                             fp_t const * const newx, fp_t * const newy, size_t n)
{
    size_t i;
    for (i=0; i<2; ++i)
    {
        if (y[i] <= 0.)
        {
            FAIL(GRTCODE_VALUE_ERR, "Cannot calculate the angstrom exponent because"
                 " y[%zu] <= 0 (%e)", i, y[i]);
        }
    }
    fp_t const alpha = -1.*angstrom_exponent(y[1], y[0], x[0], x[1]);
    for (i=0; i<n; ++i)
    {
        newy[i] = y[0]*pow(x[0]/newx[i], alpha);
    }
    return GRTCODE_SUCCESS;
}

int constant_extrapolation(fp_t const * const x, fp_t const * const y,
                           fp_t const * const newx, fp_t * const newy, size_t n)
{
    (void)x;
    (void)newx;
    size_t i;
    for (i=0; i<n; ++i)
    {
        newy[i] = y[0];
    }
    return GRTCODE_SUCCESS;
}

int linear_sample(fp_t const * const x, fp_t const * const y, fp_t const * const newx,
                  fp_t * const newy, size_t n)
{
    fp_t const slope = (y[1] - y[0])/(x[1] - x[0]);
    fp_t const intercept = y[0] - slope*x[0];
    size_t i;
    for (i=0; i<n; ++i)
    {
        newy[i] = slope*newx[i] + intercept;
    }
    return GRTCODE_SUCCESS;
}

int monotonically_increasing(fp_t const * const x, size_t n)
{
    size_t i;
    for (i=1; i<n; ++i)
    {
        if (x[i] <= x[i-1])
        {
            return 0;
        }
    }
    return 1;
}

fp_t trapezoid(fp_t const * const x, fp_t const * const y)
{
    return 0.5*(y[0] + y[1])*(x[1] - x[0]);
}

int integrate2(fp_t const * const x, fp_t const * const y, size_t n, fp_t * const s,
               Area1d_t area)
{
    REQUIRE_PTR(x);
    REQUIRE_PTR(y);
    REQUIRE_PTR(s);
    REQUIRE_PTR(area);
    if (n < 2)
    {
        FAIL(GRTCODE_VALUE_ERR, "at least two x points required (%zu given).", n);
    }
    if (!monotonically_increasing(x, n))
    {
        FAIL(GRTCODE_VALUE_ERR, "x (%p) must be monotonically increasing.", (void const *)x);
    }
    fp_t total = 0.;
    size_t i;
    for (i=0; i+1<n; ++i)
    {
        total += area(x + i, y + i);
    }
    *s = total;
    return GRTCODE_SUCCESS;
}

int interpolate2(fp_t const * const x, fp_t const * const y, size_t n,
                 fp_t const * const newx, fp_t * const newy, size_t newn,
                 Sample1d_t interp, Sample1d_t extrap)
{
    REQUIRE_PTR(x);
    REQUIRE_PTR(y);
    REQUIRE_PTR(newx);
    REQUIRE_PTR(newy);
    if (n < 2)
    {
        FAIL(GRTCODE_VALUE_ERR, "at least two x points required (%zu given).", n);
    }
    if (!monotonically_increasing(x, n))
    {
        FAIL(GRTCODE_VALUE_ERR, "x (%p) must be monotonically increasing.", (void const *)x);
    }
    if (!monotonically_increasing(newx, newn))
    {
        FAIL(GRTCODE_VALUE_ERR, "newx (%p) must be monotonically increasing.", (void const *)newx);
    }
    /* Points at or below the first source abscissa. */
    size_t done = 0;
    while (done < newn && newx[done] <= x[0])
    {
        ++done;
    }
    if (done > 0 && extrap != NULL)
    {
        TRY(extrap(x, y, newx, newy, done));
    }
    /* One source interval at a time: (x[j], x[j+1]]. */
    size_t j;
    for (j=0; j+1<n && done<newn; ++j)
    {
        size_t stop = done;
        while (stop < newn && newx[stop] <= x[j+1])
        {
            ++stop;
        }
        if (stop > done)
        {
            TRY(interp(x + j, y + j, newx + done, newy + done, stop - done));
            done = stop;
        }
    }
    /* Points beyond the last source abscissa. */
    if (done < newn && extrap != NULL)
    {
        TRY(extrap(x + n - 2, y + n - 2, newx + done, newy + done, newn - done));
    }
    return GRTCODE_SUCCESS;
}

int copy_str(char * const dest, char const * const src, size_t const len)
{
    if (strlen(src) > len)
    {
        FAIL(GRTCODE_VALUE_ERR, "input string (%s) is larger than the input buffer"
             " and would be truncated.", src);
    }
    snprintf(dest, len, "%s", src);
    return GRTCODE_SUCCESS;
}

int malloc_ptr(void ** const p, size_t const num_bytes)
{
    REQUIRE_PTR(p);
    *p = malloc(num_bytes > 0 ? num_bytes : 1);
    REQUIRE_PTR(*p);
    return GRTCODE_SUCCESS;
}

int free_ptr(void ** const p)
{
    REQUIRE_PTR(p);
    REQUIRE_PTR(*p);
    free(*p);
    *p = NULL;
    return GRTCODE_SUCCESS;
}

int open_file(FILE **file, char const * const name, char const * const mode)
{
    REQUIRE_PTR(file);
    REQUIRE_PTR(name);
    REQUIRE_PTR(mode);
    *file = fopen(name, mode);
    if (*file == NULL)
    {
        FAIL(GRTCODE_IO_ERR, "failed to open file %s.", name);
    }
    return GRTCODE_SUCCESS;
}

int to_double(char const * const s, double * const d)
{
    REQUIRE_PTR(s);
    REQUIRE_PTR(d);
    char *end;
    errno = 0;
    *d = strtod(s, &end);
    if ((errno == ERANGE && (*d == HUGE_VAL || *d == -HUGE_VAL)) || (*d == 0. && errno != 0))
    {
        FAIL(GRTCODE_RANGE_ERR, "the input string %s is out of range.", s);
    }
    if (end == s)
    {
        FAIL(GRTCODE_VALUE_ERR, "invalid input string %s, expecting the string to"
             " contain a floating point number.", s);
    }
    return GRTCODE_SUCCESS;
}

int to_fp_t(double const d, fp_t * const f)
{
    REQUIRE_PTR(f);
    *f = d;
    return GRTCODE_SUCCESS;
}

int to_int(char const * const s, int * const i)
{
    REQUIRE_PTR(s);
    REQUIRE_PTR(i);
    char *end;
    errno = 0;
    long const v = strtol(s, &end, 10);
    if ((errno == ERANGE && (v == LONG_MAX || v == LONG_MIN)) || (v == 0 && errno != 0))
    {
        FAIL(GRTCODE_RANGE_ERR, "the input string %s is out of range.", s);
    }
    if (end == s)
    {
        FAIL(GRTCODE_VALUE_ERR, "invalid input string %s, expecting the string to"
             " contain an integer.", s);
    }
    if (v < INT_MIN || v > INT_MAX)
    {
        FAIL(GRTCODE_VALUE_ERR, "input string %s cannot be represented as an int.", s);
    }
    *i = (int)v;
    return GRTCODE_SUCCESS;
}
