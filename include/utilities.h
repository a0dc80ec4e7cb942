/* Host helpers that are part of the reference's public surface
 * (utilities/src/utilities.h:33-176): piecewise resampling and integration of
 * tabulated data, used by the drivers for emissivity/albedo and by the table
 * readers. */
#ifndef UTILITIES_H_
#define UTILITIES_H_
#include <stddef.h>
#include <stdint.h>
#include <stdio.h>
#include "extern.h"
#include "floating_point_type.h"

/* Area under one interval: x[0..1], y[0..1]. */
typedef fp_t (*Area1d_t)(fp_t const * const, fp_t const * const);
/* Fill newy[0..n) for points newx[0..n) from the interval x[0..1], y[0..1]. */
typedef int (*Sample1d_t)(fp_t const * const, fp_t const * const, fp_t const * const,
                          fp_t * const, size_t);

EXTERN int linear_sample(fp_t const * const x, fp_t const * const y, fp_t const * const newx,
                         fp_t * const newy, size_t n);
EXTERN int constant_extrapolation(fp_t const * const x, fp_t const * const y,
                                  fp_t const * const newx, fp_t * const newy, size_t n);
EXTERN int angstrom_exponent_sample(fp_t const * const x, fp_t const * const y,
                                    fp_t const * const newx, fp_t * const newy, size_t n);
EXTERN fp_t angstrom_exponent(fp_t tau1, fp_t tau2, fp_t lambda1, fp_t lambda2);
/* Points newx <= x[0] and > x[n-1] are passed to extrap (left untouched when
 * extrap is NULL); each source interval (x[j], x[j+1]] goes to interp. */
EXTERN int interpolate2(fp_t const * const x, fp_t const * const y, size_t n,
                        fp_t const * const newx, fp_t * const newy, size_t newn,
                        Sample1d_t interp, Sample1d_t extrap);
EXTERN int integrate2(fp_t const * const x, fp_t const * const y, size_t n, fp_t * const s,
                      Area1d_t area);
EXTERN fp_t trapezoid(fp_t const * const x, fp_t const * const y);
EXTERN int monotonically_increasing(fp_t const * const x, size_t n);
EXTERN int activate(uint64_t * const bit_field, int const index);
EXTERN int is_active(uint64_t const bit_field, int const index);
EXTERN int copy_str(char * const dest, char const * const src, size_t const len);
EXTERN int malloc_ptr(void ** const p, size_t const num_bytes);
EXTERN int free_ptr(void ** const p);
EXTERN int open_file(FILE **file, char const * const name, char const * const mode);
EXTERN int to_double(char const * const s, double * const d);
EXTERN int to_fp_t(double const d, fp_t * const f);
EXTERN int to_int(char const * const s, int * const i);
#endif
