/* Uniform wavenumber grid (reference: utilities/src/spectral_grid.h:33-83). */
#ifndef SPECTRAL_GRID_H_
#define SPECTRAL_GRID_H_
#include <stddef.h>
#include <stdint.h>
#include "device.h"
#include "extern.h"
#include "floating_point_type.h"
#include "utilities.h"

typedef struct SpectralGrid
{
    double dw; /* spacing [cm-1] */
    uint64_t n; /* number of points = ceil((wn - w0)/dw) + 1 */
    double wn; /* upper bound [cm-1] */
    double w0; /* lower bound [cm-1] */
} SpectralGrid_t;

EXTERN int create_spectral_grid(SpectralGrid_t * const grid, double const w0, double const wn,
                                double const dw);
EXTERN int compare_spectral_grids(SpectralGrid_t const * const one,
                                  SpectralGrid_t const * const two, int * const result);
EXTERN int grid_point_index(SpectralGrid_t const grid, double const w, uint64_t * const index);
/* Host array of all grid points; `device` must be HOST_ONLY (the array is a
 * host-side helper in every caller of the reference). */
EXTERN int grid_points(SpectralGrid_t const grid, fp_t **buffer, Device_t const device);
EXTERN int interpolate_to_grid(SpectralGrid_t const grid, fp_t const * const x,
                               fp_t const * const y, size_t const n, fp_t * const newy,
                               Sample1d_t interp, Sample1d_t extrap);
#endif
