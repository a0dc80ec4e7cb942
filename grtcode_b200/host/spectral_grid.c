/* Uniform spectral grid (reference: utilities/src/spectral_grid.c:31-112). */
#include <math.h>
#include <stdlib.h>
#include "errors.h"
#include "spectral_grid.h"

int create_spectral_grid(SpectralGrid_t * const grid, double const w0, double const wn,
                         double const dw)
{
    REQUIRE_PTR(grid);
    REQUIRE_RANGE(w0, MIN_WAVENUMBER, MAX_WAVENUMBER);
    REQUIRE_RANGE(wn, w0 + 1.e-12, MAX_WAVENUMBER);
    REQUIRE_RANGE(dw, MIN_RESOLUTION, MAX_RESOLUTION);
    grid->w0 = w0;
    grid->wn = wn;
    grid->dw = dw;
    grid->n = ceil((wn - w0)/dw) + 1.;
    LOG_INFO("Spectral grid properties:\n\tlower bound: %e [1/cm]\n\tupper bound: %e [1/cm]"
             "\n\tresolution: %e [1/cm]\n\ttotal size: %llu grid points", w0, wn, dw,
             (unsigned long long)grid->n);
    return GRTCODE_SUCCESS;
}

int compare_spectral_grids(SpectralGrid_t const * const one, SpectralGrid_t const * const two,
                           int * const result)
{
    REQUIRE_PTR(one);
    REQUIRE_PTR(two);
    REQUIRE_PTR(result);
    *result = (one->w0 == two->w0 && one->wn == two->wn && one->dw == two->dw) ? 1 : 0;
    return GRTCODE_SUCCESS;
}

int grid_point_index(SpectralGrid_t const grid, double const w, uint64_t * const index)
{
    REQUIRE_PTR(index);
    REQUIRE_RANGE(w, grid.w0, grid.wn);
    *index = (uint64_t)(round((w - grid.w0)/grid.dw));
    if (fabs(grid.w0 + (*index)*grid.dw - w) > grid.dw*1.e-5)
    {
        FAIL(GRTCODE_VALUE_ERR, "value %e not located on grid.", w);
    }
    return GRTCODE_SUCCESS;
}

int grid_points(SpectralGrid_t const grid, fp_t **buffer, Device_t const device)
{
    REQUIRE_PTR(buffer);
    if (device != HOST_ONLY)
    {
        FAIL(GRTCODE_VALUE_ERR, "grid_points builds a host array; pass HOST_ONLY (got %d).", device);
    }
    fp_t *w = (fp_t *)malloc(sizeof(*w)*grid.n);
    REQUIRE_PTR(w);
    uint64_t i;
    for (i=0; i<grid.n; ++i)
    {
        w[i] = grid.w0 + i*grid.dw;
    }
    *buffer = w;
    return GRTCODE_SUCCESS;
}

int interpolate_to_grid(SpectralGrid_t const grid, fp_t const * const x, fp_t const * const y,
                        size_t const n, fp_t * const newy, Sample1d_t interp, Sample1d_t extrap)
{
    fp_t *w;
    TRY(grid_points(grid, &w, HOST_ONLY));
    int const rc = interpolate2(x, y, n, w, newy, (size_t)grid.n, interp, extrap);
    free(w);
    if (rc != GRTCODE_SUCCESS)
    {
        grt_error_frame(__FILE__, __LINE__);
    }
    return rc;
}
