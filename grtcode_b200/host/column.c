/* Additions for the drivers: broadband integration, heating rates, and the
 * allocation helpers of the Fortran bindings
 * (reference: framework/src/driver.c:285-356; fortran-bindings/malloc_structs.c:29-137). */
#include <stdlib.h>
#include <string.h>
#include "errors.h"
#include "grt_sm100.h"
#include "grtcode_column.h"
#include "malloc_structs.h"

int integrate_fluxes(Device_t const device, int const num_levels,
                     SpectralGrid_t const * const grid, fp_t const * const flux_dev,
                     fp_t * const out)
{
    REQUIRE_PTR(grid);
    REQUIRE_PTR(flux_dev);
    REQUIRE_PTR(out);
    REQUIRE_GPU(device);
    GPU(grt_integrate_flux(device, num_levels, grid->n, grid->dw, flux_dev, out));
    return GRTCODE_SUCCESS;
}

/* The reference computes no heating rates; this is the usual
 * flux-divergence form  dT/dt = g/cp * d(F_down - F_up)/dp. */
int heating_rates(int const num_levels, fp_t const * const pressure, fp_t const * const up,
                  fp_t const * const down, fp_t * const rate)
{
    REQUIRE_PTR(pressure);
    REQUIRE_PTR(up);
    REQUIRE_PTR(down);
    REQUIRE_PTR(rate);
    REQUIRE_RANGE(num_levels, MIN_NUM_LEVELS, MAX_NUM_LEVELS);
    double const gravity = 9.80665; /* [m s-2] */
    double const cp = 1004.64; /* [J kg-1 K-1] */
    double const seconds_per_day = 86400.;
    int i;
    for (i=0; i<num_levels-1; ++i)
    {
        double const net_top = down[i] - up[i];
        double const net_bottom = down[i+1] - up[i+1];
        double const dp = (pressure[i+1] - pressure[i])*100.; /* mb -> Pa */
        rate[i] = (dp != 0.) ? gravity/cp*(net_top - net_bottom)/dp*seconds_per_day : 0.;
    }
    return GRTCODE_SUCCESS;
}

enum StructTypes
{
    GRID = 0,
    OPTICS,
    GAS_OPTICS,
    SOLAR_FLUX
};

int malloc_struct(void **p, int type)
{
    size_t s;
    switch (type)
    {
        case GRID: s = sizeof(SpectralGrid_t); break;
        case OPTICS: s = sizeof(Optics_t); break;
        case GAS_OPTICS: s = sizeof(GasOptics_t); break;
        case SOLAR_FLUX: s = sizeof(SolarFlux_t); break;
        default: FAIL(GRTCODE_VALUE_ERR, "unrecognized structure type %d.", type);
    }
    REQUIRE_PTR(p);
    if (*p != NULL)
    {
        FAIL(GRTCODE_NON_NULL_ERR, "pointer at address %p is not null.", (void *)p);
    }
    *p = calloc(1, s);
    REQUIRE_PTR(*p);
    return GRTCODE_SUCCESS;
}

int free_struct(void **p)
{
    TRY(free_ptr(p));
    return GRTCODE_SUCCESS;
}

int spectral_grid_properties(SpectralGrid_t const * const grid, double * const w0,
                             uint64_t * const n, double * const dw)
{
    REQUIRE_PTR(grid);
    if (w0 != NULL) *w0 = grid->w0;
    if (n != NULL) *n = grid->n;
    if (dw != NULL) *dw = grid->dw;
    return GRTCODE_SUCCESS;
}

int solar_flux_properties(SolarFlux_t const * const solar, fp_t * const flux)
{
    REQUIRE_PTR(solar);
    REQUIRE_PTR(flux);
    memcpy(flux, solar->incident_flux, sizeof(fp_t)*solar->n);
    return GRTCODE_SUCCESS;
}
