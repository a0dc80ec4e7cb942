/* liblongwave host side (reference: longwave/src/longwave.c:27-64, 312-353). */
#include <string.h>
#include "errors.h"
#include "grt_sm100.h"
#include "grtcode_column.h"
#include "longwave.h"

int create_longwave(Longwave_t * const lw, int const num_levels,
                    SpectralGrid_t const * const grid, Device_t const * const device)
{
    REQUIRE_PTR(lw);
    REQUIRE_PTR(grid);
    REQUIRE_PTR(device);
    REQUIRE_GPU(*device);
    REQUIRE_RANGE(num_levels, MIN_NUM_LEVELS, MAX_NUM_LEVELS);
    memset(lw, 0, sizeof(*lw));
    lw->num_levels = num_levels;
    lw->grid = *grid;
    lw->device = *device;
    GPU(grt_malloc(lw->device, (void **)&lw->emissivity, sizeof(fp_t)*grid->n));
    /* Spectral flux buffers are allocated on first use: the fused column path
       never needs them. */
    return GRTCODE_SUCCESS;
}

int destroy_longwave(Longwave_t * const lw)
{
    REQUIRE_PTR(lw);
    GPU(grt_free(lw->device, lw->emissivity));
    if (lw->flux_up != NULL) GPU(grt_free(lw->device, lw->flux_up));
    if (lw->flux_down != NULL) GPU(grt_free(lw->device, lw->flux_down));
    lw->emissivity = lw->flux_up = lw->flux_down = NULL;
    return GRTCODE_SUCCESS;
}

static int check_against_optics(Longwave_t const *lw, Optics_t const *optics)
{
    REQUIRE_EQUAL(lw->device, optics->device);
    REQUIRE_EQUAL(lw->num_levels, optics->num_layers + 1);
    int same_grids;
    TRY(compare_spectral_grids(&lw->grid, &optics->grid, &same_grids));
    REQUIRE_EQUAL(same_grids, 1);
    return GRTCODE_SUCCESS;
}

int calculate_lw_fluxes(Longwave_t * const lw, Optics_t const * const optics, fp_t const T_surf,
                        fp_t * const T_layers, fp_t * const T_levels, fp_t * const emis,
                        fp_t * const flux_up, fp_t * const flux_down)
{
    REQUIRE_PTR(lw);
    REQUIRE_PTR(optics);
    REQUIRE_PTR(T_layers);
    REQUIRE_PTR(T_levels);
    REQUIRE_PTR(emis);
    REQUIRE_PTR(flux_up);
    REQUIRE_PTR(flux_down);
    TRY(check_against_optics(lw, optics));
    uint64_t const n = lw->grid.n;
    size_t const bytes = sizeof(fp_t)*n*(size_t)lw->num_levels;
    if (lw->flux_up == NULL)
    {
        GPU(grt_malloc(lw->device, (void **)&lw->flux_up, bytes));
        GPU(grt_malloc(lw->device, (void **)&lw->flux_down, bytes));
    }
    GPU(grt_copy_to_device(lw->device, lw->emissivity, emis, sizeof(fp_t)*n));
    GPU(grt_lw_fluxes(lw->device, lw->num_levels, lw->grid.w0, lw->grid.dw, n, T_surf, T_layers,
                      T_levels, optics->tau, optics->omega, lw->emissivity, lw->flux_up,
                      lw->flux_down));
    GPU(grt_copy_to_host(lw->device, flux_up, lw->flux_up, bytes));
    GPU(grt_copy_to_host(lw->device, flux_down, lw->flux_down, bytes));
    return GRTCODE_SUCCESS;
}

int calculate_lw_column_window(Longwave_t * const lw, Optics_t const * const gas,
                               fp_t * const pressure, fp_t const T_surf, fp_t * const T_layers,
                               fp_t * const T_levels, fp_t * const emis, uint64_t const first,
                               uint64_t const last, fp_t * const up, fp_t * const down)
{
    REQUIRE_PTR(lw);
    REQUIRE_PTR(gas);
    REQUIRE_PTR(pressure);
    REQUIRE_PTR(T_layers);
    REQUIRE_PTR(T_levels);
    REQUIRE_PTR(up);
    REQUIRE_PTR(down);
    TRY(check_against_optics(lw, gas));
    if (first > last || last >= lw->grid.n)
    {
        FAIL(GRTCODE_RANGE_ERR, "integration window [%llu, %llu] outside the grid of %llu points.",
             (unsigned long long)first, (unsigned long long)last, (unsigned long long)lw->grid.n);
    }
    fp_t const mbtoatm = 0.000986923f; /* rayleigh.c:104 */
    fp_t p_atm[MAX_NUM_LEVELS];
    int i;
    for (i=0; i<lw->num_levels; ++i)
    {
        p_atm[i] = pressure[i]*mbtoatm;
    }
    if (emis != NULL)
    {
        GPU(grt_copy_to_device(lw->device, lw->emissivity, emis, sizeof(fp_t)*lw->grid.n));
    }
    GPU(grt_lw_column(lw->device, lw->num_levels, lw->grid.w0, lw->grid.dw, lw->grid.n, p_atm,
                      T_surf, T_layers, T_levels, gas->tau, lw->emissivity, first, last, up, down));
    return GRTCODE_SUCCESS;
}

int calculate_lw_column(Longwave_t * const lw, Optics_t const * const gas, fp_t * const pressure,
                        fp_t const T_surf, fp_t * const T_layers, fp_t * const T_levels,
                        fp_t * const emis, fp_t * const up, fp_t * const down)
{
    REQUIRE_PTR(lw);
    int const rc = calculate_lw_column_window(lw, gas, pressure, T_surf, T_layers, T_levels, emis,
                                              0, lw->grid.n - 1, up, down);
    if (rc != GRTCODE_SUCCESS)
    {
        grt_error_frame(__FILE__, __LINE__);
    }
    return rc;
}
