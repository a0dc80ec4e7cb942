/* libshortwave host side: solver objects, Rayleigh optics and the solar
 * spectrum (reference: shortwave/src/shortwave.c:27-64, 506-547;
 * rayleigh.c:100-144; solar_flux.c:27-99). */
#include <stdlib.h>
#include <string.h>
#include "errors.h"
#include "grt_sm100.h"
#include "grtcode_column.h"
#include "parse_csv.h"
#include "rayleigh.h"
#include "shortwave.h"
#include "solar_flux.h"

int create_shortwave(Shortwave_t * const sw, int const num_levels,
                     SpectralGrid_t const * const grid, Device_t const * const device)
{
    REQUIRE_PTR(sw);
    REQUIRE_PTR(grid);
    REQUIRE_PTR(device);
    REQUIRE_GPU(*device);
    REQUIRE_RANGE(num_levels, MIN_NUM_LEVELS, MAX_NUM_LEVELS);
    memset(sw, 0, sizeof(*sw));
    sw->num_levels = num_levels;
    sw->grid = *grid;
    sw->device = *device;
    GPU(grt_malloc(sw->device, (void **)&sw->solar_flux, sizeof(fp_t)*grid->n));
    GPU(grt_malloc(sw->device, (void **)&sw->sfc_alpha_dir, sizeof(fp_t)*grid->n));
    GPU(grt_malloc(sw->device, (void **)&sw->sfc_alpha_dif, sizeof(fp_t)*grid->n));
    return GRTCODE_SUCCESS;
}

int destroy_shortwave(Shortwave_t * const sw)
{
    REQUIRE_PTR(sw);
    GPU(grt_free(sw->device, sw->solar_flux));
    GPU(grt_free(sw->device, sw->sfc_alpha_dir));
    GPU(grt_free(sw->device, sw->sfc_alpha_dif));
    if (sw->flux_up != NULL) GPU(grt_free(sw->device, sw->flux_up));
    if (sw->flux_down != NULL) GPU(grt_free(sw->device, sw->flux_down));
    memset(sw, 0, sizeof(*sw));
    return GRTCODE_SUCCESS;
}

static int check_against_optics(Shortwave_t const *sw, Optics_t const *optics)
{
    REQUIRE_EQUAL(sw->device, optics->device);
    REQUIRE_EQUAL(sw->num_levels, optics->num_layers + 1);
    int same_grids;
    TRY(compare_spectral_grids(&sw->grid, &optics->grid, &same_grids));
    REQUIRE_EQUAL(same_grids, 1);
    return GRTCODE_SUCCESS;
}

static int upload_surface(Shortwave_t *sw, fp_t const *alb_dir, fp_t const *alb_dif,
                          fp_t const *solar)
{
    size_t const bytes = sizeof(fp_t)*sw->grid.n;
    GPU(grt_copy_to_device(sw->device, sw->solar_flux, solar, bytes));
    GPU(grt_copy_to_device(sw->device, sw->sfc_alpha_dir, alb_dir, bytes));
    if (alb_dif == alb_dir)
    {
        GPU(grt_copy_on_device(sw->device, sw->sfc_alpha_dif, sw->sfc_alpha_dir, bytes));
    }
    else
    {
        GPU(grt_copy_to_device(sw->device, sw->sfc_alpha_dif, alb_dif, bytes));
    }
    return GRTCODE_SUCCESS;
}

int calculate_sw_fluxes(Shortwave_t * const sw, Optics_t const * const optics, fp_t const mu_dir,
                        fp_t const mu_dif, fp_t * const sfc_alpha_dir, fp_t * const sfc_alpha_dif,
                        fp_t const total_solar_irradiance, fp_t * const solar_flux,
                        fp_t * const flux_up, fp_t * const flux_down)
{
    REQUIRE_PTR(sw);
    REQUIRE_PTR(optics);
    REQUIRE_PTR(sfc_alpha_dir);
    REQUIRE_PTR(sfc_alpha_dif);
    REQUIRE_PTR(solar_flux);
    REQUIRE_PTR(flux_up);
    REQUIRE_PTR(flux_down);
    TRY(check_against_optics(sw, optics));
    size_t const bytes = sizeof(fp_t)*sw->grid.n*(size_t)sw->num_levels;
    if (sw->flux_up == NULL)
    {
        GPU(grt_malloc(sw->device, (void **)&sw->flux_up, bytes));
        GPU(grt_malloc(sw->device, (void **)&sw->flux_down, bytes));
    }
    TRY(upload_surface(sw, sfc_alpha_dir, sfc_alpha_dif, solar_flux));
    GPU(grt_sw_fluxes(sw->device, sw->num_levels, sw->grid.n, optics->omega, optics->g,
                      optics->tau, mu_dir, mu_dif, sw->sfc_alpha_dir, sw->sfc_alpha_dif,
                      total_solar_irradiance, sw->solar_flux, sw->flux_up, sw->flux_down));
    GPU(grt_copy_to_host(sw->device, flux_up, sw->flux_up, bytes));
    GPU(grt_copy_to_host(sw->device, flux_down, sw->flux_down, bytes));
    return GRTCODE_SUCCESS;
}

int calculate_sw_column_window(Shortwave_t * const sw, Optics_t const * const gas,
                               fp_t * const pressure, fp_t const mu_dir, fp_t const mu_dif,
                               fp_t * const sfc_alpha_dir, fp_t * const sfc_alpha_dif,
                               fp_t const total_solar_irradiance, fp_t * const solar_flux,
                               uint64_t const first, uint64_t const last, fp_t * const up,
                               fp_t * const down)
{
    REQUIRE_PTR(sw);
    REQUIRE_PTR(gas);
    REQUIRE_PTR(pressure);
    REQUIRE_PTR(up);
    REQUIRE_PTR(down);
    TRY(check_against_optics(sw, gas));
    if (first > last || last >= sw->grid.n)
    {
        FAIL(GRTCODE_RANGE_ERR, "integration window [%llu, %llu] outside the grid of %llu points.",
             (unsigned long long)first, (unsigned long long)last, (unsigned long long)sw->grid.n);
    }
    fp_t const mbtoatm = 0.000986923f;
    fp_t p_atm[MAX_NUM_LEVELS];
    int i;
    for (i=0; i<sw->num_levels; ++i)
    {
        p_atm[i] = pressure[i]*mbtoatm;
    }
    if (sfc_alpha_dir != NULL && sfc_alpha_dif != NULL && solar_flux != NULL)
    {
        TRY(upload_surface(sw, sfc_alpha_dir, sfc_alpha_dif, solar_flux));
    }
    GPU(grt_sw_column(sw->device, sw->num_levels, sw->grid.w0, sw->grid.dw, sw->grid.n, p_atm,
                      gas->tau, mu_dir, mu_dif, sw->sfc_alpha_dir, sw->sfc_alpha_dif,
                      total_solar_irradiance, sw->solar_flux, first, last, up, down));
    return GRTCODE_SUCCESS;
}

int calculate_sw_column(Shortwave_t * const sw, Optics_t const * const gas, fp_t * const pressure,
                        fp_t const mu_dir, fp_t const mu_dif, fp_t * const sfc_alpha_dir,
                        fp_t * const sfc_alpha_dif, fp_t const total_solar_irradiance,
                        fp_t * const solar_flux, fp_t * const up, fp_t * const down)
{
    REQUIRE_PTR(sw);
    int const rc = calculate_sw_column_window(sw, gas, pressure, mu_dir, mu_dif, sfc_alpha_dir,
                                              sfc_alpha_dif, total_solar_irradiance, solar_flux,
                                              0, sw->grid.n - 1, up, down);
    if (rc != GRTCODE_SUCCESS)
    {
        grt_error_frame(__FILE__, __LINE__);
    }
    return rc;
}

int rayleigh_scattering(Optics_t * const optics, fp_t * const pressure)
{
    REQUIRE_PTR(optics);
    REQUIRE_PTR(pressure);
    REQUIRE_GPU(optics->device);
    fp_t const mbtoatm = 0.000986923f;
    fp_t p_atm[MAX_NUM_LEVELS];
    int i;
    for (i=0; i<optics->num_layers+1; ++i)
    {
        p_atm[i] = pressure[i]*mbtoatm;
    }
    GPU(grt_rayleigh(optics->device, optics->num_layers, optics->grid.w0, optics->grid.dw,
                     optics->grid.n, p_atm, optics->omega, optics->g, optics->tau));
    return GRTCODE_SUCCESS;
}

int create_solar_flux(SolarFlux_t * const solar_flux, SpectralGrid_t const * const grid,
                      char const * const filepath)
{
    REQUIRE_PTR(solar_flux);
    REQUIRE_PTR(grid);
    REQUIRE_PTR(filepath);
    LOG_INFO("Reading in solar flux values from file %s.", filepath);
    fp_t *c = NULL;
    TRY(grt_table_on_grid(filepath, 2, grid->w0, grid->dw, grid->n, &c));
    /* Normalise to unit integral with the trapezoid rule on the grid points
       (solar_flux.c:73-84). */
    fp_t *w;
    int rc = grid_points(*grid, &w, HOST_ONLY);
    fp_t total = 0.;
    if (rc == GRTCODE_SUCCESS)
    {
        rc = integrate2(w, c, grid->n, &total, trapezoid);
        free(w);
    }
    if (rc != GRTCODE_SUCCESS)
    {
        free(c);
        grt_error_frame(__FILE__, __LINE__);
        return rc;
    }
    uint64_t j;
    for (j=0; j<grid->n; ++j)
    {
        c[j] /= total;
    }
    solar_flux->grid = *grid;
    solar_flux->incident_flux = c;
    solar_flux->n = grid->n;
    return GRTCODE_SUCCESS;
}

int destroy_solar_flux(SolarFlux_t * const solar_flux)
{
    REQUIRE_PTR(solar_flux);
    REQUIRE_PTR(solar_flux->incident_flux);
    free(solar_flux->incident_flux);
    solar_flux->incident_flux = NULL;
    return GRTCODE_SUCCESS;
}
