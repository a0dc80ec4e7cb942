/* libshortwave: delta-Eddington two-stream + adding shortwave solver
 * (reference: shortwave/src/shortwave.h:28-69). */
#ifndef SHORTWAVE_H_
#define SHORTWAVE_H_
#include <stdint.h>
#include "grtcode_utilities.h"

typedef struct Shortwave
{
    int num_levels;
    SpectralGrid_t grid;
    Device_t device;
    fp_t *solar_flux; /* (wavenumber), device */
    fp_t *sfc_alpha_dir; /* (wavenumber), device */
    fp_t *sfc_alpha_dif; /* (wavenumber), device */
    fp_t *flux_up; /* (level, wavenumber), device */
    fp_t *flux_down; /* (level, wavenumber), device */
} Shortwave_t;

EXTERN int create_shortwave(Shortwave_t * const sw, int const num_levels,
                            SpectralGrid_t const * const grid, Device_t const * const device);
EXTERN int destroy_shortwave(Shortwave_t * const sw);
/* All array arguments are host memory; fluxes [W cm m-2] (level, wavenumber). */
EXTERN int calculate_sw_fluxes(Shortwave_t * const sw, Optics_t const * const optics,
                               fp_t const mu_dir, fp_t const mu_dif, fp_t * const sfc_alpha_dir,
                               fp_t * const sfc_alpha_dif, fp_t const total_solar_irradiance,
                               fp_t * const solar_flux, fp_t * const flux_up,
                               fp_t * const flux_down);
#endif
