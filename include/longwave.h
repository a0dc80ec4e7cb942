/* liblongwave: four-stream no-scattering longwave solver
 * (reference: longwave/src/longwave.h:29-68). */
#ifndef LONGWAVE_H_
#define LONGWAVE_H_
#include <stdint.h>
#include "grtcode_utilities.h"

typedef struct Longwave
{
    int num_levels;
    SpectralGrid_t grid;
    Device_t device;
    fp_t *layer_temperature; /* unused placeholder (kept for layout familiarity) */
    fp_t *level_temperature; /* unused placeholder */
    fp_t *emissivity; /* (wavenumber), device */
    fp_t *flux_up; /* (level, wavenumber), device */
    fp_t *flux_down; /* (level, wavenumber), device */
} Longwave_t;

EXTERN int create_longwave(Longwave_t * const lw, int const num_levels,
                           SpectralGrid_t const * const grid, Device_t const * const device);
EXTERN int destroy_longwave(Longwave_t * const lw);
/* All array arguments are host memory; fluxes [W cm m-2] (level, wavenumber). */
EXTERN int calculate_lw_fluxes(Longwave_t * const lw, Optics_t const * const optics,
                               fp_t const T_surf, fp_t * const T_layers, fp_t * const T_levels,
                               fp_t * const emis, fp_t * const flux_up, fp_t * const flux_down);
#endif
