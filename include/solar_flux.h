/* Incident solar spectrum normalised to unit integral over the grid
 * (reference: shortwave/src/solar_flux.h:27-46). */
#ifndef SOLAR_FLUX_H_
#define SOLAR_FLUX_H_
#include "grtcode_utilities.h"

typedef struct SolarFlux
{
    SpectralGrid_t grid;
    fp_t *incident_flux; /* (wavenumber), host */
    uint64_t n;
} SolarFlux_t;

EXTERN int create_solar_flux(SolarFlux_t * const solar_flux, SpectralGrid_t const * const grid,
                             char const * const filepath);
EXTERN int destroy_solar_flux(SolarFlux_t * const solar_flux);
#endif
