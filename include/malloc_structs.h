/* Opaque allocation helpers for the Fortran bindings
 * (reference: fortran-bindings/malloc_structs.c:29-137; bound by
 * fortran-bindings/grtcode_fortran.F90). */
#ifndef MALLOC_STRUCTS_H_
#define MALLOC_STRUCTS_H_
#include "gas_optics.h"
#include "solar_flux.h"

/* type: 0 = SpectralGrid_t, 1 = Optics_t, 2 = GasOptics_t, 3 = SolarFlux_t. */
EXTERN int malloc_struct(void **p, int type);
EXTERN int free_struct(void **p);
EXTERN int spectral_grid_properties(SpectralGrid_t const * const grid, double * const w0,
                                    uint64_t * const n, double * const dw);
EXTERN int solar_flux_properties(SolarFlux_t const * const solar, fp_t * const flux);
#endif
