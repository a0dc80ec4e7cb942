/* Optical properties of a column on a spectral grid
 * (reference: utilities/src/optics.h:30-87).  The three arrays live in the
 * memory of GPU `device`, laid out [layer][wavenumber]. */
#ifndef OPTICS_H_
#define OPTICS_H_
#include "device.h"
#include "extern.h"
#include "floating_point_type.h"
#include "spectral_grid.h"

typedef struct Optics
{
    Device_t device;
    fp_t *g; /* asymmetry factor (layer, wavenumber), device memory */
    SpectralGrid_t grid;
    int num_layers;
    fp_t *omega; /* single-scatter albedo (layer, wavenumber), device memory */
    fp_t *tau; /* optical depth (layer, wavenumber), device memory */
} Optics_t;

EXTERN int create_optics(Optics_t * const optics, int const num_layers,
                         SpectralGrid_t const * const grid, Device_t const * const device);
EXTERN int destroy_optics(Optics_t * const optics);
/* result is created by the call (tau = sum, omega and g tau-weighted). */
EXTERN int add_optics(Optics_t const * const * const optics, int const num_optics,
                      Optics_t * const result);
EXTERN int optics_compatible(Optics_t const * const one, Optics_t const * const two,
                             int * const result);
EXTERN int sample_optics(Optics_t * const dest, Optics_t const * const source,
                         double const * const w0, double const * const wn);
/* Host arrays (layer, wavenumber) copied into the object. */
EXTERN int update_optics(Optics_t * const optics, fp_t const * const tau,
                         fp_t const * const omega, fp_t const * const g);
/* Device -> host copies of the three arrays; NULL pointers are skipped
 * (reference: fortran-bindings/malloc_structs.c:78-97). */
EXTERN int optical_properties(Optics_t const * const optics, fp_t * const tau,
                              fp_t * const omega, fp_t * const g);
#endif
