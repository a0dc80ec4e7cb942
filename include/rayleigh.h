/* Rayleigh scattering optics (reference: shortwave/src/rayleigh.h:28). */
#ifndef RAYLEIGH_H_
#define RAYLEIGH_H_
#include "grtcode_utilities.h"

/* pressure [mb] on levels (host); fills optics (omega = 1, g = 0, tau) on its device. */
EXTERN int rayleigh_scattering(Optics_t * const optics, fp_t * const pressure);
#endif
