/* Umbrella header (reference: utilities/src/grtcode_utilities.h). */
#ifndef GRTCODE_UTILITIES_H_
#define GRTCODE_UTILITIES_H_
#include "device.h"
#include "extern.h"
#include "floating_point_type.h"
#include "optics.h"
#include "return_codes.h"
#include "spectral_grid.h"
#include "utilities.h"
#include "verbosity.h"
#endif
