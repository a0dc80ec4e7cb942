/* Logging level and error text (reference: utilities/src/verbosity.h:28-52). */
#ifndef VERBOSITY_H_
#define VERBOSITY_H_
#include "extern.h"

enum grtcode_verbosity_levels
{
    GRTCODE_NONE = 0,
    GRTCODE_ERROR,
    GRTCODE_WARN,
    GRTCODE_INFO
};

/* Copies the accumulated error message + backtrace ("No errors." for code 0). */
EXTERN int grtcode_errstr(int const code, char * const buffer, int const buffer_size);
EXTERN void grtcode_set_verbosity(int const level);
EXTERN int grtcode_verbosity(void);
#endif
