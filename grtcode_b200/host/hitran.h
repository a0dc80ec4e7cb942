/* HITRAN-2012 line list reader (reference: gas-optics/src/parse_HITRAN_file.h:28-44,
 * parse_HITRAN_file-internal.h:27-41). */
#ifndef GRT_HITRAN_H_
#define GRT_HITRAN_H_
#include <stdint.h>
#include "floating_point_type.h"

typedef struct LineParams
{
    uint64_t num_lines;
    int *iso; /* isotopologue number (1-based; '0' -> 10, 'A' -> 11, ...) */
    fp_t *vnn; /* transition wavenumber [cm-1] */
    fp_t *snn; /* line strength scaled by Q(296)/(exp(c2 E/296)(1 - exp(c2 v/296))) */
    fp_t *yair; /* air-broadened half-width [cm-1 atm-1] (float-rounded) */
    fp_t *yself; /* self-broadened half-width [cm-1 atm-1] (float-rounded) */
    fp_t *en; /* lower-state energy [cm-1] (float-rounded) */
    fp_t *n; /* temperature exponent of the air width (float-rounded) */
    fp_t *d; /* pressure shift [cm-1 atm-1] (float-rounded) */
} LineParams_t;

/* Keeps records of molecule mol_id with w0 <= v <= wn (all when both < 0). */
int parse_hitran_file(LineParams_t * const line_params, char const * const filepath,
                      int const mol_id, double const w0, double const wn);
void free_line_params(LineParams_t * const line_params);
/* Drops the parsed-file tables kept between parse_hitran_file() calls (hitran.c). */
void grt_hitran_cache_clear(void);
/* The device that converts the records from now on (-1: the host threads). */
void grt_hitran_use_device(int const device);

#endif
