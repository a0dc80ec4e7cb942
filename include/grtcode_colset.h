/* ADDITION to the reference API (no counterpart there): dependency-free binary files for
 * the rewired drivers.  The reference's applications read their atmospheres from netCDF
 * (rfmip-irf/src/rfmip-irf.c:162-468, circ/src/circ.c:128-430, era5/src/era5.c:130-640)
 * into an Atmosphere_t (framework/src/driver.h:114-166) and write fluxes back to netCDF;
 * no netCDF library exists on the target machines of this build.  A column set carries
 * the same arrays with the same meaning and units (pressure in mb, abundances in ppmv,
 * cosine of the solar zenith angle), for clear-sky aerosol-free columns with grey surfaces.
 *
 * File layout "GRTCOLS1" (little endian):
 *   char[8] magic; int32 num_times, num_columns, num_levels, num_molecules, num_cfcs,
 *   num_cia_species, num_cias; int32 molecules[num_molecules]; int32 cfcs[num_cfcs];
 *   int32 cia_species[num_cia_species]; int32 cia_pairs[num_cias][2];
 *   char h2o_ctm[1024], o3_ctm[1024] ("none" = off); char cfc_path[num_cfcs][1024];
 *   char cia_path[num_cias][1024];
 *   float64 level_pressure[T][C][levels], level_temperature[T][C][levels],
 *   layer_temperature[T][C][layers], surface_temperature[T][C], cos_zenith[T][C],
 *   total_solar_irradiance[T][C], surface_albedo[T][C], surface_emissivity[T][C],
 *   ppmv[num_molecules][T][C][levels], cfc_ppmv[num_cfcs][T][C][levels],
 *   cia_ppmv[num_cia_species][T][C][levels].
 * File layout "GRTFLUX1": char[8] magic; int32 num_times, num_columns, num_levels, pad;
 *   float64 rlu, rld, rsu, rsd [T][C][levels] (the reference's rlucsaf, rldcsaf, rsucsaf,
 *   rsdcsaf on every level, W m-2); hr_lw, hr_sw [T][C][layers] (K day-1).
 * Python mirror: grtcode_b200/column_io.py. */
#ifndef GRTCODE_COLSET_H_
#define GRTCODE_COLSET_H_
#include <stdint.h>
#include "extern.h"
#include "floating_point_type.h"

#define COLSET_PATH_LEN 1024
#define COLSET_MAX_SPECIES 64

typedef struct ColumnSet
{
    int num_times;
    int num_columns;
    int num_levels;
    int num_molecules;
    int num_cfcs;
    int num_cia_species;
    int num_cias;
    int molecules[COLSET_MAX_SPECIES];      /* HITRAN ids (molecules.h) */
    int cfcs[COLSET_MAX_SPECIES];           /* GRTCODE CFC ids (cfcs.h) */
    int cia_species[COLSET_MAX_SPECIES];    /* GRTCODE CIA species ids */
    int cia_pairs[COLSET_MAX_SPECIES][2];
    char h2o_ctm[COLSET_PATH_LEN];
    char o3_ctm[COLSET_PATH_LEN];
    char (*cfc_path)[COLSET_PATH_LEN];
    char (*cia_path)[COLSET_PATH_LEN];
    fp_t *level_pressure;                   /* [mb] (time, column, level) */
    fp_t *level_temperature;                /* [K] */
    fp_t *layer_temperature;                /* [K] (time, column, layer) */
    fp_t *surface_temperature;              /* [K] (time, column) */
    fp_t *cos_zenith;                       /* cosine of the solar zenith angle */
    fp_t *total_solar_irradiance;           /* [W m-2] */
    fp_t *surface_albedo;                   /* grey */
    fp_t *surface_emissivity;               /* grey */
    fp_t *ppmv;                             /* (molecule, time, column, level) */
    fp_t *cfc_ppmv;                         /* (cfc, time, column, level) */
    fp_t *cia_ppmv;                         /* (species, time, column, level) */
} ColumnSet_t;

typedef struct FluxSet
{
    int num_times;
    int num_columns;
    int num_levels;
    fp_t *rlu;                              /* (time, column, level) [W m-2] */
    fp_t *rld;
    fp_t *rsu;
    fp_t *rsd;
    fp_t *hr_lw;                            /* (time, column, layer) [K day-1] */
    fp_t *hr_sw;
} FluxSet_t;

EXTERN int read_column_set(char const * const path, ColumnSet_t * const cs);
EXTERN int destroy_column_set(ColumnSet_t * const cs);
EXTERN int create_flux_set(FluxSet_t * const fs, int const num_times, int const num_columns,
                           int const num_levels);
EXTERN int write_flux_set(char const * const path, FluxSet_t const * const fs);
EXTERN int read_flux_set(char const * const path, FluxSet_t * const fs);
EXTERN int destroy_flux_set(FluxSet_t * const fs);
#endif
