/* libgas_optics: line-by-line gas optics (reference: gas-optics/src/gas_optics.h:39-180).
 *
 * Same entry points, argument meaning and error behaviour as the reference;
 * the work is done by hand-written sm_100a CUDA kernels (include/grt_sm100.h).
 * Differences a caller can observe:
 *   - `device` must be a GPU id; HOST_ONLY returns GRTCODE_VALUE_ERR;
 *   - optical_depth_method: all three methods run on the GPU (NULL -> wavenumber_sweep as in
 *     the reference, gas_optics.c:73-76); where the reference's binned methods read
 *     uninitialised memory the result is the empty range;
 *   - GasOptics_t members differ (the struct is library-owned state; callers
 *     allocate it, e.g. through malloc_struct, and never touch the fields). */
#ifndef GAS_OPTICS_H_
#define GAS_OPTICS_H_
#include <stdint.h>
#include "cfcs.h"
#include "collision_induced_absorption.h"
#include "grtcode_utilities.h"
#include "molecules.h"

#define DIR_PATH_LEN 1024

typedef struct GasOptics
{
    Device_t device;
    int num_levels;
    int num_layers;
    int num_molecules;
    uint64_t molecule_bit_field;
    int molecule_ids[NUM_MOLS]; /* in the order they were added */
    uint64_t num_lines[NUM_MOLS]; /* lines kept per added molecule */
    int num_cfcs;
    uint64_t cfc_bit_field;
    int num_cias;
    uint64_t cia_bit_field;
    int cia_pairs[MAX_NUM_CIAS][2];
    char h2o_ctm_dir[DIR_PATH_LEN];
    int use_h2o_ctm;
    char o3_ctm_file[DIR_PATH_LEN];
    int use_o3_ctm;
    SpectralGrid_t grid;
    char hitran_path[DIR_PATH_LEN];
    double wcutoff; /* stored, unused by the kernels -- as in the reference (kernels.c:417) */
    int optical_depth_method;
    fp_t *x; /* mole fraction (molecule, level), host */
    fp_t *x_cfc; /* (CFC, level), host */
    fp_t *x_cia; /* (CIA species, level), host */
    void *sm100; /* device-side state (grt_gas) */
} GasOptics_t;

enum OpticalDepthMethod
{
    wavenumber_sweep,
    line_sweep,
    line_sample
};

EXTERN int create_gas_optics(GasOptics_t * const gas_optics, int const num_levels,
                             SpectralGrid_t const * const grid, Device_t const * const device,
                             char const * const hitran_path, char const * const h2o_ctm_dir,
                             char const * const o3_ctm_file, double const * const wcutoff,
                             int const * const optical_depth_method);
EXTERN int destroy_gas_optics(GasOptics_t * const gas_optics);
EXTERN int add_molecule(GasOptics_t * const gas_optics, int const molecule_id,
                        double const * const min_line_center,
                        double const * const max_line_center);
EXTERN int set_molecule_ppmv(GasOptics_t * const gas_optics, int const molecule_id,
                             fp_t const * const ppmv);
EXTERN int add_cfc(GasOptics_t * const gas_optics, int const cfc_id,
                   char const * const filepath);
EXTERN int set_cfc_ppmv(GasOptics_t * const gas_optics, int const cfc_id,
                        fp_t const * const ppmv);
EXTERN int add_cia(GasOptics_t * const gas_optics, int const species1, int const species2,
                   char const * const filepath);
EXTERN int set_cia_ppmv(GasOptics_t * const gas_optics, int const cia_id,
                        fp_t const * const ppmv);
/* pressure [mb] and temperature [K] on levels (host); optics->tau (device) is overwritten. */
EXTERN int calculate_optical_depth(GasOptics_t * const gas_optics, fp_t * const pressure,
                                   fp_t * const temperature, Optics_t * const optics);
EXTERN int get_num_molecules(GasOptics_t const * const gas_optics, int * const n);
#endif
