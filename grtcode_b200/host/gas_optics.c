/* libgas_optics host side: bookkeeping, file parsing and table resampling in
 * C99; all arithmetic on the spectral grid is delegated to the sm_100a layer
 * (include/grt_sm100.h).  Public behaviour follows
 * gas-optics/src/gas_optics.c:51-464 of the reference. */
#include <stdlib.h>
#include <string.h>
#include "errors.h"
#include "gas_optics.h"
#include "grt_sm100.h"
#include "grtcode_column.h"
#include "hitran.h"
#include "parse_csv.h"
#include "tables.h"
#include "tips.h"

static double const MIN_CUTOFF = 1.;
static double const MAX_CUTOFF = 50.;
static double const DEFAULT_CUTOFF = 25.;
static uint64_t const MAX_NUM_LINES = 600000; /* per molecule, gas_optics.c:46 */

static int molecule_index(int const molecule_id, int * const index)
{
    if (molecule_id < H2O || molecule_id > NUM_MOLS)
    {
        FAIL(GRTCODE_VALUE_ERR, "unrecognized molecule id %d.", molecule_id);
    }
    *index = molecule_id - 1;
    return GRTCODE_SUCCESS;
}

int create_gas_optics(GasOptics_t * const gas_optics, int const num_levels,
                      SpectralGrid_t const * const grid, Device_t const * const device,
                      char const * const hitran_path, char const * const h2o_ctm_dir,
                      char const * const o3_ctm_file, double const * const wcutoff,
                      int const * const optical_depth_method)
{
    REQUIRE_PTR(gas_optics);
    REQUIRE_RANGE(num_levels, MIN_NUM_LEVELS, MAX_NUM_LEVELS);
    REQUIRE_PTR(grid);
    REQUIRE_PTR(device);
    REQUIRE_GPU(*device);
    REQUIRE_PTR(hitran_path);
    memset(gas_optics, 0, sizeof(*gas_optics));
    gas_optics->num_levels = num_levels;
    gas_optics->num_layers = num_levels - 1;
    gas_optics->grid = *grid;
    gas_optics->device = *device;
    grt_hitran_use_device(*device);
    TRY(copy_str(gas_optics->hitran_path, hitran_path, DIR_PATH_LEN));
    LOG_INFO("Using HITRAN database file %s.", gas_optics->hitran_path);
    gas_optics->wcutoff = DEFAULT_CUTOFF;
    if (wcutoff != NULL)
    {
        REQUIRE_RANGE(*wcutoff, MIN_CUTOFF, MAX_CUTOFF);
        gas_optics->wcutoff = *wcutoff;
    }
    gas_optics->optical_depth_method = wavenumber_sweep;
    if (optical_depth_method != NULL)
    {
        REQUIRE_RANGE(*optical_depth_method, wavenumber_sweep, line_sample);
        gas_optics->optical_depth_method = *optical_depth_method;
    }
    if (h2o_ctm_dir != NULL && strcmp(h2o_ctm_dir, "none") != 0)
    {
        gas_optics->use_h2o_ctm = 1;
        TRY(copy_str(gas_optics->h2o_ctm_dir, h2o_ctm_dir, DIR_PATH_LEN));
    }
    if (o3_ctm_file != NULL && strcmp(o3_ctm_file, "none") != 0)
    {
        gas_optics->use_o3_ctm = 1;
        TRY(copy_str(gas_optics->o3_ctm_file, o3_ctm_file, DIR_PATH_LEN));
    }
    gas_optics->x = (fp_t *)calloc((size_t)num_levels*NUM_MOLS, sizeof(fp_t));
    gas_optics->x_cfc = (fp_t *)calloc((size_t)num_levels*NUM_CFCS, sizeof(fp_t));
    gas_optics->x_cia = (fp_t *)calloc((size_t)num_levels*NUM_CIAS, sizeof(fp_t));
    if (gas_optics->x == NULL || gas_optics->x_cfc == NULL || gas_optics->x_cia == NULL)
    {
        FAIL(GRTCODE_NULL_ERR, "out of memory for %d levels.", num_levels);
    }
    grt_gas *dev = NULL;
    GPU(grt_gas_create(&dev, *device, num_levels, grid->w0, grid->dw, grid->n));
    gas_optics->sm100 = dev;
    return GRTCODE_SUCCESS;
}

int destroy_gas_optics(GasOptics_t * const gas_optics)
{
    REQUIRE_PTR(gas_optics);
    free(gas_optics->x);
    free(gas_optics->x_cfc);
    free(gas_optics->x_cia);
    gas_optics->x = gas_optics->x_cfc = gas_optics->x_cia = NULL;
    if (gas_optics->sm100 != NULL)
    {
        GPU(grt_gas_destroy((grt_gas *)gas_optics->sm100));
        gas_optics->sm100 = NULL;
    }
    return GRTCODE_SUCCESS;
}

/* Reads <dir>/<name> and checks it; the samples go to the device, which resamples them to the
 * grid (grt_sm100.h: grt_gas_*_samples). */
static int table_samples(char const *dir, char const *name, int cols, fp_t **table, int *rows)
{
    char path[DIR_PATH_LEN + 64];
    if (dir != NULL)
    {
        snprintf(path, sizeof(path), "%s/%s", dir, name);
    }
    else
    {
        snprintf(path, sizeof(path), "%s", name);
    }
    TRY(grt_table_samples(path, cols, table, rows));
    return GRTCODE_SUCCESS;
}

/* Reads <dir>/<name>, checks its column count and resamples column 1 to the grid on the HOST
 * (kept for the tests that compare the device resampler with it). */
int grt_continuum_table_host(GasOptics_t const *go, char const *dir, char const *name,
                             int cols, fp_t **out);
int grt_continuum_table_host(GasOptics_t const *go, char const *dir, char const *name,
                             int cols, fp_t **out)
{
    char path[DIR_PATH_LEN + 64];
    if (dir != NULL)
    {
        snprintf(path, sizeof(path), "%s/%s", dir, name);
    }
    else
    {
        snprintf(path, sizeof(path), "%s", name);
    }
    TRY(grt_table_on_grid(path, cols, go->grid.w0, go->grid.dw, go->grid.n, out));
    return GRTCODE_SUCCESS;
}

int add_molecule(GasOptics_t * const gas_optics, int const molecule_id,
                 double const * const min_line_center, double const * const max_line_center)
{
    REQUIRE_PTR(gas_optics);
    int id;
    TRY(molecule_index(molecule_id, &id));
    if (is_active(gas_optics->molecule_bit_field, id))
    {
        FAIL(GRTCODE_VALUE_ERR, "molecule %d has already been added.", molecule_id);
    }
    double w0 = gas_optics->grid.w0;
    if (min_line_center != NULL)
    {
        REQUIRE_RANGE(*min_line_center, MIN_WAVENUMBER, MAX_WAVENUMBER);
        w0 = *min_line_center;
    }
    double wn = gas_optics->grid.wn;
    if (max_line_center != NULL)
    {
        REQUIRE_RANGE(*max_line_center, MIN_WAVENUMBER, MAX_WAVENUMBER);
        wn = *max_line_center;
    }
    if (wn < w0)
    {
        FAIL(GRTCODE_RANGE_ERR, "value (%e) less than minimum allowed (%e).", wn, w0);
    }
    char const *name;
    double mass;
    int num_iso;
    if (!grt_molecule_info(molecule_id, &name, &mass, &num_iso))
    {
        FAIL(GRTCODE_VALUE_ERR, "unrecognized molecule id %d.", molecule_id);
    }
    LineParams_t lp;
    TRY(parse_hitran_file(&lp, gas_optics->hitran_path, molecule_id, w0, wn));
    LOG_MESG("Using %s (%llu lines in range %e - %e [1/cm]).", name,
             (unsigned long long)lp.num_lines, w0, wn);
    if (lp.num_lines > MAX_NUM_LINES - 1)
    {
        unsigned long long const got = lp.num_lines;
        free_line_params(&lp);
        FAIL(GRTCODE_RANGE_ERR, "value (%e) greater than maximum allowed (%e).", (double)got,
             (double)(MAX_NUM_LINES - 1));
    }

    /* tips2017.c is absent from the reference mount: unless an external table supplied this
       molecule's partition sums (GRT_TIPS_TABLE, host/tips.c) they are a substitute that only
       reproduces the reference's test pins.  Say so once per molecule and process. */
    {
        static unsigned char warned[64];
        if (grt_tips_is_substitute(molecule_id) && molecule_id < 64 && !warned[molecule_id])
        {
            warned[molecule_id] = 1;
            LOG_WARN("partition sums Q(T) of %s are a SUBSTITUTE for TIPS-2017 (its source is "
                     "not part of this build): line strengths are not HITRAN-accurate; set "
                     "GRT_TIPS_TABLE=<file of 'mol iso T Q' rows> to supply real values.", name);
        }
    }

    /* Partition-sum tables of every isotopologue, shared verbatim with the device. */
    double *q = (double *)malloc(sizeof(double)*(size_t)(num_iso > 0 ? num_iso : 1)*GRT_TIPS_NT);
    if (q == NULL)
    {
        free_line_params(&lp);
        FAIL(GRTCODE_NULL_ERR, "out of memory for partition sums of %s.", name);
    }
    int k;
    for (k=0; k<num_iso; ++k)
    {
        memcpy(q + (size_t)k*GRT_TIPS_NT, grt_tips_table(molecule_id, k + 1),
               sizeof(double)*GRT_TIPS_NT);
    }
    int const rc = grt_gas_add_molecule((grt_gas *)gas_optics->sm100, molecule_id, mass, num_iso,
                                        lp.num_lines, lp.iso, lp.vnn, lp.snn, lp.yair, lp.yself,
                                        lp.en, lp.n, lp.d, q, GRT_TIPS_TMIN, GRT_TIPS_NT);
    free(q);
    uint64_t const kept = lp.num_lines;
    free_line_params(&lp);
    if (rc != 0)
    {
        return grt_gpu_failure(__FILE__, __LINE__);
    }
    int const index = gas_optics->num_molecules++;
    gas_optics->molecule_ids[index] = molecule_id;
    gas_optics->num_lines[index] = kept;
    TRY(activate(&gas_optics->molecule_bit_field, id));

    if (molecule_id == H2O && gas_optics->use_h2o_ctm)
    {
        /* water_vapor_continuum.c:49-64: two 2-column and two 4-column tables. */
        LOG_MESG("Using the %s continuum.", name);
        /* S296, CKDS, F296, CKDF: the argument order of launch.c:165-170 */
        static char const * const files[4] = {"296MTCKD25_S.csv", "CKDS.csv", "296MTCKD25_F.csv",
                                              "CKDF.csv"};
        static int const cols[4] = {2, 4, 2, 4};
        fp_t *t[4] = {NULL, NULL, NULL, NULL};
        fp_t const *x[4];
        fp_t const *y[4];
        uint64_t m[4];
        int r = GRTCODE_SUCCESS;
        for (k=0; k<4 && r == GRTCODE_SUCCESS; ++k)
        {
            int rows = 0;
            r = table_samples(gas_optics->h2o_ctm_dir, files[k], cols[k], &t[k], &rows);
            if (r == GRTCODE_SUCCESS)
            {
                x[k] = t[k];
                y[k] = t[k] + rows;        /* first value column only */
                m[k] = (uint64_t)rows;
            }
        }
        int g = 0;
        if (r == GRTCODE_SUCCESS)
        {
            g = grt_gas_set_h2o_continuum_samples((grt_gas *)gas_optics->sm100, x, y, m);
        }
        for (k=0; k<4; ++k) free(t[k]);
        if (r != GRTCODE_SUCCESS)
        {
            grt_error_frame(__FILE__, __LINE__);
            return r;
        }
        if (g != 0) return grt_gpu_failure(__FILE__, __LINE__);
    }
    if (molecule_id == O3 && gas_optics->use_o3_ctm)
    {
        LOG_MESG("Using the %s continuum.", name);
        fp_t *t = NULL;
        int rows = 0;
        TRY(table_samples(NULL, gas_optics->o3_ctm_file, 2, &t, &rows));
        int const g = grt_gas_set_o3_continuum_samples((grt_gas *)gas_optics->sm100, t, t + rows,
                                                       (uint64_t)rows);
        free(t);
        if (g != 0) return grt_gpu_failure(__FILE__, __LINE__);
    }
    return GRTCODE_SUCCESS;
}

/* ppmv -> mole fraction into row `row` of a (species, level) host array. */
static void store_fraction(fp_t *dest, int row, int num_levels, fp_t const *ppmv)
{
    int i;
    for (i=0; i<num_levels; ++i)
    {
        dest[(size_t)row*num_levels + i] = ppmv[i]*1.e-6;
    }
}

int set_molecule_ppmv(GasOptics_t * const gas_optics, int const molecule_id,
                      fp_t const * const ppmv)
{
    REQUIRE_PTR(gas_optics);
    REQUIRE_PTR(ppmv);
    int index;
    TRY(molecule_index(molecule_id, &index));
    if (!is_active(gas_optics->molecule_bit_field, index))
    {
        LOG_WARN("molecule %d is not being used.", molecule_id);
        return GRTCODE_SUCCESS;
    }
    store_fraction(gas_optics->x, index, gas_optics->num_levels, ppmv);
    return GRTCODE_SUCCESS;
}

int add_cfc(GasOptics_t * const gas_optics, int const cfc_id, char const * const filepath)
{
    REQUIRE_PTR(gas_optics);
    REQUIRE_PTR(filepath);
    REQUIRE_RANGE(cfc_id, 0, NUM_CFCS - 1);
    if (is_active(gas_optics->cfc_bit_field, cfc_id))
    {
        FAIL(GRTCODE_VALUE_ERR, "cfc %d has already been added.", cfc_id);
    }
    fp_t *t = NULL;
    int rows = 0;
    TRY(table_samples(NULL, filepath, 2, &t, &rows));
    int const g = grt_gas_add_cfc_samples((grt_gas *)gas_optics->sm100, cfc_id, t, t + rows,
                                          (uint64_t)rows);
    free(t);
    if (g != 0) return grt_gpu_failure(__FILE__, __LINE__);
    gas_optics->num_cfcs++;
    TRY(activate(&gas_optics->cfc_bit_field, cfc_id));
    LOG_MESG("Using CFC %s.", grt_cfc_name(cfc_id));
    return GRTCODE_SUCCESS;
}

int set_cfc_ppmv(GasOptics_t * const gas_optics, int const cfc_id, fp_t const * const ppmv)
{
    REQUIRE_PTR(gas_optics);
    REQUIRE_PTR(ppmv);
    REQUIRE_RANGE(cfc_id, 0, NUM_CFCS - 1);
    if (!is_active(gas_optics->cfc_bit_field, cfc_id))
    {
        LOG_WARN("CFC %d is not being used.", cfc_id);
        return GRTCODE_SUCCESS;
    }
    store_fraction(gas_optics->x_cfc, cfc_id, gas_optics->num_levels, ppmv);
    return GRTCODE_SUCCESS;
}

int add_cia(GasOptics_t * const gas_optics, int const species1, int const species2,
            char const * const filepath)
{
    REQUIRE_PTR(gas_optics);
    REQUIRE_PTR(filepath);
    REQUIRE_RANGE(species1, 0, NUM_CIAS - 1);
    REQUIRE_RANGE(species2, 0, NUM_CIAS - 1);
    int i;
    for (i=0; i<gas_optics->num_cias; ++i)
    {
        /* Same pair test as the reference (gas_optics.c:379): the sum of the two ids. */
        if (gas_optics->cia_pairs[i][0] + gas_optics->cia_pairs[i][1] == species1 + species2)
        {
            FAIL(GRTCODE_VALUE_ERR, "CIA with %s and %s is already active.",
                 grt_cia_name(gas_optics->cia_pairs[i][0]),
                 grt_cia_name(gas_optics->cia_pairs[i][1]));
        }
    }
    if (gas_optics->num_cias >= MAX_NUM_CIAS)
    {
        FAIL(GRTCODE_RANGE_ERR, "at most %d CIA pairs are supported.", MAX_NUM_CIAS);
    }
    fp_t *t = NULL;
    int rows = 0;
    TRY(table_samples(NULL, filepath, 2, &t, &rows));
    int const g = grt_gas_add_cia_samples((grt_gas *)gas_optics->sm100, species1, species2, t,
                                          t + rows, (uint64_t)rows);
    free(t);
    if (g != 0) return grt_gpu_failure(__FILE__, __LINE__);
    int const index = gas_optics->num_cias++;
    gas_optics->cia_pairs[index][0] = species1;
    gas_optics->cia_pairs[index][1] = species2;
    TRY(activate(&gas_optics->cia_bit_field, species1));
    TRY(activate(&gas_optics->cia_bit_field, species2));
    LOG_INFO("Using collision-induced absorption between %s and %s.", grt_cia_name(species1),
             grt_cia_name(species2));
    return GRTCODE_SUCCESS;
}

int set_cia_ppmv(GasOptics_t * const gas_optics, int const cia_id, fp_t const * const ppmv)
{
    REQUIRE_PTR(gas_optics);
    REQUIRE_PTR(ppmv);
    REQUIRE_RANGE(cia_id, 0, NUM_CIAS - 1);
    if (!is_active(gas_optics->cia_bit_field, cia_id))
    {
        LOG_WARN("CIA %d is not being used.", cia_id);
        return GRTCODE_SUCCESS;
    }
    store_fraction(gas_optics->x_cia, cia_id, gas_optics->num_levels, ppmv);
    return GRTCODE_SUCCESS;
}

int calculate_optical_depth(GasOptics_t * const gas_optics, fp_t * const pressure,
                            fp_t * const temperature, Optics_t * const optics)
{
    REQUIRE_PTR(gas_optics);
    REQUIRE_PTR(pressure);
    REQUIRE_PTR(temperature);
    REQUIRE_PTR(optics);
    REQUIRE_EQUAL(gas_optics->device, optics->device);
    REQUIRE_EQUAL(gas_optics->num_layers, optics->num_layers);
    int same_grids;
    TRY(compare_spectral_grids(&gas_optics->grid, &optics->grid, &same_grids));
    REQUIRE_EQUAL(same_grids, 1);
    fp_t const mbtoatm = 0.000986923f;
    fp_t p[MAX_NUM_LEVELS];
    int i;
    for (i=0; i<gas_optics->num_levels; ++i)
    {
        p[i] = pressure[i]*mbtoatm;
    }
    GPU(grt_gas_optical_depth((grt_gas *)gas_optics->sm100, p, temperature, gas_optics->x,
                              gas_optics->x_cfc, gas_optics->x_cia,
                              gas_optics->optical_depth_method, optics->tau));
    return GRTCODE_SUCCESS;
}

int get_num_molecules(GasOptics_t const * const gas_optics, int * const n)
{
    REQUIRE_PTR(gas_optics);
    REQUIRE_PTR(n);
    *n = gas_optics->num_molecules;
    return GRTCODE_SUCCESS;
}

int gas_optics_stats(GasOptics_t const * const gas_optics, uint64_t out[8])
{
    REQUIRE_PTR(gas_optics);
    REQUIRE_PTR(out);
    GPU(grt_gas_last_stats((grt_gas *)gas_optics->sm100, out));
    return GRTCODE_SUCCESS;
}

int gas_optics_patch_counts(GasOptics_t const * const gas_optics, uint64_t out[2])
{
    REQUIRE_PTR(gas_optics);
    REQUIRE_PTR(out);
    GPU(grt_gas_patch_counts((grt_gas const *)gas_optics->sm100, out));
    return GRTCODE_SUCCESS;
}

int gas_optics_debug_records(GasOptics_t const * const gas_optics, int const layer,
                             uint64_t const first, uint64_t const count, float * const repwid,
                             float * const y, double * const centre)
{
    REQUIRE_PTR(gas_optics);
    GPU(grt_gas_debug_records((grt_gas *)gas_optics->sm100, layer, first, count, repwid, y,
                              centre));
    return GRTCODE_SUCCESS;
}

int gas_optics_debug_line_order(GasOptics_t const * const gas_optics, uint64_t const count,
                                uint64_t * const order)
{
    REQUIRE_PTR(gas_optics);
    REQUIRE_PTR(order);
    GPU(grt_gas_debug_line_order((grt_gas const *)gas_optics->sm100, count, order));
    return GRTCODE_SUCCESS;
}

int gas_optics_class_histogram(GasOptics_t const * const gas_optics, uint64_t out[8])
{
    REQUIRE_PTR(gas_optics);
    REQUIRE_PTR(out);
    GPU(grt_gas_class_histogram((grt_gas *)gas_optics->sm100, out));
    return GRTCODE_SUCCESS;
}

int gas_optics_debug_table(GasOptics_t const * const gas_optics, int const kind, int const index,
                           double * const out)
{
    REQUIRE_PTR(gas_optics);
    REQUIRE_PTR(out);
    GPU(grt_gas_debug_table((grt_gas *)gas_optics->sm100, kind, index, out));
    return GRTCODE_SUCCESS;
}
