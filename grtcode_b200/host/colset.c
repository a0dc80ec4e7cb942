/* Column-set / flux-set files of the headless drivers (include/grtcode_colset.h): the
 * dependency-free stand-in for the netCDF readers and writers of the reference's
 * applications (rfmip-irf/src/rfmip-irf.c:162-468, circ/src/circ.c:128-430). */
#include <stdlib.h>
#include <string.h>
#include "errors.h"
#include "grtcode_colset.h"

static char const colset_magic[8] = {'G', 'R', 'T', 'C', 'O', 'L', 'S', '1'};
static char const flux_magic[8] = {'G', 'R', 'T', 'F', 'L', 'U', 'X', '1'};

static int read_exact(FILE *f, void *buf, size_t bytes, char const *path, char const *what)
{
    if (bytes > 0 && fread(buf, 1, bytes, f) != bytes)
    {
        FAIL(GRTCODE_IO_ERR, "file %s is truncated (reading %s).", path, what);
    }
    return GRTCODE_SUCCESS;
}

static int read_doubles(FILE *f, fp_t **out, size_t count, char const *path, char const *what)
{
    *out = (fp_t *)malloc(sizeof(fp_t)*(count > 0 ? count : 1));
    REQUIRE_PTR(*out);
    TRY(read_exact(f, *out, sizeof(fp_t)*count, path, what));
    return GRTCODE_SUCCESS;
}

int read_column_set(char const * const path, ColumnSet_t * const cs)
{
    REQUIRE_PTR(path);
    REQUIRE_PTR(cs);
    memset(cs, 0, sizeof(*cs));
    FILE *f = fopen(path, "rb");
    if (f == NULL)
    {
        FAIL(GRTCODE_IO_ERR, "cannot open column set %s.", path);
    }
    char magic[8];
    int32_t head[7];
    int rc = read_exact(f, magic, 8, path, "magic");
    if (rc == GRTCODE_SUCCESS && memcmp(magic, colset_magic, 8) != 0)
    {
        fclose(f);
        FAIL(GRTCODE_VALUE_ERR, "%s is not a GRTCOLS1 column set.", path);
    }
    if (rc == GRTCODE_SUCCESS) rc = read_exact(f, head, sizeof(head), path, "header");
    if (rc != GRTCODE_SUCCESS)
    {
        fclose(f);
        return rc;
    }
    cs->num_times = head[0];
    cs->num_columns = head[1];
    cs->num_levels = head[2];
    cs->num_molecules = head[3];
    cs->num_cfcs = head[4];
    cs->num_cia_species = head[5];
    cs->num_cias = head[6];
    if (cs->num_times < 1 || cs->num_columns < 1 || cs->num_levels < MIN_NUM_LEVELS ||
        cs->num_levels > MAX_NUM_LEVELS || cs->num_molecules < 0 ||
        cs->num_molecules > COLSET_MAX_SPECIES || cs->num_cfcs < 0 ||
        cs->num_cfcs > COLSET_MAX_SPECIES || cs->num_cia_species < 0 ||
        cs->num_cia_species > COLSET_MAX_SPECIES || cs->num_cias < 0 ||
        cs->num_cias > COLSET_MAX_SPECIES)
    {
        fclose(f);
        FAIL(GRTCODE_RANGE_ERR, "column set %s has an invalid header (%d times, %d columns, %d "
             "levels, %d molecules, %d CFCs, %d CIA species, %d CIA pairs).", path, cs->num_times,
             cs->num_columns, cs->num_levels, cs->num_molecules, cs->num_cfcs,
             cs->num_cia_species, cs->num_cias);
    }
    int32_t ids[2*COLSET_MAX_SPECIES];
    int i;
#define READ_OR_CLOSE(call) \
    do { \
        rc = (call); \
        if (rc != GRTCODE_SUCCESS) \
        { \
            fclose(f); \
            destroy_column_set(cs); \
            grt_error_frame(__FILE__, __LINE__); \
            return rc; \
        } \
    } while (0)
    READ_OR_CLOSE(read_exact(f, ids, sizeof(int32_t)*cs->num_molecules, path, "molecule ids"));
    for (i=0; i<cs->num_molecules; ++i) cs->molecules[i] = ids[i];
    READ_OR_CLOSE(read_exact(f, ids, sizeof(int32_t)*cs->num_cfcs, path, "CFC ids"));
    for (i=0; i<cs->num_cfcs; ++i) cs->cfcs[i] = ids[i];
    READ_OR_CLOSE(read_exact(f, ids, sizeof(int32_t)*cs->num_cia_species, path, "CIA species"));
    for (i=0; i<cs->num_cia_species; ++i) cs->cia_species[i] = ids[i];
    READ_OR_CLOSE(read_exact(f, ids, sizeof(int32_t)*2*cs->num_cias, path, "CIA pairs"));
    for (i=0; i<cs->num_cias; ++i)
    {
        cs->cia_pairs[i][0] = ids[2*i];
        cs->cia_pairs[i][1] = ids[2*i+1];
    }
    READ_OR_CLOSE(read_exact(f, cs->h2o_ctm, COLSET_PATH_LEN, path, "h2o_ctm"));
    READ_OR_CLOSE(read_exact(f, cs->o3_ctm, COLSET_PATH_LEN, path, "o3_ctm"));
    cs->h2o_ctm[COLSET_PATH_LEN-1] = '\0';
    cs->o3_ctm[COLSET_PATH_LEN-1] = '\0';
    cs->cfc_path = malloc((size_t)COLSET_PATH_LEN*(cs->num_cfcs > 0 ? cs->num_cfcs : 1));
    cs->cia_path = malloc((size_t)COLSET_PATH_LEN*(cs->num_cias > 0 ? cs->num_cias : 1));
    if (cs->cfc_path == NULL || cs->cia_path == NULL)
    {
        fclose(f);
        destroy_column_set(cs);
        FAIL(GRTCODE_NULL_ERR, "out of memory reading %s.", path);
    }
    READ_OR_CLOSE(read_exact(f, cs->cfc_path, (size_t)COLSET_PATH_LEN*cs->num_cfcs, path,
                             "CFC paths"));
    READ_OR_CLOSE(read_exact(f, cs->cia_path, (size_t)COLSET_PATH_LEN*cs->num_cias, path,
                             "CIA paths"));
    for (i=0; i<cs->num_cfcs; ++i) cs->cfc_path[i][COLSET_PATH_LEN-1] = '\0';
    for (i=0; i<cs->num_cias; ++i) cs->cia_path[i][COLSET_PATH_LEN-1] = '\0';
    size_t const ncol = (size_t)cs->num_times*(size_t)cs->num_columns;
    size_t const lev = (size_t)cs->num_levels;
    READ_OR_CLOSE(read_doubles(f, &cs->level_pressure, ncol*lev, path, "level_pressure"));
    READ_OR_CLOSE(read_doubles(f, &cs->level_temperature, ncol*lev, path, "level_temperature"));
    READ_OR_CLOSE(read_doubles(f, &cs->layer_temperature, ncol*(lev - 1), path,
                               "layer_temperature"));
    READ_OR_CLOSE(read_doubles(f, &cs->surface_temperature, ncol, path, "surface_temperature"));
    READ_OR_CLOSE(read_doubles(f, &cs->cos_zenith, ncol, path, "cos_zenith"));
    READ_OR_CLOSE(read_doubles(f, &cs->total_solar_irradiance, ncol, path,
                               "total_solar_irradiance"));
    READ_OR_CLOSE(read_doubles(f, &cs->surface_albedo, ncol, path, "surface_albedo"));
    READ_OR_CLOSE(read_doubles(f, &cs->surface_emissivity, ncol, path, "surface_emissivity"));
    READ_OR_CLOSE(read_doubles(f, &cs->ppmv, (size_t)cs->num_molecules*ncol*lev, path, "ppmv"));
    READ_OR_CLOSE(read_doubles(f, &cs->cfc_ppmv, (size_t)cs->num_cfcs*ncol*lev, path, "cfc_ppmv"));
    READ_OR_CLOSE(read_doubles(f, &cs->cia_ppmv, (size_t)cs->num_cia_species*ncol*lev, path,
                               "cia_ppmv"));
#undef READ_OR_CLOSE
    fclose(f);
    return GRTCODE_SUCCESS;
}

int destroy_column_set(ColumnSet_t * const cs)
{
    REQUIRE_PTR(cs);
    free(cs->cfc_path);
    free(cs->cia_path);
    free(cs->level_pressure);
    free(cs->level_temperature);
    free(cs->layer_temperature);
    free(cs->surface_temperature);
    free(cs->cos_zenith);
    free(cs->total_solar_irradiance);
    free(cs->surface_albedo);
    free(cs->surface_emissivity);
    free(cs->ppmv);
    free(cs->cfc_ppmv);
    free(cs->cia_ppmv);
    memset(cs, 0, sizeof(*cs));
    return GRTCODE_SUCCESS;
}

int create_flux_set(FluxSet_t * const fs, int const num_times, int const num_columns,
                    int const num_levels)
{
    REQUIRE_PTR(fs);
    REQUIRE_RANGE(num_times, 1, 1 << 30);
    REQUIRE_RANGE(num_columns, 1, 1 << 30);
    REQUIRE_RANGE(num_levels, MIN_NUM_LEVELS, MAX_NUM_LEVELS);
    memset(fs, 0, sizeof(*fs));
    fs->num_times = num_times;
    fs->num_columns = num_columns;
    fs->num_levels = num_levels;
    size_t const n = (size_t)num_times*(size_t)num_columns;
    fs->rlu = (fp_t *)calloc(n*num_levels, sizeof(fp_t));
    fs->rld = (fp_t *)calloc(n*num_levels, sizeof(fp_t));
    fs->rsu = (fp_t *)calloc(n*num_levels, sizeof(fp_t));
    fs->rsd = (fp_t *)calloc(n*num_levels, sizeof(fp_t));
    fs->hr_lw = (fp_t *)calloc(n*(num_levels - 1), sizeof(fp_t));
    fs->hr_sw = (fp_t *)calloc(n*(num_levels - 1), sizeof(fp_t));
    if (!fs->rlu || !fs->rld || !fs->rsu || !fs->rsd || !fs->hr_lw || !fs->hr_sw)
    {
        destroy_flux_set(fs);
        FAIL(GRTCODE_NULL_ERR, "out of memory for a %d x %d x %d flux set.", num_times,
             num_columns, num_levels);
    }
    return GRTCODE_SUCCESS;
}

int write_flux_set(char const * const path, FluxSet_t const * const fs)
{
    REQUIRE_PTR(path);
    REQUIRE_PTR(fs);
    FILE *f = fopen(path, "wb");
    if (f == NULL)
    {
        FAIL(GRTCODE_IO_ERR, "cannot create flux file %s.", path);
    }
    int32_t const head[4] = {fs->num_times, fs->num_columns, fs->num_levels, 0};
    size_t const n = (size_t)fs->num_times*(size_t)fs->num_columns;
    size_t const lev = (size_t)fs->num_levels;
    int ok = fwrite(flux_magic, 1, 8, f) == 8 && fwrite(head, sizeof(head), 1, f) == 1;
    ok = ok && fwrite(fs->rlu, sizeof(fp_t), n*lev, f) == n*lev;
    ok = ok && fwrite(fs->rld, sizeof(fp_t), n*lev, f) == n*lev;
    ok = ok && fwrite(fs->rsu, sizeof(fp_t), n*lev, f) == n*lev;
    ok = ok && fwrite(fs->rsd, sizeof(fp_t), n*lev, f) == n*lev;
    ok = ok && fwrite(fs->hr_lw, sizeof(fp_t), n*(lev - 1), f) == n*(lev - 1);
    ok = ok && fwrite(fs->hr_sw, sizeof(fp_t), n*(lev - 1), f) == n*(lev - 1);
    if (fclose(f) != 0 || !ok)
    {
        FAIL(GRTCODE_IO_ERR, "short write to flux file %s.", path);
    }
    return GRTCODE_SUCCESS;
}

int read_flux_set(char const * const path, FluxSet_t * const fs)
{
    REQUIRE_PTR(path);
    REQUIRE_PTR(fs);
    FILE *f = fopen(path, "rb");
    if (f == NULL)
    {
        FAIL(GRTCODE_IO_ERR, "cannot open flux file %s.", path);
    }
    char magic[8];
    int32_t head[4];
    int rc = read_exact(f, magic, 8, path, "magic");
    if (rc == GRTCODE_SUCCESS) rc = read_exact(f, head, sizeof(head), path, "header");
    if (rc == GRTCODE_SUCCESS && memcmp(magic, flux_magic, 8) != 0)
    {
        fclose(f);
        FAIL(GRTCODE_VALUE_ERR, "%s is not a GRTFLUX1 flux file.", path);
    }
    if (rc == GRTCODE_SUCCESS) rc = create_flux_set(fs, head[0], head[1], head[2]);
    if (rc != GRTCODE_SUCCESS)
    {
        fclose(f);
        grt_error_frame(__FILE__, __LINE__);
        return rc;
    }
    size_t const n = (size_t)fs->num_times*(size_t)fs->num_columns;
    size_t const lev = (size_t)fs->num_levels;
    fp_t *arrays[6] = {fs->rlu, fs->rld, fs->rsu, fs->rsd, fs->hr_lw, fs->hr_sw};
    int k;
    for (k=0; k<6 && rc == GRTCODE_SUCCESS; ++k)
    {
        rc = read_exact(f, arrays[k], sizeof(fp_t)*n*(k < 4 ? lev : lev - 1), path, "fluxes");
    }
    fclose(f);
    if (rc != GRTCODE_SUCCESS)
    {
        destroy_flux_set(fs);
        grt_error_frame(__FILE__, __LINE__);
    }
    return rc;
}

int destroy_flux_set(FluxSet_t * const fs)
{
    REQUIRE_PTR(fs);
    free(fs->rlu);
    free(fs->rld);
    free(fs->rsu);
    free(fs->rsd);
    free(fs->hr_lw);
    free(fs->hr_sw);
    memset(fs, 0, sizeof(*fs));
    return GRTCODE_SUCCESS;
}
