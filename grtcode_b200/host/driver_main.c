/* grt_lbl -- headless clear-sky line-by-line driver on the sm_100a library.
 *
 * The reference's drivers (rfmip-irf, circ, era5) are framework/src/driver.c plus an
 * application file that reads a netCDF atmosphere and writes netCDF fluxes
 * (rfmip-irf/src/rfmip-irf.c:162-468, circ/src/circ.c:128-430).  No netCDF library exists
 * on this machine, so this driver keeps driver.c's structure -- one GasOptics_t + solver per
 * band, the (time, column) loop of driver.c:691-743, LONGWAVE_PASS then SHORTWAVE_PASS when
 * the sun is up (driver.c:704-716), clear-sky aerosol-free outputs (RLU/RLD/RSU/RSD CSAF,
 * driver.c:387-424) -- and replaces the two application hooks by a dependency-free binary
 * column set (include/grtcode_colset.h) whose arrays carry the semantics of Atmosphere_t
 * (framework/src/driver.h:114-166).  Fluxes are always spectrally integrated on the GPU
 * (the reference's -integrated mode, run-rfmip-irf.sh:119) and written on every level,
 * together with the layer heating rates.  Clouds and aerosols are out of scope.
 *
 * usage: grt_lbl hitran_file solar_flux column_set [-o out] [-d gpu] [-r-lw dw] [-r-sw dw]
 *                [-w-lw w0] [-W-lw wn] [-w-sw w0] [-W-sw wn] [-x first] [-X last]
 *                [-ranks N -rank r] [-method 0|1|2] [-dry-run] [-v]
 * Options are the reference driver's (driver.c:885-904; -x/-X: rfmip-irf.c:186-189);
 * -ranks/-rank select the block partition cols[r*C/N, (r+1)*C/N) for one process per GPU.
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include "gas_optics.h"
#include "grtcode_colset.h"
#include "grtcode_column.h"
#include "longwave.h"
#include "shortwave.h"
#include "solar_flux.h"
#include "verbosity.h"

#define CHECK(call) \
    do { \
        int rc__ = (call); \
        if (rc__ != GRTCODE_SUCCESS) \
        { \
            char buf__[4096]; \
            grtcode_errstr(rc__, buf__, (int)sizeof(buf__)); \
            fprintf(stderr, "[%s:%d] %s failed (%d):\n%s\n", __FILE__, __LINE__, #call, rc__, buf__); \
            return EXIT_FAILURE; \
        } \
    } while (0)

static char const *opt(int argc, char **argv, char const *name, char const *alt)
{
    int i;
    for (i=1; i<argc-1; ++i)
    {
        if (strcmp(argv[i], name) == 0 || (alt != NULL && strcmp(argv[i], alt) == 0))
        {
            return argv[i+1];
        }
    }
    return NULL;
}

static int flag(int argc, char **argv, char const *name)
{
    int i;
    for (i=1; i<argc; ++i)
    {
        if (strcmp(argv[i], name) == 0) return 1;
    }
    return 0;
}

static double now(void)
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9*(double)ts.tv_nsec;
}

/* One band: the objects driver() creates per grid (driver.c:614-672). */
typedef struct Band
{
    SpectralGrid_t grid;
    GasOptics_t lbl;
    Optics_t gas;
} Band_t;

static int create_band(Band_t *b, ColumnSet_t const *cs, double w0, double wn, double dw,
                       Device_t const *device, char const *hitran, int method)
{
    CHECK(create_spectral_grid(&b->grid, w0, wn, dw));
    CHECK(create_gas_optics(&b->lbl, cs->num_levels, &b->grid, device, hitran, cs->h2o_ctm,
                            cs->o3_ctm, NULL, &method));
    int i;
    for (i=0; i<cs->num_molecules; ++i)
    {
        CHECK(add_molecule(&b->lbl, cs->molecules[i], NULL, NULL));
    }
    for (i=0; i<cs->num_cfcs; ++i)
    {
        CHECK(add_cfc(&b->lbl, cs->cfcs[i], cs->cfc_path[i]));
    }
    for (i=0; i<cs->num_cias; ++i)
    {
        CHECK(add_cia(&b->lbl, cs->cia_pairs[i][0], cs->cia_pairs[i][1], cs->cia_path[i]));
    }
    CHECK(create_optics(&b->gas, cs->num_levels - 1, &b->grid, device));
    return EXIT_SUCCESS;
}

/* calculate_gas_optics (driver.c:247-281): abundances of the column, then optical depths. */
static int column_gas_optics(Band_t *b, ColumnSet_t const *cs, size_t col)
{
    int i;
    size_t const L = (size_t)cs->num_levels;
    size_t const ncol = (size_t)cs->num_times*(size_t)cs->num_columns;
    for (i=0; i<cs->num_molecules; ++i)
    {
        CHECK(set_molecule_ppmv(&b->lbl, cs->molecules[i], cs->ppmv + ((size_t)i*ncol + col)*L));
    }
    for (i=0; i<cs->num_cfcs; ++i)
    {
        CHECK(set_cfc_ppmv(&b->lbl, cs->cfcs[i], cs->cfc_ppmv + ((size_t)i*ncol + col)*L));
    }
    for (i=0; i<cs->num_cia_species; ++i)
    {
        CHECK(set_cia_ppmv(&b->lbl, cs->cia_species[i], cs->cia_ppmv + ((size_t)i*ncol + col)*L));
    }
    CHECK(calculate_optical_depth(&b->lbl, cs->level_pressure + col*L,
                                  cs->level_temperature + col*L, &b->gas));
    return EXIT_SUCCESS;
}

int main(int argc, char **argv)
{
    if (argc < 4 || flag(argc, argv, "-h") || flag(argc, argv, "--help"))
    {
        fprintf(stderr, "usage: %s hitran_file solar_flux column_set [-o out] [-d gpu] "
                "[-r-lw dw] [-r-sw dw] [-w-lw w0] [-W-lw wn] [-w-sw w0] [-W-sw wn] "
                "[-x first] [-X last] [-ranks N -rank r] [-method 0|1|2] [-dry-run] [-v]\n",
                argv[0]);
        return argc < 4 ? EXIT_FAILURE : EXIT_SUCCESS;
    }
    char const *hitran = argv[1];
    char const *solar_path = argv[2];
    char const *colset_path = argv[3];
    char const *s;
    grtcode_set_verbosity(flag(argc, argv, "-v") ? GRTCODE_INFO : GRTCODE_WARN);

    ColumnSet_t cs;
    CHECK(read_column_set(colset_path, &cs));
    int const levels = cs.num_levels;
    int const layers = levels - 1;
    size_t const ncol = (size_t)cs.num_times*(size_t)cs.num_columns;

    /* spectral grids: defaults of driver.c:921-931 */
    double lw_w0 = (s = opt(argc, argv, "-w-lw", "--lw-lower-bound")) ? atof(s) : 1.;
    double lw_wn = (s = opt(argc, argv, "-W-lw", "--lw-upper-bound")) ? atof(s) : 3250.;
    double lw_dw = (s = opt(argc, argv, "-r-lw", "--lw-resolution")) ? atof(s) : 0.1;
    double sw_w0 = (s = opt(argc, argv, "-w-sw", "--sw-lower-bound")) ? atof(s) : 1.;
    double sw_wn = (s = opt(argc, argv, "-W-sw", "--sw-upper-bound")) ? atof(s) : 50000.;
    double sw_dw = (s = opt(argc, argv, "-r-sw", "--sw-resolution")) ? atof(s) : 1.;
    int const method = (s = opt(argc, argv, "-method", NULL)) ? atoi(s) : line_sample;

    /* columns of this process: -x/-X (rfmip-irf.c:186-189) or the block partition */
    long first = (s = opt(argc, argv, "-x", NULL)) ? atol(s) : 0;
    long last = (s = opt(argc, argv, "-X", NULL)) ? atol(s) : (long)cs.num_columns - 1;
    if ((s = opt(argc, argv, "-ranks", NULL)) != NULL)
    {
        long const ranks = atol(s);
        long const rank = (s = opt(argc, argv, "-rank", NULL)) ? atol(s) : 0;
        if (ranks < 1 || rank < 0 || rank >= ranks)
        {
            fprintf(stderr, "bad -ranks/-rank\n");
            return EXIT_FAILURE;
        }
        first = (rank*(long)cs.num_columns)/ranks;
        last = ((rank + 1)*(long)cs.num_columns)/ranks - 1;
    }
    if (first < 0 || last >= cs.num_columns || first > last + 1)
    {
        fprintf(stderr, "column range [%ld, %ld] outside [0, %d]\n", first, last,
                cs.num_columns - 1);
        return EXIT_FAILURE;
    }
    printf("column set %s: %d times x %d columns x %d levels, %d molecules, %d CFCs, %d CIA pairs\n",
           colset_path, cs.num_times, cs.num_columns, levels, cs.num_molecules, cs.num_cfcs,
           cs.num_cias);
    printf("columns [%ld, %ld]; LW %g-%g @ %g, SW %g-%g @ %g cm-1; method %d\n", first, last,
           lw_w0, lw_wn, lw_dw, sw_w0, sw_wn, sw_dw, method);
    if (flag(argc, argv, "-dry-run"))
    {
        size_t const c = (size_t)first;
        printf("first column: p_sfc %.6f mb, T_sfc %.4f K, mu0 %.6f, TSI %.3f, albedo %.4f, "
               "emissivity %.4f\n", cs.level_pressure[c*levels + levels - 1],
               cs.surface_temperature[c], cs.cos_zenith[c], cs.total_solar_irradiance[c],
               cs.surface_albedo[c], cs.surface_emissivity[c]);
        destroy_column_set(&cs);
        return EXIT_SUCCESS;
    }

    Device_t device;
    int dev_id = (s = opt(argc, argv, "-d", "--device")) ? atoi(s) : 0;
    CHECK(create_device(&device, &dev_id));

    double const t_setup0 = now();
    Band_t lw, sw;
    if (create_band(&lw, &cs, lw_w0, lw_wn, lw_dw, &device, hitran, method) != EXIT_SUCCESS ||
        create_band(&sw, &cs, sw_w0, sw_wn, sw_dw, &device, hitran, method) != EXIT_SUCCESS)
    {
        return EXIT_FAILURE;
    }
    SolarFlux_t solar;
    CHECK(create_solar_flux(&solar, &sw.grid, solar_path));
    Longwave_t longwave;
    CHECK(create_longwave(&longwave, levels, &lw.grid, &device));
    Shortwave_t shortwave;
    CHECK(create_shortwave(&shortwave, levels, &sw.grid, &device));
    fp_t *emissivity = (fp_t *)malloc(sizeof(fp_t)*lw.grid.n);
    fp_t *albedo = (fp_t *)malloc(sizeof(fp_t)*sw.grid.n);
    if (emissivity == NULL || albedo == NULL) return EXIT_FAILURE;
    double const t_setup = now() - t_setup0;

    FluxSet_t out;
    CHECK(create_flux_set(&out, cs.num_times, cs.num_columns, levels));

    /* the (time, column) loop of driver.c:691-743 */
    fp_t last_emis = -1., last_alb = -1.;
    int solar_resident = 0;
    double const t_run0 = now();
    long done = 0;
    int t;
    for (t=0; t<cs.num_times; ++t)
    {
        long i;
        for (i=first; i<=last; ++i)
        {
            size_t const col = (size_t)t*(size_t)cs.num_columns + (size_t)i;
            fp_t *p = cs.level_pressure + col*levels;
            fp_t *up = out.rlu + col*levels;
            fp_t *down = out.rld + col*levels;
            /* LONGWAVE_PASS; the grey surface arrays are re-uploaded only when they change */
            if (column_gas_optics(&lw, &cs, col) != EXIT_SUCCESS) return EXIT_FAILURE;
            fp_t *emis = NULL;
            if (cs.surface_emissivity[col] != last_emis)
            {
                uint64_t k;
                for (k=0; k<lw.grid.n; ++k) emissivity[k] = cs.surface_emissivity[col];
                last_emis = cs.surface_emissivity[col];
                emis = emissivity;
            }
            CHECK(calculate_lw_column(&longwave, &lw.gas, p, cs.surface_temperature[col],
                                      cs.layer_temperature + col*layers,
                                      cs.level_temperature + col*levels, emis, up, down));
            CHECK(heating_rates(levels, p, up, down, out.hr_lw + col*layers));
            /* SHORTWAVE_PASS when the sun is above the horizon (driver.c:709) */
            if (cs.cos_zenith[col] > 0.)
            {
                if (column_gas_optics(&sw, &cs, col) != EXIT_SUCCESS) return EXIT_FAILURE;
                fp_t *alb = NULL;
                fp_t *sol = NULL;
                if (cs.surface_albedo[col] != last_alb || !solar_resident)
                {
                    uint64_t k;
                    for (k=0; k<sw.grid.n; ++k) albedo[k] = cs.surface_albedo[col];
                    last_alb = cs.surface_albedo[col];
                    alb = albedo;
                    sol = solar.incident_flux;
                    solar_resident = 1;
                }
                up = out.rsu + col*levels;
                down = out.rsd + col*levels;
                CHECK(calculate_sw_column(&shortwave, &sw.gas, p, cs.cos_zenith[col], 0.5, alb,
                                          alb, cs.total_solar_irradiance[col], sol, up, down));
                CHECK(heating_rates(levels, p, up, down, out.hr_sw + col*layers));
            }
            ++done;
        }
    }
    double const t_run = now() - t_run0;

    char const *out_path = (s = opt(argc, argv, "-o", NULL)) ? s : "output.grtflux";
    CHECK(write_flux_set(out_path, &out));
    size_t const c0 = (size_t)first;
    printf("setup %.2f s; %ld columns in %.3f s = %.3f columns/s (%.1f ms per column)\n", t_setup,
           done, t_run, done > 0 ? (double)done/t_run : 0., done > 0 ? 1e3*t_run/(double)done : 0.);
    if (done > 0)
    {
        printf("column %ld: rlutcsaf %.6f rldscsaf %.6f rsutcsaf %.6f rsdscsaf %.6f W m-2\n", first,
               out.rlu[c0*levels], out.rld[c0*levels + levels - 1], out.rsu[c0*levels],
               out.rsd[c0*levels + levels - 1]);
    }
    printf("wrote %s\n", out_path);

    CHECK(destroy_gas_optics(&lw.lbl));
    CHECK(destroy_gas_optics(&sw.lbl));
    CHECK(destroy_optics(&lw.gas));
    CHECK(destroy_optics(&sw.gas));
    CHECK(destroy_solar_flux(&solar));
    CHECK(destroy_longwave(&longwave));
    CHECK(destroy_shortwave(&shortwave));
    free(emissivity);
    free(albedo);
    destroy_flux_set(&out);
    destroy_column_set(&cs);
    (void)ncol;
    return EXIT_SUCCESS;
}
