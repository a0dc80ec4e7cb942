/* Source-level drop-in check: a C99 caller written against the reference's public API
 * (gas_optics.h, longwave.h, shortwave.h, ...) compiled with THIS repository's include/
 * and linked with libgrtcode.so.  The call sequences follow the reference's own unit tests
 * (gas-optics/test/test_gas_optics.c:18-140, longwave/test/test_longwave.c:52-209,
 * shortwave/test/test_shortwave.c:57-215), which check return codes only; the expected
 * error codes for misuse are the reference's (gas_optics.c:235-239, longwave.c:326-329).
 *
 * usage: test_api_dropin host
 *        test_api_dropin gpu <hitran.par> <h2o_ctm_dir> <o3_ctm.csv> <cfc.csv> <cia.csv> <solar.csv>
 * exit status = number of failed expectations. */
#include <float.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "gas_optics.h"
#include "grtcode_utilities.h"
#include "longwave.h"
#include "rayleigh.h"
#include "shortwave.h"
#include "solar_flux.h"

static int failures = 0;

#define EXPECT_RC(call, want) do { \
    int const rc_ = (call); \
    if (rc_ != (want)) { \
        char buf_[4096]; \
        grtcode_errstr(rc_, buf_, (int)sizeof(buf_)); \
        fprintf(stderr, "%s:%d: %s returned %d, expected %d\n%s\n", __FILE__, __LINE__, #call, \
                rc_, (int)(want), buf_); \
        ++failures; \
    } } while (0)

#define EXPECT(cond) do { \
    if (!(cond)) { fprintf(stderr, "%s:%d: %s is false\n", __FILE__, __LINE__, #cond); ++failures; } \
    } while (0)

static int all_finite(fp_t const *a, size_t n)
{
    size_t i;
    for (i=0; i<n; ++i) if (!isfinite(a[i])) return 0;
    return 1;
}

static void host_section(void)
{
    SpectralGrid_t grid;
    EXPECT_RC(create_spectral_grid(&grid, 1., 3000., 1.), GRTCODE_SUCCESS);
    EXPECT(grid.n == 3000);
    uint64_t index = 0;
    EXPECT_RC(grid_point_index(grid, 11., &index), GRTCODE_SUCCESS);
    EXPECT(index == 10);
    /* no CPU code path in this library: every constructor refuses HOST_ONLY */
    Device_t const host = HOST_ONLY;
    Optics_t optics;
    EXPECT_RC(create_optics(&optics, 6, &grid, &host), GRTCODE_VALUE_ERR);
    Longwave_t lw;
    EXPECT_RC(create_longwave(&lw, 7, &grid, &host), GRTCODE_VALUE_ERR);
    Shortwave_t sw;
    EXPECT_RC(create_shortwave(&sw, 7, &grid, &host), GRTCODE_VALUE_ERR);
    GasOptics_t gas;
    int const method = line_sample;
    EXPECT_RC(create_gas_optics(&gas, 7, &grid, &host, "", "none", "none", NULL, &method),
              GRTCODE_VALUE_ERR);
    char buf[4096];
    EXPECT_RC(grtcode_errstr(GRTCODE_VALUE_ERR, buf, (int)sizeof(buf)), GRTCODE_SUCCESS);
    EXPECT(strlen(buf) > 0);
    grtcode_set_verbosity(GRTCODE_WARN);
    EXPECT(grtcode_verbosity() == GRTCODE_WARN);
}

static void fill(fp_t *a, size_t n, fp_t v)
{
    size_t i;
    for (i=0; i<n; ++i) a[i] = v;
}

/* test_longwave.c / test_shortwave.c: simple, optically thick, optically thin, strong
 * absorption on the reference's own grid (1-50000 cm-1 at 10 cm-1). */
static void solver_section(Device_t const device)
{
    enum { levels = 8, layers = levels - 1 };
    SpectralGrid_t grid;
    EXPECT_RC(create_spectral_grid(&grid, 1., 50000., 10.), GRTCODE_SUCCESS);
    size_t const n = (size_t)grid.n;
    Optics_t optics;
    EXPECT_RC(create_optics(&optics, layers, &grid, &device), GRTCODE_SUCCESS);
    fp_t *tau = malloc(sizeof(fp_t)*layers*n);
    fp_t *omega = malloc(sizeof(fp_t)*layers*n);
    fp_t *g = malloc(sizeof(fp_t)*layers*n);
    fp_t *up = malloc(sizeof(fp_t)*levels*n);
    fp_t *down = malloc(sizeof(fp_t)*levels*n);
    fp_t *emis = malloc(sizeof(fp_t)*n);
    fp_t *albedo = malloc(sizeof(fp_t)*n);
    fp_t *solar = malloc(sizeof(fp_t)*n);
    fp_t t_layers[layers], t_levels[levels];
    int i;
    for (i=0; i<layers; ++i)
    {
        t_layers[i] = 250. + 1.*i;
        t_levels[i] = t_layers[i] - 0.5;
    }
    t_levels[layers] = t_layers[layers-1] + 0.5;
    fp_t const t_surf = t_levels[layers] + 2.;
    fill(emis, n, 0.98);
    fill(albedo, n, 0.35);
    fill(solar, n, 1./(10.*(double)(n - 1)));
    Longwave_t lw;
    EXPECT_RC(create_longwave(&lw, levels, &grid, &device), GRTCODE_SUCCESS);
    Shortwave_t sw;
    EXPECT_RC(create_shortwave(&sw, levels, &grid, &device), GRTCODE_SUCCESS);
    fp_t const cases[4] = {-1., 1.e12, 1.e-12, log(DBL_MAX) - 12.};
    int c;
    srand(1);
    for (c=0; c<4; ++c)
    {
        size_t k;
        for (k=0; k<layers*n; ++k)
        {
            tau[k] = cases[c] < 0. ? rand()/((double)RAND_MAX) : cases[c];
            omega[k] = 0.;
            g[k] = 0.;
        }
        EXPECT_RC(update_optics(&optics, tau, omega, g), GRTCODE_SUCCESS);
        EXPECT_RC(calculate_lw_fluxes(&lw, &optics, t_surf, t_layers, t_levels, emis, up, down),
                  GRTCODE_SUCCESS);
        EXPECT(all_finite(up, levels*n) && all_finite(down, levels*n));
        EXPECT(down[0] == 0.);                      /* nothing comes in at the top */
        for (k=0; k<layers*n; ++k)
        {
            omega[k] = 0.85*(rand()/((double)RAND_MAX));
            g[k] = 0.85*(rand()/((double)RAND_MAX));
        }
        EXPECT_RC(update_optics(&optics, tau, omega, g), GRTCODE_SUCCESS);
        EXPECT_RC(calculate_sw_fluxes(&sw, &optics, 0.8, 0.5, albedo, albedo, 1361., solar, up,
                                      down), GRTCODE_SUCCESS);
        EXPECT(all_finite(up, levels*n) && all_finite(down, levels*n));
    }
    /* an Optics_t on another grid is refused (longwave.c:326-329, shortwave.c:520-523) */
    SpectralGrid_t other;
    EXPECT_RC(create_spectral_grid(&other, 1., 3000., 1.), GRTCODE_SUCCESS);
    Optics_t mismatched;
    EXPECT_RC(create_optics(&mismatched, layers, &other, &device), GRTCODE_SUCCESS);
    EXPECT_RC(calculate_lw_fluxes(&lw, &mismatched, t_surf, t_layers, t_levels, emis, up, down),
              GRTCODE_VALUE_ERR);
    EXPECT_RC(calculate_sw_fluxes(&sw, &mismatched, 0.8, 0.5, albedo, albedo, 1361., solar, up,
                                  down), GRTCODE_VALUE_ERR);
    EXPECT_RC(calculate_lw_fluxes(&lw, NULL, t_surf, t_layers, t_levels, emis, up, down),
              GRTCODE_NULL_ERR);
    EXPECT_RC(destroy_optics(&mismatched), GRTCODE_SUCCESS);
    EXPECT_RC(destroy_shortwave(&sw), GRTCODE_SUCCESS);
    EXPECT_RC(destroy_longwave(&lw), GRTCODE_SUCCESS);
    EXPECT_RC(destroy_optics(&optics), GRTCODE_SUCCESS);
    free(tau); free(omega); free(g); free(up); free(down); free(emis); free(albedo); free(solar);
}

/* test_gas_optics.c, with real (synthetic) input files, then the column_calculation chain of
 * framework/src/driver.c:360-424. */
static void gas_section(Device_t const device, char **files)
{
    enum { levels = 7, layers = levels - 1 };
    SpectralGrid_t grid;
    EXPECT_RC(create_spectral_grid(&grid, 500., 700., 0.1), GRTCODE_SUCCESS);
    size_t const n = (size_t)grid.n;
    GasOptics_t gas;
    double const wcutoff = 25.;
    int const method = line_sample;
    EXPECT_RC(create_gas_optics(&gas, levels, &grid, &device, files[0], files[1], files[2],
                                &wcutoff, &method), GRTCODE_SUCCESS);
    EXPECT_RC(add_molecule(&gas, H2O, NULL, NULL), GRTCODE_SUCCESS);
    EXPECT_RC(add_molecule(&gas, H2O, NULL, NULL), GRTCODE_VALUE_ERR);   /* gas_optics.c:235-239 */
    EXPECT_RC(add_molecule(&gas, 999, NULL, NULL), GRTCODE_VALUE_ERR);
    double const lo = 550., hi = 650.;
    EXPECT_RC(add_molecule(&gas, CO2, &lo, &hi), GRTCODE_SUCCESS);
    int num = 0;
    EXPECT_RC(get_num_molecules(&gas, &num), GRTCODE_SUCCESS);
    EXPECT(num == 2);
    EXPECT_RC(add_cfc(&gas, CFC11, files[3]), GRTCODE_SUCCESS);
    EXPECT_RC(add_cfc(&gas, CFC11, files[3]), GRTCODE_VALUE_ERR);
    EXPECT_RC(add_cia(&gas, CIA_N2, CIA_N2, files[4]), GRTCODE_SUCCESS);
    fp_t pressure[levels], temperature[levels], h2o[levels], co2[levels], cfc[levels], n2[levels];
    int i;
    for (i=0; i<levels; ++i)
    {
        pressure[i] = 1. + 1012.*i/(levels - 1.);
        temperature[i] = 210. + 80.*i/(levels - 1.);
        h2o[i] = 5. + 15000.*i/(levels - 1.);
        co2[i] = 400.;
        cfc[i] = 2.3e-4;
        n2[i] = 781000.;
    }
    EXPECT_RC(set_molecule_ppmv(&gas, H2O, h2o), GRTCODE_SUCCESS);
    EXPECT_RC(set_molecule_ppmv(&gas, CO2, co2), GRTCODE_SUCCESS);
    /* an inactive species warns and succeeds (gas_optics.c:301-306) */
    grtcode_set_verbosity(GRTCODE_NONE);
    EXPECT_RC(set_molecule_ppmv(&gas, O3, co2), GRTCODE_SUCCESS);
    grtcode_set_verbosity(GRTCODE_WARN);
    EXPECT_RC(set_cfc_ppmv(&gas, CFC11, cfc), GRTCODE_SUCCESS);
    EXPECT_RC(set_cia_ppmv(&gas, CIA_N2, n2), GRTCODE_SUCCESS);
    Optics_t optics;
    EXPECT_RC(create_optics(&optics, layers, &grid, &device), GRTCODE_SUCCESS);
    EXPECT_RC(calculate_optical_depth(&gas, pressure, temperature, &optics), GRTCODE_SUCCESS);
    fp_t *tau = malloc(sizeof(fp_t)*layers*n);
    EXPECT_RC(optical_properties(&optics, tau, NULL, NULL), GRTCODE_SUCCESS);
    EXPECT(all_finite(tau, layers*n));
    size_t k;
    double sum = 0.;
    for (k=0; k<layers*n; ++k)
    {
        if (tau[k] < 0.) { ++failures; fprintf(stderr, "negative optical depth at %zu\n", k); break; }
        sum += tau[k];
    }
    EXPECT(sum > 0.);
    /* layers / grid of the Optics_t must match the GasOptics_t (gas_optics.c:440-444) */
    Optics_t wrong;
    EXPECT_RC(create_optics(&wrong, layers - 1, &grid, &device), GRTCODE_SUCCESS);
    EXPECT_RC(calculate_optical_depth(&gas, pressure, temperature, &wrong), GRTCODE_VALUE_ERR);
    EXPECT_RC(destroy_optics(&wrong), GRTCODE_SUCCESS);
    /* the rest of column_calculation: Rayleigh, add_optics, fluxes */
    Optics_t rayleigh, total;
    EXPECT_RC(create_optics(&rayleigh, layers, &grid, &device), GRTCODE_SUCCESS);
    EXPECT_RC(rayleigh_scattering(&rayleigh, pressure), GRTCODE_SUCCESS);
    Optics_t const *parts[2] = {&optics, &rayleigh};
    EXPECT_RC(add_optics(parts, 2, &total), GRTCODE_SUCCESS);
    int compatible = 0;
    EXPECT_RC(optics_compatible(&optics, &total, &compatible), GRTCODE_SUCCESS);
    EXPECT(compatible == 1);
    SolarFlux_t solar;
    EXPECT_RC(create_solar_flux(&solar, &grid, files[5]), GRTCODE_SUCCESS);
    EXPECT(solar.n == grid.n);
    fp_t *up = malloc(sizeof(fp_t)*levels*n);
    fp_t *down = malloc(sizeof(fp_t)*levels*n);
    fp_t *albedo = malloc(sizeof(fp_t)*n);
    fill(albedo, n, 0.2);
    Shortwave_t sw;
    EXPECT_RC(create_shortwave(&sw, levels, &grid, &device), GRTCODE_SUCCESS);
    EXPECT_RC(calculate_sw_fluxes(&sw, &total, 0.7, 0.5, albedo, albedo, 1361.,
                                  solar.incident_flux, up, down), GRTCODE_SUCCESS);
    EXPECT(all_finite(up, levels*n) && all_finite(down, levels*n));
    EXPECT(down[n/2] > 0.);
    EXPECT_RC(destroy_shortwave(&sw), GRTCODE_SUCCESS);
    EXPECT_RC(destroy_solar_flux(&solar), GRTCODE_SUCCESS);
    EXPECT_RC(destroy_optics(&total), GRTCODE_SUCCESS);
    EXPECT_RC(destroy_optics(&rayleigh), GRTCODE_SUCCESS);
    EXPECT_RC(destroy_optics(&optics), GRTCODE_SUCCESS);
    EXPECT_RC(destroy_gas_optics(&gas), GRTCODE_SUCCESS);
    free(tau); free(up); free(down); free(albedo);
}

int main(int argc, char **argv)
{
    if (argc < 2)
    {
        fprintf(stderr, "usage: %s host | gpu <six input files>\n", argv[0]);
        return 99;
    }
    host_section();
    if (strcmp(argv[1], "gpu") == 0)
    {
        if (argc < 8) return 98;
        Device_t device;
        EXPECT_RC(create_device(&device, NULL), GRTCODE_SUCCESS);
        solver_section(device);
        gas_section(device, argv + 2);
    }
    printf("%d failed expectations\n", failures);
    return failures > 99 ? 99 : failures;
}
