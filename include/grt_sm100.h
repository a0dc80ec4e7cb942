/* grt_sm100.h -- thin C ABI of the sm_100a CUDA layer (libgrt_sm100.so).
 *
 * Everything the C99 host library (include/gas_optics.h, longwave.h, shortwave.h,
 * rayleigh.h, optics.h) needs from the GPU goes through these entry points:
 * plain pointers and sizes only, no CUDA or C++ types.  Each group cites the
 * reference interface it replaces (paths under the NOAA-GFDL/GRTCODE tree).
 *
 * Conventions: every function returns 0 on success or GRT_SM100_ERR (= 12,
 * GRTCODE_GPU_ERR of utilities/src/return_codes.h:39); grt_last_error() gives
 * the message.  Pointers named *_dev are device pointers obtained from
 * grt_malloc(); all others are host pointers.  All work of one device is
 * ordered on one library-owned CUDA stream; calls that return data to the host
 * synchronise that stream before returning (grt_gas_optical_depth only enqueues its
 * sweep kernels: tau_dev is complete for every later call on the library stream).
 */
#ifndef GRT_SM100_H_
#define GRT_SM100_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Working precision of every data array below (fp_t of the host library:
 * utilities/src/floating_point_type.h:23-29): double, or float when the stack is built
 * with -DSINGLE_PRECISION (libgrtcode_sp.so + libgrt_sm100_sp.so).  Scalars stay double. */
#ifdef SINGLE_PRECISION
typedef float grt_real;
#else
typedef double grt_real;
#endif

#define GRT_SM100_ERR 12
#define GRT_MAX_LAYERS 200   /* utilities/src/grtcode_config.h:52 */
#define GRT_MAX_MOLS 53      /* gas-optics/src/molecules.h:86 */
#define GRT_MAX_CFCS 21      /* gas-optics/src/cfcs.h:55 */
#define GRT_MAX_CIAS 3       /* gas-optics/src/collision_induced_absorption.h:34 */
#define GRT_NUM_CIA_SPECIES 2

/* ---- device + memory: replaces the gmalloc/gfree/gmemset/gmemcpy macros of
 * utilities/src/debug.h:311-348 and get_num_gpus of utilities/src/device.c:26.
 * grt_malloc / grt_free are ordered on the library stream (cudaMallocAsync / cudaFreeAsync from
 * the device's pool, freed blocks kept for reuse): a pointer is valid for every later call of
 * this library at once; code that touches it on ANOTHER stream must first synchronise with the
 * library stream (any call that returns data to the host does). */
int grt_device_count(int *count);
int grt_device_name(int device, char *buf, int len);
int grt_malloc(int device, void **ptr_dev, size_t bytes);
int grt_free(int device, void *ptr_dev);
int grt_memset(int device, void *ptr_dev, int value, size_t bytes);
int grt_copy_to_device(int device, void *dst_dev, void const *src, size_t bytes);
int grt_copy_to_host(int device, void *dst, void const *src_dev, size_t bytes);
int grt_copy_on_device(int device, void *dst_dev, void const *src_dev, size_t bytes);
int grt_synchronize(int device);
/* The library's cudaStream_t for `device`, as an opaque pointer, so that a
 * caller can record CUDA events on the stream the kernels run on. */
int grt_stream_handle(int device, void **stream);
/* Pinned host buffers for the batched drivers (no reference counterpart). */
int grt_host_alloc(void **ptr, size_t bytes);
int grt_host_free(void *ptr);
/* Copy of the last error message of the calling thread. */
int grt_last_error(char *buf, int len);

/* ---- gas optics: replaces launch() of gas-optics/src/launch.c:40-226 and the
 * device halves of create_gas_optics / add_molecule / add_cfc / add_cia
 * (gas-optics/src/gas_optics.c:51-176, 219-279, 307-399). */
typedef struct grt_gas grt_gas;

int grt_gas_create(grt_gas **gas, int device, int num_levels, double w0, double dw,
                   uint64_t num_wpoints);
int grt_gas_destroy(grt_gas *gas);

/* Upload one molecule's HITRAN lines (host SoA as filled by parse_hitran_file,
 * gas-optics/src/parse_HITRAN_file.c:224-413; snn already scaled to 296 K) and
 * its partition-sum table q_table[num_iso][nt] on temperatures tmin + k [K]
 * (replaces inittips_d, gas-optics/src/tips2017.h:29). */
int grt_gas_add_molecule(grt_gas *gas, int mol_id, double mass_g, int num_iso,
                         uint64_t num_lines, int const *iso, grt_real const *vnn,
                         grt_real const *snn, grt_real const *yair, grt_real const *yself,
                         grt_real const *en, grt_real const *n_air, grt_real const *delta,
                         double const *q_table, int tmin, int nt);
/* Continuum / cross-section tables already resampled to the spectral grid
 * (water_vapor_continuum.c:32-122, ozone_continuum.c:31, cfcs.c:30,
 * collision_induced_absorption.c:29); each array has num_wpoints entries. */
int grt_gas_set_h2o_continuum(grt_gas *gas, grt_real const *s296, grt_real const *ckds,
                              grt_real const *f296, grt_real const *ckdf);
int grt_gas_set_o3_continuum(grt_gas *gas, grt_real const *cross_section);
int grt_gas_add_cfc(grt_gas *gas, int cfc_id, grt_real const *cross_section);
int grt_gas_add_cia(grt_gas *gas, int species0, int species1, grt_real const *cross_section);

/* The same tables given by their SAMPLES (x ascending, at least two; as read from the CSV
 * files) and resampled to the grid ON THE DEVICE by one thread per grid point with the
 * reference's rule (interpolate_to_grid / interpolate2 / linear_sample,
 * utilities/src/spectral_grid.c:87-112, utilities.c:149-246: linear inside (x_j, x_{j+1}],
 * 0 at or below x_0 and beyond x_m-1): the setup path of SURVEY section 8 f.3 -- no
 * num_wpoints-sized host array and no num_wpoints-sized copy per table.  h2o: the four
 * tables in the order S296, CKDS, F296, CKDF. */
int grt_gas_set_h2o_continuum_samples(grt_gas *gas, grt_real const *const x[4],
                                      grt_real const *const y[4], uint64_t const m[4]);
int grt_gas_set_o3_continuum_samples(grt_gas *gas, grt_real const *x, grt_real const *y,
                                     uint64_t m);
int grt_gas_add_cfc_samples(grt_gas *gas, int cfc_id, grt_real const *x, grt_real const *y,
                            uint64_t m);
int grt_gas_add_cia_samples(grt_gas *gas, int species0, int species1, grt_real const *x,
                            grt_real const *y, uint64_t m);
/* Diagnostic read-back of a device table (tests): kind 0-3 = H2O S296 / CKDS / F296 / CKDF,
 * 4 = O3, 5 = CFC number `index` (add order), 6 = CIA pair number `index`; out has
 * num_wpoints doubles. */
int grt_gas_debug_table(grt_gas *gas, int kind, int index, double *out);

/* Setup on the device: the fixed-width HITRAN records of `text`
 * (`size` bytes on the host; record k starts at start[k]) converted to numbers by one thread per
 * record, into host arrays of num_records entries -- the conversions of
 * gas-optics/src/parse_HITRAN_file.c:266-370 (strtod per field, float rounding of the widths,
 * E'', n and delta).  status[k]: 0 = converted, 3 = not decided here (unusual text, a malformed
 * field, a molecule id that is not two characters of digits): the caller converts that row
 * with strtod itself.  No reference counterpart: the reference parses on the host only. */
int grt_hitran_convert(int device, char const *text, uint64_t size, uint64_t const *start,
                       uint64_t num_records, short *mol, int *iso, grt_real *vnn, grt_real *snn,
                       grt_real *yair, grt_real *yself, grt_real *en, grt_real *n, grt_real *d,
                       signed char *status);

/* Test hook of the decimal conversion above: `count` fields of `width` (<= 31) characters,
 * back to back in `text`; out[k] = the double, status[k] = 0 or 3 as above. */
int grt_decimal_fields(int device, char const *text, int width, uint64_t count, double *out,
                       signed char *status);

/* One column: level pressure [atm] and temperature [K], mole fractions
 * x_mol[GRT_MAX_MOLS][num_levels] (row = HITRAN id - 1), x_cfc[GRT_MAX_CFCS][num_levels]
 * (row = cfc id), x_cia[GRT_NUM_CIA_SPECIES][num_levels]; writes (overwrites)
 * tau_dev[num_layers][num_wpoints].  method follows enum OpticalDepthMethod
 * (gas-optics/src/gas_optics.h:88-93): wavenumber_sweep (0), line_sweep (1), line_sample (2). */
int grt_gas_optical_depth(grt_gas *gas, grt_real const *p_atm, grt_real const *t,
                          grt_real const *x_mol, grt_real const *x_cfc, grt_real const *x_cia,
                          int method, grt_real *tau_dev);

/* Counters of the last grt_gas_optical_depth call (bench/roofline bookkeeping):
 * out[0] = (line,layer) records inside the grid, out[1] = Voigt samples the
 * reference would evaluate for them, out[2] = kernels launched,
 * out[3] = records taking the pure-Lorentz branch; device durations in ns measured
 * with CUDA events on the library stream: out[4] = column + line preprocessing,
 * out[5] = far-wing sweep kernel, out[6] = line-core kernel; out[7] = upper bound of
 * the samples the line-core kernel evaluates. */
int grt_gas_last_stats(grt_gas *gas, uint64_t out[8]);

/* K1 evaluates exp/pow with CUDA's libdevice, the reference with the host libm
 * (utilities/src/grtcode_config.h:25-47).  Records whose Y = (float)(REPWID*gamma)
 * (gas-optics/src/RFM_voigt.c:95) lies within a few double ulps of a float rounding
 * boundary are rebuilt on the host with the host libm and patched before the sweep, so
 * that every float of the line shape has the reference's bits.  out[0] = records patched
 * by the last grt_gas_optical_depth call, out[1] = since grt_gas_create.
 * (GRT_K1_PATCH=0 in the environment disables the patch: tests only.) */
int grt_gas_patch_counts(grt_gas const *gas, uint64_t out[2]);

/* Diagnostics (tests): REPWID, Y and the shifted centre of K1's records
 * [first, first + count) of one layer, in device line order; the device line order as
 * indices into the upload order (all grt_gas_add_molecule calls concatenated); line-tile
 * visits per line class of the last line_sample call when GRT_CLASS_HIST=1 is set
 * (out[0..5] = fast, region 1, far edge, near, Lorentz covered, Lorentz edge;
 * out[6] = grid points per tile). */
int grt_gas_debug_records(grt_gas *gas, int layer, uint64_t first, uint64_t count,
                          float *repwid, float *y, double *centre);
int grt_gas_debug_line_order(grt_gas const *gas, uint64_t count, uint64_t *order);
int grt_gas_class_histogram(grt_gas *gas, uint64_t out[8]);

/* ---- Rayleigh + optics mixing: replaces rayleigh_scattering
 * (shortwave/src/rayleigh.c:100-144) and add_optics (utilities/src/optics.c:84-148). */
int grt_rayleigh(int device, int num_layers, double w0, double dw, uint64_t num_wpoints,
                 grt_real const *p_atm, grt_real *omega_dev, grt_real *g_dev, grt_real *tau_dev);
int grt_add_optics(int device, uint64_t num_values, int num_optics,
                   grt_real const *const *g_dev, grt_real const *const *omega_dev,
                   grt_real const *const *tau_dev, grt_real *g_out_dev, grt_real *omega_out_dev,
                   grt_real *tau_out_dev);

/* ---- solvers: replace calculate_lw_fluxes (longwave/src/longwave.c:312-353) and
 * calculate_sw_fluxes (shortwave/src/shortwave.c:506-547).  Spectral fluxes are
 * written to device buffers [num_levels][num_wpoints]; the host library copies
 * them out. */
int grt_lw_fluxes(int device, int num_levels, double w0, double dw, uint64_t num_wpoints,
                  double t_surf, grt_real const *t_layers, grt_real const *t_levels,
                  grt_real const *tau_dev, grt_real const *omega_dev, grt_real const *emis_dev,
                  grt_real *flux_up_dev, grt_real *flux_down_dev);
int grt_sw_fluxes(int device, int num_levels, uint64_t num_wpoints, grt_real const *omega_dev,
                  grt_real const *g_dev, grt_real const *tau_dev, double mu_dir, double mu_dif,
                  grt_real const *alb_dir_dev, grt_real const *alb_dif_dev, double tsi,
                  grt_real const *solar_dev, grt_real *flux_up_dev, grt_real *flux_down_dev);

/* ---- broadband integration: replaces the serial trapezoid loop of
 * output_fluxes (framework/src/driver.c:302-319), for every level at once:
 * out[level] = sum_{i<n-1} 0.5*(f[level][i] + f[level][i+1])*dw. */
int grt_integrate_flux(int device, int num_levels, uint64_t num_wpoints, double dw,
                       grt_real const *flux_dev, grt_real *out);

/* ---- allocation-free broadband integration.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA
 * graph capture, tight serving loops, harnesses that forbid it).  The same sums as
 * grt_integrate_flux, out_dev[level] = sum_{i<n-1} 0.5*(f[level][i] + f[level][i+1])*dw, as a
 * two-stage reduction whose scratch (per-block partial sums) is a file-scope
 * `static __device__` array of GRT_NOALLOC_SCRATCH_BYTES that every GPU loading the library
 * holds until exit.  The call allocates and frees nothing and does not synchronise: it
 * enqueues two kernels on the library stream of `device`; out_dev (DEVICE memory,
 * num_levels values, caller-owned like flux_dev) is complete for any later work on that
 * stream (grt_copy_to_host synchronises).  num_levels > GRT_NOALLOC_MAX_LEVELS is rejected
 * with GRT_SM100_ERR and a message (use grt_integrate_flux).  All calls on a device share
 * the one scratch array: calls that overlap on the GPU race; calls ordered on the library
 * stream do not.  The first call per device looks up the array's device address
 * (cudaGetSymbolAddress) and, if no other call has yet, creates the library stream. */
#define GRT_NOALLOC_MAX_LEVELS (GRT_MAX_LAYERS + 1)
#define GRT_NOALLOC_BLOCKS 64
#define GRT_NOALLOC_SCRATCH_BYTES (8*GRT_NOALLOC_MAX_LEVELS*GRT_NOALLOC_BLOCKS)   /* 102912 */
int grt_integrate_flux_noalloc(int device, int num_levels, uint64_t num_wpoints, double dw,
                               grt_real const *flux_dev, grt_real *out_dev);

/* ---- fused column solvers for the drivers (no spectral arrays leave the GPU):
 * Rayleigh + optics mixing + solver + broadband integration in one kernel each;
 * replaces the chain column_calculation -> add_optics -> calculate_*_fluxes ->
 * output_fluxes (framework/src/driver.c:360-424).  up/down have num_levels entries.
 * [first, last] is the index range of the trapezoid integration (0, num_wpoints - 1 for
 * the whole grid; a sub-range when the grid carries halo points of a spectral split). */
int grt_lw_column(int device, int num_levels, double w0, double dw, uint64_t num_wpoints,
                  grt_real const *p_atm, double t_surf, grt_real const *t_layers,
                  grt_real const *t_levels, grt_real const *tau_gas_dev,
                  grt_real const *emis_dev, uint64_t first, uint64_t last, grt_real *up,
                  grt_real *down);
int grt_sw_column(int device, int num_levels, double w0, double dw, uint64_t num_wpoints,
                  grt_real const *p_atm, grt_real const *tau_gas_dev, double mu_dir,
                  double mu_dif, grt_real const *alb_dir_dev, grt_real const *alb_dif_dev,
                  double tsi, grt_real const *solar_dev, uint64_t first, uint64_t last,
                  grt_real *up, grt_real *down);

/* ---- measurement helper: sustained FP64 / FP32 FMA throughput of this GPU
 * [TFLOP/s, FMA = 2 flop], the roofline denominator of the Voigt sweep. */
int grt_measure_fma_peak(int device, double *fp64_tflops, double *fp32_tflops);

/* ---- self-test: the sweep's division and float->double shortcuts against the IEEE
 * operations on n pseudo-random operands; *mismatches must come back 0. */
int grt_selftest_arithmetic(int device, unsigned long long n, unsigned long long seed,
                            unsigned long long *mismatches);
/* The per-line Voigt evaluation of the line-core kernel (hoisted per-line polynomials,
 * packed CPF12 sums) against the sample-by-sample restatement of RFM_voigt.c:109-277 on n
 * pseudo-random (x, Y) pairs over all regions; *mismatches (bit-level) must come back 0. */
int grt_selftest_voigt(int device, unsigned long long n, unsigned long long seed,
                       unsigned long long *mismatches);

#ifdef __cplusplus
}
#endif
#endif
