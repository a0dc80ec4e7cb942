/* Optics_t objects on the GPU (reference: utilities/src/optics.c:84-357 and
 * optical_properties of fortran-bindings/malloc_structs.c:78-97). */
#include <stdlib.h>
#include "errors.h"
#include "grt_sm100.h"
#include "optics.h"

int create_optics(Optics_t * const optics, int const num_layers,
                  SpectralGrid_t const * const grid, Device_t const * const device)
{
    REQUIRE_PTR(optics);
    REQUIRE_PTR(grid);
    REQUIRE_PTR(device);
    REQUIRE_GPU(*device);
    REQUIRE_RANGE(num_layers, MIN_NUM_LAYERS, MAX_NUM_LAYERS);
    optics->num_layers = num_layers;
    optics->grid = *grid;
    optics->device = *device;
    optics->g = optics->omega = optics->tau = NULL;
    size_t const bytes = sizeof(fp_t)*(size_t)num_layers*grid->n;
    GPU(grt_malloc(*device, (void **)&optics->g, bytes));
    GPU(grt_malloc(*device, (void **)&optics->omega, bytes));
    GPU(grt_malloc(*device, (void **)&optics->tau, bytes));
    GPU(grt_memset(*device, optics->g, 0, bytes));
    GPU(grt_memset(*device, optics->omega, 0, bytes));
    GPU(grt_memset(*device, optics->tau, 0, bytes));
    return GRTCODE_SUCCESS;
}

int destroy_optics(Optics_t * const optics)
{
    REQUIRE_PTR(optics);
    GPU(grt_free(optics->device, optics->g));
    GPU(grt_free(optics->device, optics->omega));
    GPU(grt_free(optics->device, optics->tau));
    optics->g = optics->omega = optics->tau = NULL;
    return GRTCODE_SUCCESS;
}

int optics_compatible(Optics_t const * const one, Optics_t const * const two, int * const result)
{
    REQUIRE_PTR(one);
    REQUIRE_PTR(two);
    REQUIRE_PTR(result);
    int same;
    TRY(compare_spectral_grids(&one->grid, &two->grid, &same));
    *result = (one->num_layers == two->num_layers && same == 1 && one->device == two->device);
    return GRTCODE_SUCCESS;
}

int add_optics(Optics_t const * const * const optics, int const num_optics,
               Optics_t * const result)
{
    REQUIRE_PTR(optics);
    REQUIRE_PTR(result);
    if (num_optics < 1 || num_optics > 8)
    {
        FAIL(GRTCODE_RANGE_ERR, "number of optics objects (%d) outside [1, 8].", num_optics);
    }
    Optics_t const *first = optics[0];
    REQUIRE_PTR(first);
    fp_t const *g[8];
    fp_t const *omega[8];
    fp_t const *tau[8];
    int j;
    for (j=0; j<num_optics; ++j)
    {
        REQUIRE_PTR(optics[j]);
        int ok;
        TRY(optics_compatible(optics[j], first, &ok));
        if (!ok)
        {
            FAIL(GRTCODE_VALUE_ERR, "input optics objects (%p, %p) are incompatible.",
                 (void const *)first, (void const *)optics[j]);
        }
        g[j] = optics[j]->g;
        omega[j] = optics[j]->omega;
        tau[j] = optics[j]->tau;
    }
    TRY(create_optics(result, first->num_layers, &first->grid, &first->device));
    GPU(grt_add_optics(first->device, (uint64_t)first->num_layers*first->grid.n, num_optics, g,
                       omega, tau, result->g, result->omega, result->tau));
    return GRTCODE_SUCCESS;
}

int update_optics(Optics_t * const optics, fp_t const * const tau, fp_t const * const omega,
                  fp_t const * const g)
{
    REQUIRE_PTR(optics);
    REQUIRE_PTR(tau);
    REQUIRE_PTR(omega);
    REQUIRE_PTR(g);
    size_t const bytes = sizeof(fp_t)*(size_t)optics->num_layers*optics->grid.n;
    GPU(grt_copy_to_device(optics->device, optics->tau, tau, bytes));
    GPU(grt_copy_to_device(optics->device, optics->omega, omega, bytes));
    GPU(grt_copy_to_device(optics->device, optics->g, g, bytes));
    return GRTCODE_SUCCESS;
}

int optical_properties(Optics_t const * const optics, fp_t * const tau, fp_t * const omega,
                       fp_t * const g)
{
    REQUIRE_PTR(optics);
    size_t const bytes = sizeof(fp_t)*(size_t)optics->num_layers*optics->grid.n;
    if (tau != NULL)
    {
        GPU(grt_copy_to_host(optics->device, tau, optics->tau, bytes));
    }
    if (omega != NULL)
    {
        GPU(grt_copy_to_host(optics->device, omega, optics->omega, bytes));
    }
    if (g != NULL)
    {
        GPU(grt_copy_to_host(optics->device, g, optics->g, bytes));
    }
    return GRTCODE_SUCCESS;
}

/* Strided down-sampling onto a coarser grid that shares its points with the
 * source (reference: optics.c:234-340).  Done through the host: it is not on
 * the column hot path. */
int sample_optics(Optics_t * const dest, Optics_t const * const source,
                  double const * const w0, double const * const wn)
{
    REQUIRE_PTR(dest);
    REQUIRE_PTR(source);
    if (dest->device != source->device)
    {
        FAIL(GRTCODE_VALUE_ERR, "Device ids (%d, %d) must match.", dest->device, source->device);
    }
    if (dest->num_layers != source->num_layers)
    {
        FAIL(GRTCODE_VALUE_ERR, "Number of layers (%d, %d) must match.", dest->num_layers,
             source->num_layers);
    }
    double lower = dest->grid.w0;
    uint64_t lower_d = 0;
    uint64_t lower_s;
    if (w0 != NULL)
    {
        TRY(grid_point_index(dest->grid, *w0, &lower_d));
        lower = *w0;
    }
    TRY(grid_point_index(source->grid, lower, &lower_s));
    double upper = dest->grid.wn;
    uint64_t upper_d = dest->grid.n;
    uint64_t upper_s;
    if (wn != NULL)
    {
        TRY(grid_point_index(dest->grid, *wn, &upper_d));
        upper = *wn;
    }
    TRY(grid_point_index(source->grid, upper, &upper_s));
    if (upper < lower)
    {
        FAIL(GRTCODE_RANGE_ERR, "value (%e) less than minimum allowed (%e).", upper, lower);
    }
    uint64_t const n_d = upper_d - lower_d + 1;
    uint64_t const n_s = upper_s - lower_s + 1;
    if (n_d > n_s || n_d < 2 || (n_s - 1) % (n_d - 1) != 0)
    {
        FAIL(GRTCODE_VALUE_ERR, "New grid must be a subdomain of the original (%p).",
             (void const *)&source->grid);
    }
    uint64_t const factor = (n_s - 1)/(n_d - 1);
    size_t const ns = (size_t)source->num_layers*source->grid.n;
    size_t const nd = (size_t)dest->num_layers*dest->grid.n;
    fp_t *src = (fp_t *)malloc(sizeof(*src)*3*ns);
    fp_t *dst = (fp_t *)malloc(sizeof(*dst)*3*nd);
    if (src == NULL || dst == NULL)
    {
        free(src);
        free(dst);
        FAIL(GRTCODE_NULL_ERR, "out of memory in sample_optics.");
    }
    int rc = optical_properties(source, src, src + ns, src + 2*ns);
    if (rc == GRTCODE_SUCCESS)
    {
        rc = optical_properties(dest, dst, dst + nd, dst + 2*nd);
    }
    if (rc == GRTCODE_SUCCESS)
    {
        int k;
        for (k=0; k<3; ++k)
        {
            int i;
            for (i=0; i<dest->num_layers; ++i)
            {
                fp_t *d = dst + k*nd + (size_t)i*dest->grid.n + lower_d;
                fp_t const *s = src + k*ns + (size_t)i*source->grid.n + lower_s;
                uint64_t j;
                for (j=0; j<n_d && lower_d + j < dest->grid.n; ++j)
                {
                    d[j] = s[j*factor];
                }
            }
        }
        rc = update_optics(dest, dst, dst + nd, dst + 2*nd);
    }
    free(src);
    free(dst);
    if (rc != GRTCODE_SUCCESS)
    {
        grt_error_frame(__FILE__, __LINE__);
    }
    return rc;
}
