/* CSV reader + "table -> spectral grid" resampling.
 *
 * Format rules follow utilities/src/parse_csv.c:55-166: first line is a header,
 * at most 1024 characters per line and 31 per token, no blank lines, constant
 * column count.  The resampling is what get_water_vapor_continuum_coefs
 * (water_vapor_continuum.c:84-104), get_ozone_continuum_coefs
 * (ozone_continuum.c:53-73), get_cfc_cross_sections (cfcs.c:122-143) and
 * get_collision_induced_cross_sections (collision_induced_absorption.c:66-87)
 * all do: zero-filled buffer, linear_sample inside the table, no extrapolation. */
#include <stdlib.h>
#include <string.h>
#include "errors.h"
#include "parse_csv.h"
#include "spectral_grid.h"
#include "utilities.h"

#define MAX_LINE 1024
#define MAX_TOKEN 32

static int count_columns(char const *line)
{
    int n = 1;
    char const *c;
    for (c=line; *c != '\0' && *c != '\n'; ++c)
    {
        if (*c == ',')
        {
            ++n;
        }
    }
    return n;
}

int grt_read_csv(char const *path, int *num_rows, int *num_cols, fp_t **values)
{
    REQUIRE_PTR(path);
    REQUIRE_PTR(num_rows);
    REQUIRE_PTR(num_cols);
    REQUIRE_PTR(values);
    FILE *f = NULL;
    TRY(open_file(&f, path, "r"));
    LOG_INFO("Reading csv file %s.", path);
    char line[MAX_LINE + 2];
    int rows = 0;
    int cols = -1;
    while (fgets(line, sizeof(line), f) != NULL)
    {
        ++rows;
        if (line[0] == '\n')
        {
            fclose(f);
            FAIL(GRTCODE_VALUE_ERR, "line %d in file %s is blank.", rows, path);
        }
        if (strlen(line) > MAX_LINE)
        {
            fclose(f);
            FAIL(GRTCODE_VALUE_ERR, "line %d of file %s exceeds the maximum length (%d).", rows,
                 path, MAX_LINE);
        }
        int const c = count_columns(line);
        if (cols < 0)
        {
            cols = c;
        }
        else if (c != cols)
        {
            fclose(f);
            FAIL(GRTCODE_VALUE_ERR, "the number of columns (%d) on line %d of file %s differs"
                 " from the number of columns (%d) on the other lines.", c, rows, path, cols);
        }
    }
    if (rows == 0)
    {
        fclose(f);
        FAIL(GRTCODE_VALUE_ERR, "the file %s is empty.", path);
    }
    if (--rows == 0)
    {
        fclose(f);
        FAIL(GRTCODE_VALUE_ERR, "the file %s only contains headers, no data.", path);
    }
    fp_t *v = (fp_t *)malloc(sizeof(*v)*(size_t)rows*(size_t)cols);
    if (v == NULL)
    {
        fclose(f);
        FAIL(GRTCODE_NULL_ERR, "out of memory reading %s.", path);
    }
    rewind(f);
    if (fgets(line, sizeof(line), f) == NULL) /* header */
    {
        free(v);
        fclose(f);
        FAIL(GRTCODE_IO_ERR, "failed to read %s.", path);
    }
    int r;
    for (r=0; r<rows; ++r)
    {
        if (fgets(line, sizeof(line), f) == NULL)
        {
            free(v);
            fclose(f);
            FAIL(GRTCODE_IO_ERR, "failed to read %s.", path);
        }
        char *save = NULL;
        char *tok = strtok_r(line, ",", &save);
        int c;
        for (c=0; c<cols; ++c)
        {
            size_t const len = tok != NULL ? strlen(tok) : 0;
            if (len < 1 || len > MAX_TOKEN - 1)
            {
                free(v);
                fclose(f);
                FAIL(GRTCODE_RANGE_ERR, "bad token length %zu on line %d of file %s.", len, r + 2,
                     path);
            }
            double d;
            int const rc = to_double(tok, &d);
            if (rc != GRTCODE_SUCCESS)
            {
                free(v);
                fclose(f);
                grt_error_frame(__FILE__, __LINE__);
                return rc;
            }
            v[(size_t)c*rows + r] = d;
            tok = strtok_r(NULL, ",", &save);
        }
    }
    if (fclose(f) != 0)
    {
        free(v);
        FAIL(GRTCODE_VALUE_ERR, "failed to close csv file %s.", path);
    }
    *num_rows = rows;
    *num_cols = cols;
    *values = v;
    return GRTCODE_SUCCESS;
}

/* The samples of a two-column table (column 0 = x, column 1 = y) for the device resampler:
 * reads the file, checks its column count and the conditions interpolate2 checks
 * (utilities.c:157-166: at least two points, x strictly increasing).  *table is the
 * column-major buffer of grt_read_csv (free it); x = *table, y = *table + *rows. */
int grt_table_samples(char const *path, int expected_cols, fp_t **table, int *rows)
{
    int cols;
    fp_t *raw = NULL;
    TRY(grt_read_csv(path, rows, &cols, &raw));
    if (cols != expected_cols)
    {
        free(raw);
        FAIL(GRTCODE_VALUE_ERR, "The number of columns (%d) in file %s does not match the"
             " expected number (%d).", cols, path, expected_cols);
    }
    if (*rows < 2)
    {
        free(raw);
        FAIL(GRTCODE_VALUE_ERR, "at least two x points required (%d given).", *rows);
    }
    if (!monotonically_increasing(raw, (size_t)*rows))
    {
        free(raw);
        FAIL(GRTCODE_VALUE_ERR, "x (%p) must be monotonically increasing.", (void const *)raw);
    }
    *table = raw;
    return GRTCODE_SUCCESS;
}

int grt_table_on_grid(char const *path, int expected_cols, double w0, double dw,
                      unsigned long long n, fp_t **on_grid)
{
    int rows;
    int cols;
    fp_t *raw = NULL;
    TRY(grt_read_csv(path, &rows, &cols, &raw));
    if (cols != expected_cols)
    {
        free(raw);
        FAIL(GRTCODE_VALUE_ERR, "The number of columns (%d) in file %s does not match the"
             " expected number (%d).", cols, path, expected_cols);
    }
    fp_t *w = (fp_t *)malloc(sizeof(*w)*n);
    fp_t *out = (fp_t *)calloc(n, sizeof(*out));
    if (w == NULL || out == NULL)
    {
        free(raw);
        free(w);
        free(out);
        FAIL(GRTCODE_NULL_ERR, "out of memory resampling %s.", path);
    }
    unsigned long long i;
    for (i=0; i<n; ++i)
    {
        w[i] = w0 + i*dw;
    }
    int const rc = interpolate2(raw, raw + rows, (size_t)rows, w, out, (size_t)n, linear_sample,
                                NULL);
    free(raw);
    free(w);
    if (rc != GRTCODE_SUCCESS)
    {
        free(out);
        grt_error_frame(__FILE__, __LINE__);
        return rc;
    }
    *on_grid = out;
    return GRTCODE_SUCCESS;
}
