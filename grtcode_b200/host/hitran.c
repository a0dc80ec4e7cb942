/* HITRAN-2012 160-character record parser.
 *
 * Behaviour follows gas-optics/src/parse_HITRAN_file.c:224-413:
 *   - field widths 2,1,12,10,10,5,5,10,4,8 then 100 characters that are ignored
 *     (:79-100); records shorter than 160 or longer than 162 characters are errors;
 *   - the 1-character isotopologue code maps '0' -> 10 and 'A'.. -> 11.. (:177-194);
 *   - v and S stay double; gamma_air, gamma_self, E'', n_air and delta are rounded
 *     through float (:197-213, 331-345) -- this is part of the numerical contract;
 *   - finally S *= Q(296)/(exp(c2 E''/296)(1 - exp(c2 v/296))), c2 = -1.4387686f (:372-384).
 * The whole file is read in one pass into memory instead of twice line by line. */
#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include "errors.h"
#include "hitran.h"
#include "tips.h"
#include "utilities.h"

#define RECORD_LEN 160
#define RECORD_PAD 2

void free_line_params(LineParams_t * const lp)
{
    if (lp == NULL)
    {
        return;
    }
    free(lp->iso);
    free(lp->vnn);
    free(lp->snn);
    free(lp->yair);
    free(lp->yself);
    free(lp->en);
    free(lp->n);
    free(lp->d);
    memset(lp, 0, sizeof(*lp));
}

static int reserve(LineParams_t *lp, uint64_t capacity)
{
    size_t const c = capacity > 0 ? capacity : 1;
    lp->iso = (int *)realloc(lp->iso, sizeof(int)*c);
    lp->vnn = (fp_t *)realloc(lp->vnn, sizeof(fp_t)*c);
    lp->snn = (fp_t *)realloc(lp->snn, sizeof(fp_t)*c);
    lp->yair = (fp_t *)realloc(lp->yair, sizeof(fp_t)*c);
    lp->yself = (fp_t *)realloc(lp->yself, sizeof(fp_t)*c);
    lp->en = (fp_t *)realloc(lp->en, sizeof(fp_t)*c);
    lp->n = (fp_t *)realloc(lp->n, sizeof(fp_t)*c);
    lp->d = (fp_t *)realloc(lp->d, sizeof(fp_t)*c);
    if (!lp->iso || !lp->vnn || !lp->snn || !lp->yair || !lp->yself || !lp->en || !lp->n || !lp->d)
    {
        FAIL(GRTCODE_NULL_ERR, "out of memory for %llu spectral lines.",
             (unsigned long long)capacity);
    }
    return GRTCODE_SUCCESS;
}

/* Copies a fixed-width field and converts it. */
static int field_double(char const *rec, int offset, int width, double *out)
{
    char tmp[24];
    memcpy(tmp, rec + offset, (size_t)width);
    tmp[width] = '\0';
    TRY(to_double(tmp, out));
    return GRTCODE_SUCCESS;
}

static int field_float(char const *rec, int offset, int width, int column, fp_t *out)
{
    double d;
    TRY(field_double(rec, offset, width, &d));
    if (!(d >= -1.*FLT_MAX && d <= FLT_MAX))
    {
        FAIL(GRTCODE_VALUE_ERR, "value %e from column %d cannot be safely cast as a float.", d,
             column);
    }
    *out = (fp_t)(float)d;
    return GRTCODE_SUCCESS;
}

int parse_hitran_file(LineParams_t * const out, char const * const filepath, int const mol_id,
                      double const w0, double const wn)
{
    REQUIRE_PTR(out);
    REQUIRE_PTR(filepath);
    memset(out, 0, sizeof(*out));
    LOG_INFO("Opening and reading HITRAN line parameters from file %s.", filepath);
    FILE *fp = NULL;
    TRY(open_file(&fp, filepath, "r"));
    LineParams_t lp;
    memset(&lp, 0, sizeof(lp));
    uint64_t capacity = 1 << 16;
    int rc = reserve(&lp, capacity);
    char *buf = NULL;
    size_t buf_len = 0;
    ssize_t ll;
    uint64_t n = 0;
    size_t line_count = 0;
    while (rc == GRTCODE_SUCCESS && (ll = getline(&buf, &buf_len, fp)) != -1)
    {
        ++line_count;
        if (ll > RECORD_LEN + RECORD_PAD || ll < RECORD_LEN)
        {
            grt_error_reset();
            char m[512];
            snprintf(m, sizeof(m), "Error: Found bad record at line %zu (%zd characters, expected"
                     " %d to %d) in file %s.\nBacktrace:", line_count, ll, RECORD_LEN,
                     RECORD_LEN + RECORD_PAD, filepath);
            grt_error_append(m);
            grt_error_frame(__FILE__, __LINE__);
            rc = GRTCODE_VALUE_ERR;
            break;
        }
        char tmp[4];
        tmp[0] = buf[0];
        tmp[1] = buf[1];
        tmp[2] = '\0';
        int mol;
        rc = to_int(tmp, &mol);
        if (rc != GRTCODE_SUCCESS || mol != mol_id)
        {
            continue;
        }
        if (n == capacity)
        {
            capacity *= 2;
            rc = reserve(&lp, capacity);
            if (rc != GRTCODE_SUCCESS)
            {
                break;
            }
        }
        char const c = buf[2];
        if (c == '0')
        {
            lp.iso[n] = 10;
        }
        else if (c >= 'A' && c <= 'Z')
        {
            lp.iso[n] = c - 'A' + 11;
        }
        else
        {
            tmp[0] = c;
            tmp[1] = '\0';
            rc = to_int(tmp, &lp.iso[n]);
        }
        /* offsets: 3 v(12) | 15 S(10) | 25 A(10) | 35 g_air(5) | 40 g_self(5) |
           45 E''(10) | 55 n(4) | 59 delta(8) */
        if (rc == GRTCODE_SUCCESS) rc = field_double(buf, 3, 12, &lp.vnn[n]);
        if (rc == GRTCODE_SUCCESS) rc = field_double(buf, 15, 10, &lp.snn[n]);
        if (rc == GRTCODE_SUCCESS) rc = field_float(buf, 35, 5, 5, &lp.yair[n]);
        if (rc == GRTCODE_SUCCESS) rc = field_float(buf, 40, 5, 6, &lp.yself[n]);
        if (rc == GRTCODE_SUCCESS) rc = field_float(buf, 45, 10, 7, &lp.en[n]);
        if (rc == GRTCODE_SUCCESS) rc = field_float(buf, 55, 4, 8, &lp.n[n]);
        if (rc == GRTCODE_SUCCESS) rc = field_float(buf, 59, 8, 9, &lp.d[n]);
        if (rc != GRTCODE_SUCCESS)
        {
            grt_error_frame(__FILE__, __LINE__);
            break;
        }
        if ((w0 < 0 && wn < 0) || (lp.vnn[n] >= w0 && lp.vnn[n] <= wn))
        {
            ++n;
        }
    }
    free(buf);
    if (fclose(fp) != 0 && rc == GRTCODE_SUCCESS)
    {
        free_line_params(&lp);
        FAIL(GRTCODE_IO_ERR, "error closing file %s.", filepath);
    }
    if (rc != GRTCODE_SUCCESS)
    {
        free_line_params(&lp);
        return rc;
    }

    /* Scale the raw intensities (parse_HITRAN_file.c:372-384). */
    fp_t const tref = 296.f;
    fp_t const c2 = -1.4387686f;
    uint64_t i;
    for (i=0; i<n; ++i)
    {
        lp.snn[i] *= grt_tips_q(mol_id, tref, lp.iso[i])/(exp(c2*lp.en[i]/tref)*
                     (1.f - exp(c2*lp.vnn[i]/tref)));
    }
    lp.num_lines = n;
    *out = lp;
    return GRTCODE_SUCCESS;
}
