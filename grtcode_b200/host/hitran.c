/* HITRAN-2012 160-character record parser.
 *
 * Behaviour follows gas-optics/src/parse_HITRAN_file.c:224-413:
 *   - field widths 2,1,12,10,10,5,5,10,4,8 then 100 characters that are ignored
 *     (:79-100); records shorter than 160 or longer than 162 characters are errors;
 *   - the 1-character isotopologue code maps '0' -> 10 and 'A'.. -> 11.. (:177-194);
 *   - v and S stay double; gamma_air, gamma_self, E'', n_air and delta are rounded
 *     through float (:197-213, 331-345) -- this is part of the numerical contract;
 *   - finally S *= Q(296)/(exp(c2 E''/296)(1 - exp(c2 v/296))), c2 = -1.4387686f (:372-384).
 * The file is read and converted once per process, see HitranTable_t below. */
#include <float.h>
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>
#include <unistd.h>
#include "errors.h"
#include "grt_sm100.h"
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

/* Isotopologue code (parse_HITRAN_file.c:177-194). */
static int parse_iso(char const c, int *iso)
{
    if (c == '0')
    {
        *iso = 10;
        return GRTCODE_SUCCESS;
    }
    if (c >= 'A' && c <= 'Z')
    {
        *iso = c - 'A' + 11;
        return GRTCODE_SUCCESS;
    }
    char tmp[2];
    tmp[0] = c;
    tmp[1] = '\0';
    return to_int(tmp, iso);
}

/* One record -> one row.  offsets: 3 v(12) | 15 S(10) | 25 A(10) | 35 g_air(5) |
   40 g_self(5) | 45 E''(10) | 55 n(4) | 59 delta(8). */
static int parse_record(char const *rec, int *iso, fp_t *vnn, fp_t *snn, fp_t *yair,
                        fp_t *yself, fp_t *en, fp_t *n, fp_t *d)
{
    int rc = parse_iso(rec[2], iso);
    /* v and S: lp.vnn[n] = (fp_t)(val.d) (parse_HITRAN_file.c:319-324) */
    double dv = 0., ds = 0.;
    if (rc == GRTCODE_SUCCESS) rc = field_double(rec, 3, 12, &dv);
    if (rc == GRTCODE_SUCCESS) rc = field_double(rec, 15, 10, &ds);
    *vnn = (fp_t)dv;
    *snn = (fp_t)ds;
    if (rc == GRTCODE_SUCCESS) rc = field_float(rec, 35, 5, 5, yair);
    if (rc == GRTCODE_SUCCESS) rc = field_float(rec, 40, 5, 6, yself);
    if (rc == GRTCODE_SUCCESS) rc = field_float(rec, 45, 10, 7, en);
    if (rc == GRTCODE_SUCCESS) rc = field_float(rec, 55, 4, 8, n);
    if (rc == GRTCODE_SUCCESS) rc = field_float(rec, 59, 8, 9, d);
    return rc;
}

/* The reference reads the whole file once per add_molecule (parse_HITRAN_file.c:224-413:
 * 14 passes for seven molecules on two grids).  Here a file is read and converted ONCE, by
 * all host cores, into a table of every record; parse_hitran_file() then filters that
 * table by molecule and wavenumber range.  The table is keyed by (path, size, mtime); the
 * two most recently used files are kept (one per spectral band). */
enum { ROW_OK = 0, ROW_BAD_MOLECULE = 1, ROW_BAD_FIELD = 2, ROW_HOST = 3 };

/* Where the records are converted: on the device a GasOptics_t was created on
 * (grt_hitran_use_device, called by create_gas_optics; csrc/hitran_parse.cu), or -- for the
 * host-only parse_hitran_file() of a process that has created none, and with GRT_PARSE=host --
 * by the host threads below. */
static int parse_device = -1;

void grt_hitran_use_device(int const device)
{
    parse_device = device;
}

typedef struct HitranTable
{
    char *path;
    off_t size;
    struct timespec mtime;
    uint64_t num_records;
    char *text;               /* the file, kept while rows may have to be re-read for errors */
    uint64_t *start;          /* offset of every record in text */
    signed char *status;
    short *mol;
    int *iso;
    fp_t *vnn, *snn, *yair, *yself, *en, *n, *d;
    uint64_t stamp;
} HitranTable_t;

#define TABLE_SLOTS 2
static HitranTable_t tables[TABLE_SLOTS];
static uint64_t table_clock = 0;

static void free_table(HitranTable_t *t)
{
    free(t->path); free(t->text); free(t->start); free(t->status); free(t->mol); free(t->iso);
    free(t->vnn); free(t->snn); free(t->yair); free(t->yself); free(t->en); free(t->n);
    free(t->d);
    memset(t, 0, sizeof(*t));
}

void grt_hitran_cache_clear(void)
{
    int i;
    for (i=0; i<TABLE_SLOTS; ++i)
    {
        free_table(&tables[i]);
    }
}

typedef struct ParseJob
{
    HitranTable_t *t;
    uint64_t first, last;
} ParseJob_t;

/* Worker: converts rows [first, last).  No error text here (the error buffer is global):
 * a failing row is only marked and re-read serially when a caller asks for it. */
static void *parse_rows(void *arg)
{
    ParseJob_t const *job = (ParseJob_t const *)arg;
    HitranTable_t *t = job->t;
    uint64_t i;
    for (i=job->first; i<job->last; ++i)
    {
        char const *rec = t->text + t->start[i];
        char tmp[3];
        tmp[0] = rec[0];
        tmp[1] = rec[1];
        tmp[2] = '\0';
        char *end = NULL;
        long const mol = strtol(tmp, &end, 10);
        if (end == tmp || *end != '\0')
        {
            t->status[i] = ROW_BAD_MOLECULE;
            t->mol[i] = -1;
            continue;
        }
        t->mol[i] = (short)mol;
        t->status[i] = ROW_OK;
    }
    return NULL;
}

static void *convert_rows(void *arg)
{
    ParseJob_t const *job = (ParseJob_t const *)arg;
    HitranTable_t *t = job->t;
    uint64_t i;
    for (i=job->first; i<job->last; ++i)
    {
        if (t->status[i] != ROW_OK)
        {
            continue;
        }
        if (parse_record(t->text + t->start[i], &t->iso[i], &t->vnn[i], &t->snn[i], &t->yair[i],
                         &t->yself[i], &t->en[i], &t->n[i], &t->d[i]) != GRTCODE_SUCCESS)
        {
            t->status[i] = ROW_BAD_FIELD;
        }
    }
    return NULL;
}

static int worker_count(uint64_t const rows)
{
    long n = sysconf(_SC_NPROCESSORS_ONLN);
    char const *env = getenv("GRT_PARSE_THREADS");
    if (env != NULL && atoi(env) > 0)
    {
        n = atoi(env);
    }
    if (n > 64) n = 64;
    if (n < 1) n = 1;
    if ((uint64_t)n > rows/4096 + 1) n = (long)(rows/4096 + 1);
    return (int)n;
}

static void run_parallel(HitranTable_t *t, void *(*fn)(void *))
{
    int const nthreads = worker_count(t->num_records);
    pthread_t tid[64];
    ParseJob_t job[64];
    int started[64];
    int k;
    for (k=0; k<nthreads; ++k)
    {
        job[k].t = t;
        job[k].first = t->num_records*(uint64_t)k/(uint64_t)nthreads;
        job[k].last = t->num_records*(uint64_t)(k + 1)/(uint64_t)nthreads;
        started[k] = (k > 0) && (pthread_create(&tid[k], NULL, fn, &job[k]) == 0);
    }
    for (k=0; k<nthreads; ++k)
    {
        if (!started[k])
        {
            fn(&job[k]);       /* the caller's share, and any share whose thread failed to start */
        }
    }
    for (k=0; k<nthreads; ++k)
    {
        if (started[k])
        {
            pthread_join(tid[k], NULL);
        }
    }
}

/* Reads and converts a file; bad record lengths are reported exactly as the reference does. */
static int load_table(HitranTable_t *t, char const *filepath, struct stat const *st)
{
    FILE *fp = NULL;
    TRY(open_file(&fp, filepath, "r"));
    size_t const size = (size_t)st->st_size;
    t->text = (char *)malloc(size + 1);
    if (t->text == NULL)
    {
        fclose(fp);
        FAIL(GRTCODE_NULL_ERR, "out of memory reading %s.", filepath);
    }
    size_t const got = fread(t->text, 1, size, fp);
    if (fclose(fp) != 0 || got != size)
    {
        FAIL(GRTCODE_IO_ERR, "error reading file %s.", filepath);
    }
    t->text[size] = '\0';
    /* index the records */
    uint64_t capacity = size/(RECORD_LEN + 1) + 2;
    t->start = (uint64_t *)malloc(sizeof(uint64_t)*capacity);
    if (t->start == NULL)
    {
        FAIL(GRTCODE_NULL_ERR, "out of memory indexing %s.", filepath);
    }
    uint64_t n = 0;
    size_t pos = 0;
    while (pos < size)
    {
        char const *nl = (char const *)memchr(t->text + pos, '\n', size - pos);
        size_t const ll = nl != NULL ? (size_t)(nl - (t->text + pos)) + 1 : size - pos;
        if (ll > RECORD_LEN + RECORD_PAD || ll < RECORD_LEN)
        {
            grt_error_reset();
            char m[512];
            snprintf(m, sizeof(m), "Error: Found bad record at line %llu (%zu characters, expected"
                     " %d to %d) in file %s.\nBacktrace:", (unsigned long long)(n + 1), ll,
                     RECORD_LEN, RECORD_LEN + RECORD_PAD, filepath);
            grt_error_append(m);
            grt_error_frame(__FILE__, __LINE__);
            return GRTCODE_VALUE_ERR;
        }
        if (n == capacity)
        {
            capacity *= 2;
            uint64_t *grown = (uint64_t *)realloc(t->start, sizeof(uint64_t)*capacity);
            if (grown == NULL)
            {
                FAIL(GRTCODE_NULL_ERR, "out of memory indexing %s.", filepath);
            }
            t->start = grown;
        }
        t->start[n++] = pos;
        pos += ll;
    }
    t->num_records = n;
    size_t const c = n > 0 ? n : 1;
    t->status = (signed char *)malloc(c);
    t->mol = (short *)malloc(sizeof(short)*c);
    t->iso = (int *)malloc(sizeof(int)*c);
    t->vnn = (fp_t *)malloc(sizeof(fp_t)*c);
    t->snn = (fp_t *)malloc(sizeof(fp_t)*c);
    t->yair = (fp_t *)malloc(sizeof(fp_t)*c);
    t->yself = (fp_t *)malloc(sizeof(fp_t)*c);
    t->en = (fp_t *)malloc(sizeof(fp_t)*c);
    t->n = (fp_t *)malloc(sizeof(fp_t)*c);
    t->d = (fp_t *)malloc(sizeof(fp_t)*c);
    if (!t->status || !t->mol || !t->iso || !t->vnn || !t->snn || !t->yair || !t->yself ||
        !t->en || !t->n || !t->d)
    {
        FAIL(GRTCODE_NULL_ERR, "out of memory for %llu spectral lines.", (unsigned long long)n);
    }
    char const *where = getenv("GRT_PARSE");
    if (parse_device < 0 || n == 0 || (where != NULL && strcmp(where, "host") == 0))
    {
        run_parallel(t, parse_rows);
        run_parallel(t, convert_rows);
        return GRTCODE_SUCCESS;
    }
    GPU(grt_hitran_convert(parse_device, t->text, size, t->start, n, t->mol, t->iso, t->vnn,
                           t->snn, t->yair, t->yself, t->en, t->n, t->d, t->status));
    /* Rows the device did not decide (unusual text, malformed fields): strtod, as above. */
    uint64_t i;
    for (i=0; i<n; ++i)
    {
        if (t->status[i] == ROW_HOST)
        {
            ParseJob_t one;
            one.t = t;
            one.first = i;
            one.last = i + 1;
            parse_rows(&one);
            convert_rows(&one);
        }
    }
    return GRTCODE_SUCCESS;
}

static int find_table(char const *filepath, HitranTable_t **found)
{
    struct stat st;
    if (stat(filepath, &st) != 0)
    {
        /* let open_file produce the reference's message */
        FILE *fp = NULL;
        TRY(open_file(&fp, filepath, "r"));
        fclose(fp);
        FAIL(GRTCODE_IO_ERR, "cannot stat file %s.", filepath);
    }
    int i;
    int victim = 0;
    for (i=0; i<TABLE_SLOTS; ++i)
    {
        HitranTable_t *t = &tables[i];
        if (t->path != NULL && strcmp(t->path, filepath) == 0 && t->size == st.st_size &&
            t->mtime.tv_sec == st.st_mtim.tv_sec && t->mtime.tv_nsec == st.st_mtim.tv_nsec)
        {
            t->stamp = ++table_clock;
            *found = t;
            return GRTCODE_SUCCESS;
        }
        if (tables[i].stamp < tables[victim].stamp)
        {
            victim = i;
        }
    }
    HitranTable_t *t = &tables[victim];
    free_table(t);
    int const rc = load_table(t, filepath, &st);
    if (rc != GRTCODE_SUCCESS)
    {
        free_table(t);
        return rc;
    }
    t->path = strdup(filepath);
    t->size = st.st_size;
    t->mtime = st.st_mtim;
    t->stamp = ++table_clock;
    *found = t;
    return GRTCODE_SUCCESS;
}

int parse_hitran_file(LineParams_t * const out, char const * const filepath, int const mol_id,
                      double const w0, double const wn)
{
    REQUIRE_PTR(out);
    REQUIRE_PTR(filepath);
    memset(out, 0, sizeof(*out));
    LOG_INFO("Opening and reading HITRAN line parameters from file %s.", filepath);
    HitranTable_t *t = NULL;
    TRY(find_table(filepath, &t));
    /* first pass: count, and stop at the first row the reference would have failed on */
    uint64_t i;
    uint64_t n = 0;
    for (i=0; i<t->num_records; ++i)
    {
        if (t->status[i] == ROW_BAD_MOLECULE)
        {
            char tmp[3];
            int mol;
            tmp[0] = t->text[t->start[i]];
            tmp[1] = t->text[t->start[i] + 1];
            tmp[2] = '\0';
            TRY(to_int(tmp, &mol));           /* produces the conversion error */
            FAIL(GRTCODE_VALUE_ERR, "bad molecule id in record %llu.", (unsigned long long)(i + 1));
        }
        if (t->mol[i] != mol_id)
        {
            continue;
        }
        if (t->status[i] == ROW_BAD_FIELD)
        {
            int iso;
            fp_t f[7];
            TRY(parse_record(t->text + t->start[i], &iso, &f[0], &f[1], &f[2], &f[3], &f[4], &f[5],
                             &f[6]));
            FAIL(GRTCODE_VALUE_ERR, "bad field in record %llu.", (unsigned long long)(i + 1));
        }
        if ((w0 < 0 && wn < 0) || (t->vnn[i] >= w0 && t->vnn[i] <= wn))
        {
            ++n;
        }
    }
    LineParams_t lp;
    memset(&lp, 0, sizeof(lp));
    TRY(reserve(&lp, n));
    uint64_t k = 0;
    fp_t const tref = 296.f;
    fp_t const c2 = -1.4387686f;
    for (i=0; i<t->num_records; ++i)
    {
        if (t->mol[i] != mol_id ||
            !((w0 < 0 && wn < 0) || (t->vnn[i] >= w0 && t->vnn[i] <= wn)))
        {
            continue;
        }
        lp.iso[k] = t->iso[i];
        lp.vnn[k] = t->vnn[i];
        lp.yair[k] = t->yair[i];
        lp.yself[k] = t->yself[i];
        lp.en[k] = t->en[i];
        lp.n[k] = t->n[i];
        lp.d[k] = t->d[i];
        /* Scale the raw intensity (parse_HITRAN_file.c:372-384). */
        lp.snn[k] = t->snn[i]*((fp_t)grt_tips_q(mol_id, tref, lp.iso[k])/
                    (GRT_EXP(c2*lp.en[k]/tref)*(1.f - GRT_EXP(c2*lp.vnn[k]/tref))));
        ++k;
    }
    lp.num_lines = n;
    *out = lp;
    return GRTCODE_SUCCESS;
}
