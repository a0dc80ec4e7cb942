/* Substitute total internal partition sums -- see tips.h.
 *
 * Q(mol, T, iso) = linear interpolation of a 1 K table.  The table is built
 * once per process from
 *     base(T) = q296(mol) * isofac(iso) * (T/296)^beta(mol)
 * multiplied by a correction c(T) that is piecewise linear in T through the
 * ratios anchor/base at the temperatures where the reference's unit tests pin
 * a value (constant outside the anchored range).  Finally the two table
 * entries that bracket an anchor are adjusted so that the interpolated value
 * AT the anchor temperature equals the anchor.
 *
 * Call sites this serves in the reference: parse_HITRAN_file.c:382
 * (Q at 296 K), kernels.c:62 (1/Q per layer), gas_optics.c:169 (inittips_d).
 */
#include <math.h>
#include <pthread.h>
#include <stddef.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "molecule_table.h"
#include "tips.h"

typedef struct Anchor
{
    int mol;
    int iso;
    double t;
    double q;
} Anchor_t;

/* Pinned values.  H2O isotopologue 1 at the five test_kernels.c layer
 * temperatures: derived from the line-strength goldens (test_kernels.c:208-217),
 * which pin Q to ~8 digits; at 275.234324 K from test_tips2017.c:34-65.
 * H2O isotopologues 2-9: 1/q_ref of test_kernels.c:180-190 (6 digits). */
static Anchor_t const anchors[] = {
    {1, 1, 230.92, 120.58294649}, {1, 1, 236.24, 124.73834586},
    {1, 1, 241.93, 129.23429516}, {1, 1, 275.15, 156.53744172},
    {1, 1, 275.234324, 156.6091754}, {1, 1, 288.99, 168.44124037},
#define H2O_ISO(iso, a, b, c, d, e) \
    {1, iso, 230.92, 1./a}, {1, iso, 236.24, 1./b}, {1, iso, 241.93, 1./c}, \
    {1, iso, 275.15, 1./d}, {1, iso, 288.99, 1./e}
    H2O_ISO(2, 0.00822385, 0.00794978, 0.00767377, 0.00633500, 0.00588716),
    H2O_ISO(3, 0.00137605, 0.00133023, 0.00128397, 0.00105998, 0.00098513),
    H2O_ISO(4, 0.00167721, 0.00162114, 0.00156452, 0.00129043, 0.00119881),
    H2O_ISO(5, 0.00165653, 0.00160113, 0.00154519, 0.00127448, 0.00118398),
    H2O_ISO(6, 0.00027761, 0.00026832, 0.00025894, 0.00021353, 0.00019835),
    H2O_ISO(7, 0.00141451, 0.00136700, 0.00131902, 0.00108669, 0.00100896),
    H2O_ISO(8, 0.00139350, 0.00134669, 0.00129942, 0.00107048, 0.00099389),
    H2O_ISO(9, 0.00023390, 0.00022604, 0.00021811, 0.00017968, 0.00016683),
#undef H2O_ISO
    {2, 1, 275.234324, 261.25798746},
    {3, 1, 275.234324, 3087.3115616000005},
    {4, 1, 275.234324, 4524.7762498},
    {6, 1, 275.234324, 528.2642260800001}
};
#define NUM_ANCHORS ((int)(sizeof(anchors)/sizeof(anchors[0])))

typedef struct MolInfo
{
    int num_iso;
    double q296;
    double beta;
} MolInfo_t;

static MolInfo_t const mols[GRT_TIPS_NUM_MOLS + 1] = {
    {0, 1., 0.},
#define X(id, name, mass, niso, q296, beta) {niso, q296, beta},
    GRT_MOLECULE_TABLE(X)
#undef X
};

static double table[GRT_TIPS_NUM_MOLS + 1][GRT_TIPS_MAX_ISO][GRT_TIPS_NT];
static pthread_once_t once = PTHREAD_ONCE_INIT;

/* The interpolant of one table row at T (inside the table range), for the loader. */
static double grt_tips_q_unlocked(double const *tab, double T)
{
    double const tmax = (double)(GRT_TIPS_TMIN + GRT_TIPS_NT - 1);
    if (T < (double)GRT_TIPS_TMIN) T = (double)GRT_TIPS_TMIN;
    if (T >= tmax) return tab[GRT_TIPS_NT-1];
    int const it = (int)T;
    double const f = T - (double)it;
    int const idx = it - GRT_TIPS_TMIN;
    return tab[idx] + f*(tab[idx+1] - tab[idx]);
}


/* Synthetic ratio Q_iso/Q_main for isotopologues that have no anchors. */
static double isofac(int const iso)
{
    return 1. + 0.5*(iso - 1);
}


static double base(int const mol, int const iso, double const t)
{
    return mols[mol].q296*isofac(iso)*pow(t/296., mols[mol].beta);
}


static void build(void)
{
    int mol;
    for (mol=1; mol<=GRT_TIPS_NUM_MOLS; ++mol)
    {
        int iso;
        for (iso=1; iso<=GRT_TIPS_MAX_ISO; ++iso)
        {
            /* Anchors of this (mol, iso); the list above is sorted by T. */
            Anchor_t const *a[8];
            int na = 0;
            int k;
            for (k=0; k<NUM_ANCHORS; ++k)
            {
                if (anchors[k].mol == mol && anchors[k].iso == iso && na < 8)
                {
                    a[na++] = &anchors[k];
                }
            }
            double *tab = table[mol][iso-1];
            int i;
            for (i=0; i<GRT_TIPS_NT; ++i)
            {
                double const t = (double)(GRT_TIPS_TMIN + i);
                double c = 1.;
                if (na > 0)
                {
                    if (t <= a[0]->t)
                    {
                        c = a[0]->q/base(mol, iso, a[0]->t);
                    }
                    else if (t >= a[na-1]->t)
                    {
                        c = a[na-1]->q/base(mol, iso, a[na-1]->t);
                    }
                    else
                    {
                        for (k=0; k<na-1; ++k)
                        {
                            if (t >= a[k]->t && t <= a[k+1]->t)
                            {
                                double const r0 = a[k]->q/base(mol, iso, a[k]->t);
                                double const r1 = a[k+1]->q/base(mol, iso, a[k+1]->t);
                                double const f = (t - a[k]->t)/(a[k+1]->t - a[k]->t);
                                c = r0 + f*(r1 - r0);
                                break;
                            }
                        }
                    }
                }
                tab[i] = base(mol, iso, t)*c;
            }

            /* Make the interpolant pass through each anchor. */
            for (k=0; k<na; ++k)
            {
                int const j = (int)floor(a[k]->t) - GRT_TIPS_TMIN;
                double const f = a[k]->t - floor(a[k]->t);
                if (k + 1 < na && (int)floor(a[k+1]->t) - GRT_TIPS_TMIN == j)
                {
                    /* Two anchors inside one 1 K interval: the straight line
                       through both fixes the two bracketing entries. */
                    double const f2 = a[k+1]->t - floor(a[k+1]->t);
                    double const slope = (a[k+1]->q - a[k]->q)/(f2 - f);
                    tab[j] = a[k]->q - slope*f;
                    tab[j+1] = tab[j] + slope;
                    ++k;
                }
                else
                {
                    double const v = tab[j] + f*(tab[j+1] - tab[j]);
                    double const s = a[k]->q/v;
                    tab[j] *= s;
                    tab[j+1] *= s;
                }
            }
        }
    }
}


/* ---- external partition sums ---------------------------------------------------------
 * GRT_TIPS_TABLE=<file> in the environment replaces the substitute for every (molecule,
 * isotopologue) the file lists: text rows "mol iso T Q" (blank- or comma-separated, '#'
 * starts a comment), T ascending within a (mol, iso) group, at least two rows per group.
 * The 1 K table is filled by linear interpolation between the given points; outside the
 * given range the substitute continues, scaled to meet the table at its ends.  The file is
 * read once per process, inside the same table build the oracle's shim uses
 * (oracle/tips_shim.c links this file), so both sides keep evaluating one function. */
static int external_groups[GRT_TIPS_NUM_MOLS + 1];
static char external_path[1024];
static int external_rows = -1;        /* -1: no file; >= 0: rows used */

typedef struct TipsRow
{
    int mol;
    int iso;
    double t;
    double q;
} TipsRow_t;

static void apply_group(TipsRow_t const *r, int const n)
{
    if (n < 2)
    {
        return;
    }
    int const mol = r[0].mol;
    int const iso = r[0].iso;
    double *tab = table[mol][iso-1];
    double const lo_ratio = r[0].q/grt_tips_q_unlocked(tab, r[0].t);
    double const hi_ratio = r[n-1].q/grt_tips_q_unlocked(tab, r[n-1].t);
    int i;
    int k = 0;
    for (i=0; i<GRT_TIPS_NT; ++i)
    {
        double const t = (double)(GRT_TIPS_TMIN + i);
        if (t <= r[0].t)
        {
            tab[i] *= lo_ratio;
        }
        else if (t >= r[n-1].t)
        {
            tab[i] *= hi_ratio;
        }
        else
        {
            while (k + 2 < n && r[k+1].t < t)
            {
                ++k;
            }
            double const f = (t - r[k].t)/(r[k+1].t - r[k].t);
            tab[i] = r[k].q + f*(r[k+1].q - r[k].q);
        }
    }
    external_groups[mol] += 1;
}

static void load_external(char const *path)
{
    FILE *f = fopen(path, "r");
    external_rows = 0;
    snprintf(external_path, sizeof(external_path), "%s", path);
    if (f == NULL)
    {
        fprintf(stderr, "[grtcode] GRT_TIPS_TABLE=%s cannot be opened: keeping the substitute "
                "partition sums.\n", path);
        return;
    }
    size_t cap = 4096, n = 0;
    TipsRow_t *rows = (TipsRow_t *)malloc(sizeof(TipsRow_t)*cap);
    char line[512];
    while (rows != NULL && fgets(line, sizeof(line), f) != NULL)
    {
        char *c;
        for (c=line; *c; ++c)
        {
            if (*c == ',' || *c == ';' || *c == '\t') *c = ' ';
            if (*c == '#') { *c = '\0'; break; }
        }
        TipsRow_t r;
        if (sscanf(line, "%d %d %lf %lf", &r.mol, &r.iso, &r.t, &r.q) != 4)
        {
            continue;
        }
        if (r.mol < 1 || r.mol > GRT_TIPS_NUM_MOLS || r.iso < 1 || r.iso > GRT_TIPS_MAX_ISO ||
            !(r.q > 0.) || !(r.t > 0.))
        {
            continue;
        }
        if (n == cap)
        {
            cap *= 2;
            TipsRow_t *grown = (TipsRow_t *)realloc(rows, sizeof(TipsRow_t)*cap);
            if (grown == NULL) break;
            rows = grown;
        }
        rows[n++] = r;
    }
    fclose(f);
    size_t start = 0, i;
    for (i=1; i<=n; ++i)
    {
        int const boundary = i == n || rows[i].mol != rows[start].mol ||
                             rows[i].iso != rows[start].iso || !(rows[i].t > rows[i-1].t);
        if (boundary)
        {
            apply_group(rows + start, (int)(i - start));
            if (i - start >= 2) external_rows += (int)(i - start);
            start = i;
        }
    }
    free(rows);
}

static void build_all(void)
{
    build();
    char const *path = getenv("GRT_TIPS_TABLE");
    if (path != NULL && path[0] != '\0')
    {
        load_external(path);
    }
}

/* 1 if Q of this molecule is the built-in substitute (no external group was loaded). */
int grt_tips_is_substitute(int const mol)
{
    pthread_once(&once, build_all);
    if (mol < 1 || mol > GRT_TIPS_NUM_MOLS)
    {
        return 1;
    }
    return external_groups[mol] == 0;
}

/* Path of the external table in use (empty string if none) and the rows taken from it. */
char const *grt_tips_external(int * const rows)
{
    pthread_once(&once, build_all);
    if (rows != NULL) *rows = external_rows < 0 ? 0 : external_rows;
    return external_rows < 0 ? "" : external_path;
}

double const *grt_tips_table(int const mol, int const iso)
{
    pthread_once(&once, build_all);
    if (mol < 1 || mol > GRT_TIPS_NUM_MOLS || iso < 1 || iso > GRT_TIPS_MAX_ISO)
    {
        return NULL;
    }
    return table[mol][iso-1];
}


int grt_tips_num_iso(int const mol)
{
    if (mol < 1 || mol > GRT_TIPS_NUM_MOLS)
    {
        return 0;
    }
    return mols[mol].num_iso;
}


/* Evaluation order is part of the contract with the device version
   (csrc/column_prep.cu): idx = (int)T - TMIN; f = T - (double)(int)T;
   q = tab[idx] + f*(tab[idx+1] - tab[idx]), each operation rounded once. */
double grt_tips_q(int const mol, double const T, int const iso)
{
    double const *tab = grt_tips_table(mol, iso);
    if (tab == NULL)
    {
        return 1.;
    }
    double t = T;
    double const tmin = (double)GRT_TIPS_TMIN;
    double const tmax = (double)(GRT_TIPS_TMIN + GRT_TIPS_NT - 1);
    if (!(t >= tmin))
    {
        t = tmin;
    }
    if (t >= tmax)
    {
        return tab[GRT_TIPS_NT-1];
    }
    int const it = (int)t;
    double const f = t - (double)it;
    int const idx = it - GRT_TIPS_TMIN;
    double const d = tab[idx+1] - tab[idx];
    double const fd = f*d;
    return tab[idx] + fd;
}
