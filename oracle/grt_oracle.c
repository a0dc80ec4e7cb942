/* TEST INFRASTRUCTURE -- CPU restatement ("port") of the reference's hot path.
 *
 * Plain C99, scalar loops (OpenMP over the outer loop only), no shared code
 * with the CUDA product.  Every function restates one reference routine and
 * cites it (paths under NOAA-GFDL/GRTCODE).  Parity status: PINNED -- this
 * file is validated against the compiled reference itself
 * (oracle/_ref/libgrtcode_ref.so, tests/test_oracle.py) and against the
 * known-answer vectors of the reference's unit tests (tests/golden/).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may
 * load the library built from this file.
 *
 * Floating-point contract: the reference mixes float and double on purpose
 * (SURVEY.md F5/F6, appendix A); compile with -ffp-contract=off.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include "../grtcode_b200/host/tips.h"

#define MAXL 201

/* utilities/src/curtis_godson.c:25-36, 59-70, 92-105 */
void ora_layer_quantities(int nlay, double const *p, double const *t, double *n, double *pavg,
                          double *tavg)
{
    for (int i=0; i<nlay; ++i)
    {
        double dp = p[i] - p[i+1];
        dp = dp >= 0.f ? dp : -1.f*dp;
        n[i] = 2.147822334314468e+25*dp;
        pavg[i] = 0.5f*(p[i] + p[i+1]);
        tavg[i] = 0.5f*(t[i] + t[i+1]);
    }
}

void ora_partial_pressure(int nlay, double const *p, double const *x, double const *n,
                          double *ps, double *ns)
{
    double const third = 1.f/3.f;
    double const sixth = 1.f/6.f;
    for (int i=0; i<nlay; ++i)
    {
        ps[i] = third*(x[i]*p[i] + x[i+1]*p[i+1]) + sixth*(x[i]*p[i+1] + x[i+1]*p[i]);
        ns[i] = n[i]*0.5f*(x[i] + x[i+1]);
    }
}

/* gas-optics/src/kernels.c:34-131: centres, strengths, Lorentz and Doppler widths
 * for one molecule; outputs are [nlay][nlines]. */
void ora_line_params(int mol, double mass, int num_iso, uint64_t nlines, int nlay,
                     int const *iso, double const *v0, double const *s0, double const *yair,
                     double const *yself, double const *en, double const *nair,
                     double const *delta, double const *pavg, double const *tavg,
                     double const *ps, double *vnn, double *snn, double *gamma, double *alpha)
{
    double const c2 = -1.4387686f;
    double const tref = 296.f;
    double const sqrt_ln2 = 0.83255461115f;
    double const kb = 1.380658E-16;
    double const c = 2.99792458E10;
    for (int i=0; i<nlay; ++i)
    {
        double q[32];
        for (int k=0; k<num_iso && k<32; ++k)
        {
            q[k] = 1.f/grt_tips_q(mol, tavg[i], k+1);       /* kernels.c:62 */
        }
        for (uint64_t j=0; j<nlines; ++j)
        {
            uint64_t const o = (uint64_t)i*nlines + j;
            vnn[o] = v0[j] + delta[j]*pavg[i];                               /* :44 */
            snn[o] = s0[j]*exp(c2*en[j]/tavg[i])*(1.f - exp(c2*v0[j]/tavg[i]))*q[iso[j]-1]; /* :83 */
            gamma[o] = pow(tref/tavg[i], nair[j])*(yair[j]*(pavg[i]-ps[i]) + yself[j]*ps[i]); /* :105 */
            alpha[o] = sqrt_ln2*vnn[o]*sqrt((2.f*kb*tavg[i])/(mass*c*c));       /* :127 */
        }
    }
}

/* gas-optics/src/RFM_voigt.c:85-281 (Humlicek/RFM Voigt, float internals). */
void ora_voigt(double DWNO, uint64_t N, double WRES, double WNOADJ, double WIDADJ,
               double DOPADJ, double *K)
{
    float const RSQRPI = 0.56418958f;
    float const SQRLN2 = 0.832554611f;
    float const REPWID = SQRLN2/DOPADJ;
    float const Y = REPWID*WIDADJ;
    float const YQ = Y*Y;
    if (Y >= 70.55f)
    {
        for (uint64_t i=0; i<N; ++i)
        {
            float const XI = (DWNO + i*WRES - WNOADJ)*REPWID;
            K[i] = REPWID*Y/(3.14159265358979323846*(XI*XI + YQ));
        }
        return;
    }
    float const YRRTPI = Y*RSQRPI;
    float const XLIM0 = sqrt(15100.0f + Y*(40.0f - Y*3.6f));
    float XLIM1 = (Y >= 8.425f) ? 0.0f : (float)sqrt(164.0f - Y*(4.3f + Y*1.8f));
    float XLIM2 = 6.8f - Y;
    float const XLIM3 = 2.4f*Y;
    float const XLIM4 = 18.1f*Y + 1.65f;
    if (Y <= 0.000001f)
    {
        XLIM1 = XLIM0;
        XLIM2 = XLIM0;
    }
    float const Y0 = 1.5f, Y0PY0 = 3.f, Y0Q = 2.25f;
    float const YPY0 = Y + Y0;
    float const YPY0Q = YPY0*YPY0;
    static float const C[6] = {1.0117281f, -0.75197147f, 0.012557727f, 0.010022008f,
                               -0.00024206814f, 0.00000050084806f};
    static float const S[6] = {1.393237f, 0.23115241f, -0.15535147f, 0.0062183662f,
                               0.000091908299f, -0.00000062752596f};
    static float const T[6] = {0.31424038f, 0.94778839f, 1.5976826f, 2.2795071f, 3.0206370f,
                               3.8897249f};
    for (uint64_t i=0; i<N; ++i)
    {
        float const XI = (DWNO + i*WRES - WNOADJ)*REPWID;
        float const ABX = fabs(XI);
        float const XQ = ABX*ABX;
        if (ABX >= XLIM0)
        {
            K[i] = YRRTPI/(XQ + YQ);
        }
        else if (ABX >= XLIM1)
        {
            float const A0 = YQ + 0.5;
            float const D0 = A0*A0;
            float const D2 = YQ + YQ - 1.0;
            float const D = RSQRPI/(D0 + XQ*(D2 + XQ));
            K[i] = D*Y*(A0 + XQ);
        }
        else if (ABX >= XLIM2)
        {
            float const H0 = 0.5625f + YQ*(4.5f + YQ*(10.5f + YQ*(6.0f + YQ)));
            float const H2 = -4.5f + YQ*(9.0f + YQ*(6.0f + YQ*4.0f));
            float const H4 = 10.5f - YQ*(6.0f - YQ*6.0f);
            float const H6 = -6.0f + YQ*4.0f;
            float const E0 = 1.875f + YQ*(8.25f + YQ*(5.5f + YQ));
            float const E2 = 5.25f + YQ*(1.0f + YQ*3.0f);
            float const E4 = 0.75f*H6;
            float const D = RSQRPI/(H0 + XQ*(H2 + XQ*(H4 + XQ*(H6 + XQ))));
            K[i] = D*Y*(E0 + XQ*(E2 + XQ*(E4 + XQ)));
        }
        else if (ABX < XLIM3)
        {
            float const Z0 = 272.1014f + Y*(1280.829f + Y*(2802.870f + Y*(3764.966f
                             + Y*(3447.629f + Y*(2256.981f + Y*(1074.409f + Y*(369.1989f
                             + Y*(88.26741f + Y*(13.39880f + Y)))))))));
            float const Z2 = 211.678f + Y*(902.3066f + Y*(1758.336f + Y*(2037.310f
                             + Y*(1549.675f + Y*(793.4273f + Y*(266.2987f
                             + Y*(53.59518f + Y*5.0f)))))));
            float const Z4 = 78.86585f + Y*(308.1852f + Y*(497.3014f + Y*(479.2576f
                             + Y*(269.2916f + Y*(80.39278f + Y*10.0f)))));
            float const Z6 = 22.03523f + Y*(55.02933f + Y*(92.75679f + Y*(53.59518f + Y*10.0f)));
            float const Z8 = 1.496460f + Y*(13.39880f + Y*5.0f);
            float const P0 = 153.5168f + Y*(549.3954f + Y*(919.4955f + Y*(946.8970f
                             + Y*(662.8097f + Y*(328.2151f + Y*(115.3772f + Y*(27.93941f
                             + Y*(4.264678f + Y*0.3183291f))))))));
            float const P2 = -34.16955f + Y*(-1.322256f + Y*(124.5975f + Y*(189.7730f
                             + Y*(139.4665f + Y*(56.81652f + Y*(12.79458f + Y*1.2733163f))))));
            float const P4 = 2.584042f + Y*(10.46332f + Y*(24.01655f + Y*(29.81482f
                             + Y*(12.79568f + Y*1.9099744f))));
            float const P6 = -0.07272979f + Y*(0.9377051f + Y*(4.266322f + Y*1.273316f));
            float const P8 = 0.0005480304f + Y*0.3183291f;
            float const D = 1.7724538f/(Z0 + XQ*(Z2 + XQ*(Z4 + XQ*(Z6 + XQ*(Z8 + XQ)))));
            K[i] = D*(P0 + XQ*(P2 + XQ*(P4 + XQ*(P6 + XQ*P8))));
        }
        else
        {
            K[i] = 0.0f;
            float MQ[6], MF[6], XM[6], YM[6], PQ[6], PF[6], XP[6], YP[6];
            for (int J=0; J<=5; J++)
            {
                float D = XI - T[J];
                MQ[J] = D*D;
                MF[J] = 1.0f/(MQ[J] + YPY0Q);
                XM[J] = MF[J]*D;
                YM[J] = MF[J]*YPY0;
                D = XI + T[J];
                PQ[J] = D*D;
                PF[J] = 1.0f/(PQ[J] + YPY0Q);
                XP[J] = PF[J]*D;
                YP[J] = PF[J]*YPY0;
            }
            if (ABX <= XLIM4)
            {
                for (int J=0; J<=5; J++)
                {
                    K[i] = K[i] + C[J]*(YM[J] + YP[J]) - S[J]*(XM[J] - XP[J]);
                }
            }
            else
            {
                float const YF = Y + Y0PY0;
                for (int J=0; J<=5; J++)
                {
                    K[i] = K[i]
                           + (C[J]*(MQ[J]*MF[J] - Y0*YM[J]) + S[J]*YF*XM[J])/(MQ[J] + Y0Q)
                           + (C[J]*(PQ[J]*PF[J] - Y0*YP[J]) - S[J]*YF*XP[J])/(PQ[J] + Y0Q);
                }
                K[i] = Y*K[i] + exp(-XQ);
            }
        }
        K[i] = RSQRPI*REPWID*K[i];
    }
}

/* gas-optics/src/kernels.c:410-465 (pedestal pointers NULL as in launch.c:90-98).
 * Serial accumulation per layer instead of atomics. */
void ora_line_sample(uint64_t nlines, int nlay, double const *vnn, double const *snn,
                     double const *gamma, double const *alpha, double const *ns, double w0,
                     double wres, uint64_t nw, double *tau)
{
    uint64_t const fsteps = ceil(25.f/wres);
#pragma omp parallel for schedule(dynamic)
    for (int i=0; i<nlay; ++i)
    {
        double *t = (double *)malloc(sizeof(double)*(2*fsteps + 2));
        for (uint64_t j=0; j<nlines; ++j)
        {
            uint64_t const o = (uint64_t)i*nlines + j;
            double const fcd = floor((2*((vnn[o] - w0)/wres) + 1)/2);
            if (!(fcd >= 0.) || fcd >= (double)nw)
            {
                continue;   /* uint64 wrap of a negative value in the reference: skipped */
            }
            uint64_t const fc = (uint64_t)fcd;
            uint64_t const s = (int64_t)(fc - fsteps) < 0 ? 0 : fc - fsteps;
            uint64_t const e = fc + fsteps >= nw ? nw - 1 : fc + fsteps;
            ora_voigt(s*wres + w0, e - s + 1, wres, vnn[o], gamma[o], alpha[o], t);
            for (uint64_t f=s; f<=e; ++f)
            {
                tau[(uint64_t)i*nw + f] += snn[o]*ns[i]*t[f-s];
            }
        }
        free(t);
    }
}

/* ---- the two binned methods (library default wavenumber_sweep, and line_sweep) ---------
 * Spectral bins of gas-optics/src/spectral_bin.c:30-96: bins of `width` cm-1 holding
 * ppb = floor(width/wres) + 1 grid points (the last one what is left), three interpolation
 * wavenumbers per bin (first point, midpoint, last point).
 * shape[0..4] = ppb, last_ppb, number of bins, do_interp, do_last_interp. */
#define ORA_NIP 3            /* spectral_bin-internal.h:30 */

void ora_bins_shape(uint64_t nw, double wres, double width, uint64_t *shape)
{
    uint64_t const ppb = (uint64_t)(floor(width/wres) + 1);
    uint64_t last = nw % ppb;
    if (last == 0) last = ppb;
    uint64_t n = nw/ppb;
    if (ppb != last) ++n;
    shape[0] = ppb;
    shape[1] = last;
    shape[2] = n;
    shape[3] = ppb > 3;
    shape[4] = last > 3;
}

/* spectral_bin.c:73-84: l, r = first / last grid index of each bin, w = its three
 * interpolation wavenumbers. */
void ora_bins_fill(uint64_t nw, double w0, double wres, double width, uint64_t *l, uint64_t *r,
                   double *w)
{
    uint64_t shape[5];
    ora_bins_shape(nw, wres, width, shape);
    uint64_t const ppb = shape[0], last = shape[1], n = shape[2];
    for (uint64_t i=0; i<n; ++i)
    {
        uint64_t const pts = i + 1 < n ? ppb : last;
        l[i] = i*ppb;
        r[i] = l[i] + pts - 1;
        double const first = w0 + (double)(ppb*i)*wres;
        double const final = first + (double)((int)pts - 1)*wres;
        w[ORA_NIP*i] = first;
        w[ORA_NIP*i + 2] = final;
        w[ORA_NIP*i + 1] = 0.5f*(first + final);
    }
}

/* kernels.c:135-172: per layer, the four line arrays ordered by centre; equal centres keep
 * their order (the reference's insertion sort only moves an element past strictly larger
 * ones). */
void ora_sort_lines(uint64_t nlines, int nlay, double *vnn, double *snn, double *gamma,
                    double *alpha)
{
#pragma omp parallel for schedule(dynamic)
    for (int k=0; k<nlay; ++k)
    {
        double *v = vnn + (uint64_t)k*nlines, *s = snn + (uint64_t)k*nlines;
        double *g = gamma + (uint64_t)k*nlines, *a = alpha + (uint64_t)k*nlines;
        for (uint64_t i=1; i<nlines; ++i)
        {
            double const cv = v[i], cs = s[i], cg = g[i], ca = a[i];
            uint64_t j = i;
            for (; j > 0 && v[j-1] > cv; --j)
            {
                v[j] = v[j-1]; s[j] = s[j-1]; g[j] = g[j-1]; a[j] = a[j-1];
            }
            v[j] = cv; s[j] = cs; g[j] = cg; a[j] = ca;
        }
    }
}

/* kernel_utils.c:26-76.  Returns 0 when val lies outside the array; the reference then
 * leaves left = 0, right = size - 1 and its callers ignore the error. */
static int ora_bracket(uint64_t size, double const *a, double val, uint64_t *left,
                       uint64_t *right)
{
    uint64_t lo = 0, hi = size - 1;
    if (val < a[lo] || val > a[hi])
    {
        *left = lo;
        *right = hi;
        return 0;
    }
    if (a[lo] == val) hi = lo;
    else if (a[hi] == val) lo = hi;
    else
    {
        while (hi - lo > 1)
        {
            uint64_t const mid = lo + (hi - lo)/2;
            if (a[mid] == val) { lo = hi = mid; break; }
            if (val > a[mid]) lo = mid; else hi = mid;
        }
    }
    *left = lo;
    *right = hi;
    return 1;
}

/* One line on the grid points of bin j (fine mesh) or on its three interpolation points. */
static void ora_line_on_bin(double v, double sn, double g, double a, double w, uint64_t npts,
                            double wres, double *dst, double *scratch)
{
    ora_voigt(w, npts, wres, v, g, a, scratch);
    for (uint64_t q=0; q<npts; ++q) dst[q] += sn*scratch[q];
}

/* calc_optical_depth_bin_sweep, kernels.c:177-304 (lines sorted per layer).  "Local" lines
 * (bracketed by the edges of bins j-1 .. j+1) go onto the bin's grid points, "remote" ones
 * (out to 25 bins either side) onto its three interpolation points in bins_tau.  Where the
 * reference hands bracket() an empty sub-array (kernels.c:251, 281: undefined there) the
 * range is taken as empty. */
void ora_bin_sweep(uint64_t nlines, int nlay, double const *vnn, double const *snn,
                   double const *gamma, double const *alpha, double const *ns, uint64_t nw,
                   double w0, double wres, double width, double *tau, double *bins_tau)
{
    uint64_t shape[5];
    ora_bins_shape(nw, wres, width, shape);
    uint64_t const nb = shape[2];
    uint64_t *bl = (uint64_t *)malloc(sizeof(uint64_t)*nb), *br = (uint64_t *)malloc(sizeof(uint64_t)*nb);
    double *bw = (double *)malloc(sizeof(double)*ORA_NIP*nb);
    ora_bins_fill(nw, w0, wres, width, bl, br, bw);
    uint64_t const none = (uint64_t)(-1);
#pragma omp parallel for schedule(dynamic) collapse(2)
    for (int i=0; i<nlay; ++i)
    {
        for (uint64_t j=0; j<nb; ++j)
        {
            double const *v = vnn + (uint64_t)i*nlines, *s = snn + (uint64_t)i*nlines;
            double const *g = gamma + (uint64_t)i*nlines, *a = alpha + (uint64_t)i*nlines;
            double *scratch = (double *)malloc(sizeof(double)*(shape[0] + ORA_NIP));
            uint64_t reach = 1;
            double const leftw = j > reach ? bw[ORA_NIP*(j - reach)] : bw[0];
            double const rightw = j >= (nb - 1) - reach ? bw[ORA_NIP*nb - 1]
                                                        : bw[ORA_NIP*(j + reach + 1) - 1];
            uint64_t left = 0, right = 0, tmp;
            if (leftw <= v[nlines-1] && rightw >= v[0])
            {
                ora_bracket(nlines, v, leftw, &left, &tmp);
                ora_bracket(nlines - left, v + left, rightw, &tmp, &right);
                right += left;
                for (uint64_t k=left; k<=right; ++k)
                {
                    ora_line_on_bin(v[k], s[k]*ns[i], g[k], a[k], w0 + bl[j]*wres,
                                    br[j] - bl[j] + 1, wres, tau + (uint64_t)i*nw + bl[j], scratch);
                }
            }
            else if (leftw > v[nlines-1]) left = nlines;
            else right = none;

            reach = 25;
            double const rem_w = bw[ORA_NIP*j];
            double const rem_res = bw[ORA_NIP*j + 1] - rem_w;
            double *rem = bins_tau + ((uint64_t)i*nb + j)*ORA_NIP;
            double const leftw_r = j > reach ? bw[ORA_NIP*(j - reach)] : bw[0];
            if (leftw >= v[0] && leftw_r <= v[nlines-1] && left > 0)
            {
                uint64_t left_r;
                ora_bracket(left, v, leftw_r, &left_r, &tmp);
                for (uint64_t k=left_r; k<left; ++k)
                {
                    ora_line_on_bin(v[k], s[k]*ns[i], g[k], a[k], rem_w, ORA_NIP, rem_res, rem, scratch);
                }
            }
            double const rightw_r = j >= (nb - 1) - reach ? bw[ORA_NIP*nb - 1]
                                                          : bw[ORA_NIP*(j + reach + 1) - 1];
            if (rightw <= v[nlines-1] && rightw_r >= v[0])
            {
                uint64_t const start = right == none ? 0 : right;    /* right + f */
                if (start < nlines)
                {
                    uint64_t right_r;
                    ora_bracket(nlines - start, v + start, rightw_r, &tmp, &right_r);
                    right_r += start;
                    for (uint64_t k=(right == none ? 0 : right + 1); k<=right_r; ++k)
                    {
                        ora_line_on_bin(v[k], s[k]*ns[i], g[k], a[k], rem_w, ORA_NIP, rem_res, rem,
                                        scratch);
                    }
                }
            }
            free(scratch);
        }
    }
    free(bl); free(br); free(bw);
}

/* calc_optical_depth_line_sweep, kernels.c:308-406 (no sorting): every line goes onto the
 * grid points of the bins within 1.5 cm-1 of it and onto the interpolation points of the bins
 * out to 25 cm-1.  Serial per layer instead of atomics.  A bin index equal to the number of
 * bins (a line at the very end of the grid, kernels.c:346-350: out of bounds there) is
 * dropped. */
void ora_line_sweep(uint64_t nlines, int nlay, double const *vnn, double const *snn,
                    double const *gamma, double const *alpha, double const *ns, uint64_t nw,
                    double w0, double wres, double width, double *tau, double *bins_tau)
{
    uint64_t shape[5];
    ora_bins_shape(nw, wres, width, shape);
    uint64_t const nb = shape[2];
    uint64_t *bl = (uint64_t *)malloc(sizeof(uint64_t)*nb), *br = (uint64_t *)malloc(sizeof(uint64_t)*nb);
    double *bw = (double *)malloc(sizeof(double)*ORA_NIP*nb);
    ora_bins_fill(nw, w0, wres, width, bl, br, bw);
    double const bin_width = wres*(double)shape[0];
    double const maxw = w0 + (double)nw*wres;
#pragma omp parallel for schedule(dynamic)
    for (int i=0; i<nlay; ++i)
    {
        double *scratch = (double *)malloc(sizeof(double)*(shape[0] + ORA_NIP));
        for (uint64_t j=0; j<nlines; ++j)
        {
            uint64_t const o = (uint64_t)i*nlines + j;
            double const sn = snn[o]*ns[i];
            double cut = 1.5f;
            double lo = vnn[o] - cut;
            if (lo < w0) lo = w0;
            uint64_t const left = (uint64_t)floor((lo - w0)/bin_width);
            double hi = vnn[o] + cut;
            if (hi > maxw) hi = maxw;
            uint64_t const right = (uint64_t)floor((hi - w0)/bin_width);
            for (uint64_t k=left; k<=right && k<nb; ++k)
            {
                ora_line_on_bin(vnn[o], sn, gamma[o], alpha[o], w0 + bl[k]*wres, br[k] - bl[k] + 1,
                                wres, tau + (uint64_t)i*nw + bl[k], scratch);
            }
            cut = 25.f;
            lo = vnn[o] - cut;
            if (lo < w0) lo = w0;
            uint64_t const left_r = (uint64_t)floor((lo - w0)/bin_width);
            for (uint64_t k=left_r; k<left && k<nb; ++k)
            {
                ora_line_on_bin(vnn[o], sn, gamma[o], alpha[o], bw[ORA_NIP*k], ORA_NIP,
                                bw[ORA_NIP*k + 1] - bw[ORA_NIP*k],
                                bins_tau + ((uint64_t)i*nb + k)*ORA_NIP, scratch);
            }
            hi = vnn[o] + cut;
            if (hi > maxw) hi = maxw;
            uint64_t const right_r = (uint64_t)floor((hi - w0)/bin_width);
            for (uint64_t k=right+1; k<=right_r && k<nb; ++k)
            {
                ora_line_on_bin(vnn[o], sn, gamma[o], alpha[o], bw[ORA_NIP*k], ORA_NIP,
                                bw[ORA_NIP*k + 1] - bw[ORA_NIP*k],
                                bins_tau + ((uint64_t)i*nb + k)*ORA_NIP, scratch);
            }
        }
        free(scratch);
    }
    free(bl); free(br); free(bw);
}

/* interpolate + interpolate_last_bin (kernels.c:514-581) with bin_quad_interp /
 * bin_no_interp (kernel_utils.c:81-118): the remote sums of every bin, interpolated
 * quadratically in wavenumber and clamped at zero (or copied when a bin has at most three
 * points), are added to tau. */
void ora_bins_interpolate(int nlay, uint64_t nw, double w0, double wres, double width,
                          double const *bins_tau, double *tau)
{
    uint64_t shape[5];
    ora_bins_shape(nw, wres, width, shape);
    uint64_t const nb = shape[2];
    uint64_t *bl = (uint64_t *)malloc(sizeof(uint64_t)*nb), *br = (uint64_t *)malloc(sizeof(uint64_t)*nb);
    double *bw = (double *)malloc(sizeof(double)*ORA_NIP*nb);
    ora_bins_fill(nw, w0, wres, width, bl, br, bw);
    for (int i=0; i<nlay; ++i)
    {
        for (uint64_t j=0; j<nb; ++j)
        {
            int const quad = j + 1 < nb ? (int)shape[3] : (int)shape[4];
            double const *x = bw + ORA_NIP*j;
            double const *y = bins_tau + ((uint64_t)i*nb + j)*ORA_NIP;
            double *t = tau + (uint64_t)i*nw;
            for (uint64_t f=bl[j]; f<=br[j]; ++f)
            {
                if (!quad)
                {
                    t[f] += y[f - bl[j]];
                    continue;
                }
                double const w = w0 + f*wres;
                double q = (w - x[1])*(w - x[2])*y[0]/((x[0] - x[1])*(x[0] - x[2])) +
                           (w - x[0])*(w - x[2])*y[1]/((x[1] - x[0])*(x[1] - x[2])) +
                           (w - x[0])*(w - x[1])*y[2]/((x[2] - x[0])*(x[2] - x[1]));
                if (q < 0.f) q = 0.f;
                t[f] += q;
            }
        }
    }
    free(bl); free(br); free(bw);
}

/* kernels.c:469-491 with the argument order of launch.c:165-170. */
void ora_h2o_continuum(uint64_t nw, int nlay, double *tau, double const *CS, double const *T,
                       double const *Ps, double const *N, double const *T0, double const *CF,
                       double const *P, double const *T0F)
{
    double const tref = 296.f;
    for (int i=0; i<nlay; ++i)
    {
        for (uint64_t j=0; j<nw; ++j)
        {
            tau[(uint64_t)i*nw + j] += N[i]*(tref/T[i])*((CS[j]*Ps[i]*exp(T0[j]*(tref - T[i]))) +
                                       (CF[j]*(P[i] - Ps[i])*exp(T0F[j]*(tref - T[i]))));
        }
    }
}

/* kernels.c:495-510 */
void ora_o3_continuum(uint64_t nw, int nlay, double const *xs, double const *N, double *tau)
{
    for (int i=0; i<nlay; ++i)
    {
        for (uint64_t j=0; j<nw; ++j)
        {
            tau[(uint64_t)i*nw + j] += N[i]*xs[j];
        }
    }
}

/* kernels.c:585-601 */
void ora_cfc(uint64_t nw, int nlay, double const *n, double const *x, double const *xs,
             double *tau)
{
    double const half = 0.5;
    for (int i=0; i<nlay; ++i)
    {
        for (uint64_t j=0; j<nw; ++j)
        {
            tau[(uint64_t)i*nw + j] += half*n[i]*(x[i] + x[i+1])*xs[j];
        }
    }
}

/* kernels.c:605-630; p on levels [atm], t on layers. */
void ora_cia(uint64_t nw, int nlay, double const *p, double const *t, double const *x1,
             double const *x2, double const *xs, double *tau)
{
    double const quarter = 0.25;
    double const m = 28.97/6.02214076e23;
    double const g = 980.;
    double const k = 1.38064852e-16;
    double const atmtobarye = 1.013e6;
    double const c = (atmtobarye*atmtobarye)/(k*m*g*2.);
    for (int i=0; i<nlay; ++i)
    {
        for (uint64_t j=0; j<nw; ++j)
        {
            double n = c*((p[i]*p[i] - p[i+1]*p[i+1])/t[i])*quarter*(x1[i] + x1[i+1])*
                       (x2[i] + x2[i+1]);
            n = (n >= 0) ? n : n*-1.f;
            tau[(uint64_t)i*nw + j] += n*xs[j];
        }
    }
}

/* shortwave/src/rayleigh.c:29-41, 58-69 */
void ora_rayleigh(int nlay, double w0, double dw, uint64_t nw, double const *n, double *omega,
                  double *g, double *tau)
{
    for (int i=0; i<nlay; ++i)
    {
        for (uint64_t j=0; j<nw; ++j)
        {
            double const w = w0 + j*dw;
            double const W = w*1.e-4;
            uint64_t const o = (uint64_t)i*nw + j;
            omega[o] = 1.;
            g[o] = 0.;
            tau[o] = (n[i]*1.e-20*W*W*W*W)/(0.268675*1.e5*(9.38076E2 - 10.8426*W*W));
        }
    }
}

/* utilities/src/optics.c:124-148; inputs [num][n] stacked. */
void ora_add_optics(uint64_t n, int num, double const *g_in, double const *omega_in,
                    double const *tau_in, double *g_out, double *omega_out, double *tau_out)
{
    for (uint64_t i=0; i<n; ++i)
    {
        g_out[i] = omega_out[i] = tau_out[i] = 0.;
        for (int j=0; j<num; ++j)
        {
            uint64_t const o = j*n + i;
            g_out[i] += g_in[o]*omega_in[o]*tau_in[o];
            omega_out[i] += omega_in[o]*tau_in[o];
            tau_out[i] += tau_in[o];
        }
        g_out[i] /= omega_out[i];
        omega_out[i] /= tau_out[i];
    }
}

/* longwave/src/longwave.c:68-93 */
static double planck(double T, double w)
{
    double const c1 = 1.1910429526245744e-8;
    double const c2 = 1.4387773538277202;
    double e = c2*w/T;
    if (e > 700.) e = 700.;
    e = exp(e);
    return (c1*w*w*w)/(e - 1.);
}

/* longwave.c:100-118 */
static double effective_planck(double Tc, double Te, double w, double tau)
{
    double const a = 0.193;
    double const b = 0.013;
    double const pc = planck(Tc, w);
    double const pe = planck(Te, w);
    return (pc + (a*tau + b*tau*tau)*pe)/(1. + a*tau + b*tau*tau);
}

/* longwave.c:127-262: four streams, stream-major loops as in the reference. */
void ora_lw_fluxes(int nlev, double w0, double wres, uint64_t nw, double T_surf,
                   double const *T_layers, double const *T_levels, double const *tau,
                   double const *omega, double const *emis, double *flux_up, double *flux_down)
{
    static double const c1[4] = {-14.402613260847248, -3.0302159969901132, -1.4925584280108841,
                                 -1.0746123148178333};
    static double const c2[4] = {0.07587638482015649, 0.676114979733751, 1.3726594476601073,
                                 1.0169418413757783};
    int const nlay = nlev - 1;
#pragma omp parallel for schedule(static)
    for (uint64_t k=0; k<nw; ++k)
    {
        double t[MAXL], up[MAXL], dn[MAXL], ext[MAXL];
        double const w = w0 + k*wres;
        for (int i=0; i<nlay; ++i)
        {
            t[i] = tau[(uint64_t)i*nw + k]*(1. - omega[(uint64_t)i*nw + k]);   /* :253 */
        }
        memset(up, 0, sizeof(up));
        memset(dn, 0, sizeof(dn));
        for (int j=0; j<4; ++j)
        {
            for (int i=0; i<nlay; ++i)
            {
                double e = c1[j]*t[i];
                if (e > 700.) e = 700.;
                ext[i] = exp(e);
            }
            double I_down = 0.;
            for (int i=0; i<nlay; ++i)
            {
                double const val = effective_planck(T_layers[i], T_levels[i+1], w, t[i]);
                double const p = (1. - ext[i])*val;
                I_down = p + I_down*ext[i];
                dn[i+1] += c2[j]*I_down;
            }
            double I_up = planck(T_surf, w);
            I_up = emis[k]*I_up + (1 - emis[k])*I_down;
            up[nlev-1] += c2[j]*I_up;
            for (int i=nlay-1; i>=0; --i)
            {
                double const val = effective_planck(T_layers[i], T_levels[i], w, t[i]);
                double const p = (1. - ext[i])*val;
                I_up = p + I_up*ext[i];
                up[i] += c2[j]*I_up;
            }
        }
        for (int i=0; i<nlev; ++i)
        {
            flux_up[(uint64_t)i*nw + k] = up[i];
            flux_down[(uint64_t)i*nw + k] = dn[i];
        }
    }
}

/* shortwave/src/shortwave.c:97-236: Eddington gammas + Meador-Weaver R/T. */
static void eddington(double omega, double tau, double mu, double g, double *R, double *T,
                      double *T_pure)
{
    double const gamma1 = 0.25*(7. - omega*(4. + 3.*g));
    double const gamma2 = -0.25*(1. - omega*(4. - 3.*g));
    double const gamma3 = 0.25*(2. - 3.*g*mu);
    if (omega <= 0.0)
    {
        *R = 0.;
        *T = exp(-tau/mu);
        if (T_pure) *T_pure = *T;
    }
    else
    {
        double const gamma4 = 1. - gamma3;
        double const alpha1 = gamma1*gamma4 + gamma2*gamma3;
        double const alpha2 = gamma1*gamma3 + gamma2*gamma4;
        double const k = sqrt(gamma1*gamma1 - gamma2*gamma2);
        double t = tau;
        if (1./mu > k && tau/mu > 700.)
        {
            t = 700.*mu;
        }
        else if (tau*k > 700.)
        {
            t = 700./k;
        }
        double const tp = exp(t/mu);
        if (tp <= 1.0)
        {
            *R = 0.;
            *T = 1.;
            if (T_pure) *T_pure = 1.;
        }
        else
        {
            double const tm = exp(-t/mu);
            double const tkm = exp(-t*k);
            double const tkp = exp(t*k);
            if (T_pure) *T_pure = tm;
            if (omega >= 1.)
            {
                *R = (1./(1. + gamma1*t))*(gamma1*t + (gamma3 - gamma1*mu)*(1. - tm));
                *T = 1. - *R;
            }
            else
            {
                *R = (omega/((1. - k*k*mu*mu)*((k + gamma1)*tkp + (k - gamma1)*tkm)))*
                     ((1. - k*mu)*(alpha2 + k*gamma3)*tkp - (1. + k*mu)*(alpha2 - k*gamma3)*tkm -
                      2.*k*(gamma3 - alpha2*mu)*tm);
                *T = tm*(1. - (omega/((1. - k*k*mu*mu)*((k + gamma1)*tkp + (k - gamma1)*tkm)))*
                     ((1. + k*mu)*(alpha1 + k*gamma4)*tkp - (1. - k*mu)*(alpha1 - k*gamma4)*tkm -
                      2.*k*(gamma4 + alpha1*mu)*tp));
            }
        }
    }
    if (T_pure && *T_pure > *T)
    {
        *T = *T_pure;
    }
}

/* shortwave.c:68-92 (delta-Eddington), :242-330 (adding), :339-405, :410-455. */
void ora_sw_fluxes(int nlev, uint64_t nw, double const *omega, double const *g,
                   double const *tau, double mu_dir, double mu_dif, double const *alb_dir,
                   double const *alb_dif, double tsi, double const *solar, double *flux_up,
                   double *flux_down)
{
    int const nlay = nlev - 1;
#pragma omp parallel for schedule(static)
    for (uint64_t k=0; k<nw; ++k)
    {
        double R_dir[MAXL], R_dif[MAXL], T_dir[MAXL], T_dif[MAXL], Tp[MAXL];
        double Rdd[MAXL], Rfd[MAXL], Rfu[MAXL], R[MAXL], T[MAXL];
        for (int i=0; i<nlay; ++i)
        {
            uint64_t const o = (uint64_t)i*nw + k;
            double const gs = g[o]/(g[o] + 1.);
            double const f = g[o]*g[o];
            double const os = (1. - f)*omega[o]/(1. - omega[o]*f);
            double const ts = tau[o]*(1. - omega[o]*f);
            eddington(os, ts, mu_dir, gs, &R_dir[i], &T_dir[i], &Tp[i]);
            eddington(os, ts, mu_dif, gs, &R_dif[i], &T_dif[i], NULL);
        }
        Rdd[nlev-1] = alb_dir[k];
        Rfd[nlev-1] = alb_dif[k];
        for (int i=nlay-1; i>=0; --i)
        {
            double const A = Tp[i];
            double const B = 1./(1. - R_dif[i]*Rfd[i+1]);
            Rdd[i] = R_dir[i] + (A*Rdd[i+1] + (T_dir[i] - A)*Rfd[i+1])*T_dif[i]*B;
            Rfd[i] = R_dif[i] + T_dif[i]*T_dif[i]*Rfd[i+1]*B;
        }
        Rfu[0] = R_dif[0];
        for (int i=1; i<nlay; ++i)
        {
            double const B = 1./(1. - R_dif[i]*Rfu[i-1]);
            Rfu[i] = R_dif[i] + T_dif[i]*T_dif[i]*Rfu[i-1]*B;
        }
        double dir_beam = 1.;
        double dif_beam = 0.;
        R[0] = dir_beam*Rdd[0];
        T[0] = dir_beam;
        for (int i=1; i<nlev; ++i)
        {
            if (i > 1)
            {
                double const C = 1./(1. - R_dif[i-1]*Rfu[i-2]);
                dif_beam = (dir_beam*R_dir[i-1]*Rfu[i-2] + dif_beam)*T_dif[i-1]*C +
                           dir_beam*(T_dir[i-1] - Tp[i-1]);
            }
            else
            {
                dif_beam = dir_beam*(T_dir[i-1] - Tp[i-1]);
            }
            dir_beam *= Tp[i-1];
            double const B = 1./(1. - Rfd[i]*Rfu[i-1]);
            R[i] = (dir_beam*Rdd[i] + dif_beam*Rfd[i])*B;
            T[i] = dir_beam*(1. + Rdd[i]*Rfu[i-1]*B) + dif_beam*B;
        }
        for (int i=0; i<nlev; ++i)
        {
            R[i] *= solar[k]*mu_dir;
            T[i] *= solar[k]*mu_dir;
            flux_up[(uint64_t)i*nw + k] = tsi*R[i];
            flux_down[(uint64_t)i*nw + k] = tsi*T[i];
        }
    }
}

/* framework/src/driver.c:302-319, for every level. */
void ora_broadband(int nlev, uint64_t nw, double dw, double const *flux, double *out)
{
    for (int l=0; l<nlev; ++l)
    {
        double acc = 0.;
        double const *f = flux + (uint64_t)l*nw;
        for (uint64_t i=0; i+1<nw; ++i)
        {
            acc += 0.5*(f[i] + f[i+1])*dw;
        }
        out[l] = acc;
    }
}

/* utilities/src/utilities.c:149-246 with linear_sample and no extrapolation, onto
 * the grid w0 + i*dw (spectral_grid.c:94); out must be zero-initialised. */
void ora_table_to_grid(double const *x, double const *y, uint64_t n, double w0, double dw,
                       uint64_t nw, double *out)
{
    uint64_t i = 0;
    while (i < nw && (w0 + i*dw) <= x[0]) ++i;
    for (uint64_t j=0; j+1<n && i<nw; ++j)
    {
        double const m = (y[j+1] - y[j])/(x[j+1] - x[j]);
        double const b = y[j] - m*x[j];
        while (i < nw && (w0 + i*dw) <= x[j+1])
        {
            out[i] = m*(w0 + i*dw) + b;
            ++i;
        }
    }
}

/* parse_HITRAN_file.c:372-384: scale raw intensities to the reference form. */
void ora_scale_strengths(int mol, uint64_t n, int const *iso, double const *en,
                         double const *vnn, double *snn)
{
    double const tref = 296.f;
    double const c2 = -1.4387686f;
    for (uint64_t i=0; i<n; ++i)
    {
        snn[i] *= grt_tips_q(mol, tref, iso[i])/(exp(c2*en[i]/tref)*(1.f - exp(c2*vnn[i]/tref)));
    }
}

double ora_tips_q(int mol, double t, int iso)
{
    return grt_tips_q(mol, t, iso);
}
