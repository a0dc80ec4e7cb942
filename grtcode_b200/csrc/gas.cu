// Line-by-line gas optics on sm_100a: the device side of calculate_optical_depth.
//
// Replaces launch() (gas-optics/src/launch.c:40-226) and the kernels it calls
// (gas-optics/src/kernels.c, utilities/src/curtis_godson.c).  Instead of the
// reference's ~8 launches per molecule plus one per continuum/CFC/CIA species,
// one column is four launches:
//   K0 column_prep  layer quantities for every molecule/CFC/CIA pair + 1/Q
//   K1 line_prep    per (line, layer): shifted centre, S(T), Lorentz/Doppler
//                   widths, the Voigt per-line constants -> one 32-byte record
//   K3 wing_sweep   gather: each CTA owns a tile of grid points of one layer,
//                   stages the records of all lines within +-25 cm-1 through
//                   shared memory and accumulates the far-wing / pure-Lorentz
//                   samples in registers; epilogue adds the H2O/O3 continua and
//                   every CFC/CIA cross section and stores tau once (no atomics)
//   K3c core_sweep  the few samples per line with |x| < XLIM0 (Humlicek regions
//                   1-3 and CPF12) added with double atomics
// All molecules share one merged, centre-sorted line list.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <numeric>
#include <vector>
#include "common.cuh"
#include "voigt.cuh"

namespace grt {

constexpr int kMaxIso = 18;           // largest isotopologue count (molecules.c:48)

struct Slot
{
    int mol_id;
    double mass;
    int num_iso;
};

// Per-column layer quantities, one allocation, all [.. x num_layers] row-major.
struct LayerView
{
    double *n;        // [L] integrated air number density          (curtis_godson.c:25)
    double *pavg;     // [L]                                         (curtis_godson.c:59)
    double *tavg;     // [L]
    double *ps;       // [slots][L] partial pressure                 (curtis_godson.c:92)
    double *ns;       // [slots][L] species number density
    double *qinv;     // [slots][L][kMaxIso] 1/Q                     (kernels.c:52)
    double *cfc;      // [cfcs][L]  0.5*n*(x_i + x_{i+1})            (kernels.c:597)
    double *cia;      // [cias][L]  |c (p_i^2 - p_{i+1}^2)/T ...|    (kernels.c:623)
};

// Base line parameters on the device (merged over molecules, sorted by v0).
struct LineView
{
    double const *v0;
    double const *delta;
    double const *s0;
    double const *en;
    double const *nair;
    double const *yair;
    double const *yself;
    uint16_t const *slotiso;   // slot << 8 | iso
    uint64_t num_lines;
};

struct Tables
{
    double const *h2o[4];      // S296, CKDS, F296, CKDF (argument order of launch.c:165-170)
    double const *o3;
    double const *cfc[GRT_MAX_CFCS];
    double const *cia[GRT_MAX_CIAS];
    int h2o_slot;              // slot of H2O or -1
    int o3_slot;
    int num_cfcs;
    int num_cias;
};

struct ColumnInputs
{
    double const *p;           // [levels] atm
    double const *t;           // [levels]
    double const *x_mol;       // [GRT_MAX_MOLS][levels]
    double const *x_cfc;       // [GRT_MAX_CFCS][levels]
    double const *x_cia;       // [2][levels]
};

struct PrepMeta
{
    Slot slots[GRT_MAX_MOLS];
    int num_slots;
    int cfc_id[GRT_MAX_CFCS];
    int cia_id[GRT_MAX_CIAS][2];
    int num_cfcs;
    int num_cias;
    double const *tips;        // [slots][kMaxIso][nt]
    int tips_tmin;
    int tips_nt;
};


// Q(T) exactly as grtcode_b200/host/tips.c:grt_tips_q evaluates it.
__host__ __device__ inline double tips_q(double const *tab, int const tmin, int const nt, double const T)
{
    double t = T;
    double const lo = (double)tmin;
    double const hi = (double)(tmin + nt - 1);
    if (!(t >= lo)) t = lo;
    if (t >= hi) return tab[nt-1];
    int const it = (int)t;
    // plain operators: this translation unit is compiled -fmad=false (device) and
    // -ffp-contract=off (host), so nothing is fused on either side
    double const f = t - (double)it;
    int const idx = it - tmin;
    double const d = tab[idx+1] - tab[idx];
    double const fd = f*d;
    return tab[idx] + fd;
}


// Layer quantities of curtis_godson.c:25-120 and kernels.c:52-66, one function per value so
// that the host can re-evaluate them for the few records it patches (see patch_near_ties).
// T is the reference's fp_t: double, or float in the single build, where every operation below
// is a float operation as it is in the reference compiled with -DSINGLE_PRECISION.
template <typename T>
__host__ __device__ inline T layer_density(T const p0, T const p1)
{
    T dp = p0 - p1;
    dp = dp >= (T)0 ? dp : -dp;
    return (T)2.147822334314468e+25*dp;
}

template <typename T>
__host__ __device__ inline T layer_mean(T const a0, T const a1)
{
    return (T)0.5f*(a0 + a1);
}

template <typename T>
__host__ __device__ inline T layer_partial_pressure(T const x0, T const x1, T const p0, T const p1)
{
    T const third = (T)(1.f/3.f);
    T const sixth = (T)(1.f/6.f);
    return third*(x0*p0 + x1*p1) + sixth*(x0*p1 + x1*p0);
}

template <typename T>
__host__ __device__ inline T layer_species_density(T const n, T const x0, T const x1)
{
    return n*(T)0.5f*(x0 + x1);
}

// q = 1.f/Q(mol, T, iso) (kernels.c:62); Q is the substitute table evaluated in double
// (host/tips.c) and returned in fp_t (oracle/tips_shim.c does the same for the reference).
template <typename T>
__host__ __device__ inline T layer_qinv(double const *tab, int const tmin, int const nt, T const t)
{
    return (T)1.f/(T)tips_q(tab, tmin, nt, (double)t);
}


// K0: one thread per layer.
__global__ void column_prep_kernel(int const num_layers, ColumnInputs const in,
                                   PrepMeta const meta, LayerView const out, double const cia_c)
{
    typedef real_t T;
    int const i = blockIdx.x*blockDim.x + threadIdx.x;
    if (i >= num_layers) return;
    int const L = num_layers;
    int const levels = L + 1;
    T const p0 = (T)in.p[i];
    T const p1 = (T)in.p[i+1];
    T const n = layer_density<T>(p0, p1);
    T const pavg = layer_mean<T>(p0, p1);
    T const tavg = layer_mean<T>((T)in.t[i], (T)in.t[i+1]);
    out.n[i] = (double)n;
    out.pavg[i] = (double)pavg;
    out.tavg[i] = (double)tavg;
    for (int s=0; s<meta.num_slots; ++s)
    {
        double const *x = in.x_mol + (size_t)(meta.slots[s].mol_id - 1)*levels;
        T const x0 = (T)x[i];
        T const x1 = (T)x[i+1];
        out.ps[s*L+i] = (double)layer_partial_pressure<T>(x0, x1, p0, p1);
        out.ns[s*L+i] = (double)layer_species_density<T>(n, x0, x1);
        for (int k=0; k<meta.slots[s].num_iso; ++k)
        {
            double const *tab = meta.tips + ((size_t)s*kMaxIso + k)*meta.tips_nt;
            out.qinv[((size_t)s*L + i)*kMaxIso + k] =
                (double)layer_qinv<T>(tab, meta.tips_tmin, meta.tips_nt, tavg);
        }
    }
    for (int c=0; c<meta.num_cfcs; ++c)
    {
        double const *x = in.x_cfc + (size_t)meta.cfc_id[c]*levels;
        out.cfc[c*L+i] = (double)((T)0.5*n*((T)x[i] + (T)x[i+1]));
    }
    for (int c=0; c<meta.num_cias; ++c)
    {
        // (the reference's single build overflows here: its constant c = 7.9e46 exceeds
        // FLT_MAX, kernels.c:610-615; this build forms the factor in double in both
        // precisions)
        double const *x1 = in.x_cia + (size_t)meta.cia_id[c][0]*levels;
        double const *x2 = in.x_cia + (size_t)meta.cia_id[c][1]*levels;
        double const pd0 = (double)p0, pd1 = (double)p1;
        double v = cia_c*((pd0*pd0 - pd1*pd1)/(double)tavg)*0.25*(x1[i] + x1[i+1])*(x2[i] + x2[i+1]);
        v = (v >= 0) ? v : v*-1.;
        out.cia[c*L+i] = v;
    }
}


}  // namespace grt

#include "sweep.cuh"
#ifndef GRT_SINGLE
#include "binned.cuh"
#include <cub/device/device_segmented_radix_sort.cuh>
#include <cub/device/device_radix_sort.cuh>
#endif

namespace grt {

struct PrepStats
{
    unsigned long long records;   // (line, layer) pairs inside the grid
    unsigned long long samples;   // Voigt samples the reference evaluates for them
    unsigned long long lorentz;   // records on the pure-Lorentz branch
    unsigned long long core;      // samples handled by the core kernel (upper bound)
};

// What one (line, layer) pair adds to PrepStats.
struct RecordCounts
{
    unsigned long long samples;
    unsigned long long core;
    unsigned rec;
    unsigned lor;
};

// Base parameters of one line and the layer quantities of its molecule, in the build's fp_t.
template <typename T>
struct LineInputs
{
    T v0, delta, s0, en, nair, yair, yself;   // parse_HITRAN_file.h:28-44
    T t, P, Ps, Ns, qinv, mass;
};

// Records K1 flags for re-evaluation on the host (see patch_near_ties): the product
// REPWID*gamma lies within kTieUlps double ulps of a float rounding boundary, so a last-bit
// difference between this GPU's pow() and the host libm's -- the one the reference calls --
// could round Y = (float)(REPWID*gamma) (RFM_voigt.c:95) the other way and move every sample
// of the line by 6e-8 relative.  libdevice pow is within 2 ulp and glibc within 1 ulp of the
// true value; the multiplications that follow add one rounding each.  (Double build only: the
// single build's tolerance is 1e-4.)
constexpr unsigned kTieUlps = 64;
constexpr unsigned kTieInline = 254;       // flagged indices copied back with the counter

struct TieList
{
    unsigned count;
    unsigned pad;
    unsigned long long idx[kTieInline];    // layer*num_lines + line
};

__host__ __device__ inline bool near_float_tie(double const g)
{
#ifdef __CUDA_ARCH__
    unsigned const lo = (unsigned)__double2loint(g) & 0x1fffffffu;
#else
    unsigned long long bits;
    memcpy(&bits, &g, sizeof(bits));
    unsigned const lo = (unsigned)bits & 0x1fffffffu;
#endif
    // the float rounding of a normal double is decided by its low 29 mantissa bits
    // against the half-way pattern 0x10000000
    unsigned const d = lo >= 0x10000000u ? lo - 0x10000000u : 0x10000000u - lo;
    return d <= kTieUlps;
}

__host__ __device__ inline float sqrt_fp(float const x) { return sqrtf(x); }
__host__ __device__ inline double sqrt_fp(double const x) { return sqrt(x); }

// One 64-byte record from the line's base parameters: formulas and operation order of
// calc_line_centers / calc_line_strengths / calc_lorentz_hw / calc_doppler_hw
// (kernels.c:34-131) followed by the per-line part of rfm_voigt_line_shape
// (RFM_voigt.c:94-109) and the window of calc_optical_depth_line_sample
// (kernels.c:417,431-437), in the reference's fp_t (T).  The three transcendental values come
// from the caller: ex_en = EXP(c2*en/T), ex_v = EXP(c2*v0/T), pw = POW(296.f/T, n) --
// libdevice's in K1, the host libm's in patch_near_ties.  Returns true when Y sits near a
// float tie (double build).
template <typename T>
__host__ __device__ inline bool build_record(LineInputs<T> const &in, T const ex_en,
                                             T const ex_v, T const pw, T const w0,
                                             T const wres, uint64_t const nw,
                                             int const fsteps, LineRec &r, RecordCounts &cnt)
{
    T const v = in.v0 + in.delta*in.P;
    T const S = in.s0*ex_en*((T)1.f - ex_v)*in.qinv;
    T const gamma = pw*(in.yair*(in.P - in.Ps) + in.yself*in.Ps);
    T const kb = (T)1.380658E-16;
    T const cl = (T)2.99792458E10;
    T const alpha = (T)0.83255461115f*v*sqrt_fp(((T)2.f*kb*in.t)/(in.mass*cl*cl));

    float const repwid = (float)((T)kSqrln2/alpha);
    T const yd = (T)repwid*gamma;
    float const y = (float)yd;
    r.repwid = repwid;
    r.y = y;
    T const sn = S*in.Ns;
    float const yq = y*y;
    r.yq = yq;
    r.yrrtpi = y*kRsqrpi;
    r.a0 = (float)((double)yq + 0.5);                 // RFM_voigt.c:177-179
    r.d0 = r.a0*r.a0;
    r.d2 = (float)((double)(yq + yq) - 1.0);
    voigt_limits(y, &r.xlim0, &r.xlim1);
    // Grid points farther than hw0 from fc have |x| >= XLIM0: |w_f - v| >=
    // (|f - fc| - 1/2) wres, and 1.5 grid steps of slack cover the float rounding of x.
    // Single build: the grid wavenumber itself is a float sum, off by up to ~3 float ulps of
    // the wavenumber (0.01 cm-1 at 50000 cm-1): that many more grid steps of slack.
    double const step = (double)repwid*(double)wres;          // x per grid point
    double slack = 2.;
    if (sizeof(T) == sizeof(float))
    {
        double const av = (double)(v >= (T)0 ? v : -v) + 25.;
        slack += ceil(3.*av*1.1920929e-07/(double)wres) + 1.;
    }
    double const h0 = ceil((double)r.xlim0/step) + slack;
    // Conservative 16-bit companions for chunk classification near the centre: every
    // point with |f - fc| < hw0in has |x| < XLIM0, every point with |f - fc| > hw1
    // has |x| >= XLIM1 (so chunks between them are pure Humlicek region 1).
    double h0in = floor((double)r.xlim0/step) - slack;
    double h1 = ceil((double)r.xlim1/step) + slack;
    if (!(h0in >= 0.)) h0in = 0.;
    if (!(h0in <= 65535.)) h0in = 65535.;
    if (!(h1 >= 0.)) h1 = 65535.;
    if (!(h1 <= 65535.)) h1 = 65535.;
    r.hw1_hw0in = (unsigned)h1 | ((unsigned)h0in << 16);
    double cline;
    if (y >= kLorentzY)
    {
        r.hw0 = kHwLorentz;
        cline = (double)sn*(double)(repwid*y)/3.14159265358979323846;
    }
    else
    {
        r.hw0 = h0 < (double)kHwMax ? (int)h0 : kHwMax;
        cline = (double)(sn*(T)(kRsqrpi*repwid));
        // Vanishing (or NaN) Y: far-wing values may be zero or subnormal; let
        // the line-core kernel evaluate the whole window with the full routine.
        if (!(y >= 1e-18f)) r.hw0 = kHwCoreOnly;
    }
#ifdef GRT_SINGLE
    r.v = (float)v;
    r.c = (float)cline;
    r.cy = r.c*r.yrrtpi;
    r.pad_a = 0.f;
    // a float c*Y/sqrt(pi) that underflows would drop the line from the far wing: such a line
    // (optical depth below 1e-38 per sample) goes through the core kernel's full routine
    if (r.hw0 >= 0 && !(r.cy >= 1.1754944e-38f) && r.c > 0.f) r.hw0 = kHwCoreOnly;
#else
    r.v = (double)v;
    r.c = cline;
#endif
    cnt.samples = 0;
    cnt.core = 0;
    cnt.rec = 0;
    cnt.lor = 0;
    double const fcd = floor((double)(((T)2*((v - w0)/wres) + (T)1)/(T)2));
    if (fcd >= 0. && fcd < (double)nw)
    {
        int const fc = (int)fcd;
        r.fc = fc;
        int const s = fc - fsteps < 0 ? 0 : fc - fsteps;
        int const e = (uint64_t)(fc + fsteps) >= nw ? (int)(nw - 1) : fc + fsteps;
        cnt.samples = (unsigned long long)(e - s + 1);
        cnt.rec = 1;
        if (y >= kLorentzY)
        {
            cnt.lor = 1;
        }
        else
        {
            double const h = (double)r.xlim1/((double)repwid*(double)wres);
            cnt.core = (unsigned long long)fmin(2.*h + 1., (double)(e - s + 1));
            if (r.hw0 == kHwCoreOnly) cnt.core = cnt.samples;
        }
    }
    else
    {
        r.fc = kSkip;
    }
    return sizeof(T) == sizeof(double) && near_float_tie((double)yd);
}


// K1: one thread per (line, layer).
__global__ void __launch_bounds__(256)
line_prep_kernel(LineView const lines, int const num_layers, LayerView const lay,
                 PrepMeta const meta, double const w0, double const wres, uint64_t const nw,
                 int const fsteps, LineRec *recs, PrepStats *stats, double *vkey,
                 unsigned *vidx, TieList *ties, unsigned long long *tie_overflow,
                 unsigned const tie_capacity, int *hmax)
{
    uint64_t const j = (uint64_t)blockIdx.x*blockDim.x + threadIdx.x;
    int const i = blockIdx.y;
    int my_h = -1;                       // core half-width of this record (core_gather_kernel)
    unsigned long long my_samples = 0;
    unsigned int my_rec = 0;
    unsigned int my_lor = 0;
    unsigned long long my_core = 0;
    if (j < lines.num_lines)
    {
        int const L = num_layers;
        unsigned const si = lines.slotiso[j];
        int const slot = si >> 8;
        int iso = si & 0xff;
        if (iso < 1) iso = 1;
        if (iso > kMaxIso) iso = kMaxIso;
        typedef real_t T;
        LineInputs<T> in;
        in.t = (T)lay.tavg[i];
        in.P = (T)lay.pavg[i];
        in.Ps = (T)lay.ps[slot*L+i];
        in.Ns = (T)lay.ns[slot*L+i];
        in.qinv = (T)lay.qinv[((size_t)slot*L + i)*kMaxIso + iso - 1];
        in.mass = (T)meta.slots[slot].mass;
        in.v0 = (T)lines.v0[j];
        in.delta = (T)lines.delta[j];
        in.s0 = (T)lines.s0[j];
        in.en = (T)lines.en[j];
        in.nair = (T)lines.nair[j];
        in.yair = (T)lines.yair[j];
        in.yself = (T)lines.yself[j];
        T const c2 = (T)(-1.4387686f);
        LineRec r;
        RecordCounts cnt;
#ifdef GRT_SINGLE
        bool const tie = build_record<T>(in, expf(c2*in.en/in.t), expf(c2*in.v0/in.t),
                                         powf(296.f/in.t, in.nair), (T)w0, (T)wres, nw, fsteps,
                                         r, cnt);
#else
        bool const tie = build_record<T>(in, exp(c2*in.en/in.t), exp(c2*in.v0/in.t),
                                         pow(296./in.t, in.nair), w0, wres, nw, fsteps, r, cnt);
#endif
        my_samples = cnt.samples;
        my_core = cnt.core;
        my_rec = cnt.rec;
        my_lor = cnt.lor;
        size_t const gi = (size_t)i*lines.num_lines + j;
        recs[gi] = r;
        if (r.fc != kSkip && r.hw0 != kHwLorentz && (r.hw0 == kHwCoreOnly || r.xlim1 > 0.f))
        {
            my_h = core_half_width(r.repwid, r.xlim1, r.hw0 == kHwCoreOnly, wres, fsteps);
        }
        if (tie && ties != nullptr && r.fc != kSkip)
        {
            unsigned const k = atomicAdd(&ties->count, 1u);
            if (k < kTieInline) ties->idx[k] = gi;
            else if (k - kTieInline < tie_capacity) tie_overflow[k - kTieInline] = gi;
        }
        if (vkey != nullptr)
        {
            // sort keys of the binned methods (sort_lines, kernels.c:135-172)
            vkey[gi] = (double)r.v;
            vidx[gi] = (unsigned)gi;
        }
    }
    if (hmax != nullptr)
    {
        my_h = __reduce_max_sync(0xffffffffu, my_h);
        if ((threadIdx.x & 31) == 0 && my_h >= 0) atomicMax(&hmax[i], my_h);
    }
    // Block-level statistics (bench bookkeeping only).
    __shared__ unsigned long long sh[4];
    if (threadIdx.x < 4) sh[threadIdx.x] = 0;
    __syncthreads();
    for (int o=16; o>0; o>>=1)
    {
        my_samples += __shfl_down_sync(0xffffffffu, my_samples, o);
        my_core += __shfl_down_sync(0xffffffffu, my_core, o);
        my_rec += __shfl_down_sync(0xffffffffu, my_rec, o);
        my_lor += __shfl_down_sync(0xffffffffu, my_lor, o);
    }
    if ((threadIdx.x & 31) == 0)
    {
        atomicAdd(&sh[0], (unsigned long long)my_rec);
        atomicAdd(&sh[1], my_samples);
        atomicAdd(&sh[2], (unsigned long long)my_lor);
        atomicAdd(&sh[3], my_core);
    }
    __syncthreads();
    if (threadIdx.x == 0)
    {
        atomicAdd(&stats->records, sh[0]);
        atomicAdd(&stats->samples, sh[1]);
        atomicAdd(&stats->lorentz, sh[2]);
        atomicAdd(&stats->core, sh[3]);
    }
}

// Patched records from the host replace K1's.
struct RecordPatch
{
    unsigned long long idx;
    unsigned long long pad;
    LineRec rec;
};

__global__ void patch_records_kernel(RecordPatch const *patch, unsigned const count, LineRec *recs)
{
    unsigned const k = blockIdx.x*blockDim.x + threadIdx.x;
    if (k < count) recs[patch[k].idx] = patch[k].rec;
}

// Setup on the device: a continuum / cross-section table given by its m samples (x, y) is
// resampled to the spectral grid by one thread per grid point -- the reference's
// interpolate_to_grid / interpolate2 / linear_sample (spectral_grid.c:87-112,
// utilities.c:149-246): grid wavenumber w_i = w0 + i*dw; a point in (x_j, x_{j+1}] gets
// slope*w + intercept with slope = (y_{j+1} - y_j)/(x_{j+1} - x_j), intercept = y_j -
// slope*x_j; points at or below x_0 or beyond x_{m-1} stay 0 (no extrapolation).  The same
// expressions in the same order and type (fp_t) as the host code, so the same bits.
__global__ void __launch_bounds__(256)
resample_table_kernel(real_t const *x, real_t const *y, unsigned const m, double const w0,
                      double const dw, uint64_t const n, double *out)
{
    uint64_t const i = (uint64_t)blockIdx.x*blockDim.x + threadIdx.x;
    if (i >= n) return;
    typedef real_t T;
    T const w = (T)(w0 + (double)i*dw);
    T v = (T)0;
    if (w > x[0] && w <= x[m-1])
    {
        unsigned lo = 0, hi = m - 1;             // x[lo] < w <= x[hi]
        while (hi - lo > 1)
        {
            unsigned const mid = (lo + hi)/2;
            if (x[mid] < w) lo = mid; else hi = mid;
        }
        T const slope = (y[lo+1] - y[lo])/(x[lo+1] - x[lo]);
        T const intercept = y[lo] - slope*x[lo];
        v = slope*w + intercept;
    }
    out[i] = (double)v;
}


}  // namespace grt


// ------------------------------------------------------------------ host side
using namespace grt;

struct grt_gas
{
    int device = 0;
    int num_levels = 0;
    int num_layers = 0;
    double w0 = 0.;
    double dw = 0.;
    uint64_t nw = 0;
    cudaStream_t stream = nullptr;

    // merged host copy of the line list (kept until finalised)
    std::vector<double> h_v0, h_delta, h_s0, h_en, h_nair, h_yair, h_yself;
    std::vector<uint16_t> h_slotiso;
    std::vector<double> h_sorted_v0;     // centres in device order (tile cost estimate)
    std::vector<double> h_packed;        // the 7 base-parameter arrays in device order
    std::vector<uint16_t> h_si;          // slot << 8 | iso in device order
    std::vector<uint64_t> h_order;       // device position -> upload index
    std::vector<double> h_inputs;        // column inputs of the current call, widened
    bool dirty = true;
    double max_abs_delta = 0.;

    PrepMeta meta{};
    std::vector<double> h_tips;          // [slots][kMaxIso][nt]
    Tables tab{};

    // device buffers
    double *d_lines = nullptr;           // 7 arrays of num_lines
    uint16_t *d_slotiso = nullptr;
    unsigned *d_core_order = nullptr;    // line_sample: records molecule by molecule (core kernel)
    uint64_t num_lines = 0;
    double *d_tips = nullptr;
    double *d_layer = nullptr;           // LayerView backing store
    double *d_inputs = nullptr;          // p, t, x_mol, x_cfc, x_cia
    LineRec *d_recs = nullptr;
    size_t recs_capacity = 0;
    PrepStats *d_stats = nullptr;
    int *d_tile_order = nullptr;         // tiles by line count, heaviest first
    unsigned int *d_work_counter = nullptr;
    int *d_hmax = nullptr;               // [layers]: largest core half-width (core gather)
    int *d_first_at = nullptr;           // core gather: first record at or after every 64th point
    int num_granules = 0;
    int num_tiles = 0;
    unsigned *d_tile_first = nullptr;    // (inside the d_tile_order allocation)
    unsigned *d_tile_last = nullptr;
    int tiles_pts = 0;                   // points per tile the work list was built for
    double tiles_margin = 0.;            // margin [cm-1] it was built with
    int order_method = -1;               // line order the device list was built for
    std::vector<unsigned> slot_off;      // molecule segments of the molecule-major order
    double *d_vkey = nullptr;            // binned methods: centres, sorted centres, permutation
    double *d_vsorted = nullptr;
    unsigned *d_vidx = nullptr;
    unsigned *d_perm = nullptr;
    bool tiles_spectral = false;         // tile order of the sweep: spectral + tile-major items
    LineRec *d_sorted_recs = nullptr;    // binned methods: records in sorted order
    int *d_seg_begin = nullptr;
    int *d_seg_end = nullptr;
    double2 *d_wing_a = nullptr;         // binned methods: 32-byte far-wing records, sorted order
    float4 *d_wing_b = nullptr;
    unsigned short *d_segkey = nullptr;  // [2][layers*lines] segment ids of the second sort pass
    void *d_sort_temp = nullptr;
    size_t sort_temp_bytes = 0;
    size_t binned_capacity = 0;
    unsigned long long *d_evals = nullptr;
    void *d_ranges = nullptr;            // BinRanges32[layers][bins][slots]
    size_t ranges_capacity = 0;
    std::vector<double *> owned_tables;
    uint64_t stats[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    cudaEvent_t ev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
    bool timings_pending = false;        // ev[1..3] recorded, durations not read yet
    int last_method = 2;

    // near-tie records re-evaluated with the host libm (patch_near_ties)
    TieList *d_ties = nullptr;
    unsigned long long *d_tie_overflow = nullptr;
    struct HostBlock                     // pinned: what the one mid-call sync brings back
    {
        TieList ties;
        PrepStats stats;
    } *h_block = nullptr;
    RecordPatch *h_patch = nullptr;      // pinned
    RecordPatch *d_patch = nullptr;
    unsigned long long *h_tie_overflow = nullptr;   // pinned
    uint64_t patched_total = 0;          // records patched since creation
    uint64_t patched_last = 0;           // ... by the last call
    unsigned long long *d_class_counts = nullptr;   // GRT_CLASS_HIST=1: visits per line class
    unsigned long long h_class_counts[8] = {0, 0, 0, 0, 0, 0, 0, 0};
};

constexpr unsigned kTieCapacity = 1u << 16;   // flagged records beyond the inline list


// Tables live on the device as double in both builds (a float table widens exactly).
static int upload_table(grt_gas *g, real_t const *host_in, double const **dev_out)
{
    double *d = nullptr;
    std::vector<double> wide(host_in, host_in + g->nw);
    double const *host = wide.data();
    GRT_CUDA(cudaMalloc(&d, sizeof(double)*g->nw));
    GRT_CUDA(cudaMemcpyAsync(d, host, sizeof(double)*g->nw, cudaMemcpyHostToDevice, g->stream));
    GRT_CUDA(cudaStreamSynchronize(g->stream));
    g->owned_tables.push_back(d);
    *dev_out = d;
    return 0;
}


// A table from its samples, resampled on the device (no n-point host array, no n-point copy).
static int resample_table(grt_gas *g, real_t const *x, real_t const *y, uint64_t m,
                          double const **dev_out)
{
    if (x == nullptr || y == nullptr) return fail("null table samples");
    if (m < 2 || m > 0x7fffffffull) return fail("a table needs at least two samples (%llu given)",
                                                 (unsigned long long)m);
    real_t *xy = nullptr;
    double *d = nullptr;
    GRT_CUDA(cudaMalloc(&xy, sizeof(real_t)*2*m));
    GRT_CUDA(cudaMalloc(&d, sizeof(double)*g->nw));
    GRT_CUDA(cudaMemcpyAsync(xy, x, sizeof(real_t)*m, cudaMemcpyHostToDevice, g->stream));
    GRT_CUDA(cudaMemcpyAsync(xy + m, y, sizeof(real_t)*m, cudaMemcpyHostToDevice, g->stream));
    resample_table_kernel<<<(unsigned)((g->nw + 255)/256), 256, 0, g->stream>>>(
        xy, xy + m, (unsigned)m, g->w0, g->dw, g->nw, d);
    GRT_CUDA(cudaGetLastError());
    GRT_CUDA(cudaStreamSynchronize(g->stream));
    GRT_CUDA(cudaFree(xy));
    g->owned_tables.push_back(d);
    *dev_out = d;
    return 0;
}


static int find_slot(grt_gas const *g, int mol_id)
{
    for (int s=0; s<g->meta.num_slots; ++s)
    {
        if (g->meta.slots[s].mol_id == mol_id) return s;
    }
    return -1;
}


// Sort the merged list by base centre and (re)upload it with the TIPS tables.
static int finalize_lines(grt_gas *g, int const method)
{
    GRT_CUDA(cudaSetDevice(g->device));
    uint64_t const n = g->h_v0.size();
    std::vector<uint64_t> order(n);
    std::iota(order.begin(), order.end(), (uint64_t)0);
    if (method == 2)
    {
        // line_sample: one list over all molecules, sorted by centre (tile ranges)
        std::stable_sort(order.begin(), order.end(),
                         [&](uint64_t a, uint64_t b) { return g->h_v0[a] < g->h_v0[b]; });
    }
    else
    {
        // binned methods: molecule-major, file order inside a molecule (the order
        // sort_lines starts from, kernels.c:135)
        std::stable_sort(order.begin(), order.end(), [&](uint64_t a, uint64_t b)
                         { return (g->h_slotiso[a] >> 8) < (g->h_slotiso[b] >> 8); });
    }
    g->slot_off.assign((size_t)g->meta.num_slots + 1, 0u);
    for (uint64_t k=0; k<n; ++k) g->slot_off[(size_t)(g->h_slotiso[k] >> 8) + 1]++;
    for (int sl=0; sl<g->meta.num_slots; ++sl) g->slot_off[sl+1] += g->slot_off[sl];
    g->order_method = method;
    std::vector<double> packed(7*n);
    std::vector<uint16_t> si(n);
    double maxd = 0.;
    for (uint64_t k=0; k<n; ++k)
    {
        uint64_t const o = order[k];
        packed[0*n+k] = g->h_v0[o];
        packed[1*n+k] = g->h_delta[o];
        packed[2*n+k] = g->h_s0[o];
        packed[3*n+k] = g->h_en[o];
        packed[4*n+k] = g->h_nair[o];
        packed[5*n+k] = g->h_yair[o];
        packed[6*n+k] = g->h_yself[o];
        si[k] = g->h_slotiso[o];
        maxd = std::max(maxd, std::fabs(g->h_delta[o]));
    }
    g->max_abs_delta = maxd;
    g->h_sorted_v0.assign(packed.begin(), packed.begin() + n);
    if (g->d_lines) { GRT_CUDA(cudaFree(g->d_lines)); g->d_lines = nullptr; }
    if (g->d_slotiso) { GRT_CUDA(cudaFree(g->d_slotiso)); g->d_slotiso = nullptr; }
    if (g->d_tips) { GRT_CUDA(cudaFree(g->d_tips)); g->d_tips = nullptr; }
    GRT_CUDA(cudaMalloc(&g->d_lines, sizeof(double)*std::max<uint64_t>(7*n, 1)));
    GRT_CUDA(cudaMalloc(&g->d_slotiso, sizeof(uint16_t)*std::max<uint64_t>(n, 1)));
    GRT_CUDA(cudaMalloc(&g->d_tips, sizeof(double)*std::max<size_t>(g->h_tips.size(), 1)));
    if (n > 0)
    {
        GRT_CUDA(cudaMemcpyAsync(g->d_lines, packed.data(), sizeof(double)*7*n,
                                 cudaMemcpyHostToDevice, g->stream));
        GRT_CUDA(cudaMemcpyAsync(g->d_slotiso, si.data(), sizeof(uint16_t)*n,
                                 cudaMemcpyHostToDevice, g->stream));
    }
    if (!g->h_tips.empty())
    {
        GRT_CUDA(cudaMemcpyAsync(g->d_tips, g->h_tips.data(), sizeof(double)*g->h_tips.size(),
                                 cudaMemcpyHostToDevice, g->stream));
    }
    if (g->d_first_at) { GRT_CUDA(cudaFree(g->d_first_at)); g->d_first_at = nullptr; }
    if (method == 2)
    {
        // core_gather_kernel: first record at or after every 64th grid point (one point of
        // slack; the kernel widens the range by the core half-width and the pressure shift)
        int const granules = (int)((g->nw + kCoreGranule - 1)/kCoreGranule);
        std::vector<int> first_at((size_t)granules + 1);
        for (int k=0; k<granules; ++k)
        {
            double const edge = g->w0 + ((double)k*kCoreGranule - 1.)*g->dw;
            first_at[k] = (int)(std::lower_bound(packed.begin(), packed.begin() + n, edge) -
                                packed.begin());
        }
        first_at[granules] = (int)n;
        GRT_CUDA(cudaMalloc(&g->d_first_at, sizeof(int)*first_at.size()));
        GRT_CUDA(cudaMemcpyAsync(g->d_first_at, first_at.data(), sizeof(int)*first_at.size(),
                                 cudaMemcpyHostToDevice, g->stream));
        GRT_CUDA(cudaStreamSynchronize(g->stream));
        g->num_granules = granules;
    }
    if (g->d_core_order) { GRT_CUDA(cudaFree(g->d_core_order)); g->d_core_order = nullptr; }
    if (method == 2 && n > 0)
    {
        // The core kernel walks one record per thread: neighbouring threads should hold lines
        // of the same molecule at neighbouring wavenumbers (same Doppler width, so the same
        // number of core samples and the same Voigt region at the same distance), which the
        // merged centre-sorted list interleaves with the other molecules' lines.
        std::vector<unsigned> co((size_t)n);
        std::iota(co.begin(), co.end(), 0u);
        std::stable_sort(co.begin(), co.end(),
                         [&](unsigned a, unsigned b) { return (si[a] >> 8) < (si[b] >> 8); });
        GRT_CUDA(cudaMalloc(&g->d_core_order, sizeof(unsigned)*n));
        GRT_CUDA(cudaMemcpyAsync(g->d_core_order, co.data(), sizeof(unsigned)*n,
                                 cudaMemcpyHostToDevice, g->stream));
    }
    GRT_CUDA(cudaStreamSynchronize(g->stream));
    g->h_packed.swap(packed);
    g->h_order.swap(order);
    g->h_si.swap(si);
    g->num_lines = n;
    g->meta.tips = g->d_tips;
    g->dirty = false;
    g->tiles_pts = 0;
    return 0;
}


// The records K1 flagged (Y within a few double ulps of a float rounding boundary) are
// rebuilt here with the HOST libm's exp and pow -- the functions the reference itself calls
// (grtcode_config.h:25-47) -- and written over K1's.  Everything else in a record is IEEE
// arithmetic that the GPU and the host round identically, so after the patch Y, and with it
// every float the line shape is computed from, carries the bits the reference computes on
// this machine.  A handful of records per column (probability ~2.4e-7 per record).
static int patch_near_ties(grt_gas *g, unsigned const count, double const *p_atm,
                           double const *t, double const *x_mol)
{
    cudaStream_t const st = g->stream;
    uint64_t const N = g->num_lines;
    int const levels = g->num_levels;
    unsigned const extra = count > kTieInline ? count - kTieInline : 0u;
    if (extra > kTieCapacity) return fail("too many near-tie records (%u)", count);
    if (extra > 0)
    {
        GRT_CUDA(cudaMemcpyAsync(g->h_tie_overflow, g->d_tie_overflow,
                                 sizeof(unsigned long long)*extra, cudaMemcpyDeviceToHost, st));
        GRT_CUDA(cudaStreamSynchronize(st));
    }
    double const c2 = (double)(-1.4387686f);
    int const fsteps = (int)std::ceil((double)25.f/g->dw);
    double const *base = g->h_packed.data();
    for (unsigned k=0; k<count; ++k)
    {
        unsigned long long const gi = k < kTieInline ? g->h_block->ties.idx[k]
                                                     : g->h_tie_overflow[k - kTieInline];
        int const i = (int)(gi/N);
        uint64_t const j = gi - (uint64_t)i*N;
        unsigned const si = g->h_si[j];
        int const slot = (int)(si >> 8);
        int iso = (int)(si & 0xff);
        if (iso < 1) iso = 1;
        if (iso > kMaxIso) iso = kMaxIso;
        double const *x = x_mol + (size_t)(g->meta.slots[slot].mol_id - 1)*levels;
        double const p0 = p_atm[i], p1 = p_atm[i+1];
        LineInputs<double> in;
        in.t = layer_mean<double>(t[i], t[i+1]);
        in.P = layer_mean<double>(p0, p1);
        in.Ps = layer_partial_pressure<double>(x[i], x[i+1], p0, p1);
        in.Ns = layer_species_density<double>(layer_density<double>(p0, p1), x[i], x[i+1]);
        double const *tab = g->h_tips.data() + ((size_t)slot*kMaxIso + iso - 1)*g->meta.tips_nt;
        in.qinv = layer_qinv<double>(tab, g->meta.tips_tmin, g->meta.tips_nt, in.t);
        in.mass = g->meta.slots[slot].mass;
        in.v0 = base[0*N+j];
        in.delta = base[1*N+j];
        in.s0 = base[2*N+j];
        in.en = base[3*N+j];
        in.nair = base[4*N+j];
        in.yair = base[5*N+j];
        in.yself = base[6*N+j];
        RecordPatch &out = g->h_patch[k];
        RecordCounts cnt;
        out.idx = gi;
        out.pad = 0;
        build_record<double>(in, std::exp(c2*in.en/in.t), std::exp(c2*in.v0/in.t),
                             std::pow(296./in.t, in.nair), g->w0, g->dw, g->nw, fsteps, out.rec,
                             cnt);
    }
    GRT_CUDA(cudaMemcpyAsync(g->d_patch, g->h_patch, sizeof(RecordPatch)*count,
                             cudaMemcpyHostToDevice, st));
    patch_records_kernel<<<(count + 127)/128, 128, 0, st>>>(g->d_patch, count, g->d_recs);
    GRT_CUDA(cudaGetLastError());
    return 0;
}


// Work list of the persistent sweep: for every warp tile (32*PTS grid points) the range
// of records whose window can reach it (binary searches in the sorted centres, with the
// pressure-shift margin) and the tiles ordered by that count, heaviest first, so that
// the dynamic scheduler ends with the cheap tiles (longest-processing-time-first).
static int prepare_tiles(grt_gas *g, int tile_points, double margin, double shift)
{
    if (g->tiles_pts == tile_points && g->tiles_margin >= margin && g->d_tile_order != nullptr) return 0;
    margin += 0.25*shift;   // head-room so that a slightly larger surface pressure does not rebuild
    int const tile = tile_points;
    int const num_tiles = (int)((g->nw + tile - 1)/tile);
    std::vector<double> const &v0 = g->h_sorted_v0;
    std::vector<std::pair<uint64_t, int>> cost((size_t)num_tiles);
    std::vector<unsigned> first((size_t)num_tiles), last((size_t)num_tiles);
    for (int t=0; t<num_tiles; ++t)
    {
        double const lo = g->w0 + (double)t*tile*g->dw - margin;
        double const hi = g->w0 + (double)(t + 1)*tile*g->dw + margin;
        auto const a = std::lower_bound(v0.begin(), v0.end(), lo);
        auto const b = std::upper_bound(v0.begin(), v0.end(), hi);
        first[t] = (unsigned)(a - v0.begin());
        last[t] = (unsigned)(b - v0.begin());
        cost[t] = {(uint64_t)(b - a), t};
    }
    // Order of the work items.  Few items (less than 16 per warp in flight): layer groups one
    // after the other, heaviest tiles first inside a group (LPT), which trims the tail (2 %
    // on a config-1 column).  Otherwise: tiles in spectral order, all layers of a tile adjacent,
    // so that the warps in flight work on neighbouring tiles (records shared through L2) and
    // the continuum / CFC / CIA tables of a tile are read from DRAM once, not once per layer
    // (11 tables x 40 MB x 54 layers = 24 GB per config-2 shortwave launch otherwise).
    bool spectral = (size_t)num_tiles*(size_t)g->num_layers >= 16u*24u*(size_t)sm_count(g->device);
    if (char const *env = getenv("GRT_TILE_ORDER")) spectral = (env[0] == 's');
    g->tiles_spectral = spectral;
    if (!spectral)
    std::stable_sort(cost.begin(), cost.end(),
                     [](auto const &x, auto const &y) { return x.first > y.first; });
    std::vector<int> order((size_t)num_tiles);
    for (int t=0; t<num_tiles; ++t) order[t] = cost[t].second;
    if (g->d_tile_order) { GRT_CUDA(cudaFree(g->d_tile_order)); g->d_tile_order = nullptr; }
    size_t const nt = (size_t)std::max(num_tiles, 1);
    GRT_CUDA(cudaMalloc(&g->d_tile_order, sizeof(int)*3*nt));
    g->d_tile_first = reinterpret_cast<unsigned *>(g->d_tile_order + nt);
    g->d_tile_last = reinterpret_cast<unsigned *>(g->d_tile_order + 2*nt);
    GRT_CUDA(cudaMemcpyAsync(g->d_tile_order, order.data(), sizeof(int)*num_tiles,
                             cudaMemcpyHostToDevice, g->stream));
    GRT_CUDA(cudaMemcpyAsync(g->d_tile_first, first.data(), sizeof(unsigned)*num_tiles,
                             cudaMemcpyHostToDevice, g->stream));
    GRT_CUDA(cudaMemcpyAsync(g->d_tile_last, last.data(), sizeof(unsigned)*num_tiles,
                             cudaMemcpyHostToDevice, g->stream));
    GRT_CUDA(cudaStreamSynchronize(g->stream));
    g->num_tiles = num_tiles;
    g->tiles_pts = tile_points;
    g->tiles_margin = margin;
    return 0;
}

// Persistent CTAs: one wave that fills every SM; work items are pulled from the counter.
// TBL kernels carry the fl(i*wres) table behind the record rings in shared memory.
constexpr int kMaxTableBytes = 64*1024;

template <int PTS, int NL, bool TBL, int LN>
static int launch_wing_impl(SweepParams const &sp, int const sms, cudaStream_t const st)
{
    using Cfg = SweepConfig<PTS, NL, LN>;
    // the attribute is per device (a process may drive several GPUs): one flag per device id
    static bool configured[64] = {};
    int dev = 0;
    GRT_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !configured[dev])
    {
        GRT_CUDA(cudaFuncSetAttribute(wing_sweep_kernel<PTS, NL, TBL, LN>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      Cfg::kRingBytes + (TBL ? kMaxTableBytes : 0)));
        if (dev >= 0 && dev < 64) configured[dev] = true;
    }
    int const smem = Cfg::kRingBytes + (TBL ? (int)sizeof(wn_t)*sweep_table_entries<PTS, LN>(sp.fsteps) : 0);
    int per_sm = 0;
    GRT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, wing_sweep_kernel<PTS, NL, TBL, LN>,
                                                           kSweepThreads, smem));
    int const groups = (sp.num_layers + NL - 1)/NL;
    unsigned const total = (unsigned)sp.num_tiles*(unsigned)groups;
    unsigned const grid = std::min<unsigned>((total + kSweepWarps - 1)/kSweepWarps,
                                             (unsigned)(sms*std::max(per_sm, 1)));
    wing_sweep_kernel<PTS, NL, TBL, LN><<<std::max(grid, 1u), kSweepThreads, smem, st>>>(sp);
    GRT_CUDA(cudaGetLastError());
    return 0;
}

template <int PTS, int NL, int LN = 32>
static int launch_wing(SweepParams const &sp, int const sms, cudaStream_t const st)
{
    bool table = (int)sizeof(wn_t)*sweep_table_entries<PTS, LN>(sp.fsteps) <= kMaxTableBytes;
    if (char const *env = getenv("GRT_SWEEP_TABLE")) table = table && atoi(env) != 0;
    return table ? launch_wing_impl<PTS, NL, true, LN>(sp, sms, st)
                 : launch_wing_impl<PTS, NL, false, LN>(sp, sms, st);
}

#ifndef GRT_SINGLE
// Segment id (layer*slots + molecule slot) of every entry of a list of global record indices.
struct SegmentMap
{
    int num_slots;
    unsigned long long num_lines;
    unsigned slot_off[GRT_MAX_MOLS + 1];
};

__global__ void __launch_bounds__(256)
segment_keys_kernel(unsigned const *gidx, SegmentMap const map, unsigned long long const total,
                    unsigned short *seg)
{
    unsigned long long const pos = (unsigned long long)blockIdx.x*blockDim.x + threadIdx.x;
    if (pos >= total) return;
    unsigned const gi = gidx[pos];
    unsigned const layer = (unsigned)(gi/map.num_lines);
    unsigned const j = (unsigned)(gi - (unsigned long long)layer*map.num_lines);
    int lo = 0, hi = map.num_slots;          // slot_off[lo] <= j < slot_off[hi]
    while (hi - lo > 1)
    {
        int const mid = (lo + hi)/2;
        if (j >= map.slot_off[mid]) lo = mid; else hi = mid;
    }
    seg[pos] = (unsigned short)(layer*(unsigned)map.num_slots + (unsigned)lo);
}

// After the sort: gidx[p] = global index of the record at sorted position p.  Writes the
// layer-relative permutation, the sorted centres and the records in sorted order (so that the
// gather reads them with unit stride).  perm may alias gidx.
__global__ void __launch_bounds__(256)
sort_finish_kernel(unsigned const *gidx, double const *vkey, LineRec const *recs,
                   unsigned long long const num_lines, unsigned long long const total,
                   unsigned *perm, double *vsorted, LineRec *sorted_recs, double2 *wing_a,
                   float4 *wing_b)
{
    unsigned long long const pos = (unsigned long long)blockIdx.x*blockDim.x + threadIdx.x;
    if (pos >= total) return;
    unsigned const gi = gidx[pos];
    unsigned long long const layer = pos/num_lines;
    perm[pos] = (unsigned)(gi - layer*num_lines);
    vsorted[pos] = vkey[gi];
    uint4 const *src = reinterpret_cast<uint4 const *>(recs + gi);
    uint4 *dst = reinterpret_cast<uint4 *>(sorted_recs + pos);
    uint4 const a = src[0], b = src[1], c = src[2], d = src[3];
    dst[0] = a; dst[1] = b; dst[2] = c; dst[3] = d;
    wing_record(recs[gi], wing_a + pos, wing_b + pos);
}

static int sort_temp(grt_gas *g, size_t const bytes)
{
    if (bytes > g->sort_temp_bytes)
    {
        cudaFree(g->d_sort_temp);
        g->d_sort_temp = nullptr;
        GRT_CUDA(cudaMalloc(&g->d_sort_temp, bytes));
        g->sort_temp_bytes = bytes;
    }
    return 0;
}

// wavenumber_sweep / line_sweep after K0 + K1: per-(molecule, layer) sort of the shifted
// centres (sort_lines, kernels.c:135-172: stable, ascending), then the gather kernel of
// binned.cuh.  The radix sort is CUB's (stable); keys are the centres K1 wrote.
static int binned_optical_depth(grt_gas *g, int const method, LayerView const &lay,
                                double *tau_dev, int launches)
{
    cudaStream_t const st = g->stream;
    int const L = g->num_layers;
    uint64_t const N = g->num_lines;
    GRT_CUDA(cudaEventRecord(g->ev[1], st));
    if (g->d_evals == nullptr) GRT_CUDA(cudaMalloc(&g->d_evals, sizeof(unsigned long long)));
    GRT_CUDA(cudaMemsetAsync(g->d_evals, 0, sizeof(unsigned long long), st));
    int const slots = g->meta.num_slots;
    if (N > 0)
    {
        if ((size_t)N*L > (size_t)INT32_MAX) return fail("too many (line, layer) pairs for the sort");
        int const segs = slots*L;
        bool segmented = segs > 65535;
        if (char const *env = getenv("GRT_BIN_SORT")) segmented = segmented || env[0] == 's';
        if (segmented)
        {
            // One CUB segment per (layer, molecule): a thread block per segment, which leaves
            // most of the GPU idle when there are a few hundred segments of 1e5 lines.
            std::vector<int> begin((size_t)segs), end((size_t)segs);
            for (int i=0; i<L; ++i)
            {
                for (int sl=0; sl<slots; ++sl)
                {
                    begin[(size_t)i*slots + sl] = (int)((size_t)i*N + g->slot_off[sl]);
                    end[(size_t)i*slots + sl] = (int)((size_t)i*N + g->slot_off[sl+1]);
                }
            }
            cudaFree(g->d_seg_begin); cudaFree(g->d_seg_end);
            g->d_seg_begin = g->d_seg_end = nullptr;
            GRT_CUDA(cudaMalloc(&g->d_seg_begin, sizeof(int)*segs));
            GRT_CUDA(cudaMalloc(&g->d_seg_end, sizeof(int)*segs));
            GRT_CUDA(cudaMemcpyAsync(g->d_seg_begin, begin.data(), sizeof(int)*segs,
                                     cudaMemcpyHostToDevice, st));
            GRT_CUDA(cudaMemcpyAsync(g->d_seg_end, end.data(), sizeof(int)*segs,
                                     cudaMemcpyHostToDevice, st));
            GRT_CUDA(cudaStreamSynchronize(st));   // the host vectors go out of scope below
            size_t bytes = 0;
            GRT_CUDA(cub::DeviceSegmentedRadixSort::SortPairs(
                nullptr, bytes, g->d_vkey, g->d_vsorted, g->d_vidx, g->d_perm, (int)(N*L), segs,
                g->d_seg_begin, g->d_seg_end, 0, 64, st));
            GRT_TRY(sort_temp(g, bytes));
            GRT_CUDA(cub::DeviceSegmentedRadixSort::SortPairs(
                g->d_sort_temp, bytes, g->d_vkey, g->d_vsorted, g->d_vidx, g->d_perm, (int)(N*L), segs,
                g->d_seg_begin, g->d_seg_end, 0, 64, st));
            ++launches;
            // d_perm holds global indices in segment order
            sort_finish_kernel<<<(unsigned)((N*L + 255)/256), 256, 0, st>>>(
                g->d_perm, g->d_vkey, g->d_recs, N, N*L, g->d_perm, g->d_vsorted, g->d_sorted_recs,
                g->d_wing_a, g->d_wing_b);
            ++launches;
        }
        else
        {
            // Two device-wide passes, every SM busy: (1) all (line, layer) pairs by centre,
            // (2) stable sort by segment id = layer*slots + molecule slot (radix sorts are
            // stable, so each segment keeps the centre order, ties in file order as the
            // reference's insertion sort leaves them).  Segments are contiguous and in
            // segment-id order in the record layout, so position p of the result belongs to
            // layer p / N.
            SegmentMap map;
            map.num_slots = slots;
            map.num_lines = N;
            for (int sl=0; sl<=GRT_MAX_MOLS; ++sl)
            {
                map.slot_off[sl] = sl < (int)g->slot_off.size() ? g->slot_off[sl] : (unsigned)N;
            }
            int seg_bits = 1;
            while ((1 << seg_bits) < segs) ++seg_bits;
            unsigned short *seg_in = g->d_segkey;
            unsigned short *seg_out = g->d_segkey + (size_t)N*L;
            size_t bytes_a = 0, bytes_b = 0;
            GRT_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, bytes_a, g->d_vkey, g->d_vsorted,
                                                     g->d_vidx, g->d_perm, (int)(N*L), 0, 64, st));
            GRT_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, bytes_b, seg_in, seg_out, g->d_perm,
                                                     g->d_vidx, (int)(N*L), 0, seg_bits, st));
            size_t bytes = std::max(bytes_a, bytes_b);
            GRT_TRY(sort_temp(g, bytes));
            GRT_CUDA(cub::DeviceRadixSort::SortPairs(g->d_sort_temp, bytes, g->d_vkey, g->d_vsorted,
                                                     g->d_vidx, g->d_perm, (int)(N*L), 0, 64, st));
            segment_keys_kernel<<<(unsigned)((N*L + 255)/256), 256, 0, st>>>(g->d_perm, map, N*L,
                                                                             seg_in);
            GRT_CUDA(cub::DeviceRadixSort::SortPairs(g->d_sort_temp, bytes, seg_in, seg_out,
                                                     g->d_perm, g->d_vidx, (int)(N*L), 0, seg_bits,
                                                     st));
            sort_finish_kernel<<<(unsigned)((N*L + 255)/256), 256, 0, st>>>(
                g->d_vidx, g->d_vkey, g->d_recs, N, N*L, g->d_perm, g->d_vsorted, g->d_sorted_recs,
                g->d_wing_a, g->d_wing_b);
            launches += 4;
        }
    }
    GRT_CUDA(cudaEventRecord(g->ev[2], st));
    BinParams bp;
    bp.recs = g->d_recs;
    bp.vs = g->d_vsorted;
    bp.perm = g->d_perm;
    bp.sorted_recs = g->d_sorted_recs;
    bp.wing_a = g->d_wing_a;
    bp.wing_b = g->d_wing_b;
    bp.num_lines = N;
    bp.num_slots = N > 0 ? slots : 0;
    for (int sl=0; sl<=GRT_MAX_MOLS; ++sl)
    {
        bp.slot_off[sl] = sl < (int)g->slot_off.size() ? g->slot_off[sl] : (unsigned)N;
    }
    bp.num_layers = L;
    bp.method = method;
    bp.bins = make_bins(g->w0, g->nw, g->dw, 1.);      // bin width of gas_optics.c:73
    bp.lay = lay;
    bp.tab = g->tab;
    bp.tau = tau_dev;
    bp.evals = g->d_evals;
    if (bp.bins.n > 0x7fffffffull) return fail("too many spectral bins");
    size_t const smem = sizeof(double)*(size_t)bp.bins.ppb*kBinGroup*(bp.bins.ppb <= kBinLanePerLine ? 33 : 1);
    // Up to 48 KB is available by default; beyond that the kernel opts in to the large dynamic
    // shared memory of sm_100 (per device, as the attribute is): 200 KB holds the sums of
    // eight bins down to a resolution of 0.00032 cm-1, finer than the reference's own lower
    // limit (MIN_RESOLUTION 0.001, grtcode_config.h:58).
    constexpr size_t kBinSmemMax = 200*1024;
    if (smem > kBinSmemMax)
    {
        return fail("spectral resolution %g cm-1 too fine for the binned methods (%zu bytes of "
                    "shared memory per block, limit %zu): use line_sample", g->dw, smem,
                    kBinSmemMax);
    }
    if (smem > 40*1024)
    {
        static bool configured[64] = {};
        int dev = 0;
        GRT_CUDA(cudaGetDevice(&dev));
        if (dev < 0 || dev >= 64 || !configured[dev])
        {
            GRT_CUDA(cudaFuncSetAttribute(bin_gather_kernel,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)kBinSmemMax));
            if (dev >= 0 && dev < 64) configured[dev] = true;
        }
    }
    size_t const nranges = (size_t)L*bp.bins.n*(size_t)std::max(bp.num_slots, 1);
    if (nranges > g->ranges_capacity)
    {
        cudaFree(g->d_ranges);
        g->d_ranges = nullptr;
        GRT_CUDA(cudaMalloc(&g->d_ranges, sizeof(BinRanges32)*nranges));
        g->ranges_capacity = nranges;
    }
    if (bp.num_slots > 0)
    {
        bin_ranges_kernel<<<(unsigned)((nranges + 127)/128), 128, 0, st>>>(
            bp, (BinRanges32 *)g->d_ranges);
        ++launches;
    }
    dim3 grid((unsigned)((bp.bins.n + kBinGroup - 1)/kBinGroup), (unsigned)L);
    bin_gather_kernel<<<grid, kBinThreads, smem, st>>>(bp, (BinRanges32 const *)g->d_ranges);
    GRT_CUDA(cudaGetLastError());
    ++launches;
    GRT_CUDA(cudaEventRecord(g->ev[3], st));
    PrepStats const &hs = g->h_block->stats;
    g->stats[0] = (uint64_t)N*(uint64_t)L;
    g->stats[1] = 0;                 // evaluations: read with the timings (grt_gas_last_stats)
    g->stats[2] = (uint64_t)launches;
    g->stats[3] = hs.lorentz;
    g->stats[7] = 0;
    g->timings_pending = true;
    return 0;
}
#else
static int binned_optical_depth(grt_gas *, int, grt::LayerView const &, double *, int)
{
    return fail("binned methods are not part of the single-precision build");
}
#endif

extern "C" {

int grt_gas_create(grt_gas **gas, int device, int num_levels, double w0, double dw,
                   uint64_t num_wpoints)
{
    if (gas == nullptr) return fail("null gas handle");
    if (num_levels < 2 || num_levels > GRT_MAX_LAYERS + 1)
    {
        return fail("num_levels %d outside [2, %d]", num_levels, GRT_MAX_LAYERS + 1);
    }
    if (num_wpoints < 1 || num_wpoints > (uint64_t)INT32_MAX/4)
    {
        return fail("spectral grid size %llu unsupported", (unsigned long long)num_wpoints);
    }
    cudaStream_t stream;
    GRT_TRY(stream_of(device, &stream));
    grt_gas *g = new grt_gas();
    g->device = device;
    g->num_levels = num_levels;
    g->num_layers = num_levels - 1;
    g->w0 = w0;
    g->dw = dw;
    g->nw = num_wpoints;
    g->stream = stream;
    g->tab.h2o_slot = -1;
    g->tab.o3_slot = -1;
    g->meta.tips_tmin = 0;
    g->meta.tips_nt = 0;
    int const L = g->num_layers;
    size_t const layer_doubles = (size_t)L*(3 + 2*GRT_MAX_MOLS + (size_t)GRT_MAX_MOLS*kMaxIso +
                                            GRT_MAX_CFCS + GRT_MAX_CIAS);
    size_t const input_doubles = (size_t)num_levels*(2 + GRT_MAX_MOLS + GRT_MAX_CFCS +
                                                     GRT_NUM_CIA_SPECIES);
    GRT_CUDA(cudaMalloc(&g->d_layer, sizeof(double)*layer_doubles));
    GRT_CUDA(cudaMalloc(&g->d_inputs, sizeof(double)*input_doubles));
    GRT_CUDA(cudaMalloc(&g->d_stats, sizeof(PrepStats)));
    GRT_CUDA(cudaMalloc(&g->d_work_counter, sizeof(unsigned int)));
    GRT_CUDA(cudaMalloc(&g->d_hmax, sizeof(int)*GRT_MAX_LAYERS));
    GRT_CUDA(cudaMalloc(&g->d_ties, sizeof(TieList)));
    GRT_CUDA(cudaMalloc(&g->d_tie_overflow, sizeof(unsigned long long)*kTieCapacity));
    GRT_CUDA(cudaMalloc(&g->d_patch, sizeof(RecordPatch)*(kTieInline + kTieCapacity)));
    GRT_CUDA(cudaMallocHost(&g->h_block, sizeof(*g->h_block)));
    GRT_CUDA(cudaMallocHost(&g->h_patch, sizeof(RecordPatch)*(kTieInline + kTieCapacity)));
    GRT_CUDA(cudaMallocHost(&g->h_tie_overflow, sizeof(unsigned long long)*kTieCapacity));
    for (int k=0; k<5; ++k) GRT_CUDA(cudaEventCreate(&g->ev[k]));
    *gas = g;
    return 0;
}

int grt_gas_destroy(grt_gas *g)
{
    if (g == nullptr) return fail("null gas handle");
    cudaSetDevice(g->device);
    cudaStreamSynchronize(g->stream);
    cudaFree(g->d_lines);
    cudaFree(g->d_slotiso);
    cudaFree(g->d_core_order);
    cudaFree(g->d_tips);
    cudaFree(g->d_layer);
    cudaFree(g->d_inputs);
    cudaFree(g->d_recs);
    cudaFree(g->d_stats);
    cudaFree(g->d_tile_order);
    cudaFree(g->d_work_counter);
    cudaFree(g->d_hmax);
    cudaFree(g->d_first_at);
    cudaFree(g->d_ties); cudaFree(g->d_tie_overflow); cudaFree(g->d_patch);
    cudaFree(g->d_class_counts);
    cudaFreeHost(g->h_block); cudaFreeHost(g->h_patch); cudaFreeHost(g->h_tie_overflow);
    cudaFree(g->d_vkey); cudaFree(g->d_vsorted); cudaFree(g->d_vidx); cudaFree(g->d_perm);
    cudaFree(g->d_sorted_recs);
    cudaFree(g->d_seg_begin); cudaFree(g->d_seg_end); cudaFree(g->d_segkey); cudaFree(g->d_wing_a); cudaFree(g->d_wing_b); cudaFree(g->d_sort_temp); cudaFree(g->d_evals); cudaFree(g->d_ranges);
    for (double *t : g->owned_tables) cudaFree(t);
    for (int k=0; k<5; ++k) if (g->ev[k]) cudaEventDestroy(g->ev[k]);
    delete g;
    return 0;
}

int grt_gas_add_molecule(grt_gas *g, int mol_id, double mass_g, int num_iso,
                         uint64_t num_lines, int const *iso, real_t const *vnn,
                         real_t const *snn, real_t const *yair, real_t const *yself,
                         real_t const *en, real_t const *n_air, real_t const *delta,
                         double const *q_table, int tmin, int nt)
{
    if (g == nullptr) return fail("null gas handle");
    if (mol_id < 1 || mol_id > GRT_MAX_MOLS) return fail("unrecognized molecule id %d", mol_id);
    if (find_slot(g, mol_id) >= 0) return fail("molecule %d has already been added", mol_id);
    if (num_iso < 0 || num_iso > kMaxIso) return fail("isotopologue count %d unsupported", num_iso);
    if (g->meta.num_slots > 0 && (tmin != g->meta.tips_tmin || nt != g->meta.tips_nt))
    {
        return fail("partition-sum tables must share one temperature grid");
    }
    int const slot = g->meta.num_slots++;
    g->meta.slots[slot] = Slot{mol_id, mass_g, num_iso};
    g->meta.tips_tmin = tmin;
    g->meta.tips_nt = nt;
    g->h_tips.resize((size_t)(slot + 1)*kMaxIso*nt, 1.);
    for (int k=0; k<num_iso; ++k)
    {
        std::copy(q_table + (size_t)k*nt, q_table + (size_t)(k + 1)*nt,
                  g->h_tips.begin() + ((size_t)slot*kMaxIso + k)*nt);
    }
    for (uint64_t k=0; k<num_lines; ++k)
    {
        g->h_v0.push_back(vnn[k]);
        g->h_delta.push_back(delta[k]);
        g->h_s0.push_back(snn[k]);
        g->h_en.push_back(en[k]);
        g->h_nair.push_back(n_air[k]);
        g->h_yair.push_back(yair[k]);
        g->h_yself.push_back(yself[k]);
        int is = iso[k];
        if (is < 0) is = 0;
        if (is > 255) is = 255;
        g->h_slotiso.push_back((uint16_t)((slot << 8) | is));
    }
    if (mol_id == 1 && g->tab.h2o[0] != nullptr) g->tab.h2o_slot = slot;
    if (mol_id == 3 && g->tab.o3 != nullptr) g->tab.o3_slot = slot;
    g->dirty = true;
    return 0;
}

int grt_gas_set_h2o_continuum(grt_gas *g, real_t const *s296, real_t const *ckds,
                              real_t const *f296, real_t const *ckdf)
{
    if (g == nullptr) return fail("null gas handle");
    GRT_CUDA(cudaSetDevice(g->device));
    GRT_TRY(upload_table(g, s296, &g->tab.h2o[0]));
    GRT_TRY(upload_table(g, ckds, &g->tab.h2o[1]));
    GRT_TRY(upload_table(g, f296, &g->tab.h2o[2]));
    GRT_TRY(upload_table(g, ckdf, &g->tab.h2o[3]));
    g->tab.h2o_slot = find_slot(g, 1);
    return 0;
}

int grt_gas_set_o3_continuum(grt_gas *g, real_t const *cross_section)
{
    if (g == nullptr) return fail("null gas handle");
    GRT_CUDA(cudaSetDevice(g->device));
    GRT_TRY(upload_table(g, cross_section, &g->tab.o3));
    g->tab.o3_slot = find_slot(g, 3);
    return 0;
}

int grt_gas_add_cfc(grt_gas *g, int cfc_id, real_t const *cross_section)
{
    if (g == nullptr) return fail("null gas handle");
    if (cfc_id < 0 || cfc_id >= GRT_MAX_CFCS) return fail("unrecognized CFC id %d", cfc_id);
    if (g->tab.num_cfcs >= GRT_MAX_CFCS) return fail("too many CFCs");
    GRT_CUDA(cudaSetDevice(g->device));
    int const k = g->tab.num_cfcs;
    GRT_TRY(upload_table(g, cross_section, &g->tab.cfc[k]));
    g->meta.cfc_id[k] = cfc_id;
    g->tab.num_cfcs = g->meta.num_cfcs = k + 1;
    return 0;
}

int grt_gas_add_cia(grt_gas *g, int species0, int species1, real_t const *cross_section)
{
    if (g == nullptr) return fail("null gas handle");
    if (species0 < 0 || species0 >= GRT_NUM_CIA_SPECIES || species1 < 0 ||
        species1 >= GRT_NUM_CIA_SPECIES)
    {
        return fail("unrecognized CIA species (%d, %d)", species0, species1);
    }
    if (g->tab.num_cias >= GRT_MAX_CIAS) return fail("too many CIA pairs");
    GRT_CUDA(cudaSetDevice(g->device));
    int const k = g->tab.num_cias;
    GRT_TRY(upload_table(g, cross_section, &g->tab.cia[k]));
    g->meta.cia_id[k][0] = species0;
    g->meta.cia_id[k][1] = species1;
    g->tab.num_cias = g->meta.num_cias = k + 1;
    return 0;
}

int grt_gas_set_h2o_continuum_samples(grt_gas *g, real_t const *const x[4],
                                      real_t const *const y[4], uint64_t const m[4])
{
    if (g == nullptr || x == nullptr || y == nullptr || m == nullptr) return fail("null argument");
    GRT_CUDA(cudaSetDevice(g->device));
    for (int k=0; k<4; ++k) GRT_TRY(resample_table(g, x[k], y[k], m[k], &g->tab.h2o[k]));
    g->tab.h2o_slot = find_slot(g, 1);
    return 0;
}

int grt_gas_set_o3_continuum_samples(grt_gas *g, real_t const *x, real_t const *y, uint64_t m)
{
    if (g == nullptr) return fail("null gas handle");
    GRT_CUDA(cudaSetDevice(g->device));
    GRT_TRY(resample_table(g, x, y, m, &g->tab.o3));
    g->tab.o3_slot = find_slot(g, 3);
    return 0;
}

int grt_gas_add_cfc_samples(grt_gas *g, int cfc_id, real_t const *x, real_t const *y, uint64_t m)
{
    if (g == nullptr) return fail("null gas handle");
    if (cfc_id < 0 || cfc_id >= GRT_MAX_CFCS) return fail("unrecognized CFC id %d", cfc_id);
    if (g->tab.num_cfcs >= GRT_MAX_CFCS) return fail("too many CFCs");
    GRT_CUDA(cudaSetDevice(g->device));
    int const k = g->tab.num_cfcs;
    GRT_TRY(resample_table(g, x, y, m, &g->tab.cfc[k]));
    g->meta.cfc_id[k] = cfc_id;
    g->tab.num_cfcs = g->meta.num_cfcs = k + 1;
    return 0;
}

int grt_gas_add_cia_samples(grt_gas *g, int species0, int species1, real_t const *x,
                            real_t const *y, uint64_t m)
{
    if (g == nullptr) return fail("null gas handle");
    if (species0 < 0 || species0 >= GRT_NUM_CIA_SPECIES || species1 < 0 ||
        species1 >= GRT_NUM_CIA_SPECIES)
    {
        return fail("unrecognized CIA species (%d, %d)", species0, species1);
    }
    if (g->tab.num_cias >= GRT_MAX_CIAS) return fail("too many CIA pairs");
    GRT_CUDA(cudaSetDevice(g->device));
    int const k = g->tab.num_cias;
    GRT_TRY(resample_table(g, x, y, m, &g->tab.cia[k]));
    g->meta.cia_id[k][0] = species0;
    g->meta.cia_id[k][1] = species1;
    g->tab.num_cias = g->meta.num_cias = k + 1;
    return 0;
}

/* Diagnostic read-back of a device table (tests): kind 0-3 = H2O S296 / CKDS / F296 / CKDF,
 * 4 = O3, 5 = CFC number `index` (in add order), 6 = CIA pair number `index`. */
int grt_gas_debug_table(grt_gas *g, int kind, int index, double *out)
{
    if (g == nullptr || out == nullptr) return fail("null argument");
    double const *src = nullptr;
    if (kind >= 0 && kind < 4) src = g->tab.h2o[kind];
    else if (kind == 4) src = g->tab.o3;
    else if (kind == 5 && index >= 0 && index < g->tab.num_cfcs) src = g->tab.cfc[index];
    else if (kind == 6 && index >= 0 && index < g->tab.num_cias) src = g->tab.cia[index];
    if (src == nullptr) return fail("no such table (kind %d, index %d)", kind, index);
    GRT_CUDA(cudaSetDevice(g->device));
    GRT_CUDA(cudaStreamSynchronize(g->stream));
    GRT_CUDA(cudaMemcpy(out, src, sizeof(double)*g->nw, cudaMemcpyDeviceToHost));
    return 0;
}

int grt_gas_optical_depth(grt_gas *g, real_t const *p_atm_in, real_t const *t_in,
                          real_t const *x_mol_in, real_t const *x_cfc_in, real_t const *x_cia_in,
                          int method, real_t *tau_dev)
{
    if (g == nullptr) return fail("null gas handle");
    if (method < 0 || method > 2) return fail("unknown optical depth method %d", method);
#ifdef GRT_SINGLE
    if (method != 2)
    {
        return fail("the single-precision build provides optical_depth_method line_sample (the "
                    "method of every reference driver, driver.c:618-624); wavenumber_sweep and "
                    "line_sweep exist in the double build");
    }
#endif
    // the column inputs, widened to double (exact) in one staging block
    {
        size_t const lv = (size_t)g->num_levels;
        g->h_inputs.resize(lv*(2 + GRT_MAX_MOLS + GRT_MAX_CFCS + GRT_NUM_CIA_SPECIES));
        double *q = g->h_inputs.data();
        for (size_t k=0; k<lv; ++k) q[k] = (double)p_atm_in[k];
        q += lv;
        for (size_t k=0; k<lv; ++k) q[k] = (double)t_in[k];
        q += lv;
        for (size_t k=0; k<lv*GRT_MAX_MOLS; ++k) q[k] = (double)x_mol_in[k];
        q += lv*GRT_MAX_MOLS;
        for (size_t k=0; k<lv*GRT_MAX_CFCS; ++k) q[k] = (double)x_cfc_in[k];
        q += lv*GRT_MAX_CFCS;
        for (size_t k=0; k<lv*GRT_NUM_CIA_SPECIES; ++k) q[k] = (double)x_cia_in[k];
    }
    double const *p_atm = g->h_inputs.data();
    double const *t = p_atm + g->num_levels;
    double const *x_mol = t + g->num_levels;
    GRT_CUDA(cudaSetDevice(g->device));
    if (g->dirty || g->order_method != method) GRT_TRY(finalize_lines(g, method));
    int const L = g->num_layers;
    int const levels = g->num_levels;
    cudaStream_t const st = g->stream;

    // Stage the column inputs.
    double *d = g->d_inputs;
    ColumnInputs in;
    in.p = d;
    in.t = d + levels;
    in.x_mol = d + 2*levels;
    in.x_cfc = in.x_mol + (size_t)GRT_MAX_MOLS*levels;
    in.x_cia = in.x_cfc + (size_t)GRT_MAX_CFCS*levels;
    // one copy: the staging block has the device layout (p, t, x_mol, x_cfc, x_cia)
    GRT_CUDA(cudaMemcpyAsync((void *)in.p, p_atm, sizeof(double)*g->h_inputs.size(),
                             cudaMemcpyHostToDevice, st));

    LayerView lay;
    double *q = g->d_layer;
    lay.n = q; q += L;
    lay.pavg = q; q += L;
    lay.tavg = q; q += L;
    lay.ps = q; q += (size_t)GRT_MAX_MOLS*L;
    lay.ns = q; q += (size_t)GRT_MAX_MOLS*L;
    lay.qinv = q; q += (size_t)GRT_MAX_MOLS*L*kMaxIso;
    lay.cfc = q; q += (size_t)GRT_MAX_CFCS*L;
    lay.cia = q;

    // CIA constant of kernels.c:611-615.
    double const m_air = 28.97/6.02214076e23;
    double const grav = 980.;
    double const kboltz = 1.38064852e-16;
    double const atmtobarye = 1.013e6;
    double const cia_c = (atmtobarye*atmtobarye)/(kboltz*m_air*grav*2.);

    int launches = 0;
    GRT_CUDA(cudaEventRecord(g->ev[0], st));
    column_prep_kernel<<<(L + 63)/64, 64, 0, st>>>(L, in, g->meta, lay, cia_c);
    ++launches;

    uint64_t const N = g->num_lines;
#ifdef GRT_SINGLE
    int const fsteps = (int)std::ceil((double)(25.f/(float)g->dw));     // kernels.c:417, fp_t = float
#else
    int const fsteps = (int)std::ceil((double)25.f/g->dw);
#endif
    GRT_CUDA(cudaMemsetAsync(g->d_stats, 0, sizeof(PrepStats), st));
    GRT_CUDA(cudaMemsetAsync(g->d_hmax, 0xff, sizeof(int)*GRT_MAX_LAYERS, st));
    GRT_CUDA(cudaMemsetAsync(g->d_ties, 0, 2*sizeof(unsigned), st));
    if (N > 0)
    {
        size_t const need = (size_t)N*L;
        if (need > g->recs_capacity)
        {
            if (g->d_recs) GRT_CUDA(cudaFree(g->d_recs));
            g->d_recs = nullptr;
            GRT_CUDA(cudaMalloc(&g->d_recs, sizeof(LineRec)*need));
            g->recs_capacity = need;
        }
        LineView lv;
        lv.v0 = g->d_lines;
        lv.delta = g->d_lines + N;
        lv.s0 = g->d_lines + 2*N;
        lv.en = g->d_lines + 3*N;
        lv.nair = g->d_lines + 4*N;
        lv.yair = g->d_lines + 5*N;
        lv.yself = g->d_lines + 6*N;
        lv.slotiso = g->d_slotiso;
        lv.num_lines = N;
        dim3 grid((unsigned)((N + 255)/256), (unsigned)L);
        if (method != 2 && need > g->binned_capacity)
        {
            cudaFree(g->d_vkey); cudaFree(g->d_vsorted); cudaFree(g->d_vidx); cudaFree(g->d_perm);
            cudaFree(g->d_sorted_recs); cudaFree(g->d_segkey); cudaFree(g->d_wing_a); cudaFree(g->d_wing_b);
            g->d_segkey = nullptr;
            g->d_wing_a = nullptr;
            g->d_wing_b = nullptr;
            g->d_vkey = g->d_vsorted = nullptr;
            g->d_vidx = g->d_perm = nullptr;
            g->d_sorted_recs = nullptr;
            GRT_CUDA(cudaMalloc(&g->d_sorted_recs, sizeof(LineRec)*need));
            GRT_CUDA(cudaMalloc(&g->d_vkey, sizeof(double)*need));
            GRT_CUDA(cudaMalloc(&g->d_vsorted, sizeof(double)*need));
            GRT_CUDA(cudaMalloc(&g->d_vidx, sizeof(unsigned)*need));
            GRT_CUDA(cudaMalloc(&g->d_perm, sizeof(unsigned)*need));
            GRT_CUDA(cudaMalloc(&g->d_segkey, sizeof(unsigned short)*2*need));
            GRT_CUDA(cudaMalloc(&g->d_wing_a, sizeof(double2)*need));
            GRT_CUDA(cudaMalloc(&g->d_wing_b, sizeof(float4)*need));
            g->binned_capacity = need;
        }
        bool patch = true;
        if (char const *env = getenv("GRT_K1_PATCH")) patch = env[0] != '0';
        line_prep_kernel<<<grid, 256, 0, st>>>(lv, L, lay, g->meta, g->w0, g->dw, g->nw, fsteps,
                                               g->d_recs, g->d_stats,
                                               method != 2 ? g->d_vkey : nullptr,
                                               method != 2 ? g->d_vidx : nullptr,
                                               patch ? g->d_ties : nullptr, g->d_tie_overflow,
                                               kTieCapacity, method == 2 ? g->d_hmax : nullptr);
        ++launches;
    }
    // The one host synchronisation of the call: K1's counters and its list of near-tie
    // records come back; the sweep kernels below are only enqueued (later calls are ordered
    // behind them on the library stream, and every call that hands data to the host
    // synchronises).
    GRT_CUDA(cudaMemcpyAsync(&g->h_block->ties, g->d_ties, sizeof(TieList), cudaMemcpyDeviceToHost, st));
    GRT_CUDA(cudaMemcpyAsync(&g->h_block->stats, g->d_stats, sizeof(PrepStats),
                             cudaMemcpyDeviceToHost, st));
    GRT_CUDA(cudaStreamSynchronize(st));
    g->patched_last = 0;
    if (N > 0 && g->h_block->ties.count > 0)
    {
        GRT_TRY(patch_near_ties(g, g->h_block->ties.count, p_atm, t, x_mol));
        g->patched_last = g->h_block->ties.count;
        g->patched_total += g->patched_last;
        ++launches;
    }
    g->last_method = method;
    g->stats[2] = 0;
    if (method != 2)
    {
        return binned_optical_depth(g, method, lay, (double *)tau_dev, launches);
    }

    GRT_CUDA(cudaEventRecord(g->ev[1], st));
    SweepParams sp;
    sp.recs = g->d_recs;
    sp.num_lines = N;
    sp.w0 = g->w0;
    sp.wres = g->dw;
    sp.nw = (int)g->nw;
    sp.fsteps = fsteps;
    double pmax = 0.;
    for (int i=0; i<levels; ++i) pmax = std::max(pmax, std::fabs(p_atm[i]));
    double const margin = (double)fsteps*g->dw + 2.*g->dw + g->max_abs_delta*pmax + 1e-6;
    sp.num_layers = L;
    sp.lay = lay;
    sp.tab = g->tab;
    sp.tau = tau_dev;
    sp.core_order = g->d_core_order;
    if (char const *env = getenv("GRT_CORE_ORDER")) { if (env[0] == '0') sp.core_order = nullptr; }
    // Points per thread (PTS): LN*PTS points per warp tile; NL layers per work item.  Wider
    // tiles amortise the per-line work better; narrower ones waste fewer lanes at window
    // ends and balance better (a window is 5001 points at 0.01 cm-1 but only 501 at 0.1,
    // where a 64-point tile beats a 128-point one by 8 % and a 256-point one by 30 %).
    int pts = (g->dw < 0.05) ? 8 : 2;
    int nl = 1;
    if (char const *env = getenv("GRT_SWEEP_PTS")) pts = atoi(env);
    if (char const *env = getenv("GRT_SWEEP_NL")) nl = atoi(env);
    if (pts != 2 && pts != 4 && pts != 8 && pts != 16) pts = 4;
    if (nl != 1 && nl != 2) nl = 1;
    if (pts == 2 || pts == 16) nl = 1;
    // Lanes per line group (sweep.cuh): 32 = one line at a time per warp; 16 / 8 = two / four
    // lines at a time on a tile of LN*PTS points.
    int ln = 32;
    if (g->dw >= 0.05)
    {
        // 501-point windows: two lines at a time on 64-point tiles, four samples per lane and
        // visit (config-1 column 24.4 -> 20.2 ms, RFMIP-shaped column 101.5 -> 93.9 ms)
        ln = 16;
        pts = 4;
    }
    if (char const *env = getenv("GRT_SWEEP_PTS")) pts = atoi(env);
    if (pts != 2 && pts != 4 && pts != 8 && pts != 16) pts = 4;
    if (char const *env = getenv("GRT_SWEEP_LANES")) ln = atoi(env);
    if (ln != 8 && ln != 16) ln = 32;
    if (ln != 32) nl = 1;
    if (ln == 16 && pts == 2) pts = 4;
    if (ln == 8 && pts < 4) pts = 4;
    GRT_TRY(prepare_tiles(g, ln*pts, margin, g->max_abs_delta*pmax));
    sp.num_tiles = g->num_tiles;
    sp.tile_order = g->d_tile_order;
    sp.tile_major = g->tiles_spectral ? 1 : 0;
    sp.tile_first = g->d_tile_first;
    sp.tile_last = g->d_tile_last;
    sp.work_counter = g->d_work_counter;
    GRT_CUDA(cudaMemsetAsync(g->d_work_counter, 0, sizeof(unsigned int), st));
    {
        int const sms = sm_count(g->device);
        int rc = 0;
        switch ((ln != 32 ? ln*1000 : 0) + pts*10 + nl)
        {
            case 16041: rc = launch_wing<4, 1, 16>(sp, sms, st); break;
            case 16081: rc = launch_wing<8, 1, 16>(sp, sms, st); break;
            case 16161: rc = launch_wing<16, 1, 16>(sp, sms, st); break;
            case 8041: rc = launch_wing<4, 1, 8>(sp, sms, st); break;
            case 8081: rc = launch_wing<8, 1, 8>(sp, sms, st); break;
            case 8161: rc = launch_wing<16, 1, 8>(sp, sms, st); break;
            case 21: rc = launch_wing<2, 1>(sp, sms, st); break;
            case 41: rc = launch_wing<4, 1>(sp, sms, st); break;
            case 42: rc = launch_wing<4, 2>(sp, sms, st); break;
            case 81: rc = launch_wing<8, 1>(sp, sms, st); break;
            case 161: rc = launch_wing<16, 1>(sp, sms, st); break;
            default: rc = launch_wing<8, 2>(sp, sms, st); break;
        }
        GRT_TRY(rc);
    }
    ++launches;
    GRT_CUDA(cudaEventRecord(g->ev[2], st));
    // GRT_CORE=gather: the line cores through core_gather_kernel -- no atomics, optical depths
    // bit-reproducible from run to run -- at three times the core kernel's time (measured,
    // profiles/r02k_core_gather_ab.txt: config 2 526 -> 600 ms per step); the default is the
    // scatter with double atomics (3 % of the samples; the wing sweep has none).
    bool gather = false;
    if (char const *env = getenv("GRT_CORE")) gather = strcmp(env, "gather") == 0;
    if (N > 0 && gather)
    {
        CoreGather cg;
        cg.first_at = g->d_first_at;
        cg.num_granules = g->num_granules;
        cg.hmax = g->d_hmax;
        cg.margin_pts = (int)std::ceil(g->max_abs_delta*pmax/g->dw) + 3;
        cg.num_tiles = (int)((g->nw + kCoreTile - 1)/kCoreTile);
        cg.work_counter = g->d_work_counter;
        // lowest pressure first: the widest Doppler cores, i.e. the most expensive tiles
        std::vector<std::pair<double, int>> by_p((size_t)L);
        for (int i=0; i<L; ++i) by_p[i] = {std::fabs(p_atm[i]) + std::fabs(p_atm[i+1]), i};
        std::stable_sort(by_p.begin(), by_p.end());
        for (int i=0; i<L; ++i) cg.layer_order[i] = (short)by_p[i].second;
        GRT_CUDA(cudaMemsetAsync(g->d_work_counter, 0, sizeof(unsigned int), st));
        unsigned const items = (unsigned)cg.num_tiles*(unsigned)L;
        unsigned const want = (items + kCoreWarps - 1)/kCoreWarps;
        unsigned const blocks = std::min(want, (unsigned)(4*sm_count(g->device)));
        core_gather_kernel<<<blocks, kCoreWarps*32, 0, st>>>(sp, cg);
        ++launches;
    }
    else if (N > 0)
    {
        dim3 grid((unsigned)((N + 127)/128), (unsigned)L);
        core_sweep_kernel<<<grid, 128, 0, st>>>(sp);
        ++launches;
    }
    GRT_CUDA(cudaEventRecord(g->ev[3], st));
    if (char const *env = getenv("GRT_CLASS_HIST"))
    {
        if (env[0] == '1' && N > 0)
        {
            if (g->d_class_counts == nullptr)
            {
                GRT_CUDA(cudaMalloc(&g->d_class_counts, 8*sizeof(unsigned long long)));
            }
            GRT_CUDA(cudaMemsetAsync(g->d_class_counts, 0, 8*sizeof(unsigned long long), st));
            dim3 grid((unsigned)((N + 255)/256), (unsigned)L);
            class_histogram_kernel<<<grid, 256, 0, st>>>(sp, ln*pts, g->d_class_counts);
        }
    }
    GRT_CUDA(cudaGetLastError());
    PrepStats const &hs = g->h_block->stats;
    g->stats[0] = hs.records;
    g->stats[1] = hs.samples;
    g->stats[2] = (uint64_t)launches;
    g->stats[3] = hs.lorentz;
    g->stats[7] = hs.core;
    g->timings_pending = true;     // stats[4..6] are read from the events on demand
    return 0;
}

int grt_gas_last_stats(grt_gas *g, uint64_t out[8])
{
    if (g == nullptr || out == nullptr) return fail("null argument");
    if (g->timings_pending)
    {
        // waits for the kernels of the last grt_gas_optical_depth call
        GRT_CUDA(cudaSetDevice(g->device));
        GRT_CUDA(cudaEventSynchronize(g->ev[3]));
        float ms[3] = {0.f, 0.f, 0.f};
        for (int k=0; k<3; ++k) GRT_CUDA(cudaEventElapsedTime(&ms[k], g->ev[k], g->ev[k+1]));
        g->stats[4] = (uint64_t)(ms[0]*1e6f);       // K0 + K1 (+ patch) [ns]
        if (g->last_method == 2)
        {
            g->stats[5] = (uint64_t)(ms[1]*1e6f);   // K3 wing sweep [ns]
            g->stats[6] = (uint64_t)(ms[2]*1e6f);   // K3c core sweep [ns]
        }
        else
        {
            g->stats[5] = (uint64_t)(ms[2]*1e6f);   // bin gather [ns]
            g->stats[6] = (uint64_t)(ms[1]*1e6f);   // sort [ns]
            unsigned long long evals = 0;
            GRT_CUDA(cudaMemcpy(&evals, g->d_evals, sizeof(evals), cudaMemcpyDeviceToHost));
            g->stats[1] = evals;
        }
        g->timings_pending = false;
    }
    for (int k=0; k<8; ++k) out[k] = g->stats[k];
    return 0;
}

/* Records the host re-evaluated with its own libm because Y sat near a float rounding
 * boundary: out[0] = by the last grt_gas_optical_depth call, out[1] = since creation. */
int grt_gas_patch_counts(grt_gas const *g, uint64_t out[2])
{
    if (g == nullptr || out == nullptr) return fail("null argument");
    out[0] = g->patched_last;
    out[1] = g->patched_total;
    return 0;
}

/* Diagnostic read-back of K1's records for one layer: REPWID and Y (RFM_voigt.c:94-95) and
 * the shifted centre of records [first, first + count) in the device line order
 * (line_sample: merged, sorted by unshifted centre, ties in upload order; binned methods:
 * molecule-major in upload order). */
int grt_gas_debug_records(grt_gas *g, int layer, uint64_t first, uint64_t count, float *repwid,
                          float *y, double *centre)
{
    if (g == nullptr) return fail("null gas handle");
    if (layer < 0 || layer >= g->num_layers || first + count > g->num_lines)
    {
        return fail("record range outside the line list");
    }
    GRT_CUDA(cudaSetDevice(g->device));
    GRT_CUDA(cudaStreamSynchronize(g->stream));
    std::vector<LineRec> h((size_t)count);
    GRT_CUDA(cudaMemcpy(h.data(), g->d_recs + (size_t)layer*g->num_lines + first,
                        sizeof(LineRec)*count, cudaMemcpyDeviceToHost));
    for (uint64_t k=0; k<count; ++k)
    {
        if (repwid) repwid[k] = h[k].repwid;
        if (y) y[k] = h[k].y;
        if (centre) centre[k] = h[k].v;
    }
    return 0;
}

/* Device order of the uploaded lines: order[k] = index, in upload order over all
 * add_molecule calls, of the line at device position k. */
int grt_gas_debug_line_order(grt_gas const *g, uint64_t count, uint64_t *order)
{
    if (g == nullptr || order == nullptr) return fail("null argument");
    if (count != g->h_order.size()) return fail("line count mismatch");
    for (uint64_t k=0; k<count; ++k) order[k] = g->h_order[k];
    return 0;
}

/* GRT_CLASS_HIST=1 in the environment: line-tile visits of the wing sweep per line class
 * during the last line_sample call: out[0..5] = fast, region 1, far edge, near, Lorentz
 * (covered), Lorentz (edge); out[6] = points per tile; out[7] unused. */
int grt_gas_class_histogram(grt_gas *g, uint64_t out[8])
{
    if (g == nullptr || out == nullptr) return fail("null argument");
    if (g->d_class_counts == nullptr) return fail("class histogram not enabled (GRT_CLASS_HIST=1)");
    GRT_CUDA(cudaSetDevice(g->device));
    GRT_CUDA(cudaStreamSynchronize(g->stream));
    GRT_CUDA(cudaMemcpy(g->h_class_counts, g->d_class_counts, sizeof(g->h_class_counts),
                        cudaMemcpyDeviceToHost));
    for (int k=0; k<8; ++k) out[k] = g->h_class_counts[k];
    out[6] = (uint64_t)g->tiles_pts;
    return 0;
}

}  // extern "C"

#if GRT_DBG_TIMERS
// Development build only: per-class SM cycles and visits of the wing sweep since the last call.
extern "C" int grt_debug_timers(unsigned long long *cycles, unsigned long long *visits, int n)
{
    unsigned long long c[grt::kTmCount], v[grt::kTmCount];
    if (cudaMemcpyFromSymbol(c, grt::g_dbg_cycles, sizeof(c)) != cudaSuccess) return 1;
    if (cudaMemcpyFromSymbol(v, grt::g_dbg_visits, sizeof(v)) != cudaSuccess) return 1;
    for (int i=0; i<n && i<grt::kTmCount; ++i) { cycles[i] = c[i]; visits[i] = v[i]; }
    unsigned long long z[grt::kTmCount] = {};
    cudaMemcpyToSymbol(grt::g_dbg_cycles, z, sizeof(z));
    cudaMemcpyToSymbol(grt::g_dbg_visits, z, sizeof(z));
    return 0;
}
#endif
