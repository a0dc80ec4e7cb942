// K3b: the two binned optical-depth methods of the reference,
//   wavenumber_sweep = calc_optical_depth_bin_sweep  (gas-optics/src/kernels.c:177-304)
//   line_sweep       = calc_optical_depth_line_sweep (gas-optics/src/kernels.c:308-406)
// followed by interpolate / interpolate_last_bin (kernels.c:514-581), on the RFM-style two
// meshes of create_spectral_bins (spectral_bin.c:30-96): 1 cm-1 bins; lines near a bin
// ("local") are evaluated on every grid point of the bin, lines within 25 bins ("remote")
// on the bin's three interpolation points, whose sum over all molecules is interpolated
// quadratically (clamped at zero) onto the grid points.
//
// The reference scatters per line (line_sweep, with atomics) or walks bins with per-layer
// sorted lines (bin_sweep).  Here both are ONE gather kernel: a block owns one (bin, layer),
// finds -- per molecule, by binary searches in the molecule's centre-sorted lines of that
// layer -- the index ranges of its local and remote lines, accumulates them in registers /
// shared memory, interpolates its own three remote sums, adds the continua, CFC and CIA
// terms and stores tau once.  No atomics on global memory, no second pass over tau.
//
// Which lines are local and which remote decides values at the 1e-3 level, so the range
// logic follows the reference literally, including bracket()'s behaviour when the value
// lies outside the array (kernel_utils.c:26-76: left = 0, right = size - 1, error ignored
// by the callers).  Reads of uninitialised or out-of-bounds memory in the reference (empty
// sub-arrays handed to bracket, bin index n reached by a line at the very end of the grid)
// have no defined result; this kernel takes the empty range there.
#pragma once
#include "voigt.cuh"

namespace grt {

constexpr int kBinThreads = 256;          // one warp per bin of a gather group
constexpr int kNip = 3;                 // spectral_bin-internal.h:30
constexpr int kBinLanePerLine = 16;     // bins of up to 16 points: local lines lane per line
constexpr int kBinLinePerWarp = 128;    // bins of up to 128 points: local lines one at a time

struct BinGeometry
{
    double w0;
    double wres;
    unsigned long long num_wpoints;
    unsigned long long n;               // number of bins
    int ppb;
    int last_ppb;
    int do_interp;
    int do_last_interp;
};

struct BinParams
{
    LineRec const *recs;                // [layers][num_lines], molecule-major line order
    double const *vs;                   // [layers][num_lines] centres sorted per (molecule, layer)
    unsigned const *perm;               // [layers][num_lines] record index of each sorted entry
    LineRec const *sorted_recs;         // [layers][num_lines] records in sorted order
    double2 const *wing_a;              // [layers][num_lines] sorted order: v, c
    float4 const *wing_b;               //   repwid, yq, yrrtpi, far-wing limit (see WingRec)
    unsigned long long num_lines;
    int num_slots;
    unsigned slot_off[GRT_MAX_MOLS + 1];
    int num_layers;
    int method;                         // 0 wavenumber_sweep, 1 line_sweep
    BinGeometry bins;
    LayerView lay;
    Tables tab;
    double *tau;                        // [layers][num_wpoints]
    unsigned long long *evals;          // Voigt evaluations (bench bookkeeping)
};

// create_spectral_bins (spectral_bin.c:40-56).
inline BinGeometry make_bins(double const w0, unsigned long long const n, double const wres,
                             double const width)
{
    BinGeometry b;
    b.w0 = w0;
    b.wres = wres;
    b.num_wpoints = n;
    b.ppb = (int)(floor(width/wres) + 1);
    b.do_interp = b.ppb > 3 ? 1 : 0;
    b.last_ppb = (int)(n % (unsigned long long)b.ppb);
    b.last_ppb = b.last_ppb == 0 ? b.ppb : b.last_ppb;
    b.do_last_interp = b.last_ppb > 3 ? 1 : 0;
    b.n = n/(unsigned long long)b.ppb;
    if (b.ppb != b.last_ppb) b.n++;
    return b;
}

// Left-most grid index of bin i and the number of grid points in it (spectral_bin.c:75-80).
__device__ __forceinline__ unsigned long long bin_left(BinGeometry const &b, unsigned long long const i)
{
    return i*(unsigned long long)b.ppb;
}

__device__ __forceinline__ int bin_points(BinGeometry const &b, unsigned long long const i)
{
    return i < b.n - 1 ? b.ppb : b.last_ppb;
}

// bins.w[i*NIP + k] (spectral_bin.c:81-83).
__device__ __forceinline__ double bin_w(BinGeometry const &b, unsigned long long const i, int const k)
{
    double const first = __dadd_rn(b.w0, __dmul_rn((double)((unsigned long long)b.ppb*i), b.wres));
    if (k == 0) return first;
    double const last = __dadd_rn(first, __dmul_rn((double)(bin_points(b, i) - 1), b.wres));
    if (k == kNip - 1) return last;
    return __dmul_rn(0.5, __dadd_rn(first, last));
}

// bins.w[idx] for a flat index.
__device__ __forceinline__ double bin_w_flat(BinGeometry const &b, unsigned long long const idx)
{
    return bin_w(b, idx/kNip, (int)(idx % kNip));
}

// bracket (kernel_utils.c:26-76), literally.  Out of range: l = 0, r = size - 1.
__device__ void bracket_dev(unsigned long long const size, double const *a, double const val,
                            unsigned long long *left, unsigned long long *right)
{
    unsigned long long l = 0;
    unsigned long long r = size - 1;
    if (val < a[l] || val > a[r])
    {
        *left = l;
        *right = r;
        return;
    }
    if (a[l] == val)
    {
        r = l;
    }
    else if (a[r] == val)
    {
        l = r;
    }
    else
    {
        while (r - l > 1)
        {
            unsigned long long const mid = l + (r - l)/2;
            if (a[mid] == val)
            {
                l = mid;
                r = mid;
                break;
            }
            else if (val > a[mid])
            {
                l = mid;
            }
            else
            {
                r = mid;
            }
        }
    }
    *left = l;
    *right = r;
}

// Index ranges [lo, hi) into one molecule's sorted lines of one layer.
struct BinRanges
{
    unsigned long long loc_lo, loc_hi;     // local lines
    unsigned long long rl_lo, rl_hi;       // remote lines below the local ones
    unsigned long long rr_lo, rr_hi;       // remote lines above
};

struct BinRanges32
{
    unsigned loc_lo, loc_hi, rl_lo, rl_hi, rr_lo, rr_hi;
};

constexpr int kBinGroup = kBinThreads/32;   // bins per gather block

// calc_optical_depth_bin_sweep's bookkeeping for bin j (kernels.c:195-300).
__device__ void ranges_bin_sweep(BinGeometry const &b, unsigned long long const j,
                                 unsigned long long const n, double const *v, BinRanges *out)
{
    BinRanges R = {0, 0, 0, 0, 0, 0};
    unsigned long long const none = ~0ull;
    unsigned long long nbin = 1;
    double const leftw = j > nbin ? bin_w_flat(b, kNip*(j - nbin)) : bin_w_flat(b, 0);
    double const rightw = j >= (b.n - 1) - nbin ? bin_w_flat(b, kNip*b.n - 1)
                                                : bin_w_flat(b, kNip*(j + nbin + 1) - 1);
    unsigned long long left = 0;
    unsigned long long right = 0;
    bool left_set = false;
    bool right_set = false;
    if (leftw <= v[n-1] && rightw >= v[0])
    {
        unsigned long long tmp;
        bracket_dev(n, v, leftw, &left, &tmp);
        bracket_dev(n - left, v + left, rightw, &tmp, &right);
        right += left;
        R.loc_lo = left;
        R.loc_hi = right + 1;
        left_set = right_set = true;
    }
    else if (leftw > v[n-1])
    {
        left = n;
        left_set = true;
    }
    else
    {
        right = none;
        right_set = true;
    }
    nbin = 25;
    double const leftw_r = j > nbin ? bin_w_flat(b, kNip*(j - nbin)) : bin_w_flat(b, 0);
    if (leftw >= v[0] && leftw_r <= v[n-1] && left_set && left >= 1)
    {
        unsigned long long left_r, tmp;
        bracket_dev(left, v, leftw_r, &left_r, &tmp);
        R.rl_lo = left_r;
        R.rl_hi = left;
    }
    double const rightw_r = j >= (b.n - 1) - nbin ? bin_w_flat(b, kNip*b.n - 1)
                                                  : bin_w_flat(b, kNip*(j + nbin + 1) - 1);
    if (rightw <= v[n-1] && rightw_r >= v[0] && right_set)
    {
        unsigned long long const start = right == none ? 0 : right;    // right + f
        if (start < n)
        {
            unsigned long long right_r, tmp;
            bracket_dev(n - start, v + start, rightw_r, &tmp, &right_r);
            right_r += start;
            R.rr_lo = right == none ? 0 : right + 1;
            R.rr_hi = right_r + 1;
        }
    }
    *out = R;
}

// calc_optical_depth_line_sweep (kernels.c:322-400) turned around: the lines whose own
// local / remote bin ranges contain bin k.  Every bound is a monotone function of the
// line centre, so each set is a contiguous range of the sorted lines.
__device__ __forceinline__ long long sweep_bin_of(double const w, double const w0, double const bw)
{
    return (long long)floor(__ddiv_rn(__dsub_rn(w, w0), bw));
}

template <typename Pred>
__device__ unsigned long long first_true(unsigned long long const n, Pred pred)
{
    unsigned long long lo = 0, hi = n;      // pred is false ... false true ... true
    while (lo < hi)
    {
        unsigned long long const mid = lo + (hi - lo)/2;
        if (pred(mid)) hi = mid; else lo = mid + 1;
    }
    return lo;
}

__device__ void ranges_line_sweep(BinGeometry const &b, unsigned long long const k,
                                  unsigned long long const n, double const *v, BinRanges *out)
{
    double const bw = __dmul_rn(b.wres, (double)b.ppb);
    double const maxw = __dadd_rn(b.w0, __dmul_rn((double)b.num_wpoints, b.wres));
    long long const kk = (long long)k;
    auto lower = [&](double const c, double const cut) -> long long
    {
        double w = __dsub_rn(c, cut);
        if (w < b.w0) w = b.w0;
        return sweep_bin_of(w, b.w0, bw);
    };
    auto upper = [&](double const c, double const cut) -> long long
    {
        double w = __dadd_rn(c, cut);
        if (w > maxw) w = maxw;
        return sweep_bin_of(w, b.w0, bw);
    };
    // local: left(v) <= k <= right(v)
    unsigned long long const i_lo = first_true(n, [&](unsigned long long i) { return upper(v[i], 1.5) >= kk; });
    unsigned long long const i_hi = first_true(n, [&](unsigned long long i) { return lower(v[i], 1.5) > kk; });
    // remote below the line's local bins: left_r(v) <= k < left(v)
    unsigned long long const j_hi = first_true(n, [&](unsigned long long i) { return lower(v[i], 25.) > kk; });
    // remote above: right(v) < k <= right_r(v)
    unsigned long long const j_lo = first_true(n, [&](unsigned long long i) { return upper(v[i], 25.) >= kk; });
    BinRanges R;
    R.loc_lo = i_lo;
    R.loc_hi = i_hi > i_lo ? i_hi : i_lo;
    R.rl_lo = i_hi;                      // lines with left(v) > k ...
    R.rl_hi = j_hi > i_hi ? j_hi : i_hi; // ... and left_r(v) <= k
    R.rr_lo = j_lo;                      // lines with right_r(v) >= k ...
    R.rr_hi = i_lo > j_lo ? i_lo : j_lo; // ... and right(v) < k
    *out = R;
}

// The 32 bytes of a record that a far-wing or pure-Lorentz sample needs, as two arrays of
// 16-byte vectors in sorted order: a warp whose lanes take consecutive lines then loads two
// contiguous 512-byte rows instead of gathering 48 bytes out of 32 different 64-byte records
// (which kept the L1 data pipe at 85 % of its wavefront peak and the issue rate at 57 %).
// limit = XLIM0 for a Voigt record (|x| >= limit is the far wing, RFM_voigt.c:166-171),
// +inf for a record flagged core-only (never the shortcut), NaN for a pure-Lorentz record.
__device__ __forceinline__ void wing_record(LineRec const &r, double2 *a, float4 *b)
{
    float limit = r.xlim0;
    if (r.hw0 == kHwCoreOnly) limit = __int_as_float(0x7f800000);
    if (r.hw0 == kHwLorentz) limit = __int_as_float(0x7fc00000);
    *a = make_double2(r.v, r.c);
    *b = make_float4(r.repwid, r.yq, r.yrrtpi, limit);
}

// acc += S*N_s*K at one point of one line: rfm_voigt_line_shape for a single abscissa
// (RFM_voigt.c:97-106 pure Lorentz, :109-278 Voigt) with the scale of kernels.c:459 in c.
// t2 = fl(w + fl(i*wres)), the grid wavenumber as RFM_voigt.c:165 forms it (the same for
// every line, so the callers compute it once per point).  `full` is only dereferenced for the
// samples inside |x| < XLIM0.
__device__ __forceinline__ void line_point(double &acc, double2 const A, float4 const B,
                                           LineRec const *full, double const t2)
{
    // float -> double by integer ops (exact for the positive normal values here): the
    // conversion unit shares the XU pipe with the reciprocal of the division
    float const xi = abscissa_from(t2, A.x, widen_positive(B.x));
    float const ax = fabsf(xi);
    if (ax >= B.w)
    {
        // far wing (RFM_voigt.c:168-171): almost every remote sample
        acc = fma(A.y, widen_positive(div_exact(B.z, ax*ax + B.y)), acc);
    }
    else if (B.w != B.w)
    {
        acc += A.y/widen_positive(xi*xi + B.y);
    }
    else
    {
        LineRec const r = *full;
        acc = fma(r.c, voigt_core(xi, r.y, r.xlim0, r.xlim1), acc);
    }
}

// bin_quad_interp (kernel_utils.c:81-103) at grid point j.
__device__ __forceinline__ double quad_interp(double const *x, double const *y, double const w)
{
    double t = (w - x[1])*(w - x[2])*y[0]/((x[0] - x[1])*(x[0] - x[2])) +
               (w - x[0])*(w - x[2])*y[1]/((x[1] - x[0])*(x[1] - x[2])) +
               (w - x[0])*(w - x[1])*y[2]/((x[2] - x[0])*(x[2] - x[1]));
    if (t < 0.) t = 0.;
    return t;
}

// Records in sorted order, so that the gather reads them with unit stride and no
// dependent index load: out[layer][pos] = in[layer][perm[layer][pos]] (64 B each).
__global__ void __launch_bounds__(256)
permute_records_kernel(LineRec const *in, unsigned const *perm, LineRec *out,
                       unsigned long long const num_lines)
{
    unsigned long long const pos = (unsigned long long)blockIdx.x*blockDim.x + threadIdx.x;
    if (pos >= num_lines) return;
    size_t const base = (size_t)blockIdx.y*num_lines;
    uint4 const *src = reinterpret_cast<uint4 const *>(in + base + perm[base + pos]);
    uint4 *dst = reinterpret_cast<uint4 *>(out + base + pos);
    uint4 const a = src[0], b = src[1], c = src[2], d = src[3];
    dst[0] = a; dst[1] = b; dst[2] = c; dst[3] = d;
}

// Ranges of every (layer, bin, molecule), one thread each: the binary searches are serial
// chains of dependent loads, so they are spread over as many threads as there are triples
// instead of being walked by one thread of the gather block.
__global__ void __launch_bounds__(128)
bin_ranges_kernel(BinParams const p, BinRanges32 *out)
{
    unsigned long long const idx = (unsigned long long)blockIdx.x*blockDim.x + threadIdx.x;
    int const S = p.num_slots;
    unsigned long long const total = (unsigned long long)p.num_layers*p.bins.n*(unsigned long long)S;
    if (idx >= total) return;
    int const s = (int)(idx % (unsigned long long)S);
    unsigned long long const k = (idx/(unsigned long long)S) % p.bins.n;
    int const layer = (int)(idx/((unsigned long long)S*p.bins.n));
    unsigned long long const off = p.slot_off[s];
    unsigned long long const n = p.slot_off[s+1] - off;
    BinRanges R = {0, 0, 0, 0, 0, 0};
    if (n > 0)
    {
        double const *v = p.vs + (size_t)layer*p.num_lines + off;
        if (p.method == 0) ranges_bin_sweep(p.bins, k, n, v, &R);
        else ranges_line_sweep(p.bins, k, n, v, &R);
    }
    BinRanges32 r;
    r.loc_lo = (unsigned)R.loc_lo; r.loc_hi = (unsigned)R.loc_hi;
    r.rl_lo = (unsigned)R.rl_lo; r.rl_hi = (unsigned)R.rl_hi;
    r.rr_lo = (unsigned)R.rr_lo; r.rr_hi = (unsigned)R.rr_hi;
    out[idx] = r;
}

// A block owns kBinGroup consecutive bins of one layer, one warp per bin: neighbouring bins
// see almost the same remote lines (25 bins either side), so the eight warps stream the same
// records through L1 at the same time.  Remote lines: lanes take consecutive records and
// evaluate the bin's three interpolation points (three register accumulators per thread for
// the whole kernel).  Local lines: a lane per line with shuffle sums on coarse grids, a lane per
// grid point on fine ones.  The warp owns its bin's accumulators: no atomics, no block
// barrier inside the molecule loop.
#ifndef GRT_BIN_MINBLOCKS
#define GRT_BIN_MINBLOCKS 4
#endif
__global__ void __launch_bounds__(kBinThreads, GRT_BIN_MINBLOCKS)
bin_gather_kernel(BinParams const p, BinRanges32 const *ranges)
{
    extern __shared__ double sh_acc[];          // [kBinGroup][points per bin]
    // coarse grids only: per-lane partial sums of the local lines, [kBinGroup][ppb][32]
    double *sh_loc = sh_acc + kBinGroup*p.bins.ppb;
    __shared__ double sh_rem[kBinGroup][kNip];
    BinGeometry const &b = p.bins;
    unsigned long long const k0 = (unsigned long long)blockIdx.x*kBinGroup;
    int const ng = (int)((b.n - k0) < (unsigned long long)kBinGroup ? (b.n - k0) : kBinGroup);
    int const layer = blockIdx.y;
    int const tid = threadIdx.x;
    int const lane = tid & 31;
    int const g = tid >> 5;                      // the warp's bin inside the group
    int const ppb = b.ppb;
    for (int q=tid; q<kBinGroup*ppb; q+=kBinThreads) sh_acc[q] = 0.;
    if (ppb <= kBinLanePerLine)
    {
        for (int q=tid; q<kBinGroup*ppb*32; q+=kBinThreads) sh_loc[q] = 0.;
    }
    __syncthreads();
    unsigned long long evals = 0;
    double rem[kNip] = {0., 0., 0.};
    if (g < ng)
    {
        unsigned long long const k = k0 + g;
        double const rem_w = bin_w(b, k, 0);
        double const rem_res = __dsub_rn(bin_w(b, k, 1), rem_w);
        double rem_t[kNip];
#pragma unroll
        for (int l=0; l<kNip; ++l) rem_t[l] = __dadd_rn(rem_w, __dmul_rn((double)l, rem_res));
        int const npts = bin_points(b, k);
        double const loc_w = __dadd_rn(b.w0, __dmul_rn((double)bin_left(b, k), b.wres));
        for (int s=0; s<p.num_slots; ++s)
        {
            unsigned long long const off = p.slot_off[s];
            if (p.slot_off[s+1] == off) continue;
            LineRec const *recs = p.sorted_recs + (size_t)layer*p.num_lines + off;
            double2 const *wa = p.wing_a + (size_t)layer*p.num_lines + off;
            float4 const *wb = p.wing_b + (size_t)layer*p.num_lines + off;
            BinRanges32 const R = ranges[((size_t)layer*b.n + k)*p.num_slots + s];
            // remote lines on the three interpolation points
#pragma unroll 1
            for (int side=0; side<2; ++side)
            {
                unsigned const lo = side ? R.rr_lo : R.rl_lo;
                unsigned const hi = side ? R.rr_hi : R.rl_hi;
#pragma unroll 2
                for (unsigned i=lo + lane; i<hi; i+=32)
                {
                    double2 const A = wa[i];
                    float4 const B = wb[i];
#pragma unroll
                    for (int l=0; l<kNip; ++l) line_point(rem[l], A, B, recs + i, rem_t[l]);
                    evals += kNip;
                }
            }
            // local lines on the bin's grid points
            if (ppb <= kBinLanePerLine)
            {
                // coarse grid: a lane per line (consecutive lines of the sorted list sit at
                // nearly the same distance from a grid point, so the warp takes the same
                // Voigt branch); every lane keeps its own partial sum per grid point in shared
                // memory (lane-contiguous, conflict-free), summed over the lanes once per bin
                double *mine = sh_loc + (size_t)g*ppb*32 + lane;
                for (unsigned i=R.loc_lo + lane; i<R.loc_hi; i+=32)
                {
                    double2 const A = wa[i];
                    float4 const B = wb[i];
                    evals += (unsigned long long)npts;
                    for (int pt=0; pt<npts; ++pt)
                    {
                        double a = mine[pt*32];
                        line_point(a, A, B, recs + i, __dadd_rn(loc_w, __dmul_rn((double)pt, b.wres)));
                        mine[pt*32] = a;
                    }
                }
            }
            else if (npts <= kBinLinePerWarp)
            {
                // fine grid, bins of up to 128 points: line by line, a lane per grid point
                // (32 at a time).  The line's Voigt constants are formed once, and only if one
                // of its samples in this bin lies inside |x| < XLIM0 (voigt_line_eval: the
                // per-line form of RFM_voigt.c:109-278, bit-identical to voigt_core and about
                // half its instructions in the CPF12 sums); chunks wholly in the far wing
                // take the far-wing quotient directly.
                double *mine = sh_acc + g*ppb;
                for (unsigned i=R.loc_lo; i<R.loc_hi; ++i)
                {
                    double2 const A = wa[i];
                    float4 const B = wb[i];
                    double const repwid = widen_positive(B.x);
                    bool const lorentz = B.w != B.w;
                    bool have_line = false;
                    VoigtLine line;
                    evals += (unsigned long long)npts;
#pragma unroll 1
                    for (int pbase=0; pbase<npts; pbase+=32)
                    {
                        int const pt = pbase + lane;
                        bool const live = pt < npts;
                        double const loc_t = __dadd_rn(loc_w, __dmul_rn((double)(live ? pt : pbase), b.wres));
                        float const xi = abscissa_from(loc_t, A.x, repwid);
                        float const ax = fabsf(xi);
                        double k;
                        if (lorentz)
                        {
                            k = A.y/widen_positive(xi*xi + B.y);      // RFM_voigt.c:97-106
                        }
                        else if (!__any_sync(0xffffffffu, ax < B.w))
                        {
                            k = widen_positive(div_exact(B.z, ax*ax + B.y));
                        }
                        else
                        {
                            if (!have_line)
                            {
                                LineRec const *r = recs + i;
                                voigt_line_init(line, r->y, r->xlim0, r->xlim1);
                                have_line = true;
                            }
                            k = voigt_line_eval(line, xi);
                        }
                        if (live)
                        {
                            if (lorentz) mine[pt] += k;
                            else mine[pt] = fma(A.y, k, mine[pt]);
                        }
                    }
                }
                __syncwarp();
            }
            else
            {
                // finer grids still: a lane per grid point (32 at a time), every local line in turn
                for (int pbase=0; pbase<npts; pbase+=32)
                {
                    int const pt = pbase + lane;
                    if (pt >= npts) continue;
                    double const loc_t = __dadd_rn(loc_w, __dmul_rn((double)pt, b.wres));
                    double a = 0.;
#pragma unroll 2
                    for (unsigned i=R.loc_lo; i<R.loc_hi; ++i)
                    {
                        line_point(a, wa[i], wb[i], recs + i, loc_t);
                        ++evals;
                    }
                    sh_acc[g*ppb + pt] += a;
                }
            }
        }
    }
    if (ppb <= kBinLanePerLine && g < ng)
    {
        __syncwarp();
        for (int pt=0; pt<ppb; ++pt)
        {
            double a = sh_loc[((size_t)g*ppb + pt)*32 + lane];
            for (int o=16; o>0; o>>=1) a += __shfl_down_sync(0xffffffffu, a, o);
            if (lane == 0) sh_acc[g*ppb + pt] += a;
        }
    }
#pragma unroll
    for (int l=0; l<kNip; ++l)
    {
        double vsum = rem[l];
        for (int o=16; o>0; o>>=1) vsum += __shfl_down_sync(0xffffffffu, vsum, o);
        if (lane == 0) sh_rem[g][l] = vsum;
    }
    for (int o=16; o>0; o>>=1) evals += __shfl_down_sync(0xffffffffu, evals, o);
    if (lane == 0 && evals) atomicAdd(p.evals, evals);
    __syncthreads();

    // interpolate (kernels.c:514-581) + continua, CFCs, CIA (kernels.c:469-510, 585-630)
    int const L = p.num_layers;
    for (int q=tid; q<ng*ppb; q+=kBinThreads)
    {
        int const gq = q/ppb;
        int const pt = q % ppb;
        unsigned long long const k = k0 + gq;
        if (pt >= bin_points(b, k)) continue;
        unsigned long long const f = bin_left(b, k) + (unsigned long long)pt;
        double t = sh_acc[gq*ppb + pt];
        if (p.tab.h2o_slot >= 0)
        {
            double const tref = 296.;
            double const T = p.lay.tavg[layer];
            double const Ps = p.lay.ps[p.tab.h2o_slot*L + layer];
            double const N = p.lay.ns[p.tab.h2o_slot*L + layer];
            double const P = p.lay.pavg[layer];
            double const CS = p.tab.h2o[0][f];
            double const T0 = p.tab.h2o[1][f];
            double const CF = p.tab.h2o[2][f];
            double const T0F = p.tab.h2o[3][f];
            t += N*(tref/T)*((CS*Ps*exp(T0*(tref - T))) + (CF*(P - Ps)*exp(T0F*(tref - T))));
        }
        if (p.tab.o3_slot >= 0)
        {
            t += p.lay.ns[p.tab.o3_slot*L + layer]*p.tab.o3[f];
        }
        for (int c=0; c<p.tab.num_cfcs; ++c)
        {
            t += p.lay.cfc[c*L + layer]*p.tab.cfc[c][f];
        }
        for (int c=0; c<p.tab.num_cias; ++c)
        {
            t += p.lay.cia[c*L + layer]*p.tab.cia[c][f];
        }
        bool const interp = (k < b.n - 1) ? (b.do_interp != 0) : (b.do_last_interp != 0);
        double const y[kNip] = {sh_rem[gq][0], sh_rem[gq][1], sh_rem[gq][2]};
        if (interp)
        {
            double const x[kNip] = {bin_w(b, k, 0), bin_w(b, k, 1), bin_w(b, k, 2)};
            double const w = __dadd_rn(b.w0, __dmul_rn((double)f, b.wres));
            t += quad_interp(x, y, w);
        }
        else
        {
            t += y[pt];      // bin_no_interp (kernel_utils.c:106-118): at most three points
        }
        p.tau[(size_t)layer*b.num_wpoints + f] = t;
    }
}

}  // namespace grt
