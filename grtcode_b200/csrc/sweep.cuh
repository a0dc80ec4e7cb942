// K3 / K3c: the line-sample sweep (calc_optical_depth_line_sample,
// gas-optics/src/kernels.c:410-465, with rfm_voigt_line_shape inlined) as a
// gather + a small scatter.
//
// wing_sweep_kernel (gather, no atomics): persistent warps pull (tile, layer) work
// items, heaviest tiles first, from a global counter.  A warp owns 32*PTS
// consecutive grid points of one layer; lane l owns points tile*32*PTS + j*32 + l,
// j < PTS, accumulated in registers.  The records of every line within +-25 cm-1
// of the tile stream through a warp-private shared-memory slot (double-buffered
// against registers); the warp filters them with a ballot and, for every line whose
// window touches its range, evaluates
//     |x| >= XLIM0          far wing      K = Y/sqrt(pi) / (x^2 + y^2)
//     XLIM1 <= |x| < XLIM0  Humlicek region 1
//     Y >= 70.55            pure Lorentz
// Samples below XLIM1 (regions 2, 3, CPF12: a handful per line) are left to
// core_sweep_kernel, which scatters them with double atomics afterwards.
// The epilogue adds the continua / CFC / CIA terms and stores tau once.
#pragma once
#include "voigt.cuh"

namespace grt {

constexpr int kSweepThreads = 256;

struct SweepParams
{
    LineRec const *recs;       // [layers][num_lines]
    uint64_t num_lines;
    double w0;
    double wres;
    int nw;
    int fsteps;
    int num_layers;
    int num_tiles;             // warp tiles of 32*PTS grid points
    int const *tile_order;     // tiles sorted by line count, heaviest first
    unsigned const *tile_first;  // per tile: first record whose window may reach it
    unsigned const *tile_last;   // per tile: one past the last such record
    unsigned int *work_counter;
    LayerView lay;
    Tables tab;
    double *tau;               // [layers][nw]
};


// 1/d for d > 0 to ~1 ulp without the IEEE division slow path: float seed + two
// Newton steps in double.  Only used where the result feeds the double accumulator
// (pure-Lorentz branch), never before a float rounding.
__device__ __forceinline__ double fast_rcp(double const d)
{
    float r0;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r0) : "f"(__double2float_rn(d)));
    double r = (double)r0;
    double e = fma(-d, r, 1.0);
    r = fma(r, e, r);
    e = fma(-d, r, 1.0);
    r = fma(r, e, r);
    return r;
}


// One hit = one line whose window touches the warp's 32*PTS points.
template <int PTS>
__device__ __forceinline__ void accumulate_line(LineRec const &r, SweepParams const &p,
                                                int const warp_base, int const warp_last,
                                                int const lane, double (&acc)[PTS])
{
    int const F = p.fsteps;
    int const s = max(r.fc - F, 0);
    int const e = min(r.fc + F, p.nw - 1);
    double const dwno = __dadd_rn(__dmul_rn((double)s, p.wres), p.w0);
    double const repwid_d = (double)r.repwid;
    double const i0 = (double)(warp_base + lane - s);
    float const yq = r.y*r.y;
    bool const covered = (s <= warp_base) && (e >= warp_last);
    // Grid points within hf of the centre may fall below XLIM0.
    float const hf = r.xlim0/(r.repwid*(float)p.wres) + 2.f;
    bool const wing_only = covered && (r.y < kLorentzY) &&
        (((float)(warp_base - r.fc) > hf) || ((float)(r.fc - warp_last) > hf));
    if (wing_only)
    {
        // Every lane of every chunk is a far-wing sample: no predicates, PTS independent
        // dependency chains in flight.
        float const yrrtpi = r.y*kRsqrpi;
#pragma unroll
        for (int j=0; j<PTS; ++j)
        {
            float const xi = abscissa(dwno, i0 + (double)(32*j), p.wres, r.v, repwid_d);
            float const kf = div_exact(yrrtpi, xi*xi + yq);
            acc[j] = fma(r.c, widen_positive(kf), acc[j]);
        }
    }
    else if (r.y >= kLorentzY)
    {
#pragma unroll
        for (int j=0; j<PTS; ++j)
        {
            int const cb = warp_base + j*32;
            if (cb <= e && cb + 31 >= s)
            {
                int const f = cb + lane;
                float const xi = abscissa(dwno, i0 + (double)(32*j), p.wres, r.v, repwid_d);
                double const d = (double)(xi*xi + yq);
                if (f >= s && f <= e) acc[j] = fma(r.c, fast_rcp(d), acc[j]);
            }
        }
    }
    else
    {
        float const yrrtpi = r.y*kRsqrpi;
#pragma unroll
        for (int j=0; j<PTS; ++j)
        {
            int const cb = warp_base + j*32;
            if (cb <= e && cb + 31 >= s)
            {
                int const f = cb + lane;
                float const xi = abscissa(dwno, i0 + (double)(32*j), p.wres, r.v, repwid_d);
                float const ax = fabsf(xi);
                float const xq = ax*ax;
                bool const in = (f >= s) && (f <= e);
                float kf = div_exact(yrrtpi, xq + yq);
                bool const reg1 = in && (ax < r.xlim0) && (ax >= r.xlim1);
                if (reg1) kf = voigt_region1(xq, r.y, r.a0, r.d0, r.d2);
                if (in && (ax >= r.xlim1)) acc[j] = fma(r.c, widen_positive(kf), acc[j]);
            }
        }
    }
}


// Warp-granular persistent gather: every WARP pulls its own (32*PTS points, layer) work
// item from the global counter (heaviest first) and streams the records of its lines
// through a warp-private shared-memory slot: no block-wide barrier anywhere, so a warp
// never waits for a neighbour whose lines are denser.
template <int PTS>
__global__ void __launch_bounds__(kSweepThreads, (PTS <= 4) ? 3 : 2)
wing_sweep_kernel(SweepParams const p)
{
    __shared__ LineRec stage_all[kSweepThreads/32][32];
    int const lane = threadIdx.x & 31;
    int const warp = threadIdx.x >> 5;
    LineRec *stage = stage_all[warp];
    int const F = p.fsteps;
    int const L = p.num_layers;
    unsigned const total = (unsigned)p.num_tiles*(unsigned)L;

    for (;;)
    {
        unsigned item = 0;
        if (lane == 0) item = atomicAdd(p.work_counter, 1u);
        item = __shfl_sync(0xffffffffu, item, 0);
        if (item >= total) break;
        int const slot = (int)(item/(unsigned)L);
        int const layer = (int)(item%(unsigned)L);
        int const tile = p.tile_order[slot];
        unsigned const first = p.tile_first[tile];
        unsigned const last = p.tile_last[tile];
        int const warp_base = tile*(32*PTS);
        int const warp_last = warp_base + 32*PTS - 1;

        double acc[PTS];
#pragma unroll
        for (int j=0; j<PTS; ++j) acc[j] = 0.;

        int4 const *src = reinterpret_cast<int4 const *>(p.recs + (size_t)layer*p.num_lines);
        int4 *dst = reinterpret_cast<int4 *>(stage);
        // Software pipeline: the next 32 records travel global -> registers while the
        // current 32 are consumed from shared memory.
        int4 q0, q1, q2;
        {
            size_t const o = (size_t)first*3;
            size_t const lim = (size_t)last*3;
            q0 = (o + lane < lim) ? __ldg(src + o + lane) : make_int4(0, 0, 0, 0);
            q1 = (o + 32 + lane < lim) ? __ldg(src + o + 32 + lane) : make_int4(0, 0, 0, 0);
            q2 = (o + 64 + lane < lim) ? __ldg(src + o + 64 + lane) : make_int4(0, 0, 0, 0);
        }
        for (unsigned base=first; base<last; base+=32)
        {
            int const cnt = (int)min(32u, last - base);
            __syncwarp();
            dst[lane] = q0;
            dst[32 + lane] = q1;
            dst[64 + lane] = q2;
            __syncwarp();
            {
                size_t const o = (size_t)(base + 32)*3;
                size_t const lim = (size_t)last*3;
                q0 = (o + lane < lim) ? __ldg(src + o + lane) : make_int4(0, 0, 0, 0);
                q1 = (o + 32 + lane < lim) ? __ldg(src + o + 32 + lane) : make_int4(0, 0, 0, 0);
                q2 = (o + 64 + lane < lim) ? __ldg(src + o + 64 + lane) : make_int4(0, 0, 0, 0);
            }
            int const fc_l = lane < cnt ? stage[lane].fc : kSkip;
            bool const hit = (fc_l != kSkip) && (fc_l + F >= warp_base) && (fc_l - F <= warp_last);
            unsigned mask = __ballot_sync(0xffffffffu, hit);
            while (mask)
            {
                int const b = __ffs(mask) - 1;
                mask &= mask - 1;
                LineRec const r = stage[b];
                accumulate_line<PTS>(r, p, warp_base, warp_last, lane, acc);
            }
        }

        // Epilogue: continua, CFCs, CIA (kernels.c:469-510, 585-630), one store.
#pragma unroll
        for (int j=0; j<PTS; ++j)
        {
            int const f = warp_base + j*32 + lane;
            if (f >= p.nw) continue;
            double t = acc[j];
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
            p.tau[(size_t)layer*p.nw + f] = t;
        }
    }
}


// K3c: samples with |x| < XLIM1 (Humlicek regions 2, 3 and CPF12) of every Voigt-branch
// record, plus every sample of the rare records flagged xlim1 = +inf.  G lanes share
// one record and stride over its near-centre window; contributions are scattered with
// double atomics.
template <int G>
__global__ void __launch_bounds__(256)
core_sweep_kernel(SweepParams const p)
{
    uint64_t const rec_id = ((uint64_t)blockIdx.x*blockDim.x + threadIdx.x)/G;
    int const sub = threadIdx.x & (G - 1);
    int const layer = blockIdx.y;
    if (rec_id >= p.num_lines) return;
    LineRec const r = p.recs[(size_t)layer*p.num_lines + rec_id];
    if (r.fc == kSkip || r.y >= kLorentzY || r.xlim1 <= 0.f) return;
    int const F = p.fsteps;
    int const s = max(r.fc - F, 0);
    int const e = min(r.fc + F, p.nw - 1);
    double const repwid_d = (double)r.repwid;
    float xlim0 = r.xlim0;
    float xlim1 = r.xlim1;
    bool const flagged = !(xlim1 < 3.0e38f);
    if (flagged) voigt_limits(r.y, &xlim0, &xlim1);
    // Half-width of the |x| < xlim1 window in grid points, rounded up generously.
    double const hd = flagged ? (double)(2*F + 2) : (double)xlim1/(repwid_d*p.wres) + 2.;
    int const h = hd < (double)(2*F + 2) ? (int)hd : 2*F + 2;
    int const a = max(r.fc - h, s);
    int const b = min(r.fc + h, e);
    double const dwno = __dadd_rn(__dmul_rn((double)s, p.wres), p.w0);
    double *tau = p.tau + (size_t)layer*p.nw;
    for (int f=a+sub; f<=b; f+=G)
    {
        float const xi = abscissa(dwno, (double)(f - s), p.wres, r.v, repwid_d);
        if (flagged || fabsf(xi) < xlim1)
        {
            double const k = voigt_core(xi, r.y, xlim0, xlim1);
            atomicAdd(&tau[f], r.c*k);
        }
    }
}

}  // namespace grt
