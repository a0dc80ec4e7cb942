// K3 / K3c: the line-sample sweep (calc_optical_depth_line_sample,
// gas-optics/src/kernels.c:410-465, with rfm_voigt_line_shape inlined) as a
// gather + a small scatter.
//
// wing_sweep_kernel (gather, no atomics): persistent warps pull (tile, layer group)
// work items from a global counter -- layer groups in order, so that the records of
// one or two layers stay L2-resident while every tile re-reads them, and within a
// group the tiles with the most lines first.  A warp owns 32*PTS consecutive grid
// points of NL consecutive layers; lane l owns points tile*32*PTS + j*32 + l, j < PTS,
// accumulated in registers.  The 64-byte records of every line within +-25 cm-1 of the
// tile are staged through a warp-private, double-buffered shared-memory ring by 1-D
// bulk TMA copies (cp.async.bulk + mbarrier, issued by lane 0) and consumed as
// broadcast loads.  Per line the warp evaluates
//     |x| >= XLIM0          far wing      K = Y/sqrt(pi) / (x^2 + y^2)
//     XLIM1 <= |x| < XLIM0  Humlicek region 1
//     Y >= 70.55            pure Lorentz
// Samples below XLIM1 (regions 2, 3, CPF12: a handful per line) are left to
// core_sweep_kernel, which scatters them with double atomics afterwards.
// The epilogue adds the continua / CFC / CIA terms and stores tau once.
//
// The grid wavenumber of a sample, fl(fl(s*wres + w0) + fl(i*wres)) (kernels.c:438,
// RFM_voigt.c:165), depends on the line only through its window start s: it is cached
// in registers and recomputed when s changes, i.e. once per group of lines sharing a
// nearest grid point and once per layer group, not once per sample.
#pragma once
#include "voigt.cuh"

namespace grt {

constexpr int kSweepThreads = 256;
constexpr int kSweepWarps = kSweepThreads/32;

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


// ---- 1-D bulk TMA + mbarrier plumbing (PTX ISA: cp.async.bulk, mbarrier) ------------
__device__ __forceinline__ uint32_t smem_u32(void const *p)
{
    return (uint32_t)__cvta_generic_to_shared(p);
}

__device__ __forceinline__ void mbar_init(uint32_t const mbar, uint32_t const count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(mbar), "r"(count) : "memory");
}

__device__ __forceinline__ void mbar_expect_tx(uint32_t const mbar, uint32_t const bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;"
                 :: "r"(mbar), "r"(bytes) : "memory");
}

__device__ __forceinline__ void bulk_copy_g2s(uint32_t const dst, void const *src,
                                              uint32_t const bytes, uint32_t const mbar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes "
                 "[%0], [%1], %2, [%3];"
                 :: "r"(dst), "l"(src), "r"(bytes), "r"(mbar) : "memory");
}

__device__ __forceinline__ void mbar_wait(uint32_t const mbar, uint32_t const parity)
{
    uint32_t done = 0;
    unsigned spins = 0;
    do
    {
        asm volatile("{\n\t.reg .pred p;\n\t"
                     "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                     "selp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(mbar), "r"(parity) : "memory");
        if (++spins > (1u << 24)) __trap();   // a lost copy must not hang the GPU
    } while (!done);
}


// 1/d for d > 0 to ~1 ulp without the IEEE division slow path: float seed + two
// Newton steps in double.  Only used where the result feeds the double accumulator
// (pure-Lorentz branch), never before a float rounding.  d_f > 0 is a normal float.
__device__ __forceinline__ double fast_rcp_of_float(float const d_f)
{
    float r0;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r0) : "f"(d_f));
    double const d = widen_positive(d_f);
    double r = widen_positive(r0);
    double e = fma(-d, r, 1.0);
    r = fma(r, e, r);
    e = fma(-d, r, 1.0);
    r = fma(r, e, r);
    return r;
}


// Per-warp state of the sample abscissae: t2[j] = fl(dwno(s) + fl(i*wres)) for the
// lane's PTS points and the window start s it was computed for.
template <int PTS>
struct GridCache
{
    int s;
    double t2[PTS];
};

// (float)((t2 - v)*REPWID): the last two double roundings and the float rounding of
// RFM_voigt.c:165.
__device__ __forceinline__ float abscissa_from(double const t2, double const v, double const repwid)
{
    return __double2float_rn(__dmul_rn(__dsub_rn(t2, v), repwid));
}


// One line of one layer against the warp's 32*PTS points.
template <int PTS>
__device__ __forceinline__ void accumulate_line(LineRec const *rec, SweepParams const &p,
                                                int const warp_base, int const lane,
                                                GridCache<PTS> &gc, double (&acc)[PTS])
{
    // hot part of the record: three broadcast loads
    double2 const vc = *reinterpret_cast<double2 const *>(&rec->v);
    float4 const q = *reinterpret_cast<float4 const *>(&rec->repwid);
    int4 const fh = *reinterpret_cast<int4 const *>(&rec->fc);   // fc, hw0, y, xlim0
    int const fc = fh.x;
    int const hw0 = fh.y;
    int const warp_last = warp_base + 32*PTS - 1;
    int const F = p.fsteps;
    if (fc == kSkip || fc + F < warp_base || fc - F > warp_last || hw0 == kHwCoreOnly) return;
    double const v = vc.x;
    double const c = vc.y;
    double const repwid = __hiloint2double(__float_as_int(q.y), __float_as_int(q.x));
    float const yq = q.z;
    float const yrrtpi = q.w;
    int const s = max(fc - F, 0);
    int const e = min(fc + F, p.nw - 1);
    if (s != gc.s)
    {
        gc.s = s;
        double const dwno = __dadd_rn(__dmul_rn((double)s, p.wres), p.w0);
        double const i0 = (double)(warp_base + lane - s);
#pragma unroll
        for (int j=0; j<PTS; ++j)
        {
            gc.t2[j] = __dadd_rn(dwno, __dmul_rn(i0 + (double)(32*j), p.wres));
        }
    }
    bool const covered = (s <= warp_base) && (e >= warp_last);
    bool const far = (hw0 >= 0) && ((warp_base - fc > hw0) || (fc - warp_last > hw0));
    if (far)
    {
        if (covered)
        {
            // Every lane of every chunk is a far-wing sample: no predicates, PTS
            // independent dependency chains in flight.
#pragma unroll
            for (int j=0; j<PTS; ++j)
            {
                float const xi = abscissa_from(gc.t2[j], v, repwid);
                float const kf = div_exact(yrrtpi, xi*xi + yq);
                acc[j] = fma(c, widen_positive(kf), acc[j]);
            }
        }
        else
        {
            // A window end falls inside the tile (kernels.c:435-437).
#pragma unroll
            for (int j=0; j<PTS; ++j)
            {
                int const cb = warp_base + j*32;
                if (cb <= e && cb + 31 >= s)
                {
                    int const f = cb + lane;
                    float const xi = abscissa_from(gc.t2[j], v, repwid);
                    float const kf = div_exact(yrrtpi, xi*xi + yq);
                    if (f >= s && f <= e) acc[j] = fma(c, widen_positive(kf), acc[j]);
                }
            }
        }
    }
    else if (hw0 == kHwLorentz)
    {
#pragma unroll
        for (int j=0; j<PTS; ++j)
        {
            int const cb = warp_base + j*32;
            if (cb <= e && cb + 31 >= s)
            {
                int const f = cb + lane;
                float const xi = abscissa_from(gc.t2[j], v, repwid);
                double const r = fast_rcp_of_float(xi*xi + yq);
                if (f >= s && f <= e) acc[j] = fma(c, r, acc[j]);
            }
        }
    }
    else
    {
        // The tile is near the line centre: chunks may hold far-wing, region-1 and
        // line-core samples (the latter are skipped here).
        float4 const a = *reinterpret_cast<float4 const *>(&rec->xlim1);   // xlim1, a0, d0, d2
        float const y = __int_as_float(fh.z);
        float const xlim0 = __int_as_float(fh.w);
        float const xlim1 = a.x;
        float const a0 = a.y;
        float2 const d = make_float2(a.z, a.w);
#pragma unroll
        for (int j=0; j<PTS; ++j)
        {
            int const cb = warp_base + j*32;
            if (cb <= e && cb + 31 >= s)
            {
                int const f = cb + lane;
                float const xi = abscissa_from(gc.t2[j], v, repwid);
                float const ax = fabsf(xi);
                float const xq = ax*ax;
                bool const in = (f >= s) && (f <= e);
                bool const isfar = ax >= xlim0;
                bool const isr1 = !isfar && (ax >= xlim1);
                float kf;
                if (__all_sync(0xffffffffu, isfar || !in))
                {
                    kf = div_exact(yrrtpi, xq + yq);
                }
                else if (__all_sync(0xffffffffu, isr1 || !in))
                {
                    kf = voigt_region1(xq, y, a0, d.x, d.y);
                }
                else
                {
                    kf = div_exact(yrrtpi, xq + yq);
                    if (isr1) kf = voigt_region1(xq, y, a0, d.x, d.y);
                }
                if (in && (isfar || isr1)) acc[j] = fma(c, widen_positive(kf), acc[j]);
            }
        }
    }
}


template <int PTS, int NL>
struct SweepConfig
{
    static constexpr int kStageBytes = NL*32*(int)sizeof(LineRec);
    static constexpr int kSmemBytes = kSweepWarps*2*kStageBytes + kSweepWarps*2*8;
    // registers: PTS*NL accumulators + PTS cached abscissae, all double
    static constexpr int kMinBlocks = (PTS*NL <= 8) ? 2 : 1;
};


// Warp-granular persistent gather: every WARP pulls its own (32*PTS points, NL layers)
// work item from the global counter and streams the records of its lines through a
// warp-private shared-memory ring: no block-wide barrier anywhere, so a warp never
// waits for a neighbour whose lines are denser.
template <int PTS, int NL>
__global__ void __launch_bounds__(kSweepThreads, SweepConfig<PTS, NL>::kMinBlocks)
wing_sweep_kernel(SweepParams const p)
{
    using Cfg = SweepConfig<PTS, NL>;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    int const lane = threadIdx.x & 31;
    int const warp = threadIdx.x >> 5;
    unsigned char *ring = smem_raw + (size_t)warp*(2*Cfg::kStageBytes);
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem_raw + (size_t)kSweepWarps*2*Cfg::kStageBytes) +
                     warp*2;
    uint32_t const mb0 = smem_u32(bars);
    uint32_t const mb1 = smem_u32(bars + 1);
    uint32_t const sb0 = smem_u32(ring);
    uint32_t const sb1 = sb0 + Cfg::kStageBytes;
    if (lane == 0)
    {
        mbar_init(mb0, 1);
        mbar_init(mb1, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    unsigned phase = 0;        // bit k = parity the next wait on stage k expects

    int const L = p.num_layers;
    int const groups = (L + NL - 1)/NL;
    unsigned const total = (unsigned)p.num_tiles*(unsigned)groups;

    for (;;)
    {
        unsigned item = 0;
        if (lane == 0) item = atomicAdd(p.work_counter, 1u);
        item = __shfl_sync(0xffffffffu, item, 0);
        if (item >= total) break;
        int const group = (int)(item/(unsigned)p.num_tiles);
        int const slot = (int)(item%(unsigned)p.num_tiles);
        int const tile = p.tile_order[slot];
        unsigned const first = p.tile_first[tile];
        unsigned const last = p.tile_last[tile];
        int const warp_base = tile*(32*PTS);
        int const layer0 = group*NL;
        int const nl = min(NL, L - layer0);
        int const nb = (int)((last - first + 31u)/32u);

        double acc[NL][PTS];
#pragma unroll
        for (int l=0; l<NL; ++l)
        {
#pragma unroll
            for (int j=0; j<PTS; ++j) acc[l][j] = 0.;
        }
        GridCache<PTS> gc;
        gc.s = INT32_MIN;
#pragma unroll
        for (int j=0; j<PTS; ++j) gc.t2[j] = 0.;

        // Lane 0 issues the copies of batch k (32 records of each of the nl layers).
        auto issue = [&](int const k, uint32_t const sb, uint32_t const mb)
        {
            unsigned const base = first + 32u*(unsigned)k;
            uint32_t const bytes = (uint32_t)min(32u, last - base)*(uint32_t)sizeof(LineRec);
            mbar_expect_tx(mb, bytes*(uint32_t)nl);
            for (int l=0; l<nl; ++l)
            {
                bulk_copy_g2s(sb + (uint32_t)l*32u*(uint32_t)sizeof(LineRec),
                              p.recs + ((size_t)(layer0 + l)*p.num_lines + base), bytes, mb);
            }
        };
        if (lane == 0)
        {
            if (nb > 0) issue(0, sb0, mb0);
            if (nb > 1) issue(1, sb1, mb1);
        }
        for (int k=0; k<nb; ++k)
        {
            int const st = k & 1;
            mbar_wait(st ? mb1 : mb0, (phase >> st) & 1u);
            phase ^= 1u << st;
            int const cnt = (int)min(32u, last - (first + 32u*(unsigned)k));
            LineRec const *stage = reinterpret_cast<LineRec const *>(ring + st*Cfg::kStageBytes);
            for (int b=0; b<cnt; ++b)
            {
#pragma unroll
                for (int l=0; l<NL; ++l)
                {
                    if (l < nl) accumulate_line<PTS>(stage + l*32 + b, p, warp_base, lane, gc, acc[l]);
                }
            }
            __syncwarp();
            if (lane == 0 && k + 2 < nb)
            {
                // order the warp's generic-proxy reads before the async-proxy overwrite
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                issue(k + 2, st ? sb1 : sb0, st ? mb1 : mb0);
            }
        }

        // Epilogue: continua, CFCs, CIA (kernels.c:469-510, 585-630), one store.
#pragma unroll
        for (int l=0; l<NL; ++l)
        {
            if (l >= nl) continue;
            int const layer = layer0 + l;
#pragma unroll
            for (int j=0; j<PTS; ++j)
            {
                int const f = warp_base + j*32 + lane;
                if (f >= p.nw) continue;
                double t = acc[l][j];
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
}


// K3c: samples with |x| < XLIM1 (Humlicek regions 2, 3 and CPF12) of every Voigt-branch
// record, plus every sample of the rare records flagged kHwCoreOnly.  G lanes share
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
    LineRec const *rp = p.recs + (size_t)layer*p.num_lines + rec_id;
    int4 const fh = *reinterpret_cast<int4 const *>(&rp->fc);   // fc, hw0, y, xlim0
    int const fc = fh.x;
    int const hw0 = fh.y;
    if (fc == kSkip || hw0 == kHwLorentz) return;
    float const y = __int_as_float(fh.z);
    float const xlim0 = __int_as_float(fh.w);
    float const xlim1 = rp->xlim1;
    bool const flagged = hw0 == kHwCoreOnly;
    if (!flagged && xlim1 <= 0.f) return;
    double2 const vc = *reinterpret_cast<double2 const *>(&rp->v);
    double const repwid = rp->repwid;
    int const F = p.fsteps;
    int const s = max(fc - F, 0);
    int const e = min(fc + F, p.nw - 1);
    // Half-width of the |x| < xlim1 window in grid points, rounded up generously.
    double const hd = flagged ? (double)(2*F + 2) : (double)xlim1/(repwid*p.wres) + 2.;
    int const h = hd < (double)(2*F + 2) ? (int)hd : 2*F + 2;
    int const lo = max(fc - h, s);
    int const hi = min(fc + h, e);
    double const dwno = __dadd_rn(__dmul_rn((double)s, p.wres), p.w0);
    double *tau = p.tau + (size_t)layer*p.nw;
    for (int f=lo+sub; f<=hi; f+=G)
    {
        float const xi = abscissa(dwno, (double)(f - s), p.wres, vc.x, repwid);
        if (flagged || fabsf(xi) < xlim1)
        {
            double const k = voigt_core(xi, y, xlim0, xlim1);
            atomicAdd(&tau[f], vc.y*k);
        }
    }
}

}  // namespace grt
