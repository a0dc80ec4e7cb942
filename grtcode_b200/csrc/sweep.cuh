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
    int const *tile_order;     // tiles heaviest first (LPT), or in spectral order
    int tile_major;            // items run tile by tile (all layers of a tile adjacent)
    unsigned const *tile_first;  // per tile: first record whose window may reach it
    unsigned const *tile_last;   // per tile: one past the last such record
    unsigned int *work_counter;
    unsigned const *core_order;  // core kernel: thread -> record (molecule-major), or null
    LayerView lay;
    Tables tab;
    real_t *tau;               // [layers][nw] (float in the single build)
};


// ---- precision of the abscissa chain ---------------------------------------------------
// Double build: the grid wavenumber, the line centre and the product with REPWID are double
// and only the last step rounds to float (RFM_voigt.c:165 with fp_t = double).  Single build
// (-DGRT_SINGLE, the reference's --enable-single-precision: fp_t = float,
// utilities/src/floating_point_type.h:23-29): every one of those operations is a float
// operation, in the same order -- in.w = s*wres + w0 (kernels.c:438), DWNO + i*WRES - WNOADJ,
// times REPWID (RFM_voigt.c:165) -- so the float roundings of the reference's single build
// (1e-4 cm-1 steps of the wavenumber at 2000 cm-1) are reproduced bit for bit; optical depth
// is still accumulated in double and rounded to float once, at the store.
#ifdef GRT_SINGLE
typedef float wn_t;
typedef float4 RecA;           // v, c, c*Y/sqrt(pi), -   (K1 writes the same 16 bytes)
__device__ __forceinline__ float rec_v(RecA const a) { return a.x; }
// c may be zero (absent species) or subnormal: a true conversion, once per line visit
__device__ __forceinline__ double rec_c(RecA const a) { return (double)a.y; }
__device__ __forceinline__ float rec_repwid(float4 const B) { return B.x; }
#else
typedef double wn_t;
typedef double2 RecA;          // v, c
__device__ __forceinline__ double rec_v(RecA const a) { return a.x; }
__device__ __forceinline__ double rec_c(RecA const a) { return a.y; }
__device__ __forceinline__ double rec_repwid(float4 const B) { return widen_positive(B.x); }
#endif

// fl(k*wres), fl(s*wres + w0) and fl(a + b) in the build's wavenumber type.
__device__ __forceinline__ wn_t wn_index_times(int const k, wn_t const wres)
{
#ifdef GRT_SINGLE
    return __fmul_rn((float)k, wres);
#else
    return __dmul_rn((double)k, wres);
#endif
}

__device__ __forceinline__ wn_t wn_add(wn_t const a, wn_t const b)
{
#ifdef GRT_SINGLE
    return __fadd_rn(a, b);
#else
    return __dadd_rn(a, b);
#endif
}


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


// 1/d for d > 0 to 2^-45 relative without the IEEE division: float seed (2^-23) + one
// Newton step in double.  Only used where the result feeds the double accumulator
// (pure-Lorentz branch; the parity tolerance on tau is 1e-9), never before a float
// rounding.  d_f > 0 is a normal float.
__device__ __forceinline__ double fast_rcp_of_float(float const d_f)
{
    float r0;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r0) : "f"(d_f));
    double const d = widen_positive(d_f);
    double const r = widen_positive(r0);
    double const e = fma(-d, r, 1.0);
    return fma(r, e, r);
}


// Per-warp cache of the sample abscissae for kernels without the shared-memory table:
// t2[j] = fl(dwno(s) + fl(i*wres)) for the lane's PTS points, keyed by the centre index fc
// of the line it was computed for (s = max(fc - fsteps, 0), kernels.c:435).  (Empty for TBL kernels, which rebuild t2 from the table per line:
// fewer than one line in three shares its window start with its predecessor.)
template <int PTS, bool TBL>
struct GridCache
{
    int s;
    wn_t t2[PTS];
};

#ifndef GRT_TBL_CACHE
#define GRT_TBL_CACHE 1
#endif

#if !GRT_TBL_CACHE
template <int PTS>
struct GridCache<PTS, true>
{
};
#endif

template <int PTS, bool TBL>
__device__ __forceinline__ void reset_cache(GridCache<PTS, TBL> &gc)
{
    gc.s = INT32_MIN;
#pragma unroll
    for (int j=0; j<PTS; ++j) gc.t2[j] = (wn_t)0;
}

#if !GRT_TBL_CACHE
template <int PTS>
__device__ __forceinline__ void reset_cache(GridCache<PTS, true> &)
{
}
#endif

// (float)((t2 - v)*REPWID): the last two double roundings and the float rounding of
// RFM_voigt.c:165.
__device__ __forceinline__ float abscissa_from(wn_t const t2, wn_t const v, wn_t const repwid)
{
#ifdef GRT_SINGLE
    return __fmul_rn(__fsub_rn(t2, v), repwid);
#elif GRT_DBG_NOF2F     // timing experiment only (wrong results): no FP64->FP32 conversion
    return __int_as_float(__double2hiint(__dmul_rn(__dsub_rn(t2, v), repwid)));
#else
    return __double2float_rn(__dmul_rn(__dsub_rn(t2, v), repwid));
#endif
}

// What the warp knows about its work item while it walks the lines.
struct TileCtx
{
    int warp_base;
    int warp_last;
    int lane;                  // lane in the warp (owner of record `lane` of a staged batch)
    int sub;                   // lane in its line group: owns points warp_base + j*LN + sub
    unsigned gmask;            // records of a staged batch that the lane's group walks
    int fsteps;
    int nw;
    wn_t w0;
    wn_t wres;
    double lane_pt;            // (double)(warp_base + sub)
    uint32_t iw_lane;          // shared-memory address of iw[warp_base + sub] (TBL kernels)
};

// (double)k for 0 <= k < 2^31 on the FP64 pipe instead of the conversion unit.
__device__ __forceinline__ double int_to_double(int const k)
{
    return __dsub_rn(__hiloint2double(0x43300000, k), 4503599627370496.0);
}

__device__ __forceinline__ float lds_f32(uint32_t const addr)
{
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
    return v;
}

__device__ __forceinline__ double lds_f64(uint32_t const addr)
{
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr));
    return v;
}

__device__ __forceinline__ double2 lds_f64x2(uint32_t const addr)
{
    double2 v;
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "r"(addr));
    return v;
}

__device__ __forceinline__ float4 lds_f32x4(uint32_t const addr)
{
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
    return v;
}

// The grid wavenumbers of the lane's points as one line sees them (they depend on the
// line through its window start s only).  TBL: fl(i*wres) comes from the block's
// shared-memory table iw[] (i = f - s is a small integer: 0 <= i <= 2*fsteps inside a
// window), one LDS + one DADD per point.  Otherwise three FP64 operations per point,
// cached in registers across lines that share s.
template <int PTS, bool TBL, int LN>
struct LineGrid;

#if GRT_TBL_CACHE
template <int PTS, int LN>
struct LineGrid<PTS, true, LN>
{
    GridCache<PTS, true> &gc;
    __device__ __forceinline__ LineGrid(GridCache<PTS, true> &c, TileCtx const &t, int const fc)
        : gc(c)
    {
        if (fc != gc.s)
        {
            gc.s = fc;
            int const s = max(fc - t.fsteps, 0);
#ifdef GRT_SINGLE
            float const dwno = wn_add(wn_index_times(s, t.wres), t.w0);
            uint32_t const a = t.iw_lane - 4u*(uint32_t)s;
#pragma unroll
            for (int j=0; j<PTS; ++j) gc.t2[j] = wn_add(dwno, lds_f32(a + (4u*LN)*(uint32_t)j));
#else
            double const dwno = __dadd_rn(__dmul_rn(int_to_double(s), t.wres), t.w0);
            uint32_t const a = t.iw_lane - 8u*(uint32_t)s;
#pragma unroll
            for (int j=0; j<PTS; ++j) gc.t2[j] = __dadd_rn(dwno, lds_f64(a + (8u*LN)*(uint32_t)j));
#endif
        }
    }
    __device__ __forceinline__ wn_t t2(int const j) const
    {
        return gc.t2[j];
    }
};
#else
template <int PTS, int LN>
struct LineGrid<PTS, true, LN>
{
    wn_t dwno;
    uint32_t a;
    __device__ __forceinline__ LineGrid(GridCache<PTS, true> &, TileCtx const &t, int const fc)
    {
        int const s = max(fc - t.fsteps, 0);
        dwno = wn_add(wn_index_times(s, t.wres), t.w0);
        a = t.iw_lane - (uint32_t)sizeof(wn_t)*(uint32_t)s;
    }
    __device__ __forceinline__ wn_t t2(int const j) const
    {
#ifdef GRT_SINGLE
        return wn_add(dwno, lds_f32(a + (4u*LN)*(uint32_t)j));
#else
        return __dadd_rn(dwno, lds_f64(a + (8u*LN)*(uint32_t)j));
#endif
    }
};
#endif

template <int PTS, int LN>
struct LineGrid<PTS, false, LN>
{
    GridCache<PTS, false> &gc;
    __device__ __forceinline__ LineGrid(GridCache<PTS, false> &c, TileCtx const &t, int const fc)
        : gc(c)
    {
        if (fc != gc.s)
        {
            gc.s = fc;
            int const s = max(fc - t.fsteps, 0);
#ifdef GRT_SINGLE
            float const dwno = wn_add(wn_index_times(s, t.wres), t.w0);
            int const i0 = t.warp_base + t.sub - s;
#pragma unroll
            for (int j=0; j<PTS; ++j)
            {
                // out-of-window points (i < 0 on edge tiles) are predicated off by the caller
                gc.t2[j] = wn_add(dwno, wn_index_times(i0 + LN*j, t.wres));
            }
#else
            double const sd = int_to_double(s);
            double const dwno = __dadd_rn(__dmul_rn(sd, t.wres), t.w0);
            double const i0 = __dsub_rn(t.lane_pt, sd);     // exact
#pragma unroll
            for (int j=0; j<PTS; ++j)
            {
                gc.t2[j] = __dadd_rn(dwno, __dmul_rn(i0 + (double)(LN*j), t.wres));
            }
#endif
        }
    }
    __device__ __forceinline__ wn_t t2(int const j) const
    {
        return gc.t2[j];
    }
};

// How many of every four point pairs of the fast class take the XU-free conversion (the XU
// pipe and the issue slots are traded against each other; measured).
#ifndef GRT_F2F_MIX
#define GRT_F2F_MIX 0
#endif

// Line classes of the gather, decided once per (line, tile) by the lane that owns the
// line in the staged batch.
enum LineClass { kClsNone = 0, kClsFast, kClsFarEdge, kClsR1, kClsNear, kClsLorentzFull,
                 kClsLorentzEdge };

__device__ __forceinline__ f32x2 abscissa2(wn_t const t2a, wn_t const t2b, wn_t const v,
                                           wn_t const repwid)
{
    return pack2(abscissa_from(t2a, v, repwid), abscissa_from(t2b, v, repwid));
}

// a/b for two samples: the exact IEEE quotient in the double build (the reference's float
// division, bit for bit); in the single build, whose tolerance on tau is 1e-4, the reciprocal
// unit's 2^-22.
__device__ __forceinline__ f32x2 rcp2_mufu(f32x2 const q)
{
    float qx, qy, rx, ry;
    unpack2(q, qx, qy);
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rx) : "f"(qx));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(ry) : "f"(qy));
    return pack2(rx, ry);
}

__device__ __forceinline__ f32x2 div2(f32x2 const a, f32x2 const b)
{
#ifdef GRT_SINGLE
    return mul2(a, rcp2_mufu(b));
#else
    return div_exact2(a, b);
#endif
}

// Far wing K = YRRTPI/(XQ + YQ) (RFM_voigt.c:168-171) for two samples.
__device__ __forceinline__ f32x2 far_wing2(f32x2 const xq, f32x2 const yq2, f32x2 const yr2)
{
    return div2(yr2, add2(xq, yq2));
}

// Humlicek region 1 (RFM_voigt.c:172-183) for two samples, per-line constants hoisted.
__device__ __forceinline__ f32x2 region1_2(f32x2 const xq, f32x2 const y2, f32x2 const a0,
                                           f32x2 const d0, f32x2 const d2)
{
    f32x2 const den = add2(d0, mul2_unfused(xq, add2(d2, xq)));
    f32x2 const D = div2(pack2(kRsqrpi, kRsqrpi), den);
    return mul2(mul2(D, y2), add2(a0, xq));
}

// The sums a line class adds its samples to.  Double build: the tile's double accumulators,
// one DFMA per sample.  Single build: packed float partial sums of the staged batch (<= 32
// lines), flushed into the double accumulators by process_batch -- no FP64 and no conversion
// per sample in any class.
#ifdef GRT_SINGLE
template <int PTS> struct BatchAcc { typedef f32x2 type[PTS/2]; };
typedef float coef_t;
__device__ __forceinline__ coef_t rec_coef(RecA const a) { return a.y; }
#define GRT_ACC_PAIR(acc, j) acc[(j)/2]

__device__ __forceinline__ void add_pair(f32x2 &a, float const c, f32x2 const k)
{
    a = fma2(pack2(c, c), k, a);
}

__device__ __forceinline__ void add_pair_if(f32x2 &a, float const c, float const k0,
                                            float const k1, bool const p0, bool const p1)
{
    // a predicated-off sample may be any bit pattern (abscissae beyond a window end)
    a = fma2(pack2(c, c), pack2(p0 ? k0 : 0.f, p1 ? k1 : 0.f), a);
}
#else
template <int PTS> struct BatchAcc { typedef double type[PTS]; };
typedef double coef_t;
__device__ __forceinline__ coef_t rec_coef(RecA const a) { return a.y; }
#define GRT_ACC_PAIR(acc, j) acc[j], acc[(j) + 1]

__device__ __forceinline__ void add_pair(double &a0, double &a1, double const c, f32x2 const k)
{
    float k0, k1;
    unpack2(k, k0, k1);
    a0 = fma(c, widen_positive(k0), a0);
    a1 = fma(c, widen_positive(k1), a1);
}

__device__ __forceinline__ void add_pair_if(double &a0, double &a1, double const c,
                                            float const k0, float const k1, bool const p0,
                                            bool const p1)
{
    if (p0) a0 = fma(c, widen_positive(k0), a0);
    if (p1) a1 = fma(c, widen_positive(k1), a1);
}
#endif

#ifndef GRT_SEED_MIX
#define GRT_SEED_MIX 0
#endif

// Near-class visits whose window holds the whole tile skip the per-point window predicates
// (a second copy of the near body: config 2 525.9 -> 523.5 ms, profiles/r02q_ab_near_covered.txt).
#ifndef GRT_NEAR_COVERED
#define GRT_NEAR_COVERED 1
#endif

// Far-edge records share the fast class's pipelined loop (see process_batch).
#ifndef GRT_MERGE_EDGE
#define GRT_MERGE_EDGE 0
#endif

// kClsFast: the tile lies wholly inside the line's window and wholly in its far wing:
// no predicates, PTS independent dependency chains in flight, two samples per packed op.
#ifdef GRT_SINGLE
// 1/q for q >= 1 (finite) without the reciprocal unit: integer seed (12.5 % off), three
// Newton steps in packed FMAs -> 2^-22 relative.  The single build only needs K to ~1e-6
// (the tolerance on tau is 1e-4 and the reference's own float sums are no better), so the
// fast class alternates between this and MUFU.RCP to balance the XU pipe against the issue
// slots (GRT_SP_SEED_PAIRS of every four point pairs take this form).
__device__ __forceinline__ f32x2 rcp2_seeded(f32x2 const q)
{
    float qx, qy;
    unpack2(q, qx, qy);
    f32x2 r = pack2(__uint_as_float(0x7EF311C7u - __float_as_uint(qx)),
                    __uint_as_float(0x7EF311C7u - __float_as_uint(qy)));
    f32x2 const nq = pack2(-qx, -qy);
    f32x2 const one = pack2(1.0f, 1.0f);
    f32x2 e = fma2(nq, r, one);
    r = fma2(r, e, r);
    e = fma2(nq, r, one);
    r = fma2(r, e, r);
    e = fma2(nq, r, one);
    return fma2(r, e, r);
}

#ifndef GRT_SP_SEED_PAIRS
#define GRT_SP_SEED_PAIRS 2
#endif

// kClsFast, single build: float abscissa chain (bit-identical to the reference's, see wn_t),
// K*c = (c*Y/sqrt(pi))/(x*x + Y*Y) to float accuracy, float partial sums per staged batch
// (flushed into the double accumulators by process_batch): no FP64 and no conversion in the
// loop -- 4 packed float instructions + 2 reciprocals per two samples.
template <int PTS, bool TBL, int LN>
__device__ __forceinline__ void line_fast(RecA const A, float4 const B, TileCtx const &t,
                                          GridCache<PTS, TBL> &gc, f32x2 (&facc)[PTS/2])
{
    LineGrid<PTS, TBL, LN> const grid(gc, t, __float_as_int(B.w));
    f32x2 const nv2 = pack2(-A.x, -A.x);
    f32x2 const rep2 = pack2(B.x, B.x);
    f32x2 const yq2 = pack2(B.y, B.y);
    f32x2 const cy2 = pack2(A.z, A.z);
#pragma unroll
    for (int j=0; j<PTS; j+=2)
    {
        // fl(fl(t2 - v)*REPWID): a packed add feeding a packed multiply is not contracted
        f32x2 const xi = mul2(add2(pack2(grid.t2(j), grid.t2(j + 1)), nv2), rep2);
        f32x2 const q = fma2(xi, xi, yq2);
        f32x2 const r = ((j/2) % 4 < GRT_SP_SEED_PAIRS) ? rcp2_seeded(q) : rcp2_mufu(q);
        facc[j/2] = fma2(cy2, r, facc[j/2]);
    }
}
#else
template <int PTS, bool TBL, int LN>
__device__ __forceinline__ void line_fast(RecA const A, float4 const B, TileCtx const &t,
                                          GridCache<PTS, TBL> &gc, double (&acc)[PTS])
{
    LineGrid<PTS, TBL, LN> const grid(gc, t, __float_as_int(B.w));
    double const repwid = widen_positive(B.x);
    f32x2 const yq2 = pack2(B.y, B.y);
    f32x2 const yr2 = pack2(B.z, B.z);
#pragma unroll
    for (int j=0; j<PTS; j+=2)
    {
        f32x2 const xi = ((j/2) % 4 < GRT_F2F_MIX)
            ? pack2(abs_abscissa_alu(grid.t2(j), A.x, repwid), abs_abscissa_alu(grid.t2(j + 1), A.x, repwid))
            : abscissa2(grid.t2(j), grid.t2(j + 1), A.x, repwid);
        f32x2 const xq = mul2_unfused(xi, xi);
        // GRT_SEED_MIX of every four pairs take the XU-free division (voigt.cuh)
        f32x2 const k = ((j/2) % 4 < GRT_SEED_MIX) ? div_exact2_seeded(yr2, add2(xq, yq2))
                                                   : far_wing2(xq, yq2, yr2);
        add_pair(acc[j], acc[j+1], A.y, k);
    }
}

#endif

// kClsR1: the tile lies wholly inside the window and every point has
// XLIM1 <= |x| < XLIM0 (wide Doppler cores of the shortwave): region 1 only.
template <int PTS, bool TBL, int LN>
__device__ __forceinline__ void line_r1(RecA const A, float4 const B, float4 const C,
                                        float4 const D, TileCtx const &t,
                                        GridCache<PTS, TBL> &gc,
                                        typename BatchAcc<PTS>::type &acc)
{
    // C = hw0, y, xlim0, xlim1;  D = a0, d0, d2, -
    LineGrid<PTS, TBL, LN> const grid(gc, t, __float_as_int(B.w));
    wn_t const repwid = rec_repwid(B);
    coef_t const cA = rec_coef(A);
    f32x2 const y2 = pack2(C.y, C.y);
    f32x2 const a0 = pack2(D.x, D.x);
    f32x2 const d0 = pack2(D.y, D.y);
    f32x2 const d2 = pack2(D.z, D.z);
#pragma unroll
    for (int j=0; j<PTS; j+=2)
    {
        f32x2 const xi = abscissa2(grid.t2(j), grid.t2(j + 1), rec_v(A), repwid);
        add_pair(GRT_ACC_PAIR(acc, j), cA, region1_2(mul2_unfused(xi, xi), y2, a0, d0, d2));
    }
}

// Pure-Lorentz lines (Y >= 70.55, RFM_voigt.c:97-106).  FULL: the window covers the tile.
template <int PTS, bool TBL, int LN, bool FULL>
__device__ __forceinline__ void line_lorentz(RecA const A, float4 const B,
                                             TileCtx const &t, GridCache<PTS, TBL> &gc,
                                             typename BatchAcc<PTS>::type &acc)
{
    int const fc = __float_as_int(B.w);
    int const s = max(fc - t.fsteps, 0);
    int const e = min(fc + t.fsteps, t.nw - 1);
    LineGrid<PTS, TBL, LN> const grid(gc, t, fc);
    wn_t const repwid = rec_repwid(B);
    int const f0 = t.warp_base + t.sub;
#ifdef GRT_SINGLE
    float const cA = rec_coef(A);
    f32x2 const yq2 = pack2(B.y, B.y);
#pragma unroll
    for (int j=0; j<PTS; j+=2)
    {
        int const fa = f0 + LN*j;
        int const fb = fa + LN;
        f32x2 const xi = abscissa2(grid.t2(j), grid.t2(j + 1), rec_v(A), repwid);
        float k0, k1;
        unpack2(rcp2_mufu(fma2(xi, xi, yq2)), k0, k1);
        add_pair_if(acc[j/2], cA, k0, k1, FULL || ((fa >= s) && (fa <= e)),
                    FULL || ((fb >= s) && (fb <= e)));
    }
#else
    double const cA = rec_c(A);
#pragma unroll
    for (int j=0; j<PTS; ++j)
    {
        int const f = f0 + LN*j;
        float const xi = abscissa_from(grid.t2(j), rec_v(A), repwid);
        double const k = fast_rcp_of_float(xi*xi + B.y);
        if (FULL || ((f >= s) && (f <= e))) acc[j] = fma(cA, k, acc[j]);
    }
#endif
}


// kClsFarEdge: far wing only, but a window end falls inside the tile (kernels.c:435-437):
// the far-wing arithmetic with a per-point window predicate.
template <int PTS, bool TBL, int LN>
__device__ __forceinline__ void line_far_edge(RecA const A, float4 const B, TileCtx const &t,
                                              GridCache<PTS, TBL> &gc,
                                              typename BatchAcc<PTS>::type &acc)
{
    int const fc = __float_as_int(B.w);
    int const s = max(fc - t.fsteps, 0);
    int const e = min(fc + t.fsteps, t.nw - 1);
    LineGrid<PTS, TBL, LN> const grid(gc, t, fc);
    wn_t const repwid = rec_repwid(B);
    coef_t const cA = rec_coef(A);
    f32x2 const yq2 = pack2(B.y, B.y);
    f32x2 const yr2 = pack2(B.z, B.z);
    int const f0 = t.warp_base + t.sub;
#pragma unroll
    for (int j=0; j<PTS; j+=2)
    {
        // a pair of chunks wholly beyond the window end: nothing to do (warp-uniform)
        int const cb = t.warp_base + LN*j;
        if (cb + 2*LN - 1 < s || cb > e) continue;
        f32x2 const xi = abscissa2(grid.t2(j), grid.t2(j + 1), rec_v(A), repwid);
        f32x2 const k = far_wing2(mul2_unfused(xi, xi), yq2, yr2);
        float k0, k1;
        unpack2(k, k0, k1);
        int const fa = f0 + LN*j;
        int const fb = fa + LN;
        add_pair_if(GRT_ACC_PAIR(acc, j), cA, k0, k1, (fa >= s) && (fa <= e),
                    (fb >= s) && (fb <= e));
    }
}

// kClsNear: the tile overlaps the line's |x| < XLIM0 neighbourhood, so its points may be
// far-wing, Humlicek region-1 (RFM_voigt.c:172-183) or line-core samples (the latter are
// skipped here and left to core_sweep_kernel); the per-lane choice uses the reference's
// own float comparisons (RFM_voigt.c:166,172).
// Single-pair tiles (PTS = 2): far wing for both points, region 1 on top if the pair
// reaches inside XLIM0 (warp-uniform test on the record's conservative half-width).
template <int PTS, bool TBL, int LN>
__device__ __forceinline__ void line_near_single(RecA const A, float4 const B, float4 const C,
                                          float4 const D, TileCtx const &t,
                                          GridCache<PTS, TBL> &gc,
                                          typename BatchAcc<PTS>::type &acc)
{
    // C = hw0, y, xlim0, xlim1;  D = a0, d0, d2, hw1_hw0in
    int const fc = __float_as_int(B.w);
    int const hw0 = __float_as_int(C.x);
    int const s = max(fc - t.fsteps, 0);
    int const e = min(fc + t.fsteps, t.nw - 1);
    LineGrid<PTS, TBL, LN> const grid(gc, t, fc);
    wn_t const repwid = rec_repwid(B);
    coef_t const cA = rec_coef(A);
    f32x2 const yq2 = pack2(B.y, B.y);
    f32x2 const yr2 = pack2(B.z, B.z);
    int const f0 = t.warp_base + t.sub;
    f32x2 xi[PTS/2], xq[PTS/2], k[PTS/2];
#pragma unroll
    for (int j=0; j<PTS; j+=2)
    {
        xi[j/2] = abscissa2(grid.t2(j), grid.t2(j + 1), rec_v(A), repwid);
        xq[j/2] = mul2_unfused(xi[j/2], xi[j/2]);
        k[j/2] = far_wing2(xq[j/2], yq2, yr2);
    }
    f32x2 const y2 = pack2(C.y, C.y);
    f32x2 const a0 = pack2(D.x, D.x);
    f32x2 const d0 = pack2(D.y, D.y);
    f32x2 const d2 = pack2(D.z, D.z);
#pragma unroll
    for (int j=0; j<PTS; j+=2)
    {
        int const cb = t.warp_base + LN*j;            // the pair covers [cb, cb + 2*LN - 1]
        int const dlo = max(max(cb - fc, fc - (cb + 2*LN - 1)), 0);
        int const fa = f0 + LN*j;
        int const fb = fa + LN;
        bool ina = (fa >= s) && (fa <= e);
        bool inb = (fb >= s) && (fb <= e);
        float k0, k1;
        unpack2(k[j/2], k0, k1);
        if (dlo <= hw0)
        {
            float x0, x1, r0, r1;
            unpack2(xi[j/2], x0, x1);
            unpack2(region1_2(xq[j/2], y2, a0, d0, d2), r0, r1);
            float const ax0 = fabsf(x0);
            float const ax1 = fabsf(x1);
            bool const far0 = ax0 >= C.z;
            bool const far1 = ax1 >= C.z;
            k0 = far0 ? k0 : r0;
            k1 = far1 ? k1 : r1;
            ina = ina && (far0 || (ax0 >= C.w));
            inb = inb && (far1 || (ax1 >= C.w));
        }
        add_pair_if(GRT_ACC_PAIR(acc, j), cA, k0, k1, ina, inb);
    }
}

// Wider tiles: the abscissae of all PTS points are formed first, as one straight-line block
// with every dependency chain in flight; each pair of 32-point chunks is then dispatched
// warp-uniformly to far wing / region 1 / both (measured: -2.4 % on the config-2 step
// against forming them inside the branches).
template <int PTS, bool TBL, int LN, bool COVERED>
__device__ __forceinline__ void line_near_pairs(RecA const A, float4 const B, float4 const C,
                                          float4 const D, TileCtx const &t,
                                          GridCache<PTS, TBL> &gc,
                                          typename BatchAcc<PTS>::type &acc)
{
    int const fc = __float_as_int(B.w);
    int const hw0 = __float_as_int(C.x);
    int const hw0in = (int)(__float_as_uint(D.w) >> 16);
    int const s = max(fc - t.fsteps, 0);
    int const e = min(fc + t.fsteps, t.nw - 1);
    LineGrid<PTS, TBL, LN> const grid(gc, t, fc);
    wn_t const repwid = rec_repwid(B);
    coef_t const cA = rec_coef(A);
    f32x2 const yq2 = pack2(B.y, B.y);
    f32x2 const yr2 = pack2(B.z, B.z);
    f32x2 const y2 = pack2(C.y, C.y);
    f32x2 const a0 = pack2(D.x, D.x);
    f32x2 const d0 = pack2(D.y, D.y);
    f32x2 const d2 = pack2(D.z, D.z);
    int const f0 = t.warp_base + t.sub;
    f32x2 xi[PTS/2], xq[PTS/2];
#pragma unroll
    for (int j=0; j<PTS; j+=2)
    {
        xi[j/2] = abscissa2(grid.t2(j), grid.t2(j + 1), rec_v(A), repwid);
        xq[j/2] = mul2_unfused(xi[j/2], xi[j/2]);
    }
#pragma unroll
    for (int j=0; j<PTS; j+=2)
    {
        int const cb = t.warp_base + LN*j;            // the pair covers [cb, cb + 2*LN - 1]
        int const dlo = max(max(cb - fc, fc - (cb + 2*LN - 1)), 0);
        int const dhi = max(fc - cb, cb + 2*LN - 1 - fc);
        int const fa = f0 + LN*j;
        int const fb = fa + LN;
        // COVERED: the window holds the whole tile (the usual case away from window ends)
        bool const ina = COVERED || ((fa >= s) && (fa <= e));
        bool const inb = COVERED || ((fb >= s) && (fb <= e));
        float x0, x1;
        unpack2(xi[j/2], x0, x1);
        if (dlo > hw0)
        {
            float k0, k1;
            unpack2(far_wing2(xq[j/2], yq2, yr2), k0, k1);
            add_pair_if(GRT_ACC_PAIR(acc, j), cA, k0, k1, ina, inb);
        }
        else if (dhi < hw0in)
        {
            float k0, k1;
            unpack2(region1_2(xq[j/2], y2, a0, d0, d2), k0, k1);
            add_pair_if(GRT_ACC_PAIR(acc, j), cA, k0, k1, ina && (fabsf(x0) >= C.w),
                        inb && (fabsf(x1) >= C.w));
        }
        else
        {
            float kf0, kf1, kr0, kr1;
            unpack2(far_wing2(xq[j/2], yq2, yr2), kf0, kf1);
            unpack2(region1_2(xq[j/2], y2, a0, d0, d2), kr0, kr1);
            float const ax0 = fabsf(x0);
            bool const isfar0 = ax0 >= C.z;
            bool const isr10 = !isfar0 && (ax0 >= C.w);
            float const ax1 = fabsf(x1);
            bool const isfar1 = ax1 >= C.z;
            bool const isr11 = !isfar1 && (ax1 >= C.w);
            add_pair_if(GRT_ACC_PAIR(acc, j), cA, isr10 ? kr0 : kf0, isr11 ? kr1 : kf1,
                        ina && (isfar0 || isr10), inb && (isfar1 || isr11));
        }
    }
}

template <int PTS, bool TBL, int LN>
__device__ __forceinline__ void line_near(RecA const A, float4 const B, float4 const C,
                                          float4 const D, TileCtx const &t,
                                          GridCache<PTS, TBL> &gc,
                                          typename BatchAcc<PTS>::type &acc)
{
    if (PTS <= 2)
    {
        line_near_single<PTS, TBL, LN>(A, B, C, D, t, gc, acc);
        return;
    }
#if GRT_NEAR_COVERED
    int const fc = __float_as_int(B.w);
    if (fc - t.fsteps <= t.warp_base && fc + t.fsteps >= t.warp_last && fc + t.fsteps <= t.nw - 1)
    {
        line_near_pairs<PTS, TBL, LN, true>(A, B, C, D, t, gc, acc);
        return;
    }
#endif
    line_near_pairs<PTS, TBL, LN, false>(A, B, C, D, t, gc, acc);
}


__device__ __forceinline__ RecA lds_rec_a(uint32_t const addr)
{
#ifdef GRT_SINGLE
    return lds_f32x4(addr);
#else
    return lds_f64x2(addr);
#endif
}

// The four 16-byte vectors of a staged record.
template <int NV>
struct RecVec
{
    RecA A;            // v, c (single build: v, c, c*Y/sqrt(pi))
    float4 B;          // repwid, yq, yrrtpi, fc
    float4 C;          // hw0, y, xlim0, xlim1
    float4 D;          // a0, d0, d2, hw1_hw0in
    __device__ __forceinline__ void load(uint32_t const addr)
    {
        A = lds_rec_a(addr);
        B = lds_f32x4(addr + 16u);
        if (NV > 2)
        {
            C = lds_f32x4(addr + 32u);
            D = lds_f32x4(addr + 48u);
        }
    }
};

// Walk the records of one class (bit mask m over the staged batch at shared address sa).
// (Loading the next record while the current one computes, as the fast class does, was
// measured slower here: the extra live registers cost more than the load latency.)
template <int NV, class Body>
__device__ __forceinline__ void walk_class(unsigned m, uint32_t const sa, Body &&body)
{
    while (m)
    {
        int const b = __ffs(m) - 1;
        m &= m - 1;
        RecVec<NV> r;
        r.load(sa + 64u*(uint32_t)b);
        body(r);
    }
}

// Development build only (-DGRT_DBG_TIMERS=1): SM cycles and line visits per line class,
// summed over warps, read back by grt_debug_timers() (tools/class_timers.py).
#ifndef GRT_DBG_TIMERS
#define GRT_DBG_TIMERS 0
#endif
#if GRT_DBG_TIMERS
enum { kTmClassify = 0, kTmFast, kTmR1, kTmFarEdge, kTmNear, kTmLorFull, kTmLorEdge, kTmWait,
       kTmEpilogue, kTmItem, kTmCount };
__device__ unsigned long long g_dbg_cycles[kTmCount];
__device__ unsigned long long g_dbg_visits[kTmCount];
#define GRT_TM_BEGIN() long long tm_t0 = clock64()
#define GRT_TM_END(slot, visits)                                                          \
    do {                                                                                  \
        long long const tm_t1 = clock64();                                                \
        if ((threadIdx.x & 31) == 0)                                                      \
        {                                                                                 \
            atomicAdd(&g_dbg_cycles[slot], (unsigned long long)(tm_t1 - tm_t0));          \
            atomicAdd(&g_dbg_visits[slot], (unsigned long long)(visits));                 \
        }                                                                                 \
        tm_t0 = clock64();                                                                \
    } while (0)
#else
#define GRT_TM_BEGIN() do {} while (0)
#define GRT_TM_END(slot, visits) do {} while (0)
#endif

// The class of one record against one tile [warp_base, warp_last] (see LineClass).
__device__ __forceinline__ int classify_line(int const fc, int const hw0, unsigned const hh,
                                             int const warp_base, int const warp_last,
                                             int const fsteps, int const nw)
{
    bool const hit = (fc != kSkip) && (fc + fsteps >= warp_base) &&
                     (fc - fsteps <= warp_last) && (hw0 != kHwCoreOnly);
    if (!hit) return kClsNone;
    int const s = max(fc - fsteps, 0);
    int const e = min(fc + fsteps, nw - 1);
    bool const covered = (s <= warp_base) && (e >= warp_last);
    // nearest and farthest tile point from the line centre, in grid steps
    int const dlo = max(max(warp_base - fc, fc - warp_last), 0);
    int const dhi = max(fc - warp_base, warp_last - fc);
    bool const far = (hw0 >= 0) && (dlo > hw0);
    if (hw0 == kHwLorentz) return covered ? kClsLorentzFull : kClsLorentzEdge;
    if (far) return covered ? kClsFast : kClsFarEdge;
    int const hw1 = (int)(hh & 0xffffu);
    int const hw0in = (int)(hh >> 16);
    bool const r1 = covered && (dhi < hw0in) && (dlo > hw1) && (hw1 < 65535);
    return r1 ? kClsR1 : kClsNear;
}

// One staged batch (cnt <= 32 records of one layer) against the warp's tile.  Lane b
// classifies record b; the classes are then walked as bit masks, the dominant one
// (kClsFast) with the next record's vectors prefetched while the current one computes.
template <int PTS, bool TBL, int LN>
__device__ __forceinline__ void process_batch(LineRec const *stage, int const cnt,
                                              TileCtx const &t, GridCache<PTS, TBL> &gc,
                                              double (&acc)[PTS])
{
    GRT_TM_BEGIN();
    int cls = kClsNone;
    if (t.lane < cnt)
    {
        cls = classify_line(stage[t.lane].fc, stage[t.lane].hw0, stage[t.lane].hw1_hw0in,
                            t.warp_base, t.warp_last, t.fsteps, t.nw);
    }
    unsigned m = __ballot_sync(0xffffffffu, cls == kClsFast) & t.gmask;
#if GRT_MERGE_EDGE
    // Far-edge records ride in the software-pipelined loop of the fast class (the next
    // record's vectors are loaded while the current one computes) behind a warp-uniform
    // branch: on coarse grids a visit is only a few samples per lane, so the load latency a
    // plain class walk exposes per record is a large share of its cost.
    unsigned const medge = __ballot_sync(0xffffffffu, cls == kClsFarEdge) & t.gmask;
    m |= medge;
#define GRT_LINE_BODY(A, B, b_)                                                      \
    do {                                                                             \
        if ((medge >> (b_)) & 1u) line_far_edge<PTS, TBL, LN>(A, B, t, gc, GRT_FAST_ACC);     \
        else line_fast<PTS, TBL, LN>(A, B, t, gc, GRT_FAST_ACC);                     \
    } while (0)
#else
#define GRT_LINE_BODY(A, B, b_) line_fast<PTS, TBL, LN>(A, B, t, gc, GRT_FAST_ACC)
#endif
    GRT_TM_END(kTmClassify, 1);
#if GRT_DBG_TIMERS
    int tm_n = __popc(m);
#endif
    // the stage address is kept opaque so that it stays in a register
    uint32_t sa = smem_u32(stage);
    asm volatile("" : "+r"(sa));
#ifdef GRT_SINGLE
    // float partial sums of this batch, every class (see BatchAcc)
    f32x2 facc[PTS/2];
#pragma unroll
    for (int j=0; j<PTS/2; ++j) facc[j] = 0ull;
#define GRT_FAST_ACC facc
#else
#define GRT_FAST_ACC acc
#endif
    if (m)
    {
        int b = __ffs(m) - 1;
        m &= m - 1;
        // two records in flight, in alternating register sets (no copies between them)
        RecA A0 = lds_rec_a(sa + 64u*(uint32_t)b);
        float4 B0 = lds_f32x4(sa + 64u*(uint32_t)b + 16u);
        int b0 = b;
        for (;;)
        {
            bool more = m != 0;
            b = more ? __ffs(m) - 1 : b;
            m &= m - 1;
            RecA const A1 = lds_rec_a(sa + 64u*(uint32_t)b);
            float4 const B1 = lds_f32x4(sa + 64u*(uint32_t)b + 16u);
            int const b1 = b;
            GRT_LINE_BODY(A0, B0, b0);
            if (!more) break;
            more = m != 0;
            b = more ? __ffs(m) - 1 : b;
            m &= m - 1;
            A0 = lds_rec_a(sa + 64u*(uint32_t)b);
            B0 = lds_f32x4(sa + 64u*(uint32_t)b + 16u);
            b0 = b;
            GRT_LINE_BODY(A1, B1, b1);
            if (!more) break;
        }
#undef GRT_LINE_BODY
    }
    GRT_TM_END(kTmFast, tm_n);
    m = __ballot_sync(0xffffffffu, cls == kClsR1) & t.gmask;
#if GRT_DBG_TIMERS
    tm_n = __popc(m);
#endif
    walk_class<4>(m, sa, [&](RecVec<4> const &r) { line_r1<PTS, TBL, LN>(r.A, r.B, r.C, r.D, t, gc, GRT_FAST_ACC); });
    GRT_TM_END(kTmR1, tm_n);
#if !GRT_MERGE_EDGE
    m = __ballot_sync(0xffffffffu, cls == kClsFarEdge) & t.gmask;
#if GRT_DBG_TIMERS
    tm_n = __popc(m);
#endif
    walk_class<2>(m, sa, [&](RecVec<2> const &r) { line_far_edge<PTS, TBL, LN>(r.A, r.B, t, gc, GRT_FAST_ACC); });
    GRT_TM_END(kTmFarEdge, tm_n);
#endif
    m = __ballot_sync(0xffffffffu, cls == kClsNear) & t.gmask;
#if GRT_DBG_TIMERS
    tm_n = __popc(m);
#endif
    walk_class<4>(m, sa, [&](RecVec<4> const &r) { line_near<PTS, TBL, LN>(r.A, r.B, r.C, r.D, t, gc, GRT_FAST_ACC); });
    GRT_TM_END(kTmNear, tm_n);
    m = __ballot_sync(0xffffffffu, cls == kClsLorentzFull) & t.gmask;
#if GRT_DBG_TIMERS
    tm_n = __popc(m);
#endif
    walk_class<2>(m, sa, [&](RecVec<2> const &r) { line_lorentz<PTS, TBL, LN, true>(r.A, r.B, t, gc, GRT_FAST_ACC); });
    GRT_TM_END(kTmLorFull, tm_n);
    m = __ballot_sync(0xffffffffu, cls == kClsLorentzEdge) & t.gmask;
#if GRT_DBG_TIMERS
    tm_n = __popc(m);
#endif
    walk_class<2>(m, sa, [&](RecVec<2> const &r) { line_lorentz<PTS, TBL, LN, false>(r.A, r.B, t, gc, GRT_FAST_ACC); });
    GRT_TM_END(kTmLorEdge, tm_n);
#ifdef GRT_SINGLE
    // the batch's float partial sums (<= 32 lines) join the double accumulators
#pragma unroll
    for (int j=0; j<PTS; j+=2)
    {
        float f0, f1;
        unpack2(facc[j/2], f0, f1);
        acc[j] += (double)f0;
        acc[j+1] += (double)f1;
    }
#endif
#undef GRT_FAST_ACC
}


// Measurement knobs of the coarse-grid (grouped) kernels: stages of the record ring and
// resident blocks per SM the register budget is set for (0 = the general rule below).
#ifndef GRT_COARSE_STAGES
#define GRT_COARSE_STAGES 4
#endif
#ifndef GRT_COARSE_MINBLOCKS
#define GRT_COARSE_MINBLOCKS 0
#endif

template <int PTS, int NL, int LN = 32>
struct SweepConfig
{
    static constexpr int kStageBytes = NL*32*(int)sizeof(LineRec);
    // Stages of the warp's record ring.  Whole-warp lines on a fine grid spend thousands of
    // cycles on a batch and two stages hide the copy; the grouped kernels of the coarse grids
    // go through a batch in a fraction of that (narrow tiles, two or four lines at a time), so
    // they keep four batches in flight.
    static constexpr int kStages = (LN == 32) ? 2 : GRT_COARSE_STAGES;
    static constexpr int kRingBytes = kSweepWarps*kStages*kStageBytes + kSweepWarps*kStages*8;
    // registers: PTS*NL accumulators + PTS cached abscissae, all double
    static constexpr int kMinBlocks = (LN != 32 && PTS <= 4 && GRT_COARSE_MINBLOCKS > 0)
                                          ? GRT_COARSE_MINBLOCKS
                                          : ((PTS*NL <= 8) ? 3 : ((PTS*NL <= 16) ? 2 : 1));
};

// Entries of the fl(i*wres) table for a kernel with tiles of LN*PTS points: i = f - s runs
// over [0, 2*fsteps] inside a window; points of an edge tile beyond the window end read up
// to one tile further (real entries), points before the window start read below the table,
// into the record rings (any bit pattern: those samples are predicated off).
template <int PTS, int LN = 32>
__host__ __device__ constexpr int sweep_table_entries(int const fsteps)
{
    return 2*fsteps + LN*PTS + 1;
}


// Warp-granular persistent gather: every WARP pulls its own (32*PTS points, NL layers)
// work item from the global counter and streams the records of its lines through a
// warp-private shared-memory ring: no block-wide barrier after the prologue, so a warp
// never waits for a neighbour whose lines are denser.
//
// LN = lanes per line group.  LN = 32: the whole warp walks one line at a time over a tile
// of 32*PTS points.  LN = 16 / 8: the warp walks 2 / 4 lines at a time -- group g of LN
// lanes takes records [g*LN, (g+1)*LN) of each staged batch -- over a tile of LN*PTS points,
// every group holding partial sums of ALL the tile's points (summed over the groups with
// shuffles in the epilogue).  A line visit then covers PTS samples per lane on a tile 32/LN
// times narrower, which is what the 501-point windows of a 0.1 cm-1 grid need: the per-visit
// work (record loads, class walk, per-line constants) is amortised over 4-8 samples per lane
// without widening the tile beyond 64 points (wider tiles lose more to window ends).
template <int PTS, int NL, bool TBL, int LN>
__global__ void __launch_bounds__(kSweepThreads, SweepConfig<PTS, NL, LN>::kMinBlocks)
wing_sweep_kernel(SweepParams const p)
{
    static_assert(LN == 32 || LN == 16 || LN == 8, "lanes per line group");
    constexpr int kGroups = 32/LN;
    static_assert(PTS % kGroups == 0, "the groups share the tile's chunks in the epilogue");
    using Cfg = SweepConfig<PTS, NL, LN>;
    constexpr int STG = Cfg::kStages;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    int const lane = threadIdx.x & 31;
    int const warp = threadIdx.x >> 5;
    unsigned char *ring = smem_raw + (size_t)warp*(STG*Cfg::kStageBytes);
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem_raw + (size_t)kSweepWarps*STG*Cfg::kStageBytes) +
                     warp*STG;
    uint32_t const mb0 = smem_u32(bars);       // barrier of stage k: mb0 + 8k
    uint32_t const sb0 = smem_u32(ring);       // records of stage k: sb0 + k*kStageBytes
    if (lane == 0)
    {
#pragma unroll
        for (int k=0; k<STG; ++k) mbar_init(mb0 + 8u*(uint32_t)k, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    wn_t *iw = reinterpret_cast<wn_t *>(smem_raw + Cfg::kRingBytes);
    if (TBL)
    {
        // iw[k] = fl(k*wres): the product of RFM_voigt.c:165 for every sample index.
        int const entries = sweep_table_entries<PTS, LN>(p.fsteps);
        for (int k=threadIdx.x; k<entries; k+=kSweepThreads)
        {
            iw[k] = wn_index_times(k, (wn_t)p.wres);
        }
        __syncthreads();
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
        // tile-major: the layers of a tile follow each other, so its continuum / CFC / CIA
        // table values are read from DRAM once instead of once per layer
        int const group = p.tile_major ? (int)(item%(unsigned)groups) : (int)(item/(unsigned)p.num_tiles);
        int const slot = p.tile_major ? (int)(item/(unsigned)groups) : (int)(item%(unsigned)p.num_tiles);
        int const tile = p.tile_order[slot];
        unsigned const first = p.tile_first[tile];
        unsigned const last = p.tile_last[tile];
        int const warp_base = tile*(LN*PTS);
        int const layer0 = group*NL;
        int const nl = min(NL, L - layer0);
        int const nb = (int)((last - first + 31u)/32u);
#if GRT_DBG_TIMERS
        long long const ti0 = clock64();
#endif

        double acc[NL][PTS];
#pragma unroll
        for (int l=0; l<NL; ++l)
        {
#pragma unroll
            for (int j=0; j<PTS; ++j) acc[l][j] = 0.;
        }
        TileCtx tc;
        tc.warp_base = warp_base;
        tc.warp_last = warp_base + LN*PTS - 1;
        tc.lane = lane;
        tc.sub = lane & (LN - 1);
        tc.gmask = (LN == 32) ? 0xffffffffu : (((1u << (LN & 31)) - 1u) << ((lane/LN)*LN));
        tc.fsteps = p.fsteps;
        tc.nw = p.nw;
        tc.w0 = (wn_t)p.w0;
        tc.wres = (wn_t)p.wres;
        tc.lane_pt = (double)(warp_base + tc.sub);
        tc.iw_lane = smem_u32(iw) + (uint32_t)sizeof(wn_t)*(uint32_t)(warp_base + tc.sub);
        GridCache<PTS, TBL> gc;
        reset_cache(gc);

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
#pragma unroll
            for (int k=0; k<STG; ++k)
            {
                if (k < nb) issue(k, sb0 + (uint32_t)(k*Cfg::kStageBytes), mb0 + 8u*(uint32_t)k);
            }
        }
        for (int k=0; k<nb; ++k)
        {
            int const st = k & (STG - 1);
            uint32_t const mb = mb0 + 8u*(uint32_t)st;
#if GRT_DBG_TIMERS
            long long const tw0 = clock64();
#endif
            mbar_wait(mb, (phase >> st) & 1u);
#if GRT_DBG_TIMERS
            if (lane == 0)
            {
                atomicAdd(&g_dbg_cycles[kTmWait], (unsigned long long)(clock64() - tw0));
                atomicAdd(&g_dbg_visits[kTmWait], 1ull);
            }
#endif
            phase ^= 1u << st;
            int const cnt = (int)min(32u, last - (first + 32u*(unsigned)k));
            LineRec const *stage = reinterpret_cast<LineRec const *>(ring + st*Cfg::kStageBytes);
#pragma unroll
            for (int l=0; l<NL; ++l)
            {
                if (l < nl) process_batch<PTS, TBL, LN>(stage + l*32, cnt, tc, gc, acc[l]);
            }
            __syncwarp();
            if (lane == 0 && k + STG < nb)
            {
                // order the warp's generic-proxy reads before the async-proxy overwrite
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                issue(k + STG, sb0 + (uint32_t)(st*Cfg::kStageBytes), mb);
            }
        }

        // Epilogue: continua, CFCs, CIA (kernels.c:469-510, 585-630), one store.
#if GRT_DBG_TIMERS
        long long const te0 = clock64();
#endif
#pragma unroll
        for (int l=0; l<NL; ++l)
        {
            if (l >= nl) continue;
            int const layer = layer0 + l;
            if (kGroups > 1)
            {
                // every group holds partial sums of all the tile's points: sum them
                __syncwarp();
#pragma unroll
                for (int j=0; j<PTS; ++j)
                {
#pragma unroll
                    for (int off=LN; off<32; off<<=1)
                    {
                        acc[l][j] += __shfl_xor_sync(0xffffffffu, acc[l][j], off);
                    }
                }
            }
#pragma unroll
            for (int j=0; j<PTS/kGroups; ++j)
            {
                // chunk j*kGroups + g of LN points is stored by group g: 32 contiguous points
                int const f = warp_base + j*32 + lane;
                if (f >= p.nw) continue;
                double t = acc[l][j*kGroups];
#pragma unroll
                for (int gi=1; gi<kGroups; ++gi)
                {
                    if (lane/LN == gi) t = acc[l][j*kGroups + gi];
                }
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
                p.tau[(size_t)layer*p.nw + f] = (real_t)t;
            }
        }
#if GRT_DBG_TIMERS
        if (lane == 0)
        {
            long long const te1 = clock64();
            atomicAdd(&g_dbg_cycles[kTmEpilogue], (unsigned long long)(te1 - te0));
            atomicAdd(&g_dbg_visits[kTmEpilogue], 1ull);
            atomicAdd(&g_dbg_cycles[kTmItem], (unsigned long long)(te1 - ti0));
            atomicAdd(&g_dbg_visits[kTmItem], 1ull);
        }
#endif
    }
}


// Diagnostic (GRT_CLASS_HIST=1, tests): how many (line, tile) visits of the wing sweep fall
// into each line class, counted with the sweep's own classify_line() by one thread per
// record; the hot kernel itself carries no counters.
__global__ void __launch_bounds__(256)
class_histogram_kernel(SweepParams const p, int const tile_points, unsigned long long *counts)
{
    uint64_t const j = (uint64_t)blockIdx.x*blockDim.x + threadIdx.x;
    int const layer = blockIdx.y;
    unsigned long long mine[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (j < p.num_lines)
    {
        LineRec const *rp = p.recs + (size_t)layer*p.num_lines + j;
        int const fc = rp->fc;
        int const hw0 = rp->hw0;
        unsigned const hh = rp->hw1_hw0in;
        if (fc != kSkip)
        {
            int const t0 = max(fc - p.fsteps, 0)/tile_points;
            int const t1 = min(fc + p.fsteps, p.nw - 1)/tile_points;
            for (int t=t0; t<=t1; ++t)
            {
                int const cls = classify_line(fc, hw0, hh, t*tile_points,
                                              t*tile_points + tile_points - 1, p.fsteps, p.nw);
                mine[cls & 7] += 1ull;
            }
        }
    }
    for (int c=1; c<7; ++c)
    {
        unsigned long long v = mine[c];
        for (int o=16; o>0; o>>=1) v += __shfl_down_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0 && v) atomicAdd(&counts[c - 1], v);
    }
}


// Half-width of a record's |x| < xlim1 window in grid points, rounded up generously (the
// samples themselves are selected by the reference's float comparison).
__host__ __device__ __forceinline__ int core_half_width(float const repwid, float const xlim1,
                                                        bool const flagged, double const wres,
                                                        int const F)
{
    double const hd = flagged ? (double)(2*F + 2) : (double)xlim1/((double)repwid*wres) + 2.;
    return hd < (double)(2*F + 2) ? (int)hd : 2*F + 2;
}

// K3c: samples with |x| < XLIM1 (Humlicek regions 2, 3 and CPF12) of every Voigt-branch
// record, plus every sample of the rare records flagged kHwCoreOnly.  One thread per
// record walks outwards from the line centre (fc, fc-1, fc+1, ...): the records of a warp
// are neighbouring lines of one molecule (core_order), so at the same distance their |x| -- and
// with it the branch of RFM_voigt.c:172-277 they take -- mostly agree and the warp stays
// converged inside the expensive CPF12 sums.  Contributions are scattered with double
// atomics (a few per cent of all samples).
__global__ void __launch_bounds__(128)
core_sweep_kernel(SweepParams const p)
{
    uint64_t const tid = (uint64_t)blockIdx.x*blockDim.x + threadIdx.x;
    int const layer = blockIdx.y;
    if (tid >= p.num_lines) return;
    uint64_t const rec_id = p.core_order != nullptr ? p.core_order[tid] : tid;
    LineRec const *rp = p.recs + (size_t)layer*p.num_lines + rec_id;
    float4 const B = *reinterpret_cast<float4 const *>(&rp->repwid);   // repwid, yq, yrrtpi, fc
    float4 const C = *reinterpret_cast<float4 const *>(&rp->hw0);      // hw0, y, xlim0, xlim1
    int const fc = __float_as_int(B.w);
    int const hw0 = __float_as_int(C.x);
    if (fc == kSkip || hw0 == kHwLorentz) return;
    float const y = C.y;
    float const xlim0 = C.z;
    float const xlim1 = C.w;
    bool const flagged = hw0 == kHwCoreOnly;
    if (!flagged && xlim1 <= 0.f) return;
#ifdef GRT_SINGLE
    float4 const A = *reinterpret_cast<float4 const *>(&rp->v);
    wn_t const vline = A.x;
    double const cline = (double)A.y;
    wn_t const repwid = B.x;
#else
    double2 const vc = *reinterpret_cast<double2 const *>(&rp->v);
    wn_t const vline = vc.x;
    double const cline = vc.y;
    wn_t const repwid = (double)B.x;
#endif
    int const F = p.fsteps;
    int const s = max(fc - F, 0);
    int const e = min(fc + F, p.nw - 1);
    int const h = core_half_width(B.x, xlim1, flagged, p.wres, F);
    wn_t const wres = (wn_t)p.wres;
    wn_t const dwno = wn_add(wn_index_times(s, wres), (wn_t)p.w0);
    real_t *tau = p.tau + (size_t)layer*p.nw;
    VoigtLine line;
    voigt_line_init(line, y, xlim0, xlim1);
    for (int d=0; d<=h; ++d)
    {
#pragma unroll
        for (int side=0; side<2; ++side)
        {
            int const f = side ? fc + d : fc - d;
            if ((side == 1 && d == 0) || f < s || f > e) continue;
            float const xi = abscissa_from(wn_add(dwno, wn_index_times(f - s, wres)), vline, repwid);
            if (flagged || fabsf(xi) < xlim1)
            {
                double const k = voigt_line_eval(line, xi);
                atomicAdd(&tau[f], (real_t)(cline*k));
            }
        }
    }
}



// K3c as a gather (no atomics): a warp owns kCoreTile consecutive grid points of one layer,
// sums in a shared-memory tile every core sample that falls on them and adds the tile to tau
// with plain stores -- the records are sorted by centre, so the ones that can reach a tile are
// a contiguous range of the list (first_at[], one entry per 64 grid points, widened by the
// layer's largest core half-width and the largest pressure shift).  Results no longer depend
// on the order in which atomics happen to land: tau is bit-reproducible from run to run.
//   narrow records (fewer than 32 of their core points in the tile) are queued and walked 32
//     at a time, one record per lane, point by point from the low end of their windows; lanes
//     that meet on a grid point in the same step (neighbouring lines with the same nearest
//     grid point) are summed by a segmented warp scan before one of them stores;
//   wide records (Doppler cores of the shortwave at low pressure: up to a thousand points) are
//     walked by the whole warp, 32 consecutive points per step: every lane in the same Voigt
//     region, no conflicts, full lanes.
constexpr int kCoreTile = 512;
constexpr int kCoreWarps = 8;
constexpr int kCoreGranule = 64;

struct CoreGather
{
    int const *first_at;           // [num_granules + 1]: first record with v0 >= w0 + (64k - 1) wres
    int num_granules;
    int const *hmax;               // [layers]: largest core half-width of the layer's records, or -1
    int margin_pts;                // pressure shift + rounding slack, in grid points
    int num_tiles;
    unsigned *work_counter;
    short layer_order[GRT_MAX_LAYERS];   // layers by pressure, lowest (widest cores) first
};

struct CoreRec
{
    wn_t vline, repwid, dwno;
    double cline;
    float y, xlim0, xlim1;
    int s, a, b;                   // window start; first and last point of this tile to visit
    bool flagged;
};

// The record's core points inside [lo, hi]: false if there are none.
__device__ __forceinline__ bool core_rec_load(SweepParams const &p, LineRec const *rp,
                                              int const lo, int const hi, CoreRec &c)
{
    float4 const B = *reinterpret_cast<float4 const *>(&rp->repwid);   // repwid, yq, yrrtpi, fc
    float4 const C = *reinterpret_cast<float4 const *>(&rp->hw0);      // hw0, y, xlim0, xlim1
    int const fc = __float_as_int(B.w);
    int const hw0 = __float_as_int(C.x);
    if (fc == kSkip || hw0 == kHwLorentz) return false;
    c.flagged = hw0 == kHwCoreOnly;
    if (!c.flagged && C.w <= 0.f) return false;
    int const F = p.fsteps;
    int const h = core_half_width(B.x, C.w, c.flagged, p.wres, F);
    c.s = max(fc - F, 0);
    int const e = min(fc + F, p.nw - 1);
    c.a = max(max(fc - h, c.s), lo);
    c.b = min(min(fc + h, e), hi);
    if (c.a > c.b) return false;
    c.y = C.y;
    c.xlim0 = C.z;
    c.xlim1 = C.w;
#ifdef GRT_SINGLE
    float4 const A = *reinterpret_cast<float4 const *>(&rp->v);
    c.vline = A.x;
    c.cline = (double)A.y;
    c.repwid = B.x;
#else
    double2 const vc = *reinterpret_cast<double2 const *>(&rp->v);
    c.vline = vc.x;
    c.cline = vc.y;
    c.repwid = (double)B.x;
#endif
    c.dwno = wn_add(wn_index_times(c.s, (wn_t)p.wres), (wn_t)p.w0);
    return true;
}

__device__ __forceinline__ bool core_sample(CoreRec const &c, VoigtLine &line, wn_t const wres,
                                            int const f, double &val)
{
    float const xi = abscissa_from(wn_add(c.dwno, wn_index_times(f - c.s, wres)), c.vline, c.repwid);
    if (c.flagged || fabsf(xi) < c.xlim1)
    {
        val = c.cline*voigt_line_eval(line, xi);
        return true;
    }
    return false;
}

// One group of at most 32 queued narrow records, one per lane.
__device__ __forceinline__ void core_narrow_group(SweepParams const &p, LineRec const *layer_recs,
                                                  unsigned const *queue, int const n, int const lo,
                                                  int const hi, int const lane, double *acc)
{
    CoreRec c;
    c.a = 0;
    VoigtLine line;
    int len = 0;
    if (lane < n)
    {
        core_rec_load(p, layer_recs + queue[lane], lo, hi, c);
        voigt_line_init(line, c.y, c.xlim0, c.xlim1);
        len = c.b - c.a + 1;
    }
    int const maxlen = __reduce_max_sync(0xffffffffu, len);
    wn_t const wres = (wn_t)p.wres;
    for (int t=0; t<maxlen; ++t)
    {
        double val = 0.;
        bool active = false;
        int const f = c.a + t;
        if (t < len) active = core_sample(c, line, wres, f, val);
        int const key = active ? f : -1 - lane;
        // neighbouring lanes on the same grid point: segmented inclusive scan over each run
        // of adjacent lanes with one key (a run starts where the key changes)
        int const prev = __shfl_up_sync(0xffffffffu, key, 1);
        unsigned const heads = __ballot_sync(0xffffffffu, lane == 0 || prev != key);
        if (heads != 0xffffffffu)
        {
            int const start = 31 - __clz(heads & (0xffffffffu >> (31 - lane)));
#pragma unroll
            for (int off=1; off<32; off<<=1)
            {
                double const v2 = __shfl_up_sync(0xffffffffu, val, off);
                if (lane - off >= start) val += v2;
            }
        }
        int const next = __shfl_down_sync(0xffffffffu, key, 1);
        bool const writer = active && (lane == 31 || next != key);
        // runs of the same point that are not adjacent (lines of two molecules interleaved):
        // the writers take turns, lowest lane first
        unsigned const peers = __match_any_sync(0xffffffffu, writer ? key : -1 - lane);
        int const rank = __popc(peers & ((1u << lane) - 1u));
        int const rounds = __reduce_max_sync(0xffffffffu, writer ? __popc(peers) : 1);
        for (int r=0; r<rounds; ++r)
        {
            if (writer && rank == r) acc[key - lo] += val;
            __syncwarp();
        }
    }
}

// One wide record, the whole warp on it.
__device__ __forceinline__ void core_wide_record(SweepParams const &p, LineRec const *rp,
                                                 int const lo, int const hi, int const lane,
                                                 double *acc)
{
    CoreRec c;
    if (!core_rec_load(p, rp, lo, hi, c)) return;
    VoigtLine line;
    voigt_line_init(line, c.y, c.xlim0, c.xlim1);
    wn_t const wres = (wn_t)p.wres;
    for (int f0=c.a; f0<=c.b; f0+=32)
    {
        int const f = f0 + lane;
        double val;
        if (f <= c.b && core_sample(c, line, wres, f, val)) acc[f - lo] += val;
    }
    __syncwarp();
}

__global__ void __launch_bounds__(kCoreWarps*32)
core_gather_kernel(SweepParams const p, CoreGather const q)
{
    __shared__ double s_acc[kCoreWarps][kCoreTile];
    __shared__ unsigned s_queue[kCoreWarps][64];
    int const lane = threadIdx.x & 31;
    int const warp = threadIdx.x >> 5;
    double *acc = s_acc[warp];
    unsigned *queue = s_queue[warp];
    unsigned const total = (unsigned)q.num_tiles*(unsigned)p.num_layers;
    for (;;)
    {
        unsigned item = 0;
        if (lane == 0) item = atomicAdd(q.work_counter, 1u);
        item = __shfl_sync(0xffffffffu, item, 0);
        if (item >= total) break;
        int const layer = q.layer_order[item/(unsigned)q.num_tiles];
        int const tile = (int)(item%(unsigned)q.num_tiles);
        int const H = q.hmax[layer];
        if (H < 0) continue;
        int const lo = tile*kCoreTile;
        int const hi = min(lo + kCoreTile - 1, p.nw - 1);
        int const M = H + q.margin_pts;
        int const g0 = max(lo - M, 0)/kCoreGranule;
        int const g1 = min((hi + M)/kCoreGranule + 1, q.num_granules);
        unsigned const cfirst = (unsigned)q.first_at[g0];
        unsigned const clast = (unsigned)q.first_at[g1];
        LineRec const *layer_recs = p.recs + (size_t)layer*p.num_lines;
        bool touched = false;
        int qn = 0;
        for (unsigned base=cfirst; base<clast; base+=32u)
        {
            unsigned const r = base + (unsigned)lane;
            int extent = 0;
            if (r < clast)
            {
                CoreRec c;
                if (core_rec_load(p, layer_recs + r, lo, hi, c)) extent = c.b - c.a + 1;
            }
            unsigned const narrow = __ballot_sync(0xffffffffu, extent > 0 && extent < 32);
            unsigned wide = __ballot_sync(0xffffffffu, extent >= 32);
            if ((narrow | wide) && !touched)
            {
                for (int k=lane; k<kCoreTile; k+=32) acc[k] = 0.;
                touched = true;
                __syncwarp();
            }
            if (narrow)
            {
                if ((narrow >> lane) & 1u) queue[qn + __popc(narrow & ((1u << lane) - 1u))] = r;
                qn += __popc(narrow);
                __syncwarp();
                if (qn >= 32)
                {
                    core_narrow_group(p, layer_recs, queue, 32, lo, hi, lane, acc);
                    unsigned const keep = (lane < qn - 32) ? queue[32 + lane] : 0u;
                    __syncwarp();
                    if (lane < qn - 32) queue[lane] = keep;
                    qn -= 32;
                    __syncwarp();
                }
            }
            while (wide)
            {
                int const b = __ffs(wide) - 1;
                wide &= wide - 1;
                core_wide_record(p, layer_recs + (base + (unsigned)b), lo, hi, lane, acc);
            }
        }
        if (qn > 0) core_narrow_group(p, layer_recs, queue, qn, lo, hi, lane, acc);
        if (touched)
        {
            __syncwarp();
            real_t *tau = p.tau + (size_t)layer*p.nw;
            for (int k=lane; k<kCoreTile; k+=32)
            {
                double const v = acc[k];
                int const f = lo + k;
                if (f <= hi && v != 0.) tau[f] = (real_t)((double)tau[f] + v);
            }
            __syncwarp();
        }
    }
}

}  // namespace grt
