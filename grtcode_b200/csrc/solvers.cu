// Rayleigh optics, optics mixing, the longwave and shortwave column solvers and
// spectral -> broadband integration on sm_100a.
//
// Replaces rayleigh_scattering (shortwave/src/rayleigh.c:29-144), add_optics
// (utilities/src/optics.c:84-148), lw_fluxes_kernel/lw_flux/effective_planck/
// planck_law (longwave/src/longwave.c:68-309), sw_fluxes_kernel/sw_flux/
// sw_adding/eddington_mw1980/meador_weaver_1980/delta_eddington_scaling_jww1976
// (shortwave/src/shortwave.c:68-503) and the trapezoid loop of output_fluxes
// (framework/src/driver.c:302-319).
//
// One thread owns one wavenumber and walks the layers; every global access is
// layer-major with the wavenumber contiguous, so a warp reads/writes 256
// contiguous bytes per layer.  The "column" variants never materialise
// Rayleigh/total optics or spectral fluxes: they mix on the fly and reduce the
// fluxes to broadband values inside the kernel.
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include "common.cuh"

namespace grt {

constexpr int kSolverThreads = 128;

// Spectral arrays (optics in, fluxes out) carry the build's working precision (real_t: float
// in the single build); every kernel computes in double and rounds once at its stores, and the
// broadband sums of the fused kernels stay double until they are handed to the caller.
template <bool FUSED> struct FluxOut { typedef real_t type; };
template <> struct FluxOut<true> { typedef double type; };

struct LevelTemps
{
    double t_layers[GRT_MAX_LAYERS];
    double t_levels[GRT_MAX_LAYERS + 1];
};

struct LayerDensity
{
    double n[GRT_MAX_LAYERS];   // integrated air number density (curtis_godson.c:25)
};


// ---- working precision of the solver arithmetic ----------------------------------------
// `fp` is the reference's fp_t.  Every local the reference declares fp_t is an `fp` here and
// every expression is written as the reference writes it, so that C's usual arithmetic
// conversions round at the same places in both builds: in the single build an expression with
// a double literal (1., 0.25, ...) or a double variable (k, alpha1, alpha2, gamma4) is
// evaluated in double and rounded to float when it is stored in an fp, a product of two fp
// values is a float product (this TU is compiled -fmad=false: nothing is contracted).  The
// Meador-Weaver formulas are ill-conditioned where the single-scattering albedo is tiny, so
// the float roundings of gamma1..3 are visible in the fluxes: the single build has to make
// them exactly where the reference's single build does.  With fp = double nothing changes.
typedef real_t fp;
#ifdef GRT_SINGLE
#define GRT_MAX_EXP_ARG 80.f                     /* grtcode_config.h:29 */
__device__ __forceinline__ float fp_exp(float const x) { return expf(x); }     // EXP
#else
#define GRT_MAX_EXP_ARG 700.                     /* grtcode_config.h:41 */
__device__ __forceinline__ double fp_exp(double const x) { return exp(x); }
#endif
// the reference's plain exp() / sqrt(): double functions in both builds
__device__ __forceinline__ double d_exp(double const x) { return exp(x); }
__device__ __forceinline__ double d_sqrt(double const x) { return sqrt(x); }

// ---- Rayleigh (rayleigh.c:29-41) -------------------------------------------------
__device__ __forceinline__ fp rayleigh_tau(double const w, fp const n)
{
    fp const W = w*1.e-4;
    return (n*1.e-20*W*W*W*W)/(0.268675*1.e5*(9.38076E2 - 10.8426*W*W));
}

__global__ void rayleigh_kernel(int const num_layers, double const w0, double const dw,
                                uint64_t const nw, LayerDensity const dens, real_t *omega,
                                real_t *g, real_t *tau)
{
    uint64_t const i = (uint64_t)blockIdx.x*blockDim.x + threadIdx.x;
    if (i >= nw) return;
    double const w = w0 + (double)i*dw;
    for (int j=0; j<num_layers; ++j)
    {
        size_t const o = (size_t)j*nw + i;
        omega[o] = (real_t)1.;
        g[o] = (real_t)0.;
        tau[o] = rayleigh_tau(w, (fp)dens.n[j]);
    }
}


// ---- add_optics (optics.c:124-148) ------------------------------------------------
struct OpticsList
{
    real_t const *g[8];
    real_t const *omega[8];
    real_t const *tau[8];
};

__global__ void add_optics_kernel(uint64_t const n, int const num_optics, OpticsList const in,
                                  real_t *g_out, real_t *omega_out, real_t *tau_out)
{
    uint64_t const i = (uint64_t)blockIdx.x*blockDim.x + threadIdx.x;
    if (i >= n) return;
    fp g = 0., omega = 0., tau = 0.;
    for (int j=0; j<num_optics; ++j)
    {
        fp const gi = in.g[j][i];
        fp const oi = in.omega[j][i];
        fp const ti = in.tau[j][i];
        g += gi*oi*ti;
        omega += oi*ti;
        tau += ti;
    }
    g_out[i] = g/omega;
    omega_out[i] = omega/tau;
    tau_out[i] = tau;
}


// ---- longwave ----------------------------------------------------------------------
__device__ __forceinline__ fp planck(fp const T, fp const w)
{
    fp const c1 = 1.1910429526245744e-8;
    fp const c2 = 1.4387773538277202;
    fp e = c2*w/T;
    if (e > GRT_MAX_EXP_ARG) e = GRT_MAX_EXP_ARG;
    e = fp_exp(e);
    return (c1*w*w*w)/(e - 1.);
}

__device__ __forceinline__ fp effective_planck(fp const Tc, fp const Te, fp const w, fp const tau)
{
    fp const a = 0.193;
    fp const b = 0.013;
    fp const pc = planck(Tc, w);
    fp const pe = planck(Te, w);
    return (pc + (a*tau + b*tau*tau)*pe)/(1. + a*tau + b*tau*tau);
}

// Accumulates one value per thread into out[level] (block reduction + one
// atomic per block); used by the fused column kernels.
__device__ __forceinline__ void block_accumulate(double v, double *dst, double *warp_sums)
{
    for (int o=16; o>0; o>>=1) v += __shfl_down_sync(0xffffffffu, v, o);
    int const lane = threadIdx.x & 31;
    int const warp = threadIdx.x >> 5;
    __syncthreads();
    if (lane == 0) warp_sums[warp] = v;
    __syncthreads();
    if (threadIdx.x == 0)
    {
        double s = 0.;
        for (int k=0; k<(int)(blockDim.x >> 5); ++k) s += warp_sums[k];
        atomicAdd(dst, s);
    }
}

// The fused column kernels reduce one value per thread, level and direction: every warp
// parks its own sum in shared memory (no barrier per level) and the block adds them up and
// issues its atomics once, at the end of the kernel.
struct LevelSums
{
    double v[2][GRT_MAX_LAYERS + 1][kSolverThreads/32];
};

__device__ __forceinline__ void warp_accumulate(double v, LevelSums &sums, int const dir,
                                                int const level)
{
    for (int o=16; o>0; o>>=1) v += __shfl_down_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) sums.v[dir][level][threadIdx.x >> 5] = v;
}

__device__ __forceinline__ void clear_levels(LevelSums &sums)
{
    double *flat = &sums.v[0][0][0];
    for (int t=threadIdx.x; t<(int)(sizeof(LevelSums)/sizeof(double)); t+=blockDim.x) flat[t] = 0.;
    __syncthreads();
}

__device__ __forceinline__ void flush_levels(LevelSums const &sums, int const num_levels,
                                             double *flux_up, double *flux_down)
{
    __syncthreads();
    for (int t=threadIdx.x; t<2*num_levels; t+=blockDim.x)
    {
        int const dir = t/num_levels;
        int const level = t - dir*num_levels;
        double s = 0.;
        for (int k=0; k<(int)(blockDim.x >> 5); ++k) s += sums.v[dir][level][k];
        atomicAdd(dir ? &flux_down[level] : &flux_up[level], s);
    }
}

// FUSED = false: tau/omega are the caller's total optics, spectral fluxes are
// stored (drop-in calculate_lw_fluxes).  FUSED = true: tau is the gas optical
// depth; Rayleigh is mixed in on the fly (as column_calculation does through
// rayleigh_scattering + add_optics, driver.c:381-383) and the fluxes are reduced
// to broadband values with trapezoid weights.
template <bool FUSED>
__global__ void __launch_bounds__(kSolverThreads)
lw_kernel(int const num_levels, double const w0, double const dw, uint64_t const nw,
          double const t_surf, LevelTemps const temps, LayerDensity const dens,
          real_t const *tau, real_t const *omega, real_t const *emis,
          typename FluxOut<FUSED>::type *flux_up, typename FluxOut<FUSED>::type *flux_down,
          uint64_t const first, uint64_t const last)
{
    typedef typename FluxOut<FUSED>::type flux_t;
    __shared__ LevelSums sums;
    if (FUSED) clear_levels(sums);
    uint64_t const i = (uint64_t)blockIdx.x*blockDim.x + threadIdx.x;
    bool const live = i < nw;
    uint64_t const ii = live ? i : nw - 1;
    int const L = num_levels - 1;
    fp const w = w0 + (double)ii*dw;                 // longwave.c:238
    fp const c1[4] = {-14.402613260847248, -3.0302159969901132, -1.4925584280108841,
                      -1.0746123148178333};
    fp const c2[4] = {0.07587638482015649, 0.676114979733751, 1.3726594476601073,
                      1.0169418413757783};
    double weight = 0.;
    if (FUSED)
    {
        // trapezoid weights of output_fluxes over the integration window [first, last]
        weight = (live && ii >= first && ii <= last) ? ((ii == first || ii == last) ? 0.5*dw : dw)
                                                     : 0.;
        if (first == last) weight = 0.;
    }

    // tau_buf[j] = tau*(1. - omega) (longwave.c:253)
    auto layer_tau = [&](int const j) -> fp
    {
        size_t const o = (size_t)j*nw + ii;
        if (FUSED)
        {
            // add_optics of [gas, rayleigh] (optics.c:134-146)
            fp const tg = tau[o];
            fp const tr = rayleigh_tau(w0 + (double)ii*dw, (fp)dens.n[j]);
            fp const tt = tg + tr;
            fp const om = tr/tt;
            return tt*(1. - om);
        }
        return tau[o]*(1. - omega[o]);
    };

    fp I[4] = {0., 0., 0., 0.};
    if (FUSED)
    {
        // flux_down at the top of the atmosphere is zero: nothing to add.
    }
    else if (live)
    {
        flux_down[ii] = (flux_t)0.;
    }
    for (int j=0; j<L; ++j)
    {
        fp const t = layer_tau(j);
        fp const b = effective_planck((fp)temps.t_layers[j], (fp)temps.t_levels[j+1], w, t);
        fp f = 0.;
#pragma unroll
        for (int s=0; s<4; ++s)
        {
            fp const e = c1[s]*t;
            fp const ext = d_exp(e);
            fp const pp = (1. - ext)*b;
            I[s] = pp + I[s]*ext;
            f += c2[s]*I[s];
        }
        if (FUSED)
        {
            warp_accumulate((double)f*weight, sums, 1, j + 1);
        }
        else if (live)
        {
            flux_down[(size_t)(j+1)*nw + ii] = (flux_t)f;
        }
    }
    fp const bs = planck((fp)t_surf, w);
    fp const em = emis[ii];
    fp f = 0.;
#pragma unroll
    for (int s=0; s<4; ++s)
    {
        I[s] = em*bs + (1 - em)*I[s];
        f += c2[s]*I[s];
    }
    if (FUSED)
    {
        warp_accumulate((double)f*weight, sums, 0, L);
    }
    else if (live)
    {
        flux_up[(size_t)L*nw + ii] = (flux_t)f;
    }
    for (int j=L-1; j>=0; --j)
    {
        fp const t = layer_tau(j);
        fp const b = effective_planck((fp)temps.t_layers[j], (fp)temps.t_levels[j], w, t);
        f = 0.;
#pragma unroll
        for (int s=0; s<4; ++s)
        {
            fp const e = c1[s]*t;
            fp const ext = d_exp(e);
            fp const pp = (1. - ext)*b;
            I[s] = pp + I[s]*ext;
            f += c2[s]*I[s];
        }
        if (FUSED)
        {
            warp_accumulate((double)f*weight, sums, 0, j);
        }
        else if (live)
        {
            flux_up[(size_t)j*nw + ii] = (flux_t)f;
        }
    }
    if (FUSED) flush_levels(sums, num_levels, (double *)flux_up, (double *)flux_down);
}


// ---- shortwave ---------------------------------------------------------------------
// meador_weaver_1980 (shortwave.c:97-209) with the Eddington gammas of
// eddington_mw1980 (shortwave.c:214-236).  t_pure == nullptr for the diffuse beam.
// The optical depth the exponentials see (shortwave.c:137-145).
__device__ __forceinline__ fp edd_clamped_tau(fp const tau, fp const mu, double const k)
{
    fp t = tau;
    if (1./mu > k && tau/mu > GRT_MAX_EXP_ARG)
    {
        t = GRT_MAX_EXP_ARG*mu;
    }
    else if (tau*k > GRT_MAX_EXP_ARG)
    {
        t = GRT_MAX_EXP_ARG/k;
    }
    return t;
}

// One beam of meador_weaver_1980 for omega > 0, given the quantities that do not depend on
// the beam's zenith angle (gamma1, gamma2, k) and exp(-t*k), exp(t*k).
__device__ __forceinline__ void edd_beam(fp const omega, fp const mu, fp const g,
                                         fp const gamma1, fp const gamma2, double const k,
                                         fp const t, fp const tkm, fp const tkp, fp *R, fp *T,
                                         fp *T_pure)
{
    fp const gamma3 = 0.25*(2. - 3.*g*mu);
    double const gamma4 = 1. - gamma3;
    double const alpha1 = gamma1*gamma4 + gamma2*gamma3;
    double const alpha2 = gamma1*gamma3 + gamma2*gamma4;
    fp const tp = d_exp(t/mu);
    if (tp <= 1.0)
    {
        *R = 0.;
        *T = 1.;
        if (T_pure != nullptr) *T_pure = 1.;
    }
    else
    {
        fp const tm = d_exp(-t/mu);
        if (T_pure != nullptr) *T_pure = tm;
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
    if (T_pure != nullptr)
    {
        if (*T_pure > *T) *T = *T_pure;
    }
}

// eddington_mw1980 + meador_weaver_1980 (shortwave.c:97-236) for the direct (mu_d) and the
// diffuse (mu_f) beam of one layer.  The reference evaluates the two beams independently; the
// Eddington gammas 1 and 2, k and -- unless one of the two optical-depth clamps fires for one
// beam only -- exp(-t*k) and exp(t*k) do not depend on the beam, so they are computed once
// (2 of the 8 double exp and 1 of the 2 sqrt per layer): the same expressions on the same
// operands, hence the same bits.
__device__ void eddington_pair(fp const omega, fp const tau, fp const mu_d, fp const mu_f,
                               fp const g, fp *R_dir, fp *T_dir, fp *T_pure, fp *R_dif,
                               fp *T_dif)
{
    if (omega <= 0.0)
    {
        *R_dir = 0.;
        *T_dir = d_exp(-tau/mu_d);
        *T_pure = *T_dir;
        *R_dif = 0.;
        *T_dif = d_exp(-tau/mu_f);
        return;
    }
    fp const gamma1 = 0.25*(7. - omega*(4. + 3.*g));
    fp const gamma2 = -0.25*(1. - omega*(4. - 3.*g));
    double const k = d_sqrt(gamma1*gamma1 - gamma2*gamma2);
    fp const td = edd_clamped_tau(tau, mu_d, k);
    fp const tf = edd_clamped_tau(tau, mu_f, k);
    fp const tkm_d = d_exp(-td*k);
    fp const tkp_d = d_exp(td*k);
    fp tkm_f = tkm_d;
    fp tkp_f = tkp_d;
    if (tf != td)
    {
        tkm_f = d_exp(-tf*k);
        tkp_f = d_exp(tf*k);
    }
    edd_beam(omega, mu_d, g, gamma1, gamma2, k, td, tkm_d, tkp_d, R_dir, T_dir, T_pure);
    edd_beam(omega, mu_f, g, gamma1, gamma2, k, tf, tkm_f, tkp_f, R_dif, T_dif, nullptr);
}

// Resident blocks per SM the compiler must allow.  The solver is latency-bound on its
// per-thread layer arrays (local memory that spills to DRAM: L2 hit rate 21 %): measured on
// config 2, 1 / 2 / 3 / 4 / 6 / 8 / 10 blocks give 54.5 / 32.2 / 23.5 / 18.8 / 14.1 / 12.4 /
// 12.1 ms, so the register budget is capped at 48 (a few spilled words) for ten blocks.
#ifndef GRT_SW_MINBLOCKS
#define GRT_SW_MINBLOCKS 10
#endif
template <int MAXL, bool FUSED>
__global__ void __launch_bounds__(kSolverThreads, GRT_SW_MINBLOCKS)
sw_kernel(int const num_levels, double const w0, double const dw, uint64_t const nw,
          LayerDensity const dens, real_t const *omega_in, real_t const *g_in,
          real_t const *tau_in, double const mu_dir, double const mu_dif,
          real_t const *alb_dir, real_t const *alb_dif, double const tsi, real_t const *solar,
          typename FluxOut<FUSED>::type *flux_up, typename FluxOut<FUSED>::type *flux_down,
          uint64_t const first, uint64_t const last)
{
    typedef typename FluxOut<FUSED>::type flux_t;
    __shared__ LevelSums sums;
    uint64_t const i = (uint64_t)blockIdx.x*blockDim.x + threadIdx.x;
    bool const live = i < nw;
    if (!FUSED && !live) return;
    if (FUSED) clear_levels(sums);
    uint64_t const ii = live ? i : nw - 1;
    int const L = num_levels - 1;
    double const w = w0 + (double)ii*dw;
    fp const mu_d = mu_dir;                          // fp_t arguments of sw_flux
    fp const mu_f = mu_dif;
    fp const tsi_fp = tsi;

    // Layer reflectivities / transmittances and, in the same bottom-up loop, the adding
    // recursions from the surface (shortwave.c:242-260): a layer's five values are in
    // registers when Rdd / Rfd need them, so the per-thread arrays (local memory, the traffic
    // that bounds this kernel) are written once here and read once by the level loop below.
    // The layers are independent, so their order does not change any value.
    fp R_dir[MAXL], R_dif[MAXL], T_dir[MAXL], T_dif[MAXL], T_pure[MAXL];
    fp Rdd[MAXL + 1], Rfd[MAXL + 1];
    fp rdd_below = alb_dir[ii];
    fp rfd_below = alb_dif[ii];
    Rdd[L] = rdd_below;
    Rfd[L] = rfd_below;
    for (int j=L-1; j>=0; --j)
    {
        size_t const o = (size_t)j*nw + ii;
        fp omega, g, tau;
        if (FUSED)
        {
            // add_optics of [gas, rayleigh] (optics.c:134-146): g = 0*tau_R/tau_R,
            // omega = tau_R/(tau_gas + tau_R).
            fp const tg = tau_in[o];
            fp const tr = rayleigh_tau(w, (fp)dens.n[j]);
            tau = tg + tr;
            g = 0./tr;
            omega = tr/tau;
        }
        else
        {
            omega = omega_in[o];
            g = g_in[o];
            tau = tau_in[o];
        }
        // delta-Eddington scaling (shortwave.c:86-89).
        fp const gs = g/(g + 1.);
        fp const f = g*g;
        fp const os = (1. - f)*omega/(1. - omega*f);
        fp const ts = tau*(1. - omega*f);
        fp r_dir, t_dir, t_pure, r_dif, t_dif;
        eddington_pair(os, ts, mu_d, mu_f, gs, &r_dir, &t_dir, &t_pure, &r_dif, &t_dif);
        R_dir[j] = r_dir;
        T_dir[j] = t_dir;
        T_pure[j] = t_pure;
        R_dif[j] = r_dif;
        T_dif[j] = t_dif;
        // Adding (shortwave.c:242-260).
        fp const A = t_pure;
        fp const B = 1./(1. - r_dif*rfd_below);
        fp const rdd = r_dir + (A*rdd_below + (t_dir - A)*rfd_below)*t_dif*B;
        fp const rfd = r_dif + t_dif*t_dif*rfd_below*B;
        Rdd[j] = rdd;
        Rfd[j] = rfd;
        rdd_below = rdd;
        rfd_below = rfd;
    }
    // Rfu (reflectivity of the layers above a level to diffuse light from below,
    // shortwave.c:262-268) is a top-down recursion: it is advanced inside the level loop
    // below (the same operations in the same order) instead of being stored per layer.
    fp const scale = solar[ii]*mu_d;                 // flux *= solar_flux*mu_dir (shortwave.c:402)
    double weight = 0.;
    if (FUSED)
    {
        // trapezoid weights of output_fluxes over the integration window [first, last]
        weight = (live && ii >= first && ii <= last) ? ((ii == first || ii == last) ? 0.5*dw : dw)
                                                     : 0.;
        if (first == last) weight = 0.;
    }
    fp dir_beam = 1.;
    fp dif_beam = 0.;
    {
        fp const r0 = dir_beam*Rdd[0];
        fp const t0 = dir_beam;
        fp const up = tsi_fp*(r0*scale);             // shortwave.c:448
        fp const down = tsi_fp*(t0*scale);
        if (FUSED)
        {
            warp_accumulate((double)up*weight, sums, 0, 0);
            warp_accumulate((double)down*weight, sums, 1, 0);
        }
        else
        {
            flux_up[ii] = (flux_t)up;
            flux_down[ii] = (flux_t)down;
        }
    }
    fp rfu_above = 0.;          // Rfu[j-2]
    fp rfu = R_dif[0];          // Rfu[j-1]
    for (int j=1; j<num_levels; ++j)
    {
        if (j > 1)
        {
            // Rfu[j-1] from Rfu[j-2]
            fp const Bu = 1./(1. - R_dif[j-1]*rfu);
            fp const next = R_dif[j-1] + T_dif[j-1]*T_dif[j-1]*rfu*Bu;
            rfu_above = rfu;
            rfu = next;
            fp const C = 1./(1. - R_dif[j-1]*rfu_above);
            dif_beam = (dir_beam*R_dir[j-1]*rfu_above + dif_beam)*T_dif[j-1]*C +
                       dir_beam*(T_dir[j-1] - T_pure[j-1]);
        }
        else
        {
            dif_beam = dir_beam*(T_dir[j-1] - T_pure[j-1]);
        }
        dir_beam *= T_pure[j-1];
        fp const B = 1./(1. - Rfd[j]*rfu);
        fp const R = (dir_beam*Rdd[j] + dif_beam*Rfd[j])*B;
        fp const T = dir_beam*(1. + Rdd[j]*rfu*B) + dif_beam*B;
        fp const up = tsi_fp*(R*scale);
        fp const down = tsi_fp*(T*scale);
        if (FUSED)
        {
            warp_accumulate((double)up*weight, sums, 0, j);
            warp_accumulate((double)down*weight, sums, 1, j);
        }
        else
        {
            flux_up[(size_t)j*nw + ii] = (flux_t)up;
            flux_down[(size_t)j*nw + ii] = (flux_t)down;
        }
    }
    if (FUSED) flush_levels(sums, num_levels, (double *)flux_up, (double *)flux_down);
}


// ---- shortwave, staged through L2 ----------------------------------------------------
// sw_kernel above keeps seven per-thread arrays of MAXL values in local memory: with enough
// threads resident to hide its latency they spill to DRAM (13x the algorithmic traffic).  The
// expensive part of a layer -- the Eddington reflectivities and transmittances, six double
// exp and a sqrt -- does not depend on the other layers; only the adding recursions are
// sequential.  So a block of four warps takes 32 wavenumbers at a time:
//   phase A  all 128 threads evaluate the (wavenumber, layer) pairs independently (warp = one
//            layer row of 32 wavenumbers: every access 256 contiguous bytes) and park the five
//            layer values in a block-private scratch slab in global memory -- 69 KB per block
//            at 54 layers, 61 MB for 6 blocks on each SM: it lives in L2 and is overwritten in
//            place stage after stage, so it never has to reach DRAM;
//   phase B  warp 0 runs the two recursions of shortwave.c:242-330 for its 32 wavenumbers,
//            Rdd / Rfd in shared memory, while the other blocks of the SM are in phase A.
// Same expressions on the same operands as sw_kernel, hence the same bits.
constexpr int kStagedThreads = 128;
constexpr int kStagedBlocksPerSM = 6;
constexpr int kStageWidth = 32;

__host__ __device__ constexpr size_t sw_stage_doubles(int const num_layers)
{
    return (size_t)5*(size_t)num_layers*kStageWidth;
}

template <bool FUSED>
__global__ void __launch_bounds__(kStagedThreads, kStagedBlocksPerSM)
sw_staged_kernel(int const num_levels, double const w0, double const dw, uint64_t const nw,
                 LayerDensity const dens, real_t const *omega_in, real_t const *g_in,
                 real_t const *tau_in, double const mu_dir, double const mu_dif,
                 real_t const *alb_dir, real_t const *alb_dif, double const tsi,
                 real_t const *solar, typename FluxOut<FUSED>::type *flux_up,
                 typename FluxOut<FUSED>::type *flux_down, uint64_t const first,
                 uint64_t const last, double *scratch)
{
    typedef typename FluxOut<FUSED>::type flux_t;
    extern __shared__ double staged_smem[];
    int const L = num_levels - 1;
    // shared: Rdd[L+1][32], Rfd[L+1][32], then (FUSED) the block's broadband sums [2][L+1]
    double *s_rdd = staged_smem;
    double *s_rfd = staged_smem + (size_t)num_levels*kStageWidth;
    double *s_sum = staged_smem + (size_t)2*num_levels*kStageWidth;
    int const lane = threadIdx.x & 31;
    int const warp = threadIdx.x >> 5;
    if (FUSED)
    {
        for (int t=threadIdx.x; t<2*num_levels; t+=kStagedThreads) s_sum[t] = 0.;
    }
    fp const mu_d = mu_dir;                          // fp_t arguments of sw_flux
    fp const mu_f = mu_dif;
    fp const tsi_fp = tsi;
    double *slab = scratch + (size_t)blockIdx.x*sw_stage_doubles(L);
    size_t const plane = (size_t)L*kStageWidth;      // one of the five layer values, [L][32]
    uint64_t const stages = (nw + kStageWidth - 1)/kStageWidth;
    for (uint64_t stage=blockIdx.x; stage<stages; stage+=gridDim.x)
    {
        uint64_t const i = stage*kStageWidth + (uint64_t)lane;
        bool const live = i < nw;
        uint64_t const ii = live ? i : nw - 1;
        // ---- phase A: layer values (shortwave.c:86-236), one layer row per warp and turn
        {
            double const w = w0 + (double)ii*dw;
            for (int j=warp; j<L; j+=kStagedThreads/32)
            {
                size_t const o = (size_t)j*nw + ii;
                fp omega, g, tau;
                if (FUSED)
                {
                    // add_optics of [gas, rayleigh] (optics.c:134-146)
                    fp const tg = tau_in[o];
                    fp const tr = rayleigh_tau(w, (fp)dens.n[j]);
                    tau = tg + tr;
                    g = 0./tr;
                    omega = tr/tau;
                }
                else
                {
                    omega = omega_in[o];
                    g = g_in[o];
                    tau = tau_in[o];
                }
                // delta-Eddington scaling (shortwave.c:86-89)
                fp const gs = g/(g + 1.);
                fp const f = g*g;
                fp const os = (1. - f)*omega/(1. - omega*f);
                fp const ts = tau*(1. - omega*f);
                fp r_dir, t_dir, t_pure, r_dif, t_dif;
                eddington_pair(os, ts, mu_d, mu_f, gs, &r_dir, &t_dir, &t_pure, &r_dif, &t_dif);
                double *dst = slab + (size_t)j*kStageWidth + lane;
                __stcg(dst, (double)r_dir);
                __stcg(dst + plane, (double)t_dir);
                __stcg(dst + 2*plane, (double)t_pure);
                __stcg(dst + 3*plane, (double)r_dif);
                __stcg(dst + 4*plane, (double)t_dif);
            }
        }
        __syncthreads();
        // ---- phase B: adding from the surface up (shortwave.c:242-260), then the beams
        // from the top down (shortwave.c:262-330, 339-405)
        if (warp == 0)
        {
            double const *src = slab + lane;
            fp rdd_below = alb_dir[ii];
            fp rfd_below = alb_dif[ii];
            s_rdd[(size_t)L*kStageWidth + lane] = rdd_below;
            s_rfd[(size_t)L*kStageWidth + lane] = rfd_below;
            {
                size_t o = (size_t)(L - 1)*kStageWidth;
                fp r_dir = __ldcg(src + o);
                fp t_dir = __ldcg(src + o + plane);
                fp t_pure = __ldcg(src + o + 2*plane);
                fp r_dif = __ldcg(src + o + 3*plane);
                fp t_dif = __ldcg(src + o + 4*plane);
                for (int j=L-1; j>=0; --j)
                {
                    // the next layer's values are in flight while this one is added
                    fp n_r_dir = 0., n_t_dir = 0., n_t_pure = 0., n_r_dif = 0., n_t_dif = 0.;
                    if (j > 0)
                    {
                        size_t const on = (size_t)(j - 1)*kStageWidth;
                        n_r_dir = __ldcg(src + on);
                        n_t_dir = __ldcg(src + on + plane);
                        n_t_pure = __ldcg(src + on + 2*plane);
                        n_r_dif = __ldcg(src + on + 3*plane);
                        n_t_dif = __ldcg(src + on + 4*plane);
                    }
                    fp const A = t_pure;
                    fp const B = 1./(1. - r_dif*rfd_below);
                    fp const rdd = r_dir + (A*rdd_below + (t_dir - A)*rfd_below)*t_dif*B;
                    fp const rfd = r_dif + t_dif*t_dif*rfd_below*B;
                    s_rdd[(size_t)j*kStageWidth + lane] = rdd;
                    s_rfd[(size_t)j*kStageWidth + lane] = rfd;
                    rdd_below = rdd;
                    rfd_below = rfd;
                    r_dir = n_r_dir; t_dir = n_t_dir; t_pure = n_t_pure; r_dif = n_r_dif;
                    t_dif = n_t_dif;
                }
            }
            fp const scale = solar[ii]*mu_d;         // flux *= solar_flux*mu_dir (shortwave.c:402)
            double weight = 0.;
            if (FUSED)
            {
                weight = (live && ii >= first && ii <= last)
                             ? ((ii == first || ii == last) ? 0.5*dw : dw) : 0.;
                if (first == last) weight = 0.;
            }
            auto emit = [&](int const level, fp const up, fp const down)
            {
                if (FUSED)
                {
                    double u = (double)up*weight;
                    double d = (double)down*weight;
                    for (int o=16; o>0; o>>=1)
                    {
                        u += __shfl_down_sync(0xffffffffu, u, o);
                        d += __shfl_down_sync(0xffffffffu, d, o);
                    }
                    if (lane == 0)
                    {
                        s_sum[level] += u;
                        s_sum[num_levels + level] += d;
                    }
                }
                else if (live)
                {
                    flux_up[(size_t)level*nw + ii] = (flux_t)up;
                    flux_down[(size_t)level*nw + ii] = (flux_t)down;
                }
            };
            fp dir_beam = 1.;
            fp dif_beam = 0.;
            {
                fp const r0 = dir_beam*(fp)s_rdd[lane];
                fp const t0 = dir_beam;
                emit(0, tsi_fp*(r0*scale), tsi_fp*(t0*scale));      // shortwave.c:448
            }
            fp rfu_above = 0.;          // Rfu[j-2]
            fp r_dir = __ldcg(src);
            fp t_dir = __ldcg(src + plane);
            fp t_pure = __ldcg(src + 2*plane);
            fp r_dif = __ldcg(src + 3*plane);
            fp t_dif = __ldcg(src + 4*plane);
            fp rfu = r_dif;             // Rfu[j-1] = R_dif[0]
            for (int j=1; j<num_levels; ++j)
            {
                // layer j-1's values are in registers; layer j's are fetched meanwhile
                fp n_r_dir = 0., n_t_dir = 0., n_t_pure = 0., n_r_dif = 0., n_t_dif = 0.;
                if (j < L)
                {
                    size_t const on = (size_t)j*kStageWidth;
                    n_r_dir = __ldcg(src + on);
                    n_t_dir = __ldcg(src + on + plane);
                    n_t_pure = __ldcg(src + on + 2*plane);
                    n_r_dif = __ldcg(src + on + 3*plane);
                    n_t_dif = __ldcg(src + on + 4*plane);
                }
                if (j > 1)
                {
                    // Rfu[j-1] from Rfu[j-2]
                    fp const Bu = 1./(1. - r_dif*rfu);
                    fp const next = r_dif + t_dif*t_dif*rfu*Bu;
                    rfu_above = rfu;
                    rfu = next;
                    fp const C = 1./(1. - r_dif*rfu_above);
                    dif_beam = (dir_beam*r_dir*rfu_above + dif_beam)*t_dif*C +
                               dir_beam*(t_dir - t_pure);
                }
                else
                {
                    dif_beam = dir_beam*(t_dir - t_pure);
                }
                dir_beam *= t_pure;
                fp const rdd_j = s_rdd[(size_t)j*kStageWidth + lane];
                fp const rfd_j = s_rfd[(size_t)j*kStageWidth + lane];
                fp const B = 1./(1. - rfd_j*rfu);
                fp const R = (dir_beam*rdd_j + dif_beam*rfd_j)*B;
                fp const T = dir_beam*(1. + rdd_j*rfu*B) + dif_beam*B;
                emit(j, tsi_fp*(R*scale), tsi_fp*(T*scale));
                r_dir = n_r_dir; t_dir = n_t_dir; t_pure = n_t_pure; r_dif = n_r_dif;
                t_dif = n_t_dif;
            }
        }
        __syncthreads();
    }
    if (FUSED)
    {
        for (int t=threadIdx.x; t<2*num_levels; t+=kStagedThreads)
        {
            int const dir = t/num_levels;
            int const level = t - dir*num_levels;
            atomicAdd(dir ? &((double *)flux_down)[level] : &((double *)flux_up)[level], s_sum[t]);
        }
    }
}


// ---- broadband integration (driver.c:302-319) ----------------------------------------
__global__ void __launch_bounds__(256)
integrate_kernel(uint64_t const nw, double const dw, real_t const *flux, double *out)
{
    __shared__ double warp_sums[8];
    int const level = blockIdx.y;
    real_t const *f = flux + (size_t)level*nw;
    double acc = 0.;
    for (uint64_t i=(uint64_t)blockIdx.x*blockDim.x + threadIdx.x; i + 1 < nw;
         i += (uint64_t)gridDim.x*blockDim.x)
    {
        acc += 0.5*((double)f[i] + (double)f[i+1])*dw;
    }
    block_accumulate(acc, &out[level], warp_sums);
}


// ---- allocation-free broadband integration (grt_integrate_flux_noalloc) -----------------
// The same trapezoid sums as integrate_kernel for callers that must not allocate (CUDA graph
// capture, serving loops, harnesses that forbid cudaMalloc): a two-stage reduction whose
// per-block partial sums live in a file-scope device array instead of a cudaMalloc'ed
// buffer.  Every GPU that loads the library holds the array until the process exits
// (GRT_NOALLOC_SCRATCH_BYTES, about 100 KB).  All calls on a device share it: calls that
// overlap on the GPU (different streams, concurrent graphs) race.  Nothing has to start at
// zero: the first stage overwrites every slot the second stage reads, so the previous call's
// contents are never seen.
constexpr int kNoAllocBlocks = GRT_NOALLOC_BLOCKS;
constexpr int kNoAllocMaxLevels = GRT_NOALLOC_MAX_LEVELS;
static __device__ __align__(256) double g_noalloc_scratch[kNoAllocMaxLevels*kNoAllocBlocks];
static_assert(sizeof(double)*kNoAllocMaxLevels*kNoAllocBlocks == GRT_NOALLOC_SCRATCH_BYTES,
              "scratch size stated in grt_sm100.h");

// Stage 1: block (b, level) sums its grid-stride share of the level's trapezoid intervals in a
// fixed order (no atomics) and writes partial[level*gridDim.x + b].
__global__ void __launch_bounds__(256)
integrate_partial_kernel(uint64_t const nw, double const dw, real_t const *flux, double *partial)
{
    __shared__ double warp_sums[8];
    int const level = blockIdx.y;
    real_t const *f = flux + (size_t)level*nw;
    double acc = 0.;
    for (uint64_t i=(uint64_t)blockIdx.x*blockDim.x + threadIdx.x; i + 1 < nw;
         i += (uint64_t)gridDim.x*blockDim.x)
    {
        acc += 0.5*((double)f[i] + (double)f[i+1])*dw;
    }
    for (int o=16; o>0; o>>=1) acc += __shfl_down_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) warp_sums[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0)
    {
        double s = 0.;
        for (int k=0; k<(int)(blockDim.x >> 5); ++k) s += warp_sums[k];
        partial[(size_t)level*gridDim.x + blockIdx.x] = s;
    }
}

// Stage 2: one warp per level adds the level's partial sums and stores the result in the
// caller's output array.
__global__ void __launch_bounds__(32)
integrate_finish_kernel(int const num_blocks, double const *partial, real_t *out)
{
    int const level = blockIdx.x;
    double acc = 0.;
    for (int k=threadIdx.x; k<num_blocks; k+=32) acc += partial[(size_t)level*num_blocks + k];
    for (int o=16; o>0; o>>=1) acc += __shfl_down_sync(0xffffffffu, acc, o);
    if (threadIdx.x == 0) out[level] = (real_t)acc;
}

// Device address of g_noalloc_scratch on each device: in host code the array's name is only
// a stand-in for the symbol, so the address is looked up once per device, with that device
// current (stream_of() has made it so), and passed to the kernels as an ordinary pointer.
static double *g_noalloc_ptr[64];

static int noalloc_scratch(int device, double **ptr)
{
    if (device < 0 || device >= 64) return fail("device id %d out of range", device);
    if (g_noalloc_ptr[device] == nullptr)
    {
        double *p = nullptr;
        GRT_CUDA(cudaGetSymbolAddress((void **)&p, g_noalloc_scratch));
        if (p == nullptr) return fail("cudaGetSymbolAddress returned a null scratch address");
        g_noalloc_ptr[device] = p;
    }
    *ptr = g_noalloc_ptr[device];
    return 0;
}


// Per-device scratch for broadband results.
static double *g_out_buf[64];

static int out_buffer(int device, double **buf)
{
    if (g_out_buf[device] == nullptr)
    {
        GRT_CUDA(cudaMalloc(&g_out_buf[device], sizeof(double)*4*(GRT_MAX_LAYERS + 1)));
    }
    *buf = g_out_buf[device];
    return 0;
}

// n of calc_number_densities (curtis_godson.c:25-34) in the build's fp_t.
static void number_densities(int num_layers, real_t const *p_atm, LayerDensity *d)
{
    for (int i=0; i<num_layers; ++i)
    {
        real_t dp = p_atm[i] - p_atm[i+1];
        dp = dp >= (real_t)0 ? dp : -dp;
        d->n[i] = (double)((real_t)2.147822334314468e+25*dp);
    }
}

// Broadband results: double on the device, handed to the caller in the build's precision.
static int fetch_levels(cudaStream_t st, double const *buf, int num_levels, real_t *up, real_t *down)
{
    double host[2*(GRT_MAX_LAYERS + 1)];
    GRT_CUDA(cudaMemcpyAsync(host, buf, sizeof(double)*2*num_levels, cudaMemcpyDeviceToHost, st));
    GRT_CUDA(cudaStreamSynchronize(st));
    for (int i=0; i<num_levels; ++i)
    {
        up[i] = (real_t)host[i];
        if (down != nullptr) down[i] = (real_t)host[num_levels + i];
    }
    return 0;
}

}  // namespace grt

using namespace grt;

// Per-device scratch slabs of sw_staged_kernel (grow-only).
static double *g_sw_scratch[64];
static size_t g_sw_scratch_bytes[64];

static int sw_scratch(int device, size_t bytes, double **ptr)
{
    if (device < 0 || device >= 64) return fail("device id %d out of range", device);
    if (bytes > g_sw_scratch_bytes[device])
    {
        if (g_sw_scratch[device] != nullptr) GRT_CUDA(cudaFree(g_sw_scratch[device]));
        g_sw_scratch[device] = nullptr;
        g_sw_scratch_bytes[device] = 0;
        GRT_CUDA(cudaMalloc(&g_sw_scratch[device], bytes));
        g_sw_scratch_bytes[device] = bytes;
    }
    *ptr = g_sw_scratch[device];
    return 0;
}

template <bool FUSED>
static int launch_sw(int device, cudaStream_t st, int num_levels, double w0, double dw,
                     uint64_t nw, LayerDensity const &dens, real_t const *omega, real_t const *g,
                     real_t const *tau, double mu_dir, double mu_dif, real_t const *alb_dir,
                     real_t const *alb_dif, double tsi, real_t const *solar,
                     typename FluxOut<FUSED>::type *up, typename FluxOut<FUSED>::type *down,
                     uint64_t first, uint64_t last)
{
    // Measured on config 2 (profiles/r02j_sw_staged.txt): the staged kernel moves 5.4 GB of DRAM
    // traffic per launch instead of 31.5 GB but takes 16.5 ms against 11.2 ms -- with one warp in
    // four on the recursions and 24 warps per SM its in-order dependency chains leave the FP64
    // pipe half idle -- so sw_kernel stays the default; GRT_SW_KERNEL=staged selects this one.
    char const *which = getenv("GRT_SW_KERNEL");
    if (which != nullptr && strcmp(which, "staged") == 0)
    {
        static bool configured[64][2];
        int const L = num_levels - 1;
        size_t const smem = sizeof(double)*((size_t)2*num_levels*kStageWidth + 2*(size_t)num_levels);
        if (smem > 48*1024 && !configured[device][FUSED ? 1 : 0])
        {
            GRT_CUDA(cudaFuncSetAttribute(sw_staged_kernel<FUSED>,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)(sizeof(double)*((size_t)2*(GRT_MAX_LAYERS + 1)*kStageWidth +
                                                                2*(size_t)(GRT_MAX_LAYERS + 1)))));
            configured[device][FUSED ? 1 : 0] = true;
        }
        uint64_t const stages = (nw + kStageWidth - 1)/kStageWidth;
        uint64_t const resident = (uint64_t)kStagedBlocksPerSM*(uint64_t)sm_count(device);
        unsigned const blocks = (unsigned)(stages < resident ? stages : resident);
        double *scratch = nullptr;
        GRT_TRY(sw_scratch(device, sizeof(double)*sw_stage_doubles(L)*blocks, &scratch));
        sw_staged_kernel<FUSED><<<blocks, kStagedThreads, smem, st>>>(
            num_levels, w0, dw, nw, dens, omega, g, tau, mu_dir, mu_dif, alb_dir, alb_dif, tsi,
            solar, up, down, first, last, scratch);
        GRT_CUDA(cudaGetLastError());
        return 0;
    }
    unsigned const blocks = (unsigned)((nw + kSolverThreads - 1)/kSolverThreads);
    // Measurement knob: resident blocks per SM, enforced with idle dynamic shared memory (the
    // per-thread layer arrays of the resident threads are what has to fit in L2).
    size_t dyn = 0;
    if (char const *env = getenv("GRT_SW_BLOCKS_PER_SM"))
    {
        int const want = atoi(env);
        if (want >= 1 && want <= 16)
        {
            dyn = (size_t)(227*1024)/(size_t)want - (FUSED ? sizeof(LevelSums) : 0) - 1024;
            dyn &= ~(size_t)255;
            if (num_levels - 1 <= 64)
            {
                GRT_CUDA(cudaFuncSetAttribute(sw_kernel<64, FUSED>,
                                              cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn));
            }
            else
            {
                dyn = 0;
            }
        }
    }
    if (num_levels - 1 <= 64)
    {
        sw_kernel<64, FUSED><<<blocks, kSolverThreads, dyn, st>>>(
            num_levels, w0, dw, nw, dens, omega, g, tau, mu_dir, mu_dif, alb_dir, alb_dif, tsi,
            solar, up, down, first, last);
    }
    else
    {
        sw_kernel<GRT_MAX_LAYERS, FUSED><<<blocks, kSolverThreads, 0, st>>>(
            num_levels, w0, dw, nw, dens, omega, g, tau, mu_dir, mu_dif, alb_dir, alb_dif, tsi,
            solar, up, down, first, last);
    }
    GRT_CUDA(cudaGetLastError());
    return 0;
}

extern "C" {

int grt_rayleigh(int device, int num_layers, double w0, double dw, uint64_t nw,
                 real_t const *p_atm, real_t *omega_dev, real_t *g_dev, real_t *tau_dev)
{
    if (num_layers < 1 || num_layers > GRT_MAX_LAYERS) return fail("bad layer count %d", num_layers);
    cudaStream_t st;
    GRT_TRY(stream_of(device, &st));
    LayerDensity dens;
    number_densities(num_layers, p_atm, &dens);
    rayleigh_kernel<<<(unsigned)((nw + 255)/256), 256, 0, st>>>(num_layers, w0, dw, nw, dens,
                                                                omega_dev, g_dev, tau_dev);
    GRT_CUDA(cudaGetLastError());
    return 0;
}

int grt_add_optics(int device, uint64_t n, int num_optics, real_t const *const *g_dev,
                   real_t const *const *omega_dev, real_t const *const *tau_dev,
                   real_t *g_out_dev, real_t *omega_out_dev, real_t *tau_out_dev)
{
    if (num_optics < 1 || num_optics > 8) return fail("add_optics supports 1..8 objects");
    cudaStream_t st;
    GRT_TRY(stream_of(device, &st));
    OpticsList list;
    for (int j=0; j<num_optics; ++j)
    {
        list.g[j] = g_dev[j];
        list.omega[j] = omega_dev[j];
        list.tau[j] = tau_dev[j];
    }
    add_optics_kernel<<<(unsigned)((n + 255)/256), 256, 0, st>>>(n, num_optics, list, g_out_dev,
                                                                 omega_out_dev, tau_out_dev);
    GRT_CUDA(cudaGetLastError());
    return 0;
}

int grt_lw_fluxes(int device, int num_levels, double w0, double dw, uint64_t nw, double t_surf,
                  real_t const *t_layers, real_t const *t_levels, real_t const *tau_dev,
                  real_t const *omega_dev, real_t const *emis_dev, real_t *flux_up_dev,
                  real_t *flux_down_dev)
{
    if (num_levels < 2 || num_levels > GRT_MAX_LAYERS + 1) return fail("bad level count");
    cudaStream_t st;
    GRT_TRY(stream_of(device, &st));
    LevelTemps temps;
    for (int i=0; i<num_levels-1; ++i) temps.t_layers[i] = t_layers[i];
    for (int i=0; i<num_levels; ++i) temps.t_levels[i] = t_levels[i];
    LayerDensity dens{};
    unsigned const blocks = (unsigned)((nw + kSolverThreads - 1)/kSolverThreads);
    lw_kernel<false><<<blocks, kSolverThreads, 0, st>>>(num_levels, w0, dw, nw, t_surf, temps,
                                                        dens, tau_dev, omega_dev, emis_dev,
                                                        flux_up_dev, flux_down_dev, 0, nw - 1);
    GRT_CUDA(cudaGetLastError());
    return 0;
}

int grt_sw_fluxes(int device, int num_levels, uint64_t nw, real_t const *omega_dev,
                  real_t const *g_dev, real_t const *tau_dev, double mu_dir, double mu_dif,
                  real_t const *alb_dir_dev, real_t const *alb_dif_dev, double tsi,
                  real_t const *solar_dev, real_t *flux_up_dev, real_t *flux_down_dev)
{
    if (num_levels < 2 || num_levels > GRT_MAX_LAYERS + 1) return fail("bad level count");
    cudaStream_t st;
    GRT_TRY(stream_of(device, &st));
    LayerDensity dens{};
    return launch_sw<false>(device, st, num_levels, 0., 0., nw, dens, omega_dev, g_dev, tau_dev, mu_dir,
                            mu_dif, alb_dir_dev, alb_dif_dev, tsi, solar_dev, flux_up_dev,
                            flux_down_dev, 0, nw - 1);
}

int grt_integrate_flux(int device, int num_levels, uint64_t nw, double dw,
                       real_t const *flux_dev, real_t *out)
{
    if (num_levels < 1 || num_levels > GRT_MAX_LAYERS + 1) return fail("bad level count");
    cudaStream_t st;
    GRT_TRY(stream_of(device, &st));
    double *buf;
    GRT_TRY(out_buffer(device, &buf));
    GRT_CUDA(cudaMemsetAsync(buf, 0, sizeof(double)*num_levels, st));
    unsigned bx = (unsigned)((nw + 255)/256);
    if (bx > 64) bx = 64;
    if (bx < 1) bx = 1;
    integrate_kernel<<<dim3(bx, (unsigned)num_levels), 256, 0, st>>>(nw, dw, flux_dev, buf);
    GRT_CUDA(cudaGetLastError());
    double host[GRT_MAX_LAYERS + 1];
    GRT_CUDA(cudaMemcpyAsync(host, buf, sizeof(double)*num_levels, cudaMemcpyDeviceToHost, st));
    GRT_CUDA(cudaStreamSynchronize(st));
    for (int i=0; i<num_levels; ++i) out[i] = (real_t)host[i];
    return 0;
}

int grt_integrate_flux_noalloc(int device, int num_levels, uint64_t nw, double dw,
                               real_t const *flux_dev, real_t *out_dev)
{
    if (flux_dev == nullptr || out_dev == nullptr) return fail("null device pointer");
    if (num_levels < 1) return fail("bad level count %d", num_levels);
    if (num_levels > kNoAllocMaxLevels)
    {
        return fail("grt_integrate_flux_noalloc: %d levels exceed the static scratch array "
                    "(GRT_NOALLOC_MAX_LEVELS = %d); use grt_integrate_flux", num_levels,
                    kNoAllocMaxLevels);
    }
    if (nw < 1 || nw > (uint64_t)INT32_MAX) return fail("spectral grid size unsupported");
    cudaStream_t st;
    GRT_TRY(stream_of(device, &st));          // makes `device` current
    double *scratch;
    GRT_TRY(noalloc_scratch(device, &scratch));
    unsigned bx = (unsigned)((nw + 255)/256);
    if (bx > (unsigned)kNoAllocBlocks) bx = (unsigned)kNoAllocBlocks;
    if (bx < 1) bx = 1;
    integrate_partial_kernel<<<dim3(bx, (unsigned)num_levels), 256, 0, st>>>(nw, dw, flux_dev,
                                                                             scratch);
    integrate_finish_kernel<<<(unsigned)num_levels, 32, 0, st>>>((int)bx, scratch, out_dev);
    GRT_CUDA(cudaGetLastError());
    return 0;
}

int grt_lw_column(int device, int num_levels, double w0, double dw, uint64_t nw,
                  real_t const *p_atm, double t_surf, real_t const *t_layers,
                  real_t const *t_levels, real_t const *tau_gas_dev, real_t const *emis_dev,
                  uint64_t first, uint64_t last, real_t *up, real_t *down)
{
    if (first > last || last >= nw) return fail("integration window outside the grid");
    if (num_levels < 2 || num_levels > GRT_MAX_LAYERS + 1) return fail("bad level count");
    cudaStream_t st;
    GRT_TRY(stream_of(device, &st));
    LevelTemps temps;
    for (int i=0; i<num_levels-1; ++i) temps.t_layers[i] = t_layers[i];
    for (int i=0; i<num_levels; ++i) temps.t_levels[i] = t_levels[i];
    LayerDensity dens;
    number_densities(num_levels - 1, p_atm, &dens);
    double *buf;
    GRT_TRY(out_buffer(device, &buf));
    GRT_CUDA(cudaMemsetAsync(buf, 0, sizeof(double)*2*num_levels, st));
    unsigned const blocks = (unsigned)((nw + kSolverThreads - 1)/kSolverThreads);
    lw_kernel<true><<<blocks, kSolverThreads, 0, st>>>(num_levels, w0, dw, nw, t_surf, temps, dens,
                                                       tau_gas_dev, nullptr, emis_dev, buf,
                                                       buf + num_levels, first, last);
    GRT_CUDA(cudaGetLastError());
    return fetch_levels(st, buf, num_levels, up, down);
}

int grt_sw_column(int device, int num_levels, double w0, double dw, uint64_t nw,
                  real_t const *p_atm, real_t const *tau_gas_dev, double mu_dir, double mu_dif,
                  real_t const *alb_dir_dev, real_t const *alb_dif_dev, double tsi,
                  real_t const *solar_dev, uint64_t first, uint64_t last, real_t *up,
                  real_t *down)
{
    if (first > last || last >= nw) return fail("integration window outside the grid");
    if (num_levels < 2 || num_levels > GRT_MAX_LAYERS + 1) return fail("bad level count");
    cudaStream_t st;
    GRT_TRY(stream_of(device, &st));
    LayerDensity dens;
    number_densities(num_levels - 1, p_atm, &dens);
    double *buf;
    GRT_TRY(out_buffer(device, &buf));
    GRT_CUDA(cudaMemsetAsync(buf, 0, sizeof(double)*2*num_levels, st));
    GRT_TRY(launch_sw<true>(device, st, num_levels, w0, dw, nw, dens, nullptr, nullptr, tau_gas_dev,
                            mu_dir, mu_dif, alb_dir_dev, alb_dif_dev, tsi, solar_dev, buf,
                            buf + num_levels, first, last));
    return fetch_levels(st, buf, num_levels, up, down);
}

}  // extern "C"
