"""GPU parity of the column solvers on the reference's own solver test cases.

longwave/test/test_longwave.c:100-209 and shortwave/test/test_shortwave.c:103-215 run
calculate_{lw,sw}_fluxes on a 1-50000 cm-1 grid at 10 cm-1 with MIN_NUM_LEVELS levels and
(a) random optics, (b) tau = 1e12 ("optically thick"), (c) tau = 1e-12 ("optically thin"),
(d) tau = log(DBL_MAX) - 12 ("strong absorption": exp(tau) close to overflow); they only
check the return code.  Here the same inputs go through libgrtcode.so on the GPU and through
the compiled reference, and the spectral fluxes are compared point by point; the layer
counts also cover the reference's limits (grtcode_config.h:51-54: 1 and 200 layers) and the
shortwave branches the random case never reaches (omega = 0: shortwave.c:114-123,
omega = 1 conservative: :169-179, tau below the empty-layer clamp: :137-158).

Which reference: the checked build (`oracle/_ref/libgrtcode_ref.so`) range-checks its
arguments and traps floating-point exceptions inside lw_flux / sw_flux (debug.h:193-248) and
returns an error code from the layer functions when one fires -- which lw_fluxes_kernel /
sw_fluxes_kernel ignore (longwave.c:254, shortwave.c:443), copying their uninitialised
per-thread buffers out (in practice a neighbouring point's fluxes).  That happens for the
last longwave point of this grid (50001 cm-1 > MAX_WAVENUMBER, longwave.c:151), for every
conservative-scattering layer (sqrt of a rounding-negative gamma1^2 - gamma2^2 raises
FE_INVALID, shortwave.c:130) and for the optically thin case (R rounds below zero,
shortwave.c:188).  The reference's own -DFAST build (`libgrtcode_ref_fast.so`) evaluates the
same formulas without the checks, so it is the oracle at every point; the checked build is
compared as well wherever its result is defined (i.e. equals its own FAST build).
"""
import numpy as np
import pytest

from grtcode_b200 import api
from oracle import ref as R

pytestmark = pytest.mark.gpu

SPEC_RTOL = 1e-9      # spectral fluxes, same relative bound north_star puts on tau
FLUX_ATOL = 0.01      # north_star: broadband fluxes within 0.01 W m-2
W0, WN, DW = 1., 50000., 10.          # MIN_WAVENUMBER, MAX_WAVENUMBER, MAX_RESOLUTION
STRONG = float(np.log(np.finfo(np.float64).max) - 12.)


def _both_builds(run):
    """run(ref) -> (up, down) through the FAST and the checked build of the reference."""
    fast = run(R.Ref(fast=True))
    checked = run(R.Ref())
    return fast, checked


def _check(case, layers, got, fast, checked, min_defined):
    up, dn = got
    assert _rel(up, fast[0]) <= SPEC_RTOL, (case, layers, "up", _rel(up, fast[0]))
    assert _rel(dn, fast[1]) <= SPEC_RTOL, (case, layers, "down", _rel(dn, fast[1]))
    if np.isfinite(fast[0]).all() and np.isfinite(fast[1]).all():
        assert np.abs(R.broadband(up, DW) - R.broadband(fast[0], DW)).max() <= FLUX_ATOL
        assert np.abs(R.broadband(dn, DW) - R.broadband(fast[1], DW)).max() <= FLUX_ATOL
    # the checked build, wherever it wrote a result
    with np.errstate(invalid="ignore"):
        defined = np.all((checked[0] == fast[0]) & (checked[1] == fast[1]), axis=0)
    frac = float(defined.mean())
    print(f"{case}/{layers}: checked build defined on {frac:.4f} of the grid")
    assert frac >= min_defined, (case, layers, frac)
    if defined.any():
        assert _rel(up[:, defined], checked[0][:, defined]) <= SPEC_RTOL
        assert _rel(dn[:, defined], checked[1][:, defined]) <= SPEC_RTOL


def _rel(got, want):
    """Relative difference; points where both sides are the same non-finite value (or both
    NaN) count as equal, a finite/non-finite mismatch as infinite."""
    got = np.asarray(got)
    want = np.asarray(want)
    same_nonfinite = (~np.isfinite(got)) & (~np.isfinite(want)) & \
                     ((np.isnan(got) & np.isnan(want)) | (got == want))
    mismatch = np.isfinite(got) != np.isfinite(want)
    with np.errstate(invalid="ignore", divide="ignore"):
        scale = np.maximum(np.abs(got), np.abs(want))
        d = np.where(scale > 0, np.abs(got - want)/scale, 0.)
    d = np.where(same_nonfinite, 0., d)
    d = np.where(mismatch, np.inf, d)
    return float(d.max())


def _atmosphere(layers):
    # test_longwave.c:75-86 (a 1 K per layer ramp), levels half a degree either side
    t_lay = 250. + 1.*np.arange(layers)
    t_lev = np.concatenate([t_lay - 0.5, [t_lay[-1] + 0.5]])
    return t_lay, t_lev, float(t_lev[-1] + 2.)


def _tau_case(case, layers, n, rng):
    if case == "simple":
        return rng.random((layers, n))
    if case == "thick":
        return np.full((layers, n), 1.e12)
    if case == "thin":
        return np.full((layers, n), 1.e-12)
    if case == "strong":
        return np.full((layers, n), STRONG)
    if case == "decades":            # every magnitude between the clamps, layer by layer
        e = np.linspace(-14., 4., n)[None, :] + 0.37*np.arange(layers)[:, None]
        return 10.**np.minimum(e, 6.)
    raise ValueError(case)


@pytest.mark.parametrize("layers", [1, 7, 200])
@pytest.mark.parametrize("case", ["simple", "thick", "thin", "strong", "decades"])
def test_longwave_reference_cases(built, case, layers):
    r = R.Ref()
    n = int(r.grid(W0, WN, DW).n)
    rng = np.random.default_rng(11)
    tau = _tau_case(case, layers, n, rng)
    t_lay, t_lev, t_surf = _atmosphere(layers)
    emis = np.full(n, 0.98)

    def run(ref):
        ro = R.RefOptics(ref, layers, ref.grid(W0, WN, DW))
        ro.tau[:] = tau
        out = R.lw_fluxes(ref, ro, t_surf, t_lay, t_lev, emis)
        ro.close()
        return out

    fast, checked = _both_builds(run)
    g = api.spectral_grid(W0, WN, DW)
    go = api.OpticsObject(layers, g)
    go.update(tau, np.zeros_like(tau), np.zeros_like(tau))
    lw = api.LongwaveSolver(layers + 1, g)
    got = lw.fluxes(go, t_surf, t_lay, t_lev, emis)
    lw.close()
    go.close()
    # every point but the one above MAX_WAVENUMBER is defined in the checked build
    _check(case, layers, got, fast, checked, min_defined=(n - 1)/n)


def _sw_optics(case, layers, n, rng):
    """(tau, omega, g) of the shortwave cases: test_shortwave.c:70-79 scales omega and g
    by 0.85; the extra cases pin the branches of shortwave.c:114-179."""
    g = 0.85*rng.random((layers, n))
    omega = 0.85*rng.random((layers, n))
    tau = rng.random((layers, n))
    if case in ("thick", "thin", "strong", "decades"):
        tau = _tau_case(case, layers, n, rng)
    elif case == "pure_absorption":
        omega[:] = 0.
    elif case == "conservative":
        omega[:] = 1.
    elif case == "mixed_branches":
        omega[:, 0::4] = 0.
        omega[:, 1::4] = 1.
        tau[:, 2::8] = 1.e-9
        g[:, 3::8] = 0.
    elif case != "simple":
        raise ValueError(case)
    return tau, omega, g


@pytest.mark.parametrize("kernel", ["thread", "staged"])
@pytest.mark.parametrize("layers", [1, 7, 200])
@pytest.mark.parametrize("case", ["simple", "thick", "thin", "strong", "decades",
                                  "pure_absorption", "conservative", "mixed_branches"])
def test_shortwave_reference_cases(built, case, layers, kernel, monkeypatch):
    """kernel: sw_kernel (one thread per wavenumber, the default) and sw_staged_kernel
    (GRT_SW_KERNEL=staged: layer values staged through an L2-resident slab)."""
    monkeypatch.setenv("GRT_SW_KERNEL", kernel)
    r = R.Ref()
    n = int(r.grid(W0, WN, DW).n)
    rng = np.random.default_rng(13)
    tau, omega, g = _sw_optics(case, layers, n, rng)
    alb = np.full(n, 0.35)
    solar = np.full(n, 1./(DW*(n - 1)))         # unit integral, flat
    mu_dir, mu_dif, tsi = 0.8, 0.5, 1361.

    def run(ref):
        ro = R.RefOptics(ref, layers, ref.grid(W0, WN, DW))
        ro.tau[:], ro.omega[:], ro.g[:] = tau, omega, g
        out = R.sw_fluxes(ref, ro, mu_dir, mu_dif, alb, alb, tsi, solar)
        ro.close()
        return out

    fast, checked = _both_builds(run)
    grid = api.spectral_grid(W0, WN, DW)
    go = api.OpticsObject(layers, grid)
    go.update(tau, omega, g)
    sw = api.ShortwaveSolver(layers + 1, grid)
    got = sw.fluxes(go, mu_dir, mu_dif, alb, alb, tsi, solar)
    sw.close()
    go.close()
    well_defined = {"simple": 1., "thick": 1., "strong": 1., "decades": 0., "thin": 0.,
                    "pure_absorption": 1., "conservative": 0., "mixed_branches": 0.}[case]
    _check(case, layers, got, fast, checked, min_defined=well_defined)


def test_rayleigh_and_add_optics_match_the_reference(built):
    """rayleigh_scattering (rayleigh.c:100-144) and add_optics (optics.c:84-148) with three
    operands, one of them non-scattering."""
    r = R.Ref()
    layers = 9
    rg = r.grid(W0, WN, DW)
    n = int(rg.n)
    rng = np.random.default_rng(17)
    p = np.linspace(0.5, 1013., layers + 1)
    a_tau, a_om, a_g = rng.random((layers, n)), 0.9*rng.random((layers, n)), rng.random((layers, n))
    b_tau = rng.random((layers, n))

    ra, rb, rr = (R.RefOptics(r, layers, rg) for _ in range(3))
    ra.tau[:], ra.omega[:], ra.g[:] = a_tau, a_om, a_g
    rb.tau[:] = b_tau
    R.rayleigh(r, rr, p)
    want = R.add_optics(r, [ra, rb, rr])
    want_arr = (want.tau.copy(), want.omega.copy(), want.g.copy(), rr.tau.copy())
    for o in (want, ra, rb, rr):
        o.close()

    g = api.spectral_grid(W0, WN, DW)
    ga, gb, gr = (api.OpticsObject(layers, g) for _ in range(3))
    ga.update(a_tau, a_om, a_g)
    gb.update(b_tau, np.zeros_like(b_tau), np.zeros_like(b_tau))
    api.rayleigh_scattering(gr, p)
    tot = api.add_optics([ga, gb, gr])
    tau, om, gg = tot.arrays(True, True, True)
    ray_tau = gr.tau
    for o in (tot, ga, gb, gr):
        o.close()
    assert _rel(ray_tau, want_arr[3]) <= 1e-14
    assert _rel(tau, want_arr[0]) <= 1e-14
    assert _rel(om, want_arr[1]) <= 1e-13
    assert _rel(gg, want_arr[2]) <= 1e-13


def test_staged_shortwave_column_matches_the_thread_kernel(built, monkeypatch):
    """The fused column call (Rayleigh mixed in, broadband sums in the kernel) through both
    shortwave kernels: the same expressions on the same operands, so the broadband fluxes
    agree to the order of their sums."""
    layers, n = 54, 20001
    grid = api.spectral_grid(1000., 1200., 0.01)
    rng = np.random.default_rng(3)
    tau = np.exp(rng.uniform(np.log(1e-6), np.log(30.), (layers, n)))
    p = np.exp(np.linspace(np.log(0.02), np.log(1000.), layers + 1))
    alb = np.full(n, 0.2)
    solar = np.full(n, 1./(0.01*(n - 1)))
    gas = api.OpticsObject(layers, grid)
    gas.update(tau, np.zeros_like(tau), np.zeros_like(tau))
    out = {}
    for kernel in ("thread", "staged"):
        monkeypatch.setenv("GRT_SW_KERNEL", kernel)
        sw = api.ShortwaveSolver(layers + 1, grid)
        out[kernel] = [np.array(x, dtype=np.float64)
                       for x in sw.column(gas, p, 0.7, 0.5, alb, alb, 1361., solar,
                                          window=(100, n - 200))]
        sw.close()
    gas.close()
    assert out["thread"][1].max() > 100.         # the direct beam at the top: tsi*mu over the window
    for a, b in zip(out["thread"], out["staged"]):
        assert np.all(np.isfinite(a)) and a.max() > 0.
        np.testing.assert_allclose(b, a, rtol=1e-12 if api.DTYPE == np.float64 else 1e-6, atol=0.)

