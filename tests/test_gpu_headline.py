"""Parity on the headline workload's own regime.

bench.py's step is BASELINE config 2: the CIRC-1 column, seven HITRAN molecules + H2O/O3
continua + three CIA pairs + three CFCs at 0.01 cm-1 on LW 1-3250 and SW 1-50000 cm-1.  The
oracle cannot run that column in a test (6e11 Voigt samples), so the same inputs go through
both sides on windows spread over the two bands -- including > 3000, ~10000, ~30000 cm-1 and
the last cm-1 below 50000, where the Doppler cores are hundreds of grid points wide and the
wing sweep's R1 class and the core kernel do most of the work -- for all three optical-depth
methods; then BASELINE config 1 at its full size against the compiled reference, and the
headless driver against the reference's column loop.

Tolerances (north_star): tau within 1e-9 relative per wavenumber, broadband fluxes within
0.01 W m-2.
"""
import os
import subprocess

import numpy as np
import pytest

from grtcode_b200 import api, column_io, synth
from grtcode_b200 import workloads as W
from tests import parity

pytestmark = pytest.mark.gpu
TAU_RTOL = 1e-9
FLUX_ATOL = 0.01
HERE = os.path.dirname(os.path.abspath(__file__))
BIN = os.environ.get("GRT_LBL_BIN", os.path.join(os.path.dirname(HERE), "grtcode_b200", "grt_lbl"))

# (band, window centre [cm-1]): 60 cm-1 windows (wider than the 25 cm-1 cut-off on both sides,
# so that tiles wholly inside a window, window ends and grid ends all occur)
WINDOWS = [("lw", 31.), ("lw", 667.), ("lw", 1600.), ("lw", 3220.),
           ("sw", 3400.), ("sw", 5300.), ("sw", 10000.), ("sw", 14500.), ("sw", 22000.),
           ("sw", 30000.), ("sw", 41000.), ("sw", 49970.)]
WIDTH = 60.


def _window_band(band, centre):
    """A WIDTH cm-1 piece of the band with the band's mean line density per molecule."""
    w0 = max(band.w0, centre - WIDTH/2.)
    wn = min(band.wn, w0 + WIDTH)
    w0 = wn - WIDTH
    frac = WIDTH/(band.wn - band.w0)
    lines = {m: max(2, int(round(n*frac))) for m, n in band.lines.items()}
    return W.Band(f"{band.name}_{int(centre)}", w0, wn, band.dw, lines)


def _histogram(go_stats, hist):
    pts = hist["tile_points"]
    parts = [f"{k} {hist[k]}" for k in ("fast", "far_edge", "near", "region1", "lorentz",
                                         "lorentz_edge")]
    return (f"line-tile visits ({pts}-point tiles): " + ", ".join(parts) +
            f"; core samples <= {go_stats['core_samples']} of {go_stats['samples']}")


def _gpu_window(sb, col, method, solvers):
    """gpu_column with the class histogram of the sweep switched on."""
    os.environ["GRT_CLASS_HIST"] = "1"
    try:
        ds = sb.dataset
        g = api.spectral_grid(sb.w0, sb.wn, sb.dw)
        levels = len(col.p)
        go = api.GasOptics(levels, g, ds.hitran, ds.h2o_ctm_dir, ds.o3_ctm, method=method)
        for m in W.MOLECULES:
            go.add_molecule(m)
        for c, path in ds.cfc.items():
            go.add_cfc(c, path)
        for (a, b), path in ds.cia.items():
            go.add_cia(a, b, path)
        for m in W.MOLECULES:
            go.set_ppmv("mol", m, col.ppmv[m])
        for c in ds.cfc:
            go.set_ppmv("cfc", c, col.cfc_ppmv[c])
        for c in (synth.CIA_N2, synth.CIA_O2):
            go.set_ppmv("cia", c, col.cia_ppmv[c])
        gas = api.OpticsObject(levels - 1, g)
        go.optical_depth(col.p, col.t, gas)
        out = {"tau": gas.tau, "stats": go.stats()}
        if method == api.LINE_SAMPLE:
            out["hist"] = go.class_histogram()
        if solvers:
            n = int(g.n)
            if sb.name.startswith("lw"):
                lw = api.LongwaveSolver(levels, g)
                out["up"], out["dn"] = lw.column(gas, col.p, col.t_surf, col.t_layer, col.t,
                                                 np.full(n, col.emissivity))
                lw.close()
            else:
                sw = api.ShortwaveSolver(levels, g)
                alb = np.full(n, col.albedo)
                out["up"], out["dn"] = sw.column(gas, col.p, col.mu0, 0.5, alb, alb, col.tsi,
                                                 api.solar_flux(g, ds.solar))
                sw.close()
        gas.close()
        go.close()
        return out
    finally:
        del os.environ["GRT_CLASS_HIST"]


@pytest.mark.parametrize("band_name,centre", WINDOWS,
                         ids=[f"{b}{int(c)}" for b, c in WINDOWS])
def test_config2_windows_all_methods(built, tmp_path, band_name, centre):
    """The bench's inputs (workloads.circ1_column, circ1_bands line densities, all 7 molecules
    + continua + 3 CIA pairs, + 3 CFCs on the LW band) on one 60 cm-1 window at 0.01 cm-1:
    tau of the three methods and the band's broadband fluxes against oracle/_ref."""
    from oracle import ref as R
    lw, sw = W.circ1_bands(0.01, 1.)
    band = lw if band_name == "lw" else sw
    sb = _window_band(band, centre)
    W.materialise(sb, str(tmp_path), seed=500 + int(centre))
    col = W.circ1_column()
    totals = {}
    for method, name in ((api.LINE_SAMPLE, "line_sample"), (api.WAVENUMBER_SWEEP, "wavenumber_sweep"),
                         (api.LINE_SWEEP, "line_sweep")):
        want = parity.oracle_column(sb.dataset, sb.w0, sb.wn, sb.dw, col, W.MOLECULES,
                                    solvers=False, method=method)
        got = _gpu_window(sb, col, method, solvers=(method == api.LINE_SAMPLE))
        rel = parity.relative_error(got["tau"], want["tau"])
        print(f"{sb.name} [{sb.w0:g}, {sb.wn:g}] {name}: tau max rel {rel.max():.3e} "
              f"({int((rel > TAU_RTOL).sum())} points > 1e-9), {sb.num_lines} lines")
        assert rel.max() <= TAU_RTOL, name
        if method == api.LINE_SAMPLE:
            print("   " + _histogram(got["stats"], got["hist"]))
            totals = got["hist"]
            # broadband fluxes of this band on the window, reference solver chain
            r = R.Ref()
            g = r.grid(sb.w0, sb.wn, sb.dw)
            layers = len(col.p) - 1
            gas = R.RefOptics(r, layers, g)
            gas.tau[:] = want["tau"]
            ray = R.RefOptics(r, layers, g)
            R.rayleigh(r, ray, col.p)
            tot = R.add_optics(r, [gas, ray])
            n = int(g.n)
            if band_name == "lw":
                up, dn = R.lw_fluxes(r, tot, col.t_surf, col.t_layer, col.t,
                                     np.full(n, col.emissivity))
            else:
                sol = R.solar_flux(r, g, sb.dataset.solar)
                alb = np.full(n, col.albedo)
                up, dn = R.sw_fluxes(r, tot, col.mu0, 0.5, alb, alb, col.tsi, sol)
            err = max(np.abs(got["up"] - R.broadband(up, sb.dw)).max(),
                      np.abs(got["dn"] - R.broadband(dn, sb.dw)).max())
            print(f"   broadband |dF| {err:.3e} W m-2 (up TOA {got['up'][0]:.6f}, "
                  f"down surface {got['dn'][-1]:.6f})")
            assert err <= FLUX_ATOL
            for o in (tot, ray, gas):
                o.close()
    assert sum(totals[k] for k in ("fast", "far_edge", "near", "region1", "lorentz",
                                   "lorentz_edge")) > 0


def test_every_line_class_is_exercised_by_the_windows(built, tmp_path):
    """The classes of the wing sweep (fast / far edge / near / region 1 / Lorentz) over the
    twelve windows together: none is empty, i.e. the parity test above covers all of them."""
    from oracle import ref as R
    lw, sw = W.circ1_bands(0.01, 1.)
    col = W.circ1_column()
    total = {}
    for band_name, centre in WINDOWS:
        band = lw if band_name == "lw" else sw
        sb = _window_band(band, centre)
        W.materialise(sb, str(tmp_path/f"{band_name}{int(centre)}"), seed=500 + int(centre))
        got = _gpu_window(sb, col, api.LINE_SAMPLE, solvers=False)
        for k, v in got["hist"].items():
            total[k] = total.get(k, 0) + v
        total["core"] = total.get("core", 0) + got["stats"]["core_samples"]
    print({k: v for k, v in total.items() if k != "tile_points"})
    for k in ("fast", "far_edge", "near", "region1", "lorentz", "core"):
        assert total[k] > 0, k


def test_full_size_config1_against_the_compiled_reference(built, tmp_path):
    """BASELINE config 1 at its real size (60 levels, 1-3000 cm-1 at 0.1 cm-1, H2O + CO2 + O3,
    6.5e5 lines, 1.9e10 Voigt samples, continua on): tau of all 1.77e6 (layer, point) pairs and
    the LW / SW broadband fluxes against oracle/_ref -- under a minute of host time."""
    import time
    mols = (synth.H2O, synth.CO2, synth.O3)
    lines = {synth.H2O: 100_000, synth.CO2: 300_000, synth.O3: 250_000}
    w0, wn, dw, levels = 1., 3000., 0.1, 60
    ds = parity.build_dataset(str(tmp_path), w0, wn, lines, 31)
    col = synth.make_column(levels, np.random.default_rng(31), molecules=mols)
    t0 = time.perf_counter()
    want = parity.oracle_column(ds, w0, wn, dw, col, mols)
    t_ref = time.perf_counter() - t0
    got = parity.gpu_column(ds, w0, wn, dw, col, mols)
    assert got["stats"]["samples"] > 1.8e10
    rel = parity.relative_error(got["tau"], want["tau"])
    bb = lambda f: parity.R.broadband(f, dw)
    lw_err = max(np.abs(got["lw_bb_up"] - bb(want["lw_up"])).max(),
                 np.abs(got["lw_bb_dn"] - bb(want["lw_dn"])).max())
    sw_err = max(np.abs(got["sw_bb_up"] - bb(want["sw_up"])).max(),
                 np.abs(got["sw_bb_dn"] - bb(want["sw_dn"])).max())
    print(f"config 1 full size: tau max rel {rel.max():.3e} over {rel.size} values "
          f"({int((rel > TAU_RTOL).sum())} > 1e-9); LW |dF| {lw_err:.3e}, SW |dF| {sw_err:.3e} "
          f"W m-2; reference {t_ref:.1f} s; {got['stats']['samples']:.3e} samples")
    assert rel.max() <= TAU_RTOL
    assert lw_err <= FLUX_ATOL and sw_err <= FLUX_ATOL


def test_driver_matches_the_reference_column_loop(built, tmp_path):
    """grt_lbl against the reference's per-column chain (framework/src/driver.c:691-743:
    LONGWAVE_PASS, then SHORTWAVE_PASS only when cos(zenith) > 0) run column by column through
    oracle/_ref: a 2 x 4 column set with one night-side column, the -x/-X column range, and
    the heating rates formed from the reference's fluxes."""
    mols = (synth.H2O, synth.CO2, synth.O3)
    ds = synth.make_dataset(str(tmp_path/"data"), 500., 700., {m: 150 for m in mols}, seed=4,
                            cfcs=("CFC11",), cias=((0, 0), (1, 1)))
    cols = column_io.rfmip_like_columns(4, 2, 12, mols, tuple(ds.cfc.keys()))
    cols[0][2].mu0 = -0.2            # night side: no shortwave pass (driver.c:706)
    cols[1][1].mu0 = 0.
    path = str(tmp_path/"columns.grtcols")
    column_io.write_column_set(path, cols, ds, mols)
    out = str(tmp_path/"out.grtflux")
    args = [BIN, ds.hitran, ds.solar, path, "-o", out, "-w-lw", "500", "-W-lw", "700", "-r-lw",
            "0.1", "-w-sw", "500", "-W-sw", "700", "-r-sw", "0.1", "-x", "1", "-X", "3"]
    r = subprocess.run(args, capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "6 columns in" in r.stdout
    got = column_io.read_flux_set(out)
    bb = lambda f: parity.R.broadband(f, 0.1)
    worst = 0.
    for t in range(2):
        for c in range(4):
            col = cols[t][c]
            if c == 0:
                # outside -x/-X: untouched
                assert not got.rlu[t, c].any() and not got.rsd[t, c].any()
                continue
            import dataclasses
            sunlit = col if col.mu0 > 0. else dataclasses.replace(col, mu0=0.5)
            want = parity.oracle_column(ds, 500., 700., 0.1, sunlit, mols)
            errs = [np.abs(got.rlu[t, c] - bb(want["lw_up"])).max(),
                    np.abs(got.rld[t, c] - bb(want["lw_dn"])).max()]
            hr = built.heating_rates(col.p, bb(want["lw_up"]), bb(want["lw_dn"]))
            assert np.abs(got.hr_lw[t, c] - hr).max() <= 1e-6*max(1., np.abs(hr).max())
            if col.mu0 > 0.:
                errs += [np.abs(got.rsu[t, c] - bb(want["sw_up"])).max(),
                         np.abs(got.rsd[t, c] - bb(want["sw_dn"])).max()]
                hr = built.heating_rates(col.p, bb(want["sw_up"]), bb(want["sw_dn"]))
                assert np.abs(got.hr_sw[t, c] - hr).max() <= 1e-6*max(1., np.abs(hr).max())
            else:
                assert not got.rsu[t, c].any() and not got.rsd[t, c].any()
                assert not got.hr_sw[t, c].any()
            worst = max(worst, *errs)
    print(f"grt_lbl vs reference column loop: max |dF| {worst:.3e} W m-2")
    assert worst <= FLUX_ATOL
