"""Shared parity harness: the same column through the GPU library and the oracle.

The checker is the compiled reference (oracle/_ref/libgrtcode_ref.so) when it
has been built, else the C restatement (oracle/libgrt_oracle.so).
"""
from __future__ import annotations

import numpy as np

from grtcode_b200 import api, synth
from oracle import ref as R


def checker_name() -> str:
    return "reference" if R.available() else "port"


def relative_error(got: np.ndarray, want: np.ndarray) -> np.ndarray:
    """|got - want| / max(|got|, |want|) with 0/0 = 0 (the reference's own
    check_floating_point metric, testing_harness/src/test_harness.c:10-30)."""
    scale = np.maximum(np.abs(got), np.abs(want))
    diff = np.abs(got - want)
    out = np.zeros_like(diff)
    np.divide(diff, scale, out=out, where=scale > 0)
    return out


def build_dataset(root, w0, wn, lines, seed, cfcs=(), cias=()):
    return synth.make_dataset(root, w0, wn, lines, seed=seed, cfcs=cfcs, cias=cias)


def oracle_column(ds, w0, wn, dw, col, molecules, use_ctm=True, threads=None, solvers=True,
                  line_window=(None, None), method=R.LINE_SAMPLE, fast=False):
    """Reference pipeline on the host: tau_gas, then (optionally) LW + SW fluxes.
    fast: the reference's own -DFAST build (no range checks / trapped exceptions)."""
    r = R.Ref(fast=fast, threads=threads)
    g = r.grid(w0, wn, dw)
    levels = len(col.p)
    go = R.RefGasOptics(r, levels, g, ds.hitran, ds.h2o_ctm_dir if use_ctm else None,
                        ds.o3_ctm if use_ctm else None, method=method)
    for m in molecules:
        go.add_molecule(m, *line_window)
    for c, path in ds.cfc.items():
        go.add_cfc(c, path)
    for (a, b), path in ds.cia.items():
        go.add_cia(a, b, path)
    for m in molecules:
        go.set_ppmv("mol", m, col.ppmv[m])
    for c in ds.cfc:
        go.set_ppmv("cfc", c, col.cfc_ppmv[c])
    if ds.cia:
        for c in (synth.CIA_N2, synth.CIA_O2):
            go.set_ppmv("cia", c, col.cia_ppmv[c])
    gas = R.RefOptics(r, levels - 1, g)
    go.optical_depth(col.p, col.t, gas)
    out = {"tau": gas.tau.copy(), "n": int(g.n)}
    if solvers:
        ray = R.RefOptics(r, levels - 1, g)
        R.rayleigh(r, ray, col.p)
        tot = R.add_optics(r, [gas, ray])
        n = int(g.n)
        emis = np.full(n, col.emissivity)
        alb = np.full(n, col.albedo)
        up, dn = R.lw_fluxes(r, tot, col.t_surf, col.t_layer, col.t, emis)
        out["lw_up"], out["lw_dn"] = up, dn
        sol = R.solar_flux(r, g, ds.solar)
        up, dn = R.sw_fluxes(r, tot, col.mu0, 0.5, alb, alb, col.tsi, sol)
        out["sw_up"], out["sw_dn"] = up, dn
        out["solar"] = sol
        out["tau_total"] = tot.tau.copy()
        out["omega_total"] = tot.omega.copy()
        tot.close()
        ray.close()
    gas.close()
    go.close()
    return out


def gpu_column(ds, w0, wn, dw, col, molecules, use_ctm=True, solvers=True, device=0,
               line_window=(None, None), method=api.LINE_SAMPLE):
    """Same pipeline through libgrtcode.so (the reference's C API on the GPU)."""
    g = api.spectral_grid(w0, wn, dw)
    levels = len(col.p)
    go = api.GasOptics(levels, g, ds.hitran, ds.h2o_ctm_dir if use_ctm else None,
                       ds.o3_ctm if use_ctm else None, device=device, method=method)
    for m in molecules:
        go.add_molecule(m, *line_window)
    for c, path in ds.cfc.items():
        go.add_cfc(c, path)
    for (a, b), path in ds.cia.items():
        go.add_cia(a, b, path)
    for m in molecules:
        go.set_ppmv("mol", m, col.ppmv[m])
    for c in ds.cfc:
        go.set_ppmv("cfc", c, col.cfc_ppmv[c])
    if ds.cia:
        for c in (synth.CIA_N2, synth.CIA_O2):
            go.set_ppmv("cia", c, col.cia_ppmv[c])
    gas = api.OpticsObject(levels - 1, g, device)
    go.optical_depth(col.p, col.t, gas)
    out = {"tau": gas.tau, "n": int(g.n), "stats": go.stats()}
    if solvers:
        n = int(g.n)
        ray = api.OpticsObject(levels - 1, g, device)
        api.rayleigh_scattering(ray, col.p)
        tot = api.add_optics([gas, ray])
        emis = np.full(n, col.emissivity)
        alb = np.full(n, col.albedo)
        lw = api.LongwaveSolver(levels, g, device)
        out["lw_up"], out["lw_dn"] = lw.fluxes(tot, col.t_surf, col.t_layer, col.t, emis)
        out["lw_bb_up"], out["lw_bb_dn"] = lw.column(gas, col.p, col.t_surf, col.t_layer, col.t,
                                                     emis)
        sol = api.solar_flux(g, ds.solar)
        sw = api.ShortwaveSolver(levels, g, device)
        out["sw_up"], out["sw_dn"] = sw.fluxes(tot, col.mu0, 0.5, alb, alb, col.tsi, sol)
        out["sw_bb_up"], out["sw_bb_dn"] = sw.column(gas, col.p, col.mu0, 0.5, alb, alb, col.tsi,
                                                     sol)
        out["solar"] = sol
        tau_t, om_t = tot.arrays(True, True, False)
        out["tau_total"], out["omega_total"] = tau_t, om_t
        lw.close()
        sw.close()
        tot.close()
        ray.close()
    gas.close()
    go.close()
    return out


def run_column_pair(root, w0, wn, dw, num_levels, lines, seed, cfcs=(), cias=(), use_ctm=True,
                    oracle_fast=False):
    """One synthetic column through both implementations; returns error metrics."""
    ds = build_dataset(root, w0, wn, lines, seed, cfcs, cias)
    rng = np.random.default_rng(seed)
    mols = tuple(lines.keys())
    col = synth.make_column(num_levels, rng, molecules=mols, cfcs=tuple(ds.cfc.keys()))
    want = oracle_column(ds, w0, wn, dw, col, mols, use_ctm, fast=oracle_fast)
    got = gpu_column(ds, w0, wn, dw, col, mols, use_ctm)
    rel = relative_error(got["tau"], want["tau"])
    bb = lambda f: R.broadband(f, dw)
    res = {
        "checker": checker_name(),
        "tau_max_rel": float(rel.max()),
        "tau_argmax": np.unravel_index(int(rel.argmax()), rel.shape),
        "tau_num_gt_1e9": int((rel > 1e-9).sum()),
        "solar_max_rel": float(relative_error(got["solar"], want["solar"]).max()),
        "lw_spec_max_rel": float(max(relative_error(got["lw_up"], want["lw_up"]).max(),
                                     relative_error(got["lw_dn"], want["lw_dn"]).max())),
        "sw_spec_max_rel": float(max(relative_error(got["sw_up"], want["sw_up"]).max(),
                                     relative_error(got["sw_dn"], want["sw_dn"]).max())),
        "lw_max_abs": float(max(np.abs(bb(got["lw_up"]) - bb(want["lw_up"])).max(),
                                np.abs(bb(got["lw_dn"]) - bb(want["lw_dn"])).max(),
                                np.abs(got["lw_bb_up"] - bb(want["lw_up"])).max(),
                                np.abs(got["lw_bb_dn"] - bb(want["lw_dn"])).max())),
        "sw_max_abs": float(max(np.abs(bb(got["sw_up"]) - bb(want["sw_up"])).max(),
                                np.abs(bb(got["sw_dn"]) - bb(want["sw_dn"])).max(),
                                np.abs(got["sw_bb_up"] - bb(want["sw_up"])).max(),
                                np.abs(got["sw_bb_dn"] - bb(want["sw_dn"])).max())),
        "lw_up_toa": float(bb(want["lw_up"])[0]),
        "sw_dn_sfc": float(bb(want["sw_dn"])[-1]),
        "stats": got["stats"],
        "got": got,
        "want": want,
    }
    return res
