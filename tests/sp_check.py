"""TEST INFRASTRUCTURE: parity of the single-precision build, run as its own process

    GRT_PRECISION=single python -m tests.sp_check

(precision is per process: grtcode_b200/api.py and oracle/ref.py pick libgrtcode_sp.so /
oracle/_ref/libgrtcode_ref_sp.so -- the unmodified reference compiled with its own
-DSINGLE_PRECISION -- when GRT_PRECISION=single).  Prints one JSON line per case;
tests/test_gpu_single.py asserts on them.  Tolerances (north_star): optical depth within 1e-4
relative per wavenumber, broadband fluxes within 0.01 W m-2.

No CIA in these cases: the reference's single build overflows there (its constant
c = 7.9e46 exceeds FLT_MAX, kernels.c:610-615, so every CIA optical depth is inf).
"""
import json
import sys
import tempfile

import numpy as np

from grtcode_b200 import api, synth
from tests import parity

CASES = [
    # name, w0, wn, dw, levels, lines, seed, cfcs
    ("lw_window_0.1", 500., 700., 0.1, 20, {synth.H2O: 300, synth.CO2: 400, synth.O3: 200}, 3, ()),
    ("config1_shape_0.1", 1., 3000., 0.1, 60, {synth.H2O: 2000, synth.CO2: 3000, synth.O3: 2000},
     1, ("CFC11", "CFC12")),
    ("fine_0.01_grid_edges", 1000., 1030., 0.01, 12,
     {synth.H2O: 150, synth.CO2: 150, synth.CH4: 100}, 7, ()),
    ("sw_10000_0.01", 9970., 10030., 0.01, 30, {synth.H2O: 300, synth.CO2: 300, synth.O2: 50}, 11, ()),
    ("sw_30000_0.1", 29900., 30100., 0.1, 30, {synth.H2O: 300, synth.O3: 300, synth.O2: 50}, 13, ()),
    ("sw_49900_0.1", 49800., 50000., 0.1, 30, {synth.H2O: 300, synth.O3: 300}, 17, ()),
]


def _sw_detail(res, dw):
    """Where a shortwave broadband difference comes from: spectral path vs fused path, and
    the largest relative differences of the inputs the solver sees."""
    got, want = res["got"], res["want"]
    bb = lambda f: parity.R.broadband(f, dw)
    f64 = lambda a: np.asarray(a, dtype=np.float64)
    d = {"spec_up": float(np.abs(bb(got["sw_up"]) - bb(want["sw_up"])).max()),
         "spec_dn": float(np.abs(bb(got["sw_dn"]) - bb(want["sw_dn"])).max()),
         "fused_up": float(np.abs(f64(got["sw_bb_up"]) - bb(want["sw_up"])).max()),
         "fused_dn": float(np.abs(f64(got["sw_bb_dn"]) - bb(want["sw_dn"])).max()),
         "bb_dn_toa": float(bb(want["sw_dn"])[0]),
         "tau_total_max_rel": float(parity.relative_error(f64(got["tau_total"]),
                                                          f64(want["tau_total"])).max()),
         "omega_total_max_rel": float(parity.relative_error(f64(got["omega_total"]),
                                                            f64(want["omega_total"])).max())}
    rel = parity.relative_error(f64(got["sw_dn"]), f64(want["sw_dn"]))
    lev, pt = np.unravel_index(int(rel.argmax()), rel.shape)
    d["dn_worst"] = {"level": int(lev), "point": int(pt), "rel": float(rel.max()),
                     "got": float(got["sw_dn"][lev, pt]), "want": float(want["sw_dn"][lev, pt])}
    absd = np.abs(f64(got["sw_dn"]) - f64(want["sw_dn"]))
    lev, pt = np.unravel_index(int(absd.argmax()), absd.shape)
    d["dn_worst_abs"] = {"level": int(lev), "point": int(pt), "abs": float(absd.max()),
                         "got": float(got["sw_dn"][lev, pt]), "want": float(want["sw_dn"][lev, pt]),
                         "tau": float(want["tau_total"][max(lev - 1, 0), pt])}
    return d


def main():
    assert api.SINGLE and parity.R.SINGLE, "run with GRT_PRECISION=single"
    api.load()
    root = tempfile.mkdtemp(prefix="grt_sp_")
    for name, w0, wn, dw, levels, lines, seed, cfcs in CASES:
        # the reference's -DFAST single build: see oracle/Makefile (the checked build leaves
        # stale values at the points where one of its float traps fires)
        res = parity.run_column_pair(f"{root}/{name}", w0, wn, dw, levels, lines, seed, cfcs=cfcs,
                                     oracle_fast=True)
        got, want = res["got"]["tau"].astype(np.float64), res["want"]["tau"].astype(np.float64)
        rel = parity.relative_error(got, want)
        finite = bool(np.isfinite(got).all() and np.isfinite(want).all())
        out = {"case": name, "dtype": str(res["got"]["tau"].dtype), "finite": finite,
               "tau_max_rel": float(rel.max()), "tau_p999_rel": float(np.quantile(rel, 0.999)),
               "tau_num_gt_1e4": int((rel > 1e-4).sum()), "tau_points": int(rel.size),
               "lw_max_abs": res["lw_max_abs"], "sw_max_abs": res["sw_max_abs"],
               "solar_max_rel": res["solar_max_rel"], "sw_spec_max_rel": res["sw_spec_max_rel"],
               "lw_spec_max_rel": res["lw_spec_max_rel"],
               "sw_detail": _sw_detail(res, dw),
               "lw_up_toa": res["lw_up_toa"], "samples": int(res["stats"]["samples"]),
               "wing_ns": int(res["stats"]["wing_ns"])}
        print("SPCHECK " + json.dumps(out), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
