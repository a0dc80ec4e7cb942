"""TEST INFRASTRUCTURE: randomised parity soak of the GPU path against the compiled reference
(oracle/_ref): columns, bands, grids, molecule sets, line densities, methods and continuum /
CFC / CIA switches drawn from a seed, every case through both sides (tests/parity.py), the
worst optical-depth and broadband-flux differences collected.  A hunt for rare cases the fixed
test suite does not hold (line-class boundaries, window ends at grid ends, near-empty tiles,
several lines per grid point); the tolerances are north_star's: tau 1e-9 relative, fluxes
0.01 W m-2.

    python tools/parity_soak.py [--cases 120] [--seed 1] [--seconds 600] [--out FILE]
    GRT_PRECISION=single python tools/parity_soak.py ...     # the single-precision pair:
        line_sample only, no CIA (the reference's single build overflows there), tau 1e-4
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from grtcode_b200 import api, synth        # noqa: E402
from tests import parity                   # noqa: E402

TAU_TOL = 1e-4 if api.SINGLE else 1e-9
ALL_MOLS = (synth.H2O, synth.CO2, synth.O3, synth.N2O, synth.CO, synth.CH4, synth.O2)


def draw_case(rng: np.random.Generator) -> dict:
    dw = float(rng.choice([0.1, 0.1, 0.05, 0.02, 0.01, 0.01]))
    width = float(rng.choice([8., 20., 40., 60.])) if dw <= 0.02 else float(rng.choice([40., 120., 300.]))
    where = rng.random()
    if where < 0.15:
        w0 = 1.                                        # the grid's lower end (window clipping)
    elif where < 0.3:
        w0 = 50000. - width                            # the upper end
    elif where < 0.65:
        w0 = float(np.floor(rng.uniform(20., 3200.)))  # longwave
    else:
        w0 = float(np.floor(rng.uniform(3200., 49000.)))
    wn = min(w0 + width, 50000.)
    nmol = int(rng.integers(1, 5))
    mols = [int(m) for m in rng.choice(ALL_MOLS, size=nmol, replace=False)]
    density = float(rng.choice([0.2, 2., 10., 40.]))  # lines per cm-1 and molecule
    lines = {m: max(1, int(density*width*rng.uniform(0.3, 1.))) for m in mols}
    method = int(rng.choice([2, 2, 2, 0, 1]))
    if method == 0 and wn - w0 < 27.:
        # wavenumber_sweep on a grid of fewer than 26 bins is undefined in the reference:
        # `j >= (bins.n-1)-nbin` (kernels.c:272) wraps for nbin = 25 and bins.w[] is read beyond
        # its end; there is nothing to compare with (the first soak run: 19 such grids, all
        # different from the GPU's result, which extends the remote range past the grid's end)
        method = 1
    case = {"w0": w0, "wn": wn, "dw": dw, "levels": int(rng.integers(2, 62)), "lines": lines,
            "method": method, "use_ctm": bool(rng.random() < 0.7),
            "cfcs": ("CFC11", "CFC12") if rng.random() < 0.25 else (),
            "cias": ((synth.CIA_N2, synth.CIA_N2), (synth.CIA_O2, synth.CIA_N2)) if rng.random() < 0.25 else (),
            "seed": int(rng.integers(1, 1 << 30))}
    if api.SINGLE:
        case["method"] = 2
        case["cias"] = ()
    return case


def _oracle_child(path, ds, c, col, mols):
    want = parity.oracle_column(ds, c["w0"], c["wn"], c["dw"], col, mols, c["use_ctm"],
                                method=c["method"], fast=True)
    np.savez(path, **{k: v for k, v in want.items() if isinstance(v, np.ndarray)})


def run_case(root: str, c: dict, isolate: bool = True) -> dict:
    """isolate: the reference in a forked child (this tool, whose parent never enters the
    reference's OpenMP runtime); False = in this process (pytest, where earlier tests have
    started OpenMP threads that a forked child would wait for forever)."""
    ds = parity.build_dataset(root, c["w0"], c["wn"], c["lines"], c["seed"], c["cfcs"], c["cias"])
    mols = tuple(c["lines"].keys())
    col = synth.make_column(c["levels"], np.random.default_rng(c["seed"]), molecules=mols,
                            cfcs=tuple(ds.cfc.keys()))
    # the reference in a child process: its binned methods read out of bounds on some grids
    # (kernels.c:214-262 with bracket() on empty ranges) and may take the process with them
    import multiprocessing as mp
    path = os.path.join(root, "want.npz")
    exitcode = 0
    if isolate:
        child = mp.get_context("fork").Process(target=_oracle_child, args=(path, ds, c, col, mols))
        child.start()
        child.join()
        exitcode = child.exitcode
    else:
        _oracle_child(path, ds, c, col, mols)
    got = parity.gpu_column(ds, c["w0"], c["wn"], c["dw"], col, mols, c["use_ctm"],
                            method=c["method"])
    if exitcode != 0:
        return {"reference_crashed": int(exitcode), "finite": bool(np.isfinite(got["tau"]).all()),
                "tau_max_rel": 0., "tau_gt_1e9": 0, "flux_max_abs": 0., "points": 0,
                "samples": int(got["stats"]["samples"])}
    want = {k: v.astype(np.float64) for k, v in dict(np.load(path)).items()}
    got = {k: (np.asarray(v, dtype=np.float64) if isinstance(v, np.ndarray) else v)
           for k, v in got.items()}
    rel = parity.relative_error(got["tau"], want["tau"])
    bb = lambda f: parity.R.broadband(f, c["dw"])
    flux = max(float(np.abs(bb(got[k]) - bb(want[k])).max()) for k in ("lw_up", "lw_dn", "sw_up", "sw_dn"))
    fused = max(float(np.abs(got["lw_bb_up"] - bb(want["lw_up"])).max()),
                float(np.abs(got["lw_bb_dn"] - bb(want["lw_dn"])).max()),
                float(np.abs(got["sw_bb_up"] - bb(want["sw_up"])).max()),
                float(np.abs(got["sw_bb_dn"] - bb(want["sw_dn"])).max()))
    return {"tau_max_rel": float(rel.max()), "tau_gt_1e9": int((rel > TAU_TOL).sum()),
            "finite": bool(np.isfinite(got["tau"]).all()), "flux_max_abs": max(flux, fused),
            "points": int(rel.size), "samples": int(got["stats"]["samples"])}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", type=int, default=120)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--seconds", type=float, default=600.)
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    api.load()
    rng = np.random.default_rng(args.seed)
    root = tempfile.mkdtemp(prefix="grt_soak_")
    t0 = time.time()
    worst_tau, worst_flux, bad, done, points, crashed = 0., 0., [], 0, 0, []
    by_method = {0: 0, 1: 0, 2: 0}
    for k in range(args.cases):
        if time.time() - t0 > args.seconds:
            break
        c = draw_case(rng)
        try:
            r = run_case(os.path.join(root, f"c{k}"), c)
        except Exception as exc:                      # a failing case is a finding, not a crash
            bad.append({"case": c, "error": repr(exc)[:300]})
            continue
        if "reference_crashed" in r:
            crashed.append({"case": c, "signal": -r["reference_crashed"], "gpu_tau_finite": r["finite"]})
            print(f"case {k}: the reference crashed (exit {r['reference_crashed']}); GPU tau finite: "
                  f"{r['finite']}", flush=True)
            continue
        done += 1
        points += r["points"]
        by_method[c["method"]] += 1
        worst_tau = max(worst_tau, r["tau_max_rel"])
        worst_flux = max(worst_flux, r["flux_max_abs"])
        if r["tau_gt_1e9"] or not r["finite"] or r["flux_max_abs"] > 0.01:
            bad.append({"case": c, "result": r})
        print(f"case {k}: [{c['w0']:.0f}, {c['wn']:.0f}] dw {c['dw']} levels {c['levels']} method "
              f"{c['method']} lines {sum(c['lines'].values())}: tau {r['tau_max_rel']:.2e} flux "
              f"{r['flux_max_abs']:.2e}", flush=True)
    summary = {"cases": done, "by_method": by_method, "seed": args.seed, "points_compared": points,
               "worst_tau_rel": worst_tau, "worst_flux_abs_W_m2": worst_flux,
               "outside_tolerance": bad, "reference_crashed": crashed, "seconds": round(time.time() - t0, 1),
               "checker": parity.checker_name(), "precision": api.PRECISION,
               "tolerances": {"tau_rel": TAU_TOL, "flux_W_m2": 0.01}}
    print("SOAK " + json.dumps(summary, default=str))
    if args.out:
        json.dump(summary, open(args.out, "w"), indent=1, default=str)
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
