"""Generates tests/golden/near_ties.json: columns in which, on every layer, one line's
Y = (float)(REPWID*gamma) (RFM_voigt.c:95) sits within about one double ulp of a float rounding
boundary -- the case where a last-bit difference between CUDA's pow() and the host libm's
flips Y and moves every sample of that line by 6e-8 relative.

The search uses the compiled reference's own kernels (tests/k1_reference.py -> oracle/_ref):
the level temperatures are bisected, top down, until the reference's REPWID*gamma for the
chosen line straddles the boundary between two adjacent floats.  Line lists are regenerated
from their seed at test time (SHA-256 of the file recorded here).

    python tools/make_near_ties.py        # needs oracle/_ref (make -C oracle ref)
"""
from __future__ import annotations

import hashlib
import json
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from grtcode_b200 import synth                      # noqa: E402
from oracle import port as P                        # noqa: E402
from oracle import ref as R                         # noqa: E402
from tests import k1_reference as K                 # noqa: E402

CASE = dict(w0=600., wn=800., dw=0.1, levels=201, lines={synth.H2O: 300, synth.CO2: 300},
            seed=77, columns=6)


def line_params(hitran, mol, w0, wn):
    lp = P.parse_hitran(hitran, mol, w0, wn)
    return lp


def main():
    ref = R.Ref()
    c = CASE
    tmp = tempfile.mkdtemp(prefix="near_ties_")
    ds = synth.make_dataset(tmp, c["w0"], c["wn"], c["lines"], seed=c["seed"])
    sha = hashlib.sha256(open(ds.hitran, "rb").read()).hexdigest()
    mols = tuple(c["lines"])
    lp = {m: line_params(ds.hitran, m, c["w0"], c["wn"]) for m in mols}
    layers = c["levels"] - 1
    out = {"case": {**c, "lines": {str(k): v for k, v in c["lines"].items()}},
           "hitran_sha256": sha, "columns": []}
    for col_id in range(c["columns"]):
        rng = np.random.default_rng(1000 + col_id)
        col = synth.make_column(c["levels"], rng, molecules=mols)
        t = np.array(col.t, dtype=np.float64)
        chosen = []
        for i in range(layers):
            mol = mols[(i + col_id) % len(mols)]
            li = int(rng.integers(0, len(lp[mol]["v"])))
            pavg, _, ps = K.layer_quantities(ref, col.p[i:i+2], t[i:i+2], col.ppmv[mol][i:i+2])
            one = lambda a: np.array([a[li]])

            def yd_of(t_next):
                tt = np.array([t[i], t_next])
                _, tavg, _ = K.layer_quantities(ref, col.p[i:i+2], tt, col.ppmv[mol][i:i+2])
                _, gamma, alpha = K.widths(ref, mol, one(lp[mol]["v"]), one(lp[mol]["d"]),
                                           one(lp[mol]["n"]), one(lp[mol]["yair"]),
                                           one(lp[mol]["yself"]), pavg, tavg, ps)
                return float(K.repwid_y(gamma, alpha)[2][0, 0])

            lo, hi = t[i+1] - 1.5, t[i+1] + 1.5           # yd decreases with temperature
            y_hi, y_lo = yd_of(lo), yd_of(hi)
            f = np.float32(0.5*(y_hi + y_lo))
            g = np.nextafter(f, np.float32(-np.inf))
            B = 0.5*(np.float64(f) + np.float64(g))       # exact: a float tie
            if not (y_lo < B < y_hi):
                raise RuntimeError(f"column {col_id} layer {i}: no tie inside the bracket")
            while True:
                mid = 0.5*(lo + hi)
                if mid == lo or mid == hi:
                    break
                if yd_of(mid) > B:
                    lo = mid
                else:
                    hi = mid
            t[i+1] = lo if rng.random() < 0.5 else hi
            chosen.append([int(mol), li, int(K.ulps_from_float_tie(np.array([yd_of(t[i+1])]))[0])])
        dist = np.array([x[2] for x in chosen])
        print(f"column {col_id}: tie distance [double ulps] median {np.median(dist):.0f}, "
              f"max {dist.max()}, T range {t.min():.1f}-{t.max():.1f} K")
        out["columns"].append({"seed": 1000 + col_id, "t_hex": [float(x).hex() for x in t],
                               "near_tie": chosen})
    path = os.path.join(ROOT, "tests", "golden", "near_ties.json")
    json.dump(out, open(path, "w"))
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
