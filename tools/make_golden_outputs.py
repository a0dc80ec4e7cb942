"""Golden OUTPUT vectors of the reference itself for the paths its own tests leave unpinned
(Voigt / optical depths of the three methods / fluxes): the compiled reference
(oracle/_ref, the unmodified sources under /root/reference) is run on a tiny seeded synthetic
case and its results are committed as tests/golden/reference_outputs.json, so that the
restatement (oracle/) and the CUDA path can be checked where /root/reference and oracle/_ref
do not exist.  Inputs are regenerated from the seeds by grtcode_b200.synth; the SHA-256 of the
generated HITRAN file is stored to detect a drifting generator.

    python tools/make_golden_outputs.py        (needs oracle/_ref: make -C oracle ref)
"""
import hashlib
import json
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from grtcode_b200 import synth          # noqa: E402
from oracle import ref as R             # noqa: E402
from tests import parity                # noqa: E402

CASE = dict(w0=950., wn=990., dw=0.1, levels=4, lines={1: 60, 2: 80, 3: 40}, seed=17,
            cfcs=["CFC11"], cias=[[0, 0]])


VOIGT_CASES = [(1000.03, 1e-7, 9e-4), (1000.03, 5e-4, 9e-4), (1000.03, 2e-3, 9e-4),
               (1000.03, 8e-3, 9e-4), (1000.03, 0.03, 9e-4), (1000.03, 0.09, 9e-4),
               (50.01, 0.08, 7e-5), (20000.4, 0.05, 0.03), (1000.03, 0., 9e-4)]


def build(root):
    ds = synth.make_dataset(root, CASE["w0"], CASE["wn"], {int(k): v for k, v in CASE["lines"].items()},
                            seed=CASE["seed"], cfcs=tuple(CASE["cfcs"]),
                            cias=tuple(tuple(c) for c in CASE["cias"]))
    mols = tuple(int(k) for k in CASE["lines"])
    col = synth.make_column(CASE["levels"], np.random.default_rng(CASE["seed"]), molecules=mols,
                            cfcs=tuple(ds.cfc.keys()))
    return ds, mols, col


def main():
    if not R.available():
        sys.exit("oracle/_ref/libgrtcode_ref.so missing: make -C oracle ref")
    with tempfile.TemporaryDirectory() as tmp:
        ds, mols, col = build(tmp)
        out = {"source": "compiled reference (oracle/_ref/libgrtcode_ref.so, gcc -O3 -fopenmp "
                         "-ffp-contract=off, one thread) on grtcode_b200.synth inputs",
               "case": CASE,
               "hitran_sha256": hashlib.sha256(open(ds.hitran, "rb").read()).hexdigest(),
               "tau": {}}
        for name, method in (("wavenumber_sweep", 0), ("line_sweep", 1), ("line_sample", 2)):
            want = parity.oracle_column(ds, CASE["w0"], CASE["wn"], CASE["dw"], col, mols,
                                        threads=1, solvers=(method == 2), method=method)
            out["tau"][name] = [[float(x) for x in row] for row in want["tau"]]
            if method == 2:
                bb = lambda f: [float(x) for x in R.broadband(f, CASE["dw"])]
                out["broadband_W_m2"] = {k: bb(want[k]) for k in ("lw_up", "lw_dn", "sw_up", "sw_dn")}
        # rfm_voigt_line_shape itself: nine (centre, gamma, alpha) triples covering every branch
        # (far wing, regions 1-3, both CPF12 forms, pure Lorentz, Y = 0), 41 points each
        r = R.Ref(threads=1)
        out["voigt"] = []
        for v, gamma, alpha in VOIGT_CASES:
            for wres in (0.1, 0.001):
                w = v - 20*wres + 0.37*wres
                k = r.voigt(w, 41, wres, v, gamma, alpha)
                out["voigt"].append({"w": w, "n": 41, "wres": wres, "center": v, "gamma": gamma,
                                     "alpha": alpha, "K": [float(x) for x in k]})
        path = os.path.join(ROOT, "tests", "golden", "reference_outputs.json")
        json.dump(out, open(path, "w"))
        print(path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
