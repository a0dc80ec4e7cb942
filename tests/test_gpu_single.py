"""The single-precision build (the reference's --enable-single-precision: fp_t = float,
configure.ac:25-32) against the reference compiled with its own -DSINGLE_PRECISION.

Precision is per process, so the comparison runs in a subprocess (tests/sp_check.py) with
GRT_PRECISION=single.  Tolerances are north_star's: optical depth within 1e-4 relative per
wavenumber, broadband fluxes within 0.01 W m-2.
"""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu

TAU_RTOL_SINGLE = 1e-4
FLUX_ATOL = 0.01


def test_single_precision_build_matches_the_single_precision_reference(built):
    sp = os.path.join(ROOT, "oracle", "_ref", "libgrtcode_ref_sp.so")
    if not os.path.exists(sp):
        pytest.skip("oracle/_ref/libgrtcode_ref_sp.so not built")
    env = dict(os.environ, GRT_PRECISION="single")
    proc = subprocess.run([sys.executable, "-m", "tests.sp_check"], cwd=ROOT, env=env,
                          capture_output=True, text=True, timeout=900)
    assert proc.returncode == 0, proc.stdout[-3000:] + proc.stderr[-3000:]
    rows = [json.loads(l[len("SPCHECK "):]) for l in proc.stdout.splitlines()
            if l.startswith("SPCHECK ")]
    assert len(rows) >= 6
    for r in rows:
        print(r)
        assert r["dtype"] == "float32" and r["finite"], r
        assert r["tau_max_rel"] <= TAU_RTOL_SINGLE, r
        assert r["lw_max_abs"] <= FLUX_ATOL and r["sw_max_abs"] <= FLUX_ATOL, r
