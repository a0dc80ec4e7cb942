"""GPU parity: libgrtcode.so (sm_100a) against the oracle on the same inputs.

Tolerances are the ones BASELINE.json's north_star states: per-wavenumber
optical depth within 1e-9 relative (double precision), broadband fluxes within
0.01 W m-2.
"""
import numpy as np
import pytest

from grtcode_b200 import synth
from tests import parity

TAU_RTOL = 1e-9          # north_star: "per-wavenumber optical depths within relative 1e-9"
FLUX_ATOL = 0.01         # north_star: "fluxes ... within 0.01 W/m^2"

pytestmark = pytest.mark.gpu


def _report(res):
    return (f"tau max rel {res['tau_max_rel']:.3e} at {res['tau_argmax']} "
            f"({res['tau_num_gt_1e9']} points > 1e-9); lw {res['lw_max_abs']:.3e}; "
            f"sw {res['sw_max_abs']:.3e}; stats {res['stats']}")


def test_small_window_three_gases(built, tmp_path):
    res = parity.run_column_pair(str(tmp_path), 500., 700., 0.1, 20,
                                 {synth.H2O: 300, synth.CO2: 400, synth.O3: 200}, seed=3)
    print(_report(res))
    assert res["tau_max_rel"] <= TAU_RTOL, _report(res)
    assert res["solar_max_rel"] <= 1e-14
    assert res["lw_max_abs"] <= FLUX_ATOL and res["sw_max_abs"] <= FLUX_ATOL, _report(res)
    assert res["lw_spec_max_rel"] <= 1e-9, _report(res)


def test_config1_shape_with_everything(built, tmp_path):
    """BASELINE config 1 shape (60 levels, 1-3000 cm-1 @ 0.1) with fewer lines so the
    oracle finishes in seconds; continua, two CFCs and all three CIA pairs on."""
    res = parity.run_column_pair(str(tmp_path), 1., 3000., 0.1, 60,
                                 {synth.H2O: 2000, synth.CO2: 3000, synth.O3: 2000}, seed=1,
                                 cfcs=("CFC11", "CFC12"), cias=((0, 0), (1, 0), (1, 1)))
    print(_report(res))
    assert res["tau_max_rel"] <= TAU_RTOL, _report(res)
    assert res["lw_max_abs"] <= FLUX_ATOL and res["sw_max_abs"] <= FLUX_ATOL, _report(res)


def test_fine_resolution_grid_edges(built, tmp_path):
    """0.01 cm-1 (5001-point windows, PTS=8 kernel) on a narrow grid so that most
    windows are clipped at the grid ends (kernels.c:435-437)."""
    res = parity.run_column_pair(str(tmp_path), 1000., 1030., 0.01, 12,
                                 {synth.H2O: 150, synth.CO2: 150, synth.CH4: 100}, seed=7)
    print(_report(res))
    assert res["tau_max_rel"] <= TAU_RTOL, _report(res)
    assert res["lw_max_abs"] <= FLUX_ATOL and res["sw_max_abs"] <= FLUX_ATOL, _report(res)
