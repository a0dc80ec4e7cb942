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


def test_lines_outside_the_grid(built, tmp_path):
    """Line list wider than the grid, loaded with add_molecule(min, max)
    (gas_optics.h:117-119): centres beyond the grid are skipped, centres within half a
    grid step below w0 still land on point 0 (kernels.c:431-433)."""
    ds = parity.build_dataset(str(tmp_path), 400., 800., {synth.H2O: 600, synth.CO2: 800}, 5)
    rng = np.random.default_rng(5)
    mols = (synth.H2O, synth.CO2)
    col = synth.make_column(16, rng, molecules=mols)
    kw = dict(solvers=False, line_window=(400., 800.))
    want = parity.oracle_column(ds, 500., 700., 0.1, col, mols, **kw)
    got = parity.gpu_column(ds, 500., 700., 0.1, col, mols, **kw)
    rel = parity.relative_error(got["tau"], want["tau"])
    assert got["stats"]["records"] < 1400*15          # some (line, layer) pairs were skipped
    assert rel.max() <= TAU_RTOL, f"tau max rel {rel.max():.3e}"


def test_two_level_column_without_continua(built, tmp_path):
    """Smallest column the API accepts (one layer), continua off, one molecule."""
    ds = parity.build_dataset(str(tmp_path), 900., 1100., {synth.O3: 500}, 9)
    rng = np.random.default_rng(9)
    col = synth.make_column(2, rng, molecules=(synth.O3,))
    want = parity.oracle_column(ds, 900., 1100., 0.1, col, (synth.O3,), use_ctm=False)
    got = parity.gpu_column(ds, 900., 1100., 0.1, col, (synth.O3,), use_ctm=False)
    rel = parity.relative_error(got["tau"], want["tau"])
    assert rel.max() <= TAU_RTOL, f"tau max rel {rel.max():.3e}"
    bb = lambda f: parity.R.broadband(f, 0.1)
    assert np.abs(got["lw_bb_up"] - bb(want["lw_up"])).max() <= FLUX_ATOL
    assert np.abs(got["sw_bb_dn"] - bb(want["sw_dn"])).max() <= FLUX_ATOL


def test_spectral_split_reproduces_the_unsplit_column(built, tmp_path):
    """Config-4 decomposition on one GPU: three chunks (+ halo) run one after the other,
    partial broadband integrals summed as the all-reduce would; against the unsplit GPU
    run and the oracle's fluxes."""
    from grtcode_b200 import parallel as P
    from grtcode_b200 import workloads as W
    from grtcode_b200.column_engine import ColumnEngine
    mols = (synth.H2O, synth.CO2, synth.O3)
    lines = {synth.H2O: 1500, synth.CO2: 2500, synth.O3: 1500}
    w0, wn, dw, levels = 1., 3000., 0.1, 24
    ds = parity.build_dataset(str(tmp_path), w0, wn, lines, 21)
    rng = np.random.default_rng(21)
    col = synth.make_column(levels, rng, molecules=mols)
    bands = [W.Band("lw", w0, wn, dw, lines, ds), W.Band("sw", w0, wn, dw, lines, ds)]
    whole = ColumnEngine(bands, col, 0, mols)
    ref = whole.step(col, upload_surface=True)
    whole.close()
    world = 3
    total = {k: [np.zeros(levels), np.zeros(levels)] for k in ("lw", "sw")}
    for rank in range(world):
        part = P.SpectralSplitColumn(bands, col, 0, mols, world, rank)
        out = part.run(col)
        for k in total:
            total[k][0] += out[k][0]
            total[k][1] += out[k][1]
        part.close()
    for k in total:
        for d in (0, 1):
            err = np.abs(total[k][d] - ref[k][d]).max()
            scale = np.abs(ref[k][d]).max()
            print(f"{k} dir {d}: split vs unsplit |dF| {err:.3e} of {scale:.3e} W/m2")
            assert err <= 1e-9*max(scale, 1.), (k, d, err)
    want = parity.oracle_column(ds, w0, wn, dw, col, mols)
    bb = lambda f: parity.R.broadband(f, dw)
    assert np.abs(total["lw"][0] - bb(want["lw_up"])).max() <= FLUX_ATOL
    assert np.abs(total["lw"][1] - bb(want["lw_dn"])).max() <= FLUX_ATOL
    assert np.abs(total["sw"][0] - bb(want["sw_up"])).max() <= FLUX_ATOL
    assert np.abs(total["sw"][1] - bb(want["sw_dn"])).max() <= FLUX_ATOL


@pytest.mark.parametrize("method", [0, 1, None], ids=["wavenumber_sweep", "line_sweep", "default"])
@pytest.mark.parametrize("case", ["dense", "sparse"])
def test_binned_methods(built, tmp_path, method, case):
    """wavenumber_sweep (the library default, optical_depth_method = NULL) and line_sweep
    (kernels.c:177-406 + interpolate, :514-581) against the reference on the same inputs;
    "sparse" has molecules with a handful of lines, which exercises bracket()'s
    out-of-range returns (kernel_utils.c:36-41)."""
    if case == "dense":
        lines = {synth.H2O: 1500, synth.CO2: 2500, synth.O3: 1500}
        w0, wn, dw, levels = 400., 900., 0.1, 18
    else:
        lines = {synth.H2O: 40, synth.CO: 5, synth.O2: 12}
        w0, wn, dw, levels = 1000., 1400., 0.05, 10
    ds = parity.build_dataset(str(tmp_path), w0, wn, lines, 31)
    rng = np.random.default_rng(31)
    mols = tuple(lines)
    col = synth.make_column(levels, rng, molecules=mols)
    want = parity.oracle_column(ds, w0, wn, dw, col, mols, solvers=False, method=method)
    got = parity.gpu_column(ds, w0, wn, dw, col, mols, solvers=False, method=method)
    rel = parity.relative_error(got["tau"], want["tau"])
    bad = int((rel > TAU_RTOL).sum())
    where = np.unravel_index(int(rel.argmax()), rel.shape)
    print(f"method {method} {case}: tau max rel {rel.max():.3e} at {where}; {bad} points > 1e-9; "
          f"stats {got['stats']}")
    assert rel.max() <= TAU_RTOL
