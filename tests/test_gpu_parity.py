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


@pytest.mark.parametrize("method", [0, 1], ids=["wavenumber_sweep", "line_sweep"])
def test_binned_methods_at_the_finest_resolution_the_reference_allows(built, tmp_path, method):
    """0.001 cm-1 (MIN_RESOLUTION, grtcode_config.h:58): 1001 points per bin, 64 KB of
    shared memory per block -- beyond the 48 KB default, the gather kernel opts in to the
    large dynamic shared memory instead of refusing (wavenumber_sweep is the library default
    method, so a caller that passes no method must not fail here)."""
    # at least 26 bins: on a narrower grid the reference itself indexes bins.w out of bounds
    # ((bins.n-1)-nbin wraps as unsigned, kernels.c:199,238) and its result is undefined
    lines = {synth.H2O: 60, synth.CO2: 80}
    w0, wn, dw, levels = 1000., 1030., 0.001, 4
    ds = parity.build_dataset(str(tmp_path), w0, wn, lines, 41)
    mols = tuple(lines)
    col = synth.make_column(levels, np.random.default_rng(41), molecules=mols)
    want = parity.oracle_column(ds, w0, wn, dw, col, mols, solvers=False, method=method)
    got = parity.gpu_column(ds, w0, wn, dw, col, mols, solvers=False, method=method)
    rel = parity.relative_error(got["tau"], want["tau"])
    print(f"method {method} @ {dw}: tau max rel {rel.max():.3e}")
    assert rel.max() <= TAU_RTOL


# ---- BASELINE full sizes: size-independent properties (no oracle: it would take hours) ----

def _full_size_tau(ds, w0, wn, dw, col, mols, method=2, use_ctm=False):
    from grtcode_b200 import api
    got = parity.gpu_column(ds, w0, wn, dw, col, mols, use_ctm=use_ctm, solvers=False,
                            method=method)
    return got["tau"], got["stats"]


def test_full_size_config1_is_additive_over_molecules_and_line_order(built, tmp_path):
    """BASELINE config 1 at its real size (60 levels, 1-3000 cm-1 at 0.1 cm-1, 6.5e5 lines,
    1.9e10 Voigt samples).  Optical depth is a sum over lines (kernels.c:456-460), so
    (1) tau(H2O + CO2 + O3) = tau(H2O) + tau(CO2) + tau(O3), and (2) the order of the records
    in the HITRAN file does not matter -- both up to the rounding of a re-ordered double sum,
    five orders of magnitude inside the 1e-9 parity tolerance."""
    import os
    mols = (synth.H2O, synth.CO2, synth.O3)
    lines = {synth.H2O: 100_000, synth.CO2: 300_000, synth.O3: 250_000}
    w0, wn, dw, levels = 1., 3000., 0.1, 60
    ds = parity.build_dataset(str(tmp_path/"a"), w0, wn, lines, 31)
    col = synth.make_column(levels, np.random.default_rng(31), molecules=mols)
    whole, st = _full_size_tau(ds, w0, wn, dw, col, mols)
    assert st["samples"] > 1.8e10
    assert np.isfinite(whole).all() and (whole >= 0.).all()
    parts = np.zeros_like(whole)
    for m in mols:
        t, _ = _full_size_tau(ds, w0, wn, dw, col, (m,))
        parts += t
    rel = parity.relative_error(parts, whole)
    print(f"additivity: max rel {rel.max():.3e}")
    assert rel.max() <= 1e-12
    # same records, shuffled file order
    rows = open(ds.hitran).read().splitlines(keepends=True)
    np.random.default_rng(5).shuffle(rows)
    os.makedirs(tmp_path/"b", exist_ok=True)
    shuffled = str(tmp_path/"b"/"shuffled.par")
    open(shuffled, "w").writelines(rows)
    import dataclasses
    ds2 = dataclasses.replace(ds, hitran=shuffled)
    again, st2 = _full_size_tau(ds2, w0, wn, dw, col, mols)
    assert st2["samples"] == st["samples"]
    rel = parity.relative_error(again, whole)
    print(f"line order: max rel {rel.max():.3e}")
    assert rel.max() <= 1e-12


def test_full_size_fine_grid_chunks_reproduce_the_whole(built, tmp_path):
    """BASELINE config 2/4 resolution at a real line density (0.01 cm-1, 5001-point windows,
    8.8e4 lines over 325 cm-1, 54 layers): tau on a sub-grid that starts on a grid point of
    the whole grid, fed only the lines within 25 cm-1 of it (add_molecule min/max), equals
    the whole-grid tau away from the cut-off seam (kernels.c:431-439: a line's window is
    centred on its own nearest grid point, so interior points see the same samples)."""
    mols = (synth.H2O, synth.CO2, synth.CH4)
    lines = {synth.H2O: 20_000, synth.CO2: 40_000, synth.CH4: 28_000}
    w0, wn, dw, levels = 600., 925., 0.01, 55
    ds = parity.build_dataset(str(tmp_path), w0, wn, lines, 37)
    col = synth.make_column(levels, np.random.default_rng(37), molecules=mols)
    whole, st = _full_size_tau(ds, w0, wn, dw, col, mols)
    assert st["samples"] > 2e10
    a, b = 700., 760.
    got = parity.gpu_column(ds, a, b, dw, col, mols, use_ctm=False, solvers=False,
                            line_window=(a, b))
    i0 = int(round((a - w0)/dw))
    n = got["tau"].shape[1]
    # points further than 25 cm-1 (+ one step) from both ends of the sub-grid only see lines
    # that the sub-grid run also loaded
    k = int(np.ceil(25./dw)) + 2
    sub = got["tau"][:, k:n-k]
    ref = whole[:, i0+k:i0+n-k]
    rel = parity.relative_error(sub, ref)
    print(f"chunk vs whole: max rel {rel.max():.3e} over {sub.size} points")
    assert rel.max() <= 1e-9


# ---- committed outputs of the reference (tests/golden/reference_outputs.json) -------------

@pytest.mark.parametrize("name,method", [("wavenumber_sweep", 0), ("line_sweep", 1),
                                         ("line_sample", 2)])
def test_golden_reference_outputs(built, name, method):
    """The CUDA path against outputs of the compiled reference committed as a fixture
    (tools/make_golden_outputs.py), i.e. without oracle/_ref: optical depths of the three
    methods with continua, one CFC and one CIA pair, and the broadband fluxes."""
    from tests.test_oracle import _golden_case
    g, c, ds, mols, col = _golden_case()
    got = parity.gpu_column(ds, c["w0"], c["wn"], c["dw"], col, mols, solvers=(method == 2),
                            method=method)
    want = np.array(g["tau"][name])
    rel = parity.relative_error(got["tau"], want)
    assert rel.max() <= TAU_RTOL, f"{name}: tau max rel {rel.max():.3e}"
    if method == 2:
        bb = g["broadband_W_m2"]
        assert np.abs(got["lw_bb_up"] - np.array(bb["lw_up"])).max() <= FLUX_ATOL
        assert np.abs(got["lw_bb_dn"] - np.array(bb["lw_dn"])).max() <= FLUX_ATOL
        assert np.abs(got["sw_bb_up"] - np.array(bb["sw_up"])).max() <= FLUX_ATOL
        assert np.abs(got["sw_bb_dn"] - np.array(bb["sw_dn"])).max() <= FLUX_ATOL
        dw = c["dw"]
        assert np.abs(parity.R.broadband(got["lw_up"], dw) - np.array(bb["lw_up"])).max() <= 1e-9
