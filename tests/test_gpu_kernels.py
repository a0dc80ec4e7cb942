"""GPU unit tests of individual device routines."""
import pytest

pytestmark = pytest.mark.gpu


def test_division_and_widening_shortcuts_are_exact(built):
    """div_exact == IEEE float division and widen_positive == (double) on 2^28 operands
    spanning the sweep's ranges (voigt.cuh): bit-exact, zero mismatches allowed."""
    assert built.selftest_arithmetic(1 << 28, seed=12345) == 0


def test_fma_peak_is_plausible(built):
    fp64, fp32 = built.measure_fma_peak(0)
    assert 10. < fp64 < 60. and 30. < fp32 < 120.


def test_per_line_voigt_matches_the_restatement_bit_for_bit(built):
    """voigt_line_eval (hoisted constants, fp32x2 CPF12, slow-path-free division) ==
    voigt_core (RFM_voigt.c:109-277 sample by sample) on 2^28 (x, Y) pairs over all regions."""
    assert built.selftest_voigt(1 << 28, seed=4321) == 0


def _tau(tmp, w0, wn, dw, levels, lines, seed, method, env):
    """tau through the C API with the library's measurement knobs set as in `env`."""
    import os
    import numpy as np
    from grtcode_b200 import synth
    from tests import parity
    ds = parity.build_dataset(str(tmp), w0, wn, lines, seed)
    mols = tuple(lines.keys())
    col = synth.make_column(levels, np.random.default_rng(seed), molecules=mols)
    old = {k: os.environ.get(k) for k in env}
    os.environ.update(env)
    try:
        got = parity.gpu_column(ds, w0, wn, dw, col, mols, solvers=False, method=method)
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
    return got["tau"], got["stats"]


@pytest.mark.parametrize("dw,levels", [(0.1, 30), (0.01, 8)])
def test_sweep_variants_agree(built, tmp_path, dw, levels):
    """Whole-warp lines, two and four lines at a time, every tile width: the same samples
    summed in a different order (sweep.cuh, template parameters PTS and LN)."""
    import numpy as np
    from grtcode_b200 import synth
    from tests import parity
    lines = {synth.H2O: 2500, synth.CO2: 3500, synth.O3: 2000}
    span = (600., 1000.) if dw == 0.1 else (700., 760.)
    ref, st = _tau(tmp_path, *span, dw, levels, lines, 41, 2,
                   {"GRT_SWEEP_LANES": "32", "GRT_SWEEP_PTS": "2" if dw == 0.1 else "8"})
    for lanes, pts in ((32, 4), (32, 16), (16, 4), (16, 8), (16, 16), (8, 4), (8, 8), (8, 16)):
        tau, st2 = _tau(tmp_path, *span, dw, levels, lines, 41, 2,
                        {"GRT_SWEEP_LANES": str(lanes), "GRT_SWEEP_PTS": str(pts)})
        assert st2["samples"] == st["samples"]
        rel = parity.relative_error(tau, ref)
        assert rel.max() <= 1e-12, (lanes, pts, float(rel.max()))
    tau, _ = _tau(tmp_path, *span, dw, levels, lines, 41, 2, {"GRT_CORE_ORDER": "0"})
    assert parity.relative_error(tau, ref).max() <= 1e-12


@pytest.mark.parametrize("case", ["lw_0.01", "lw_0.1_dense", "sw_low_pressure"])
def test_core_gather_matches_the_scatter_kernel_and_is_reproducible(built, tmp_path, case):
    """core_gather_kernel (no atomics: a warp owns 512 grid points of a layer) against
    core_sweep_kernel (one thread per record, double atomics): the same samples in a different
    order -- and, unlike the scatter, the same bits from run to run.  The cases cover narrow
    cores with several lines per grid point (the segmented warp sum), and the wide Doppler
    cores of the shortwave at low pressure (whole-warp walk, records reaching over tiles)."""
    import numpy as np
    from grtcode_b200 import synth
    from tests import parity
    if case == "lw_0.01":
        args = (700., 760., 0.01, 12, {synth.H2O: 6000, synth.CO2: 9000, synth.O3: 5000}, 51)
    elif case == "lw_0.1_dense":
        args = (600., 700., 0.1, 20, {synth.H2O: 4000, synth.CO2: 6000, synth.N2O: 3000}, 52)
    else:
        args = (29950., 30050., 0.01, 40, {synth.H2O: 900, synth.O3: 900, synth.O2: 300}, 53)
    a, sa = _tau(tmp_path, *args, 2, {"GRT_CORE": "gather"})
    b, _ = _tau(tmp_path, *args, 2, {"GRT_CORE": "gather"})
    c, sc = _tau(tmp_path, *args, 2, {"GRT_CORE": "scatter"})
    assert a.tobytes() == b.tobytes()
    rel = parity.relative_error(a, c)
    print(f"{case}: gather vs scatter max rel {rel.max():.2e}; core {sa['core_ns']/1e6:.3f} ms "
          f"(gather) / {sc['core_ns']/1e6:.3f} ms (scatter)")
    assert rel.max() <= 1e-12


@pytest.mark.parametrize("method", [0, 1])
def test_binned_sort_paths_agree(built, tmp_path, method):
    """Two device-wide radix passes against CUB's segmented sort: both stable, so the same
    permutation and bit-identical optical depths."""
    import numpy as np
    from grtcode_b200 import synth
    lines = {synth.H2O: 3000, synth.CO2: 4000, synth.O3: 1, synth.CH4: 2500}
    a, _ = _tau(tmp_path, 500., 900., 0.1, 20, lines, 43, method, {"GRT_BIN_SORT": "d"})
    b, _ = _tau(tmp_path, 500., 900., 0.1, 20, lines, 43, method, {"GRT_BIN_SORT": "s"})
    assert np.array_equal(a, b)


def test_device_table_resampling_matches_the_host_rule(built, tmp_path):
    """Setup on the device: continuum / CFC / CIA tables are resampled to the
    spectral grid by a kernel; every table must equal, bit for bit, what the host's
    interpolate_to_grid restatement (host/parse_csv.c: grt_table_on_grid, checked against the
    reference in tests/test_host.py) gives for the same CSV: linear inside a source interval,
    0 at or below the first abscissa and beyond the last."""
    import numpy as np
    from grtcode_b200 import api, synth
    api.load()
    # grid wider than the CFC table (500-1500.5 cm-1) and finer than any table's spacing
    w0, wn, dw = 1., 3250., 0.01
    ds = synth.make_dataset(str(tmp_path), w0, wn, {synth.H2O: 20, synth.O3: 20}, seed=3,
                            cfcs=("CFC11", "CCl4"), cias=((0, 0), (1, 1)))
    g = api.spectral_grid(w0, wn, dw)
    go = api.GasOptics(4, g, ds.hitran, ds.h2o_ctm_dir, ds.o3_ctm)
    go.add_molecule(synth.H2O)
    go.add_molecule(synth.O3)
    for c, path in ds.cfc.items():
        go.add_cfc(c, path)
    for (a, b), path in ds.cia.items():
        go.add_cia(a, b, path)
    import os
    cases = [(0, 0, os.path.join(ds.h2o_ctm_dir, "296MTCKD25_S.csv"), 2),
             (1, 0, os.path.join(ds.h2o_ctm_dir, "CKDS.csv"), 4),
             (2, 0, os.path.join(ds.h2o_ctm_dir, "296MTCKD25_F.csv"), 2),
             (3, 0, os.path.join(ds.h2o_ctm_dir, "CKDF.csv"), 4),
             (4, 0, ds.o3_ctm, 2)]
    cases += [(5, k, path, 2) for k, path in enumerate(ds.cfc.values())]
    cases += [(6, k, path, 2) for k, path in enumerate(ds.cia.values())]
    for kind, index, path, cols in cases:
        got = go.debug_table(kind, index)
        want = api.table_on_grid_host(path, cols, g)
        assert got.shape == want.shape
        assert np.array_equal(got, want), (kind, index, np.abs(got - want).max())
        assert (got != 0).any()
    # the CFC table is zero outside 500-1500.5 cm-1 (no extrapolation)
    cfc = go.debug_table(5, 0)
    w = w0 + np.arange(int(g.n))*dw
    assert not cfc[w <= 500.].any() and not cfc[w > 1500.5].any()
    go.close()
