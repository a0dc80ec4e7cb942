"""The headless driver grt_lbl (grtcode_b200/host/driver_main.c) and its file formats
(include/grtcode_colset.h): the stand-in for the reference's netCDF applications
(framework/src/driver.c + rfmip-irf.c / circ.c)."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from grtcode_b200 import column_io, synth

HERE = os.path.dirname(os.path.abspath(__file__))
BIN = os.environ.get("GRT_LBL_BIN", os.path.join(os.path.dirname(HERE), "grtcode_b200", "grt_lbl"))
MOLS = (synth.H2O, synth.CO2, synth.O3)


def _colset(tmp_path, sites=3, expts=2, levels=12, cfcs=("CFC11",), cias=((0, 0), (1, 1))):
    ds = synth.make_dataset(str(tmp_path / "data"), 500., 700., {m: 150 for m in MOLS}, seed=4,
                            cfcs=cfcs, cias=cias)
    cols = column_io.rfmip_like_columns(sites, expts, levels, MOLS, tuple(ds.cfc.keys()))
    path = str(tmp_path / "columns.grtcols")
    column_io.write_column_set(path, cols, ds, MOLS)
    return ds, cols, path


def test_dry_run_parses_what_python_wrote(built, tmp_path):
    ds, cols, path = _colset(tmp_path)
    r = subprocess.run([BIN, ds.hitran, ds.solar, path, "-dry-run", "-x", "1", "-X", "2"],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "2 times x 3 columns x 12 levels, 3 molecules, 1 CFCs, 2 CIA pairs" in r.stdout
    col = cols[0][1]
    assert f"p_sfc {col.p[-1]:.6f} mb" in r.stdout
    assert f"T_sfc {col.t_surf:.4f} K" in r.stdout and f"mu0 {col.mu0:.6f}" in r.stdout


def test_block_partition_option_matches_shard_columns(built, tmp_path):
    from grtcode_b200 import parallel
    ds, cols, path = _colset(tmp_path, sites=11, expts=1)
    for rank in range(4):
        r = subprocess.run([BIN, ds.hitran, ds.solar, path, "-dry-run", "-ranks", "4", "-rank",
                            str(rank)], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        mine = parallel.shard_columns(11, 4, rank)
        assert f"columns [{mine.start}, {mine.stop - 1}]" in r.stdout


def test_malformed_inputs_fail_with_a_message(built, tmp_path):
    ds, cols, path = _colset(tmp_path)
    bad = str(tmp_path / "bad.grtcols")
    open(bad, "wb").write(b"NOTACOLS" + b"\0"*64)
    r = subprocess.run([BIN, ds.hitran, ds.solar, bad, "-dry-run"], capture_output=True, text=True)
    assert r.returncode != 0 and "not a GRTCOLS1" in r.stderr
    data = open(path, "rb").read()
    open(bad, "wb").write(data[:len(data)//2])
    r = subprocess.run([BIN, ds.hitran, ds.solar, bad, "-dry-run"], capture_output=True, text=True)
    assert r.returncode != 0 and "truncated" in r.stderr
    r = subprocess.run([BIN, ds.hitran, ds.solar, path, "-dry-run", "-x", "5"],
                       capture_output=True, text=True)
    assert r.returncode != 0 and "column range" in r.stderr


def test_flux_set_round_trip_through_the_c_library(built, tmp_path):
    """write_flux_set (C) -> column_io.read_flux_set (Python) and read_flux_set (C)."""
    api = built
    L = api.load()

    class FluxSet(C.Structure):
        _fields_ = [("num_times", C.c_int), ("num_columns", C.c_int), ("num_levels", C.c_int)] + \
                   [(n, C.POINTER(C.c_double)) for n in ("rlu", "rld", "rsu", "rsd", "hr_lw", "hr_sw")]

    fs = FluxSet()
    assert L.create_flux_set(C.byref(fs), 2, 3, 5) == 0
    rng = np.random.default_rng(0)
    want = {}
    for name in ("rlu", "rld", "rsu", "rsd", "hr_lw", "hr_sw"):
        n = 2*3*(5 if name.startswith("r") else 4)
        a = rng.normal(size=n)
        C.memmove(getattr(fs, name), a.ctypes.data, 8*n)
        want[name] = a
    path = str(tmp_path / "f.grtflux").encode()
    assert L.write_flux_set(path, C.byref(fs)) == 0
    got = column_io.read_flux_set(path.decode())
    for name, a in want.items():
        np.testing.assert_array_equal(getattr(got, name).reshape(-1), a)
    fs2 = FluxSet()
    assert L.read_flux_set(path, C.byref(fs2)) == 0
    assert (fs2.num_times, fs2.num_columns, fs2.num_levels) == (2, 3, 5)
    np.testing.assert_array_equal(np.ctypeslib.as_array(fs2.rsd, shape=(30,)), want["rsd"])
    assert L.destroy_flux_set(C.byref(fs)) == 0 and L.destroy_flux_set(C.byref(fs2)) == 0
    assert L.read_flux_set(b"/nonexistent/file", C.byref(fs2)) != 0


@pytest.mark.gpu
def test_driver_reproduces_the_python_column_loop(built, tmp_path):
    """grt_lbl on a 2 x 3 column set == the same columns through ColumnEngine (the same C
    API calls; the fused solvers reduce with one atomic per block, so the last bits depend
    on block order: 1e-12 relative), and the heating rates are those of heating_rates()."""
    from grtcode_b200 import workloads as W
    from grtcode_b200.column_engine import ColumnEngine
    ds, cols, path = _colset(tmp_path)
    out = str(tmp_path / "out.grtflux")
    r = subprocess.run([BIN, ds.hitran, ds.solar, path, "-o", out, "-w-lw", "500", "-W-lw", "700",
                        "-r-lw", "0.1", "-w-sw", "500", "-W-sw", "700", "-r-sw", "0.1"],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "6 columns in" in r.stdout
    got = column_io.read_flux_set(out)
    lines = {m: 150 for m in MOLS}
    bands = [W.Band("lw", 500., 700., 0.1, lines, ds), W.Band("sw", 500., 700., 0.1, lines, ds)]
    eng = ColumnEngine(bands, cols[0][0], 0, MOLS)
    for t in range(2):
        for c in range(3):
            res = eng.step(cols[t][c], upload_surface=True)
            np.testing.assert_allclose(got.rlu[t, c], res["lw"][0], rtol=1e-12)
            np.testing.assert_allclose(got.rld[t, c], res["lw"][1], rtol=1e-12, atol=1e-14)
            np.testing.assert_allclose(got.rsu[t, c], res["sw"][0], rtol=1e-12)
            np.testing.assert_allclose(got.rsd[t, c], res["sw"][1], rtol=1e-12)
            hr = built.heating_rates(cols[t][c].p, got.rlu[t, c], got.rld[t, c])
            np.testing.assert_array_equal(got.hr_lw[t, c], hr)
    eng.close()
