"""CPU tests of the host side of the product: the C-ABI libraries load, export every
symbol the headers declare, and the C99 host logic (no GPU calls) behaves like the
reference's."""
import ctypes as C
import glob
import os
import re
import subprocess

import numpy as np
import pytest

from grtcode_b200 import synth
from oracle import port as P

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
dp = C.POINTER(C.c_double)


def _declared(header: str) -> list[str]:
    src = open(header).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    names = re.findall(r"\b(?:EXTERN\s+)?(?:int|void|fp_t|double)\s+(\w+)\s*\(", src)
    return [n for n in names if not n.startswith("(")]


def _exported(lib: str) -> set[str]:
    out = subprocess.run(["nm", "-D", "--defined-only", lib], capture_output=True, text=True,
                         check=True).stdout
    return {line.split()[-1] for line in out.splitlines() if " T " in line}


def test_cuda_abi_exports_every_declared_symbol(built):
    want = _declared(os.path.join(ROOT, "include", "grt_sm100.h"))
    assert len(want) >= 25
    missing = set(want) - _exported(built.LIB_CUDA)
    assert not missing, missing


def test_host_library_exports_reference_api(built):
    have = _exported(built.LIB_HOST)
    for header in glob.glob(os.path.join(ROOT, "include", "*.h")):
        if header.endswith("grt_sm100.h"):
            continue
        for name in _declared(header):
            assert name in have, (os.path.basename(header), name)
    # the reference's 26 Fortran-bound symbols (fortran-bindings/grtcode_fortran.F90:127-554)
    for name in ("create_gas_optics", "destroy_gas_optics", "add_molecule", "set_molecule_ppmv",
                 "add_cfc", "set_cfc_ppmv", "add_cia", "set_cia_ppmv", "calculate_optical_depth",
                 "get_num_molecules", "create_spectral_grid", "create_optics", "destroy_optics",
                 "add_optics", "create_solar_flux", "destroy_solar_flux", "rayleigh_scattering",
                 "malloc_struct", "free_struct", "optical_properties", "spectral_grid_properties",
                 "solar_flux_properties", "grtcode_errstr", "grtcode_set_verbosity",
                 "create_device", "update_optics"):
        assert name in have, name


def test_no_product_dependency_on_oracle(built):
    """The product libraries must not link anything under oracle/."""
    for lib in (built.LIB_HOST, built.LIB_CUDA):
        out = subprocess.run(["ldd", lib], capture_output=True, text=True).stdout
        assert "oracle" not in out and "grtcode_ref" not in out


def test_spectral_grid_sizes(built):
    """n = ceil((wn - w0)/dw) + 1 (utilities/src/spectral_grid.c:61) at the BASELINE shapes."""
    for w0, wn, dw, n in ((1., 3000., 0.1, 29991), (1., 3250., 0.1, 32491),
                          (1., 50000., 0.1, 499991), (1., 3250., 0.01, 324901),
                          (1., 50000., 0.01, 4999901)):
        g = built.spectral_grid(w0, wn, dw)
        assert g.n == n and g.n == P.grid_size(w0, wn, dw)
    with pytest.raises(built.GrtcodeError) as e:
        built.spectral_grid(0.5, 10., 0.1)
    assert e.value.code == 8     # GRTCODE_RANGE_ERR


def test_host_only_is_rejected(built):
    """No CPU path: HOST_ONLY (-1) is a GRTCODE_VALUE_ERR (= 9) in every constructor."""
    L = built.load()
    g = built.spectral_grid(1., 10., 0.1)
    dev = C.c_int(-1)
    o = built.Optics()
    assert L.create_optics(C.byref(o), 5, C.byref(g), C.byref(dev)) == 9
    s = built.GasOpticsStruct()
    m = C.c_int(2)
    assert L.create_gas_optics(C.byref(s), 6, C.byref(g), C.byref(dev), b"x.par", None, None,
                               None, C.byref(m)) == 9
    lw = built.Longwave()
    assert L.create_longwave(C.byref(lw), 6, C.byref(g), C.byref(dev)) == 9
    sw = built.Shortwave()
    assert L.create_shortwave(C.byref(sw), 6, C.byref(g), C.byref(dev)) == 9
    buf = C.create_string_buffer(2048)
    L.grtcode_errstr(9, buf, 2048)
    assert b"HOST_ONLY" in buf.value and b"Backtrace" in buf.value
    L.grtcode_errstr(0, buf, 2048)
    assert buf.value == b"No errors."


def test_interpolation_rules(built):
    """interpolate2 (utilities/src/utilities.c:149-222): (x_j, x_j+1] intervals, nothing
    outside the table unless an extrapolator is given; constant_extrapolation on the
    right uses y[n-2] (the reference passes &y[n-2])."""
    L = built.load()
    fn = C.CFUNCTYPE(C.c_int, dp, dp, dp, dp, C.c_size_t)
    L.interpolate2.argtypes = [dp, dp, C.c_size_t, dp, dp, C.c_size_t, C.c_void_p, C.c_void_p]
    x = np.array([1., 2., 4.])
    y = np.array([10., 20., 0.])
    newx = np.array([0.5, 1., 1.5, 2., 3., 4., 5.])
    out = np.full(7, -1.)
    lin = C.cast(L.linear_sample, C.c_void_p)
    rc = L.interpolate2(x.ctypes.data_as(dp), y.ctypes.data_as(dp), 3, newx.ctypes.data_as(dp),
                        out.ctypes.data_as(dp), 7, lin, None)
    assert rc == 0
    assert np.allclose(out, [-1., -1., 15., 20., 10., 0., -1.])
    out[:] = -1.
    const = C.cast(L.constant_extrapolation, C.c_void_p)
    rc = L.interpolate2(x.ctypes.data_as(dp), y.ctypes.data_as(dp), 3, newx.ctypes.data_as(dp),
                        out.ctypes.data_as(dp), 7, lin, const)
    assert rc == 0
    assert np.allclose(out, [10., 10., 15., 20., 10., 0., 20.])
    bad = np.array([1., 1., 2.])
    assert L.interpolate2(bad.ctypes.data_as(dp), y.ctypes.data_as(dp), 3,
                          newx.ctypes.data_as(dp), out.ctypes.data_as(dp), 7, lin, None) == 9
    del fn


class _LineParams(C.Structure):
    _fields_ = [("num_lines", C.c_uint64), ("iso", C.POINTER(C.c_int)), ("vnn", dp), ("snn", dp),
                ("yair", dp), ("yself", dp), ("en", dp), ("n", dp), ("d", dp)]


def test_hitran_parser_matches_oracle(built, tmp_path):
    """Host C parser against the oracle's numpy restatement: filtering by molecule and
    range, float rounding of five fields, 296 K strength scaling -- bit for bit."""
    L = built.load()
    ds = synth.make_dataset(str(tmp_path), 100., 900., {1: 500, 2: 700, 6: 300}, seed=9)
    L.parse_hitran_file.argtypes = [C.POINTER(_LineParams), C.c_char_p, C.c_int, C.c_double,
                                    C.c_double]
    for mol, w0, wn in ((1, 100., 900.), (2, 300., 500.), (6, 100., 900.), (3, 100., 900.)):
        lp = _LineParams()
        assert L.parse_hitran_file(C.byref(lp), ds.hitran.encode(), mol, w0, wn) == 0
        want = P.parse_hitran(ds.hitran, mol, w0, wn)
        n = int(lp.num_lines)
        assert n == len(want["v"])
        if n == 0:
            continue
        for field, key in (("vnn", "v"), ("snn", "s"), ("yair", "yair"), ("yself", "yself"),
                           ("en", "en"), ("n", "n"), ("d", "d")):
            got = np.ctypeslib.as_array(getattr(lp, field), shape=(n,))
            assert np.array_equal(got, want[key]), (mol, field)
        assert np.array_equal(np.ctypeslib.as_array(lp.iso, shape=(n,)), want["iso"])
        assert np.all(np.diff(want["v"]) >= 0)
    # malformed record length -> GRTCODE_VALUE_ERR
    badfile = tmp_path/"bad.par"
    badfile.write_text(open(ds.hitran).readline()[:100] + "\n")
    lp = _LineParams()
    assert L.parse_hitran_file(C.byref(lp), str(badfile).encode(), 1, -1., -1.) == 9
    assert L.parse_hitran_file(C.byref(lp), b"/nonexistent.par", 1, -1., -1.) == 11   # IO_ERR


def test_hitran_table_cache_and_error_rows(built, tmp_path):
    """The parser converts a file once (all host cores) and serves every molecule from the
    table.  What a caller of the reference would see must not change: a field that cannot be
    converted fails only the molecule it belongs to (parse_HITRAN_file.c:310 skips other
    molecules before converting), a rewritten file is read again, thread count does not
    change the result."""
    import os
    L = built.load()
    L.parse_hitran_file.argtypes = [C.POINTER(_LineParams), C.c_char_p, C.c_int, C.c_double,
                                    C.c_double]
    ds = synth.make_dataset(str(tmp_path), 100., 900., {1: 6000, 2: 9000}, seed=4)
    path = ds.hitran.encode()

    def parse(mol):
        lp = _LineParams()
        rc = L.parse_hitran_file(C.byref(lp), path, mol, -1., -1.)
        if rc != 0:
            return rc, None
        n = int(lp.num_lines)
        out = {f: np.ctypeslib.as_array(getattr(lp, f), shape=(n,)).copy()
               for f in ("vnn", "snn", "yair", "yself", "en", "n", "d")}
        L.free_line_params(C.byref(lp))
        return 0, out

    os.environ["GRT_PARSE_THREADS"] = "1"
    L.grt_hitran_cache_clear()
    rc, one = parse(2)
    assert rc == 0 and len(one["vnn"]) == 9000
    os.environ["GRT_PARSE_THREADS"] = "7"
    L.grt_hitran_cache_clear()
    rc, many = parse(2)
    del os.environ["GRT_PARSE_THREADS"]
    assert rc == 0 and all(np.array_equal(one[f], many[f]) for f in one)
    # corrupt the strength field of one H2O record: H2O fails, CO2 is untouched
    rows = open(ds.hitran).read().split("\n")
    k = next(i for i, r in enumerate(rows) if r.startswith(" 1"))
    rows[k] = rows[k][:15] + "  garbage " + rows[k][25:]
    assert len(rows[k]) == 160
    st = os.stat(ds.hitran)
    open(ds.hitran, "w").write("\n".join(rows))
    os.utime(ds.hitran, ns=(st.st_atime_ns, st.st_mtime_ns + 1_000_000_000))
    rc, _ = parse(1)
    assert rc != 0
    buf = C.create_string_buffer(4096)
    L.grtcode_errstr(rc, buf, 4096)
    msg = buf.value.decode(errors="replace")
    assert "garbage" in msg or "convert" in msg.lower() or "record" in msg.lower(), msg
    rc, again = parse(2)
    assert rc == 0 and all(np.array_equal(one[f], again[f]) for f in one)
    # a non-numeric molecule field fails every request, as the reference's loop does
    rows[k] = "xx" + rows[k][2:]
    open(ds.hitran, "w").write("\n".join(rows))
    os.utime(ds.hitran, ns=(st.st_atime_ns, st.st_mtime_ns + 2_000_000_000))
    assert parse(2)[0] != 0


def test_tables_and_solar_match_oracle(built, tmp_path):
    ds = synth.make_dataset(str(tmp_path), 1., 3000., {1: 10}, seed=2, cfcs=("CFC11",),
                            cias=((0, 0),))
    g = built.spectral_grid(1., 3000., 0.1)
    sol = built.solar_flux(g, ds.solar)
    want = P.solar_flux(ds.solar, 1., 0.1, int(g.n))
    assert np.array_equal(sol, want)
    assert abs(np.sum(0.5*(sol[1:] + sol[:-1])*0.1) - 1.) < 1e-12
    L = built.load()
    L.grt_table_on_grid.argtypes = [C.c_char_p, C.c_int, C.c_double, C.c_double, C.c_ulonglong,
                                    C.POINTER(dp)]
    for path, cols in ((ds.cfc[0], 2), (os.path.join(ds.h2o_ctm_dir, "CKDS.csv"), 4),
                       (ds.o3_ctm, 2)):
        out = dp()
        assert L.grt_table_on_grid(path.encode(), cols, 1., 0.1, g.n, C.byref(out)) == 0
        got = np.ctypeslib.as_array(out, shape=(int(g.n),))
        assert np.array_equal(got, P.read_table(path, 1., 0.1, int(g.n), cols))
    out = dp()
    assert L.grt_table_on_grid(ds.o3_ctm.encode(), 4, 1., 0.1, g.n, C.byref(out)) == 9


def test_tips_host_function_is_the_oracle_function(built):
    for mol, t, iso in ((1, 296., 1), (2, 187.3, 2), (3, 250.25, 5), (7, 301.9, 1), (53, 99.5, 4)):
        assert built.tips_q(mol, t, iso) == P.tips_q(mol, t, iso)
    assert built.tips_q(1, 5., 1) == built.tips_q(1, 20., 1)       # clamped below the table


def test_heating_rates(built):
    p = np.array([100., 500., 1000.])
    up = np.array([240., 300., 400.])
    down = np.array([0., 150., 350.])
    rate = built.heating_rates(p, up, down)
    net = down - up
    want = 9.80665/1004.64*(net[:-1] - net[1:])/((p[1:] - p[:-1])*100.)*86400.
    assert np.allclose(rate, want, rtol=1e-14)


def test_synthetic_hitran_records_are_160_chars(tmp_path):
    ds = synth.make_dataset(str(tmp_path), 1., 100., {1: 50, 7: 20}, seed=4)
    lines = open(ds.hitran).read().split("\n")
    assert lines[-1] == ""
    assert all(len(l) == 160 for l in lines[:-1]) and len(lines) == 71


def test_fortran_bound_interfaces_match_the_headers():
    """Every C symbol the reference's Fortran module binds (fortran-bindings/
    grtcode_fortran.F90:127-554; fixture tests/golden/fortran_interfaces.json made by
    tools/fortran_interfaces.py) has a prototype in include/*.h with the same number of
    arguments and, position by position, the same class (by value / by reference, integer /
    real / pointer / character).  No Fortran compiler exists here, so this is the check that
    can be made; symbol export is covered by test_host_library_exports_reference_api."""
    import importlib.util
    spec = importlib.util.spec_from_file_location(
        "fortran_interfaces", os.path.join(ROOT, "tools", "fortran_interfaces.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    import json
    fx = json.load(open(os.path.join(ROOT, "tests", "golden", "fortran_interfaces.json")))["interfaces"]
    assert len(fx) >= 26
    protos = mod.parse_headers(os.path.join(ROOT, "include"))
    problems = mod.check(fx, protos)
    assert not problems, problems
    # the checker does notice a wrong argument list
    broken = dict(protos)
    broken["add_molecule"] = dict(protos["add_molecule"], args=protos["add_molecule"]["args"][:-1])
    assert mod.check(fx, broken)


def test_package_data_is_self_contained():
    """The product package reads its CIRC-1 input from its own data directory, not from
    tests/ (the two copies are the same file, both written by tools/make_golden.py)."""
    import filecmp
    from grtcode_b200 import workloads
    assert os.path.dirname(workloads.CIRC1).endswith(os.path.join("grtcode_b200", "data"))
    assert filecmp.cmp(workloads.CIRC1, os.path.join(ROOT, "tests", "golden", "circ1_profile.json"),
                       shallow=False)
    assert len(workloads.circ1_column().p) == 55


def test_external_partition_sums_replace_the_substitute(built, tmp_path):
    """GRT_TIPS_TABLE=<file of 'mol iso T Q' rows>: read once per process, so checked in a
    subprocess.  Listed (molecule, isotopologue) groups follow the file (linear between its
    points), other molecules keep the substitute, and add_molecule warns only for those."""
    import subprocess, sys, textwrap
    table = tmp_path/"tips.txt"
    rows = ["# mol iso T Q"] + [f"2 1 {t} {2.*t + 5.}" for t in range(100, 501, 50)]
    table.write_text("\n".join(rows) + "\n")
    ds = synth.make_dataset(str(tmp_path/"d"), 600., 650., {1: 5, 2: 5}, seed=2)
    code = textwrap.dedent(f"""
        import ctypes as C, sys
        sys.path.insert(0, {ROOT!r})
        from grtcode_b200 import api
        L = api.load()
        L.grt_tips_is_substitute.argtypes = [C.c_int]
        print("Q", api.tips_q(2, 225., 1), api.tips_q(2, 100., 1), api.tips_q(2, 500., 1))
        print("SUB", L.grt_tips_is_substitute(1), L.grt_tips_is_substitute(2))
        L.grtcode_set_verbosity(2)
        try:
            g = api.spectral_grid(600., 650., 0.1)
            go = api.GasOptics(5, g, {ds.hitran!r}, None, None)
            go.add_molecule(1); go.add_molecule(2)
        except Exception as e:          # no GPU here: the warning is printed before the upload
            print("EXC", type(e).__name__)
    """)
    env = dict(os.environ, GRT_TIPS_TABLE=str(table))
    out = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True)
    text = out.stdout + out.stderr
    assert "Q 455.0 205.0 1005.0" in text, text[-2000:]
    assert "SUB 1 0" in text
    # without the variable both are substitutes and the values are the built-in ones
    out2 = subprocess.run([sys.executable, "-c", code], env={k: v for k, v in os.environ.items()
                                                             if k != "GRT_TIPS_TABLE"},
                          capture_output=True, text=True)
    assert "SUB 1 1" in out2.stdout
    assert "Q 455.0" not in out2.stdout
