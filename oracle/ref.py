"""TEST INFRASTRUCTURE (oracle) -- ctypes bindings of the compiled reference.

``oracle/_ref/libgrtcode_ref.so`` is the UNMODIFIED reference C implementation
of the hot path (utilities/src, gas-optics/src, longwave/src, shortwave/src of
/root/reference) built by ``oracle/Makefile`` with ``gcc -O3 -fopenmp``, plus the
substitute partition sums (the reference's tips2017.c is missing from the mount).
Only tests/, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline legs may
import this module; the product never does.

Struct layouts mirror utilities/src/spectral_grid.h:33-39, optics.h:30-38,
longwave/src/longwave.h:29-39, shortwave/src/shortwave.h:28-38,
shortwave/src/solar_flux.h:27-32, gas-optics/src/line_shape.h:26-35 and
gas-optics/src/spectral_bin.h:29-50.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
HOST_ONLY = -1
LINE_SAMPLE = 2  # gas_optics.h:88-93
# Working precision of this process (see grtcode_b200/api.py): GRT_PRECISION=single loads the
# reference compiled with its own -DSINGLE_PRECISION (fp_t = float, configure.ac:25-32).
SINGLE = os.environ.get("GRT_PRECISION", "double") == "single"
REAL = C.c_float if SINGLE else C.c_double
DTYPE = np.float32 if SINGLE else np.float64


def lib_path(fast: bool = False) -> str:
    if SINGLE:
        return os.path.join(HERE, "_ref", "libgrtcode_ref_sp_fast.so" if fast else
                            "libgrtcode_ref_sp.so")
    return os.path.join(HERE, "_ref", "libgrtcode_ref_fast.so" if fast else "libgrtcode_ref.so")


def available(fast: bool = False) -> bool:
    return os.path.exists(lib_path(fast))


class SpectralGrid(C.Structure):
    _fields_ = [("dw", C.c_double), ("n", C.c_uint64), ("wn", C.c_double), ("w0", C.c_double)]


class Optics(C.Structure):
    _fields_ = [("device", C.c_int), ("g", C.POINTER(REAL)), ("grid", SpectralGrid),
                ("num_layers", C.c_int), ("omega", C.POINTER(REAL)),
                ("tau", C.POINTER(REAL))]


class Longwave(C.Structure):
    _fields_ = [("num_levels", C.c_int), ("grid", SpectralGrid), ("device", C.c_int),
                ("layer_temperature", C.c_void_p), ("level_temperature", C.c_void_p),
                ("emissivity", C.c_void_p), ("flux_up", C.c_void_p), ("flux_down", C.c_void_p)]


class Shortwave(C.Structure):
    _fields_ = [("num_levels", C.c_int), ("grid", SpectralGrid), ("device", C.c_int),
                ("solar_flux", C.c_void_p), ("sfc_alpha_dir", C.c_void_p),
                ("sfc_alpha_dif", C.c_void_p), ("flux_up", C.c_void_p), ("flux_down", C.c_void_p)]


class SolarFlux(C.Structure):
    _fields_ = [("grid", SpectralGrid), ("incident_flux", C.POINTER(REAL)),
                ("n", C.c_uint64)]


class LineShapeInputs(C.Structure):
    _fields_ = [("w", REAL), ("num_wpoints", C.c_uint64), ("wres", REAL),
                ("line_center", REAL), ("lorentz_hwhm", REAL),
                ("doppler_hwhm", REAL), ("eta", REAL)]


class SpectralBins(C.Structure):
    _fields_ = [("num_layers", C.c_int), ("w0", REAL), ("wres", REAL),
                ("num_wpoints", C.c_uint64), ("n", C.c_uint64), ("width", REAL),
                ("isize", C.c_uint64), ("ppb", C.c_int), ("do_interp", C.c_int),
                ("last_ppb", C.c_int), ("do_last_interp", C.c_int), ("w", C.c_void_p),
                ("tau", C.c_void_p), ("l", C.c_void_p), ("r", C.c_void_p), ("device", C.c_int)]


def _dp(a: np.ndarray):
    assert a.dtype == DTYPE and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.POINTER(REAL))


class RefError(RuntimeError):
    pass


class Ref:
    """Loaded reference library + thin helpers."""

    def __init__(self, fast: bool = False, threads: int | None = None):
        if not available(fast):
            raise FileNotFoundError(f"{lib_path(fast)} missing: run `make -C oracle ref`")
        self.lib = C.CDLL(lib_path(fast), mode=C.RTLD_LOCAL)
        self.gomp = None
        try:
            self.gomp = C.CDLL("libgomp.so.1")
            if threads is not None:
                self.gomp.omp_set_num_threads(int(threads))
        except OSError:
            pass
        L = self.lib
        L.Q.restype = REAL
        L.Q.argtypes = [C.c_int, REAL, C.c_int]
        L.rfm_voigt_line_shape.argtypes = [LineShapeInputs, C.POINTER(REAL)]
        L.calc_optical_depth_line_sample.argtypes = [
            C.c_uint64, C.c_int, C.POINTER(REAL), C.POINTER(REAL),
            C.POINTER(REAL), C.POINTER(REAL), C.POINTER(REAL), SpectralBins,
            C.POINTER(REAL), C.c_void_p, C.c_void_p]
        L.calculate_lw_fluxes.argtypes = [
            C.POINTER(Longwave), C.POINTER(Optics), REAL, C.POINTER(REAL),
            C.POINTER(REAL), C.POINTER(REAL), C.POINTER(REAL),
            C.POINTER(REAL)]
        L.calculate_sw_fluxes.argtypes = [
            C.POINTER(Shortwave), C.POINTER(Optics), REAL, REAL,
            C.POINTER(REAL), C.POINTER(REAL), REAL, C.POINTER(REAL),
            C.POINTER(REAL), C.POINTER(REAL)]
        L.create_spectral_grid.argtypes = [C.POINTER(SpectralGrid), C.c_double, C.c_double,
                                           C.c_double]
        L.add_molecule.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.add_cfc.argtypes = [C.c_void_p, C.c_int, C.c_char_p]
        L.add_cia.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_char_p]

    def set_threads(self, n: int):
        if self.gomp is not None:
            self.gomp.omp_set_num_threads(int(n))

    def max_threads(self) -> int:
        return int(self.gomp.omp_get_max_threads()) if self.gomp is not None else 1

    def check(self, rc: int, what: str = ""):
        if rc != 0:
            buf = C.create_string_buffer(4096)
            self.lib.grtcode_errstr(rc, buf, 4096)
            raise RefError(f"reference {what} failed rc={rc}: {buf.value.decode(errors='replace')}")

    def grid(self, w0: float, wn: float, dw: float) -> SpectralGrid:
        g = SpectralGrid()
        self.check(self.lib.create_spectral_grid(C.byref(g), w0, wn, dw), "create_spectral_grid")
        return g

    # -- low-level kernels (gas-optics/src/kernels.c, RFM_voigt.c) -------------
    def Q(self, mol: int, t: float, iso: int) -> float:
        return float(self.lib.Q(mol, t, iso))

    def voigt(self, w: float, n: int, wres: float, center: float, gamma: float,
              alpha: float) -> np.ndarray:
        k = np.zeros(n, dtype=DTYPE)
        v = LineShapeInputs(w, n, wres, center, gamma, alpha, 0.)
        self.check(self.lib.rfm_voigt_line_shape(v, _dp(k)), "rfm_voigt_line_shape")
        return k

    def layer_kernel(self, name: str, *args):
        self.check(getattr(self.lib, name)(*args), name)

    def line_sample(self, v, s, gamma, alpha, ns, w0, wres, n) -> np.ndarray:
        """calc_optical_depth_line_sample (kernels.c:410) on [layers, lines] arrays."""
        layers, lines = v.shape
        tau = np.zeros((layers, n), dtype=DTYPE)
        bins = SpectralBins()
        bins.num_layers = layers
        bins.w0 = w0
        bins.wres = wres
        bins.num_wpoints = n
        self.check(self.lib.calc_optical_depth_line_sample(
            lines, layers, _dp(v), _dp(s), _dp(gamma), _dp(alpha), _dp(ns), bins, _dp(tau),
            None, None), "line_sample")
        return tau


class RefOptics:
    """An Optics_t living on the host (HOST_ONLY)."""

    def __init__(self, ref: Ref, num_layers: int, grid: SpectralGrid):
        self.ref = ref
        self.o = Optics()
        dev = C.c_int(HOST_ONLY)
        ref.check(ref.lib.create_optics(C.byref(self.o), num_layers, C.byref(grid),
                                        C.byref(dev)), "create_optics")
        self.shape = (num_layers, int(grid.n))
        self._live = True

    def _arr(self, p):
        return np.ctypeslib.as_array(p, shape=self.shape)

    @property
    def tau(self):
        return self._arr(self.o.tau)

    @property
    def omega(self):
        return self._arr(self.o.omega)

    @property
    def g(self):
        return self._arr(self.o.g)

    def close(self):
        if self._live:
            self.ref.lib.destroy_optics(C.byref(self.o))
            self._live = False


class RefGasOptics:
    """create_gas_optics ... calculate_optical_depth (gas_optics.h:99-180) on the host."""

    def __init__(self, ref: Ref, num_levels: int, grid: SpectralGrid, hitran: str,
                 h2o_ctm_dir: str | None = None, o3_ctm: str | None = None,
                 method: int | None = LINE_SAMPLE):
        self.ref = ref
        self.grid = grid
        self.num_levels = num_levels
        self.buf = C.create_string_buffer(1 << 20)   # >= sizeof(GasOptics_t)
        dev = C.c_int(HOST_ONLY)
        m = C.c_int(method) if method is not None else None
        ref.check(ref.lib.create_gas_optics(
            self.buf, num_levels, C.byref(grid), C.byref(dev), hitran.encode(),
            h2o_ctm_dir.encode() if h2o_ctm_dir else None,
            o3_ctm.encode() if o3_ctm else None, None,
            C.byref(m) if m is not None else None), "create_gas_optics")
        self._live = True

    def add_molecule(self, mol: int, wmin: float | None = None, wmax: float | None = None):
        a = C.byref(C.c_double(wmin)) if wmin is not None else None
        b = C.byref(C.c_double(wmax)) if wmax is not None else None
        self.ref.check(self.ref.lib.add_molecule(self.buf, mol, a, b), "add_molecule")

    def add_cfc(self, cfc: int, path: str):
        self.ref.check(self.ref.lib.add_cfc(self.buf, cfc, path.encode()), "add_cfc")

    def add_cia(self, a: int, b: int, path: str):
        self.ref.check(self.ref.lib.add_cia(self.buf, a, b, path.encode()), "add_cia")

    def set_ppmv(self, kind: str, ident: int, ppmv: np.ndarray):
        ppmv = np.ascontiguousarray(ppmv, dtype=DTYPE)
        fn = getattr(self.ref.lib, {"mol": "set_molecule_ppmv", "cfc": "set_cfc_ppmv",
                                    "cia": "set_cia_ppmv"}[kind])
        self.ref.check(fn(self.buf, C.c_int(ident), _dp(ppmv)), f"set_{kind}_ppmv")

    def optical_depth(self, p_mb: np.ndarray, t: np.ndarray, optics: RefOptics):
        p = np.ascontiguousarray(p_mb, dtype=DTYPE)
        tt = np.ascontiguousarray(t, dtype=DTYPE)
        self.ref.check(self.ref.lib.calculate_optical_depth(
            self.buf, _dp(p), _dp(tt), C.byref(optics.o)), "calculate_optical_depth")

    def close(self):
        if self._live:
            self.ref.lib.destroy_gas_optics(self.buf)
            self._live = False


def rayleigh(ref: Ref, optics: RefOptics, p_mb: np.ndarray):
    p = np.ascontiguousarray(p_mb, dtype=DTYPE)
    ref.check(ref.lib.rayleigh_scattering(C.byref(optics.o), _dp(p)), "rayleigh_scattering")


def add_optics(ref: Ref, parts: list[RefOptics]) -> RefOptics:
    arr = (C.POINTER(Optics)*len(parts))(*[C.pointer(p.o) for p in parts])
    out = RefOptics.__new__(RefOptics)
    out.ref = ref
    out.o = Optics()
    ref.check(ref.lib.add_optics(arr, len(parts), C.byref(out.o)), "add_optics")
    out.shape = parts[0].shape
    out._live = True
    return out


def lw_fluxes(ref: Ref, optics: RefOptics, t_surf, t_layers, t_levels, emis):
    levels = optics.shape[0] + 1
    n = optics.shape[1]
    lw = Longwave()
    dev = C.c_int(HOST_ONLY)
    ref.check(ref.lib.create_longwave(C.byref(lw), levels, C.byref(optics.o.grid),
                                      C.byref(dev)), "create_longwave")
    up = np.zeros((levels, n), dtype=DTYPE)
    dn = np.zeros((levels, n), dtype=DTYPE)
    tl = np.ascontiguousarray(t_layers, dtype=DTYPE)
    tv = np.ascontiguousarray(t_levels, dtype=DTYPE)
    em = np.ascontiguousarray(emis, dtype=DTYPE)
    ref.check(ref.lib.calculate_lw_fluxes(C.byref(lw), C.byref(optics.o), float(t_surf),
                                          _dp(tl), _dp(tv), _dp(em), _dp(up), _dp(dn)),
              "calculate_lw_fluxes")
    return up, dn


def sw_fluxes(ref: Ref, optics: RefOptics, mu_dir, mu_dif, alb_dir, alb_dif, tsi, solar):
    levels = optics.shape[0] + 1
    n = optics.shape[1]
    sw = Shortwave()
    dev = C.c_int(HOST_ONLY)
    ref.check(ref.lib.create_shortwave(C.byref(sw), levels, C.byref(optics.o.grid),
                                       C.byref(dev)), "create_shortwave")
    up = np.zeros((levels, n), dtype=DTYPE)
    dn = np.zeros((levels, n), dtype=DTYPE)
    a1 = np.ascontiguousarray(alb_dir, dtype=DTYPE)
    a2 = np.ascontiguousarray(alb_dif, dtype=DTYPE)
    so = np.ascontiguousarray(solar, dtype=DTYPE)
    ref.check(ref.lib.calculate_sw_fluxes(C.byref(sw), C.byref(optics.o), float(mu_dir),
                                          float(mu_dif), _dp(a1), _dp(a2), float(tsi), _dp(so),
                                          _dp(up), _dp(dn)), "calculate_sw_fluxes")
    return up, dn


def solar_flux(ref: Ref, grid: SpectralGrid, path: str) -> np.ndarray:
    s = SolarFlux()
    ref.check(ref.lib.create_solar_flux(C.byref(s), C.byref(grid), path.encode()),
              "create_solar_flux")
    out = np.ctypeslib.as_array(s.incident_flux, shape=(int(s.n),)).copy()
    ref.lib.destroy_solar_flux(C.byref(s))
    return out


def broadband(flux: np.ndarray, dw: float) -> np.ndarray:
    """Trapezoid of framework/src/driver.c:302-319, per level (numpy pairwise sum)."""
    flux = np.asarray(flux, dtype=np.float64)
    return np.sum(0.5*(flux[:, :-1] + flux[:, 1:])*dw, axis=1)
