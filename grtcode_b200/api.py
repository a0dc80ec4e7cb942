"""ctypes bindings of ``libgrtcode.so`` -- the reference's C API on sm_100a.

Every function below is the entry point of the same name in the reference's
``libgas_optics`` / ``liblongwave`` / ``libshortwave`` / utilities
(include/*.h cite the reference headers), loaded from the in-tree shared
libraries built by ``__graft_entry__.build()``.  There is no fallback: if the
libraries are missing, or no GPU is visible, the calls raise.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
# Working precision of this PROCESS: GRT_PRECISION=single selects the libraries built with the
# reference's -DSINGLE_PRECISION (fp_t = float, configure.ac:25-32): libgrtcode_sp.so +
# libgrt_sm100_sp.so.  Every array that crosses the API then is float32.
PRECISION = os.environ.get("GRT_PRECISION", "double")
if PRECISION not in ("double", "single"):
    raise ValueError(f"GRT_PRECISION={PRECISION!r}: expected 'double' or 'single'")
SINGLE = PRECISION == "single"
REAL = C.c_float if SINGLE else C.c_double
DTYPE = np.float32 if SINGLE else np.float64
# GRT_LIB_DIR: directory of an alternative build of the two libraries (kernel variants made
# by tools/build_variant.sh for A/B measurements); default: the in-tree build.
LIB_DIR = os.environ.get("GRT_LIB_DIR") or HERE
LIB_HOST = os.path.join(LIB_DIR, "libgrtcode_sp.so" if SINGLE else "libgrtcode.so")
LIB_CUDA = os.path.join(LIB_DIR, "libgrt_sm100_sp.so" if SINGLE else "libgrt_sm100.so")

WAVENUMBER_SWEEP, LINE_SWEEP, LINE_SAMPLE = 0, 1, 2   # include/gas_optics.h
NUM_MOLS, NUM_CFCS, NUM_CIAS, MAX_NUM_CIAS, DIR_PATH_LEN = 53, 21, 2, 3, 1024


class GrtcodeError(RuntimeError):
    def __init__(self, code: int, what: str, text: str):
        super().__init__(f"{what} failed with code {code}: {text}")
        self.code = code


class SpectralGrid(C.Structure):
    _fields_ = [("dw", C.c_double), ("n", C.c_uint64), ("wn", C.c_double), ("w0", C.c_double)]


class Optics(C.Structure):
    _fields_ = [("device", C.c_int), ("g", C.c_void_p), ("grid", SpectralGrid),
                ("num_layers", C.c_int), ("omega", C.c_void_p), ("tau", C.c_void_p)]


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


class GasOpticsStruct(C.Structure):
    _fields_ = [("device", C.c_int), ("num_levels", C.c_int), ("num_layers", C.c_int),
                ("num_molecules", C.c_int), ("molecule_bit_field", C.c_uint64),
                ("molecule_ids", C.c_int*NUM_MOLS), ("num_lines", C.c_uint64*NUM_MOLS),
                ("num_cfcs", C.c_int), ("cfc_bit_field", C.c_uint64), ("num_cias", C.c_int),
                ("cia_bit_field", C.c_uint64), ("cia_pairs", (C.c_int*2)*MAX_NUM_CIAS),
                ("h2o_ctm_dir", C.c_char*DIR_PATH_LEN), ("use_h2o_ctm", C.c_int),
                ("o3_ctm_file", C.c_char*DIR_PATH_LEN), ("use_o3_ctm", C.c_int),
                ("grid", SpectralGrid), ("hitran_path", C.c_char*DIR_PATH_LEN),
                ("wcutoff", C.c_double), ("optical_depth_method", C.c_int),
                ("x", C.c_void_p), ("x_cfc", C.c_void_p), ("x_cia", C.c_void_p),
                ("sm100", C.c_void_p)]


_dbl_p = C.POINTER(REAL)          # fp_t * of the loaded build
_lib = None
_cuda = None


def libraries_present() -> bool:
    return os.path.exists(LIB_HOST) and os.path.exists(LIB_CUDA)


def load():
    """Load the two shared libraries (raises if they have not been built)."""
    global _lib, _cuda
    if _lib is not None:
        return _lib
    if not libraries_present():
        raise FileNotFoundError(
            f"{LIB_CUDA} / {LIB_HOST} not built: run `python -c 'import __graft_entry__ as g; "
            "g.build()'` (there is no CPU fallback)")
    # RTLD_LOCAL: the oracle libraries export the same public names (they are the
    # reference); neither side may bind to the other's symbols.
    _cuda = C.CDLL(LIB_CUDA, mode=C.RTLD_LOCAL)
    L = C.CDLL(LIB_HOST, mode=C.RTLD_LOCAL)
    L.create_spectral_grid.argtypes = [C.POINTER(SpectralGrid), C.c_double, C.c_double, C.c_double]
    L.create_device.argtypes = [C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.get_num_gpus.argtypes = [C.POINTER(C.c_int), C.c_int]
    L.create_optics.argtypes = [C.POINTER(Optics), C.c_int, C.POINTER(SpectralGrid),
                                C.POINTER(C.c_int)]
    L.destroy_optics.argtypes = [C.POINTER(Optics)]
    L.add_optics.argtypes = [C.POINTER(C.POINTER(Optics)), C.c_int, C.POINTER(Optics)]
    L.update_optics.argtypes = [C.POINTER(Optics), _dbl_p, _dbl_p, _dbl_p]
    L.optical_properties.argtypes = [C.POINTER(Optics), _dbl_p, _dbl_p, _dbl_p]
    L.create_gas_optics.argtypes = [C.POINTER(GasOpticsStruct), C.c_int, C.POINTER(SpectralGrid),
                                    C.POINTER(C.c_int), C.c_char_p, C.c_char_p, C.c_char_p,
                                    C.POINTER(C.c_double), C.POINTER(C.c_int)]
    L.destroy_gas_optics.argtypes = [C.POINTER(GasOpticsStruct)]
    L.add_molecule.argtypes = [C.POINTER(GasOpticsStruct), C.c_int, C.POINTER(C.c_double),
                               C.POINTER(C.c_double)]
    L.set_molecule_ppmv.argtypes = [C.POINTER(GasOpticsStruct), C.c_int, _dbl_p]
    L.add_cfc.argtypes = [C.POINTER(GasOpticsStruct), C.c_int, C.c_char_p]
    L.set_cfc_ppmv.argtypes = [C.POINTER(GasOpticsStruct), C.c_int, _dbl_p]
    L.add_cia.argtypes = [C.POINTER(GasOpticsStruct), C.c_int, C.c_int, C.c_char_p]
    L.set_cia_ppmv.argtypes = [C.POINTER(GasOpticsStruct), C.c_int, _dbl_p]
    L.calculate_optical_depth.argtypes = [C.POINTER(GasOpticsStruct), _dbl_p, _dbl_p,
                                          C.POINTER(Optics)]
    L.get_num_molecules.argtypes = [C.POINTER(GasOpticsStruct), C.POINTER(C.c_int)]
    L.gas_optics_stats.argtypes = [C.POINTER(GasOpticsStruct), C.POINTER(C.c_uint64)]
    L.gas_optics_patch_counts.argtypes = [C.POINTER(GasOpticsStruct), C.POINTER(C.c_uint64)]
    L.gas_optics_class_histogram.argtypes = [C.POINTER(GasOpticsStruct), C.POINTER(C.c_uint64)]
    L.gas_optics_debug_records.argtypes = [C.POINTER(GasOpticsStruct), C.c_int, C.c_uint64,
                                           C.c_uint64, C.POINTER(C.c_float), C.POINTER(C.c_float),
                                           C.POINTER(C.c_double)]
    L.gas_optics_debug_line_order.argtypes = [C.POINTER(GasOpticsStruct), C.c_uint64,
                                              C.POINTER(C.c_uint64)]
    L.gas_optics_debug_table.argtypes = [C.POINTER(GasOpticsStruct), C.c_int, C.c_int,
                                         C.POINTER(C.c_double)]
    L.grt_table_on_grid.argtypes = [C.c_char_p, C.c_int, C.c_double, C.c_double, C.c_ulonglong,
                                    C.POINTER(_dbl_p)]
    L.rayleigh_scattering.argtypes = [C.POINTER(Optics), _dbl_p]
    L.create_longwave.argtypes = [C.POINTER(Longwave), C.c_int, C.POINTER(SpectralGrid),
                                  C.POINTER(C.c_int)]
    L.destroy_longwave.argtypes = [C.POINTER(Longwave)]
    L.calculate_lw_fluxes.argtypes = [C.POINTER(Longwave), C.POINTER(Optics), REAL, _dbl_p,
                                      _dbl_p, _dbl_p, _dbl_p, _dbl_p]
    L.calculate_lw_column.argtypes = [C.POINTER(Longwave), C.POINTER(Optics), _dbl_p, REAL,
                                      _dbl_p, _dbl_p, _dbl_p, _dbl_p, _dbl_p]
    L.calculate_lw_column_window.argtypes = [C.POINTER(Longwave), C.POINTER(Optics), _dbl_p,
                                             REAL, _dbl_p, _dbl_p, _dbl_p, C.c_uint64,
                                             C.c_uint64, _dbl_p, _dbl_p]
    L.calculate_sw_column_window.argtypes = [C.POINTER(Shortwave), C.POINTER(Optics), _dbl_p,
                                             REAL, REAL, _dbl_p, _dbl_p, REAL,
                                             _dbl_p, C.c_uint64, C.c_uint64, _dbl_p, _dbl_p]
    L.create_shortwave.argtypes = [C.POINTER(Shortwave), C.c_int, C.POINTER(SpectralGrid),
                                   C.POINTER(C.c_int)]
    L.destroy_shortwave.argtypes = [C.POINTER(Shortwave)]
    L.calculate_sw_fluxes.argtypes = [C.POINTER(Shortwave), C.POINTER(Optics), REAL,
                                      REAL, _dbl_p, _dbl_p, REAL, _dbl_p, _dbl_p,
                                      _dbl_p]
    L.calculate_sw_column.argtypes = [C.POINTER(Shortwave), C.POINTER(Optics), _dbl_p, REAL,
                                      REAL, _dbl_p, _dbl_p, REAL, _dbl_p, _dbl_p,
                                      _dbl_p]
    L.create_solar_flux.argtypes = [C.POINTER(SolarFlux), C.POINTER(SpectralGrid), C.c_char_p]
    L.destroy_solar_flux.argtypes = [C.POINTER(SolarFlux)]
    L.integrate_fluxes.argtypes = [C.c_int, C.c_int, C.POINTER(SpectralGrid), C.c_void_p, _dbl_p]
    L.heating_rates.argtypes = [C.c_int, _dbl_p, _dbl_p, _dbl_p, _dbl_p]
    L.grtcode_errstr.argtypes = [C.c_int, C.c_char_p, C.c_int]
    L.grt_tips_q.restype = C.c_double
    L.grt_tips_q.argtypes = [C.c_int, C.c_double, C.c_int]
    _cuda.grt_measure_fma_peak.argtypes = [C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    _cuda.grt_synchronize.argtypes = [C.c_int]
    _cuda.grt_selftest_arithmetic.argtypes = [C.c_int, C.c_ulonglong, C.c_ulonglong,
                                              C.POINTER(C.c_ulonglong)]
    _cuda.grt_selftest_voigt.argtypes = [C.c_int, C.c_ulonglong, C.c_ulonglong,
                                         C.POINTER(C.c_ulonglong)]
    _cuda.grt_stream_handle.argtypes = [C.c_int, C.POINTER(C.c_void_p)]
    _cuda.grt_host_alloc.argtypes = [C.POINTER(C.c_void_p), C.c_size_t]
    _cuda.grt_host_free.argtypes = [C.c_void_p]
    _cuda.grt_copy_to_host.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_size_t]
    _lib = L
    return L


def cuda_layer():
    load()
    return _cuda


def _check(rc: int, what: str):
    if rc != 0:
        buf = C.create_string_buffer(4096)
        _lib.grtcode_errstr(rc, buf, 4096)
        raise GrtcodeError(rc, what, buf.value.decode(errors="replace").strip())


def _dp(a: np.ndarray):
    assert a.dtype == DTYPE and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(_dbl_p)


def _f64(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=DTYPE)


def num_gpus() -> int:
    n = C.c_int(0)
    _check(load().get_num_gpus(C.byref(n), 0), "get_num_gpus")
    return n.value


def spectral_grid(w0: float, wn: float, dw: float) -> SpectralGrid:
    g = SpectralGrid()
    _check(load().create_spectral_grid(C.byref(g), w0, wn, dw), "create_spectral_grid")
    return g


def tips_q(mol: int, t: float, iso: int) -> float:
    return float(load().grt_tips_q(mol, t, iso))


def stream_handle(device: int) -> int:
    """cudaStream_t of the library's stream on ``device`` (for event timing)."""
    h = C.c_void_p()
    if cuda_layer().grt_stream_handle(device, C.byref(h)) != 0:
        raise RuntimeError("grt_stream_handle failed")
    return h.value or 0


def selftest_arithmetic(n: int, seed: int = 1, device: int = 0) -> int:
    bad = C.c_ulonglong(0)
    if cuda_layer().grt_selftest_arithmetic(device, n, seed, C.byref(bad)) != 0:
        raise RuntimeError("grt_selftest_arithmetic failed")
    return int(bad.value)


def selftest_voigt(n: int, seed: int = 1, device: int = 0) -> int:
    bad = C.c_ulonglong(0)
    if cuda_layer().grt_selftest_voigt(device, n, seed, C.byref(bad)) != 0:
        raise RuntimeError("grt_selftest_voigt failed")
    return int(bad.value)


def measure_fma_peak(device: int = 0) -> tuple[float, float]:
    a, b = C.c_double(0.), C.c_double(0.)
    if cuda_layer().grt_measure_fma_peak(device, C.byref(a), C.byref(b)) != 0:
        raise RuntimeError("grt_measure_fma_peak failed")
    return a.value, b.value


class OpticsObject:
    """Optics_t on a GPU (include/optics.h)."""

    def __init__(self, num_layers: int, grid: SpectralGrid, device: int = 0, _create=True):
        self.L = load()
        self.o = Optics()
        self.shape = (num_layers, int(grid.n))
        self._live = False
        if _create:
            dev = C.c_int(device)
            _check(self.L.create_optics(C.byref(self.o), num_layers, C.byref(grid),
                                        C.byref(dev)), "create_optics")
            self._live = True

    def arrays(self, tau=True, omega=False, g=False):
        out = []
        bufs = []
        for want in (tau, omega, g):
            bufs.append(np.empty(self.shape, dtype=DTYPE) if want else None)
        _check(self.L.optical_properties(
            C.byref(self.o), *[(_dp(b) if b is not None else None) for b in bufs]),
            "optical_properties")
        out = [b for b in bufs if b is not None]
        return out[0] if len(out) == 1 else out

    @property
    def tau(self) -> np.ndarray:
        return self.arrays(True, False, False)

    def update(self, tau, omega, g):
        _check(self.L.update_optics(C.byref(self.o), _dp(_f64(tau)), _dp(_f64(omega)),
                                    _dp(_f64(g))), "update_optics")

    def close(self):
        if self._live:
            _check(self.L.destroy_optics(C.byref(self.o)), "destroy_optics")
            self._live = False

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def add_optics(parts: list[OpticsObject]) -> OpticsObject:
    L = load()
    arr = (C.POINTER(Optics)*len(parts))(*[C.pointer(p.o) for p in parts])
    grid = parts[0].o.grid
    out = OpticsObject(parts[0].shape[0], grid, _create=False)
    _check(L.add_optics(arr, len(parts), C.byref(out.o)), "add_optics")
    out._live = True
    return out


def rayleigh_scattering(optics: OpticsObject, p_mb):
    _check(load().rayleigh_scattering(C.byref(optics.o), _dp(_f64(p_mb))), "rayleigh_scattering")


class GasOptics:
    """GasOptics_t (include/gas_optics.h): create -> add_* -> set_*_ppmv -> optical depth."""

    def __init__(self, num_levels: int, grid: SpectralGrid, hitran: str,
                 h2o_ctm_dir: str | None = None, o3_ctm: str | None = None,
                 method: int | None = LINE_SAMPLE, device: int = 0,
                 wcutoff: float | None = None):
        self.L = load()
        self.s = GasOpticsStruct()
        self.grid = grid
        self.num_levels = num_levels
        dev = C.c_int(device)
        m = C.c_int(method) if method is not None else None
        wc = C.c_double(wcutoff) if wcutoff is not None else None
        self._live = False
        _check(self.L.create_gas_optics(
            C.byref(self.s), num_levels, C.byref(grid), C.byref(dev), hitran.encode(),
            h2o_ctm_dir.encode() if h2o_ctm_dir else None, o3_ctm.encode() if o3_ctm else None,
            C.byref(wc) if wc is not None else None, C.byref(m) if m is not None else None),
            "create_gas_optics")
        self._live = True

    def add_molecule(self, mol: int, wmin: float | None = None, wmax: float | None = None):
        a = C.byref(C.c_double(wmin)) if wmin is not None else None
        b = C.byref(C.c_double(wmax)) if wmax is not None else None
        _check(self.L.add_molecule(C.byref(self.s), mol, a, b), "add_molecule")

    def add_cfc(self, cfc: int, path: str):
        _check(self.L.add_cfc(C.byref(self.s), cfc, path.encode()), "add_cfc")

    def add_cia(self, a: int, b: int, path: str):
        _check(self.L.add_cia(C.byref(self.s), a, b, path.encode()), "add_cia")

    def set_ppmv(self, kind: str, ident: int, ppmv):
        fn = {"mol": self.L.set_molecule_ppmv, "cfc": self.L.set_cfc_ppmv,
              "cia": self.L.set_cia_ppmv}[kind]
        _check(fn(C.byref(self.s), ident, _dp(_f64(ppmv))), f"set_{kind}_ppmv")

    def optical_depth(self, p_mb, t, optics: OpticsObject):
        _check(self.L.calculate_optical_depth(C.byref(self.s), _dp(_f64(p_mb)), _dp(_f64(t)),
                                              C.byref(optics.o)), "calculate_optical_depth")

    def num_molecules(self) -> int:
        n = C.c_int(0)
        _check(self.L.get_num_molecules(C.byref(self.s), C.byref(n)), "get_num_molecules")
        return n.value

    def stats(self) -> dict:
        out = (C.c_uint64*8)()
        _check(self.L.gas_optics_stats(C.byref(self.s), out), "gas_optics_stats")
        return {"records": out[0], "samples": out[1], "launches": out[2], "lorentz": out[3],
                "prep_ns": out[4], "wing_ns": out[5], "core_ns": out[6], "core_samples": out[7]}

    def patch_counts(self) -> tuple[int, int]:
        """(records the host re-evaluated with its libm in the last call, since creation)."""
        out = (C.c_uint64*2)()
        _check(self.L.gas_optics_patch_counts(C.byref(self.s), out), "gas_optics_patch_counts")
        return int(out[0]), int(out[1])

    def class_histogram(self) -> dict:
        """Line-tile visits per wing-sweep line class of the last line_sample call (needs
        GRT_CLASS_HIST=1 in the environment of that call)."""
        out = (C.c_uint64*8)()
        _check(self.L.gas_optics_class_histogram(C.byref(self.s), out),
               "gas_optics_class_histogram")
        names = ("fast", "region1", "far_edge", "near", "lorentz", "lorentz_edge")
        d = {n: int(out[k]) for k, n in enumerate(names)}
        d["tile_points"] = int(out[6])
        return d

    def debug_records(self, layer: int, first: int, count: int):
        """(REPWID float32[count], Y float32[count], shifted centre float64[count]) of K1's
        records of one layer, in device line order."""
        rep = np.empty(count, dtype=np.float32)
        y = np.empty(count, dtype=np.float32)
        v = np.empty(count, dtype=np.float64)
        _check(self.L.gas_optics_debug_records(
            C.byref(self.s), layer, first, count, rep.ctypes.data_as(C.POINTER(C.c_float)),
            y.ctypes.data_as(C.POINTER(C.c_float)), v.ctypes.data_as(C.POINTER(C.c_double))),
            "gas_optics_debug_records")
        return rep, y, v

    def debug_table(self, kind: int, index: int = 0) -> np.ndarray:
        """A continuum / cross-section table as it lives on the device (float64[n]): kind
        0-3 = H2O S296 / CKDS / F296 / CKDF, 4 = O3, 5 = CFC `index`, 6 = CIA pair `index`."""
        out = np.empty(int(self.grid.n), dtype=np.float64)
        _check(self.L.gas_optics_debug_table(C.byref(self.s), kind, index,
                                             out.ctypes.data_as(C.POINTER(C.c_double))),
               "gas_optics_debug_table")
        return out

    def debug_line_order(self, count: int) -> np.ndarray:
        order = np.empty(count, dtype=np.uint64)
        _check(self.L.gas_optics_debug_line_order(
            C.byref(self.s), count, order.ctypes.data_as(C.POINTER(C.c_uint64))),
            "gas_optics_debug_line_order")
        return order

    def close(self):
        if self._live:
            _check(self.L.destroy_gas_optics(C.byref(self.s)), "destroy_gas_optics")
            self._live = False

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class LongwaveSolver:
    def __init__(self, num_levels: int, grid: SpectralGrid, device: int = 0):
        self.L = load()
        self.s = Longwave()
        self.num_levels = num_levels
        self.n = int(grid.n)
        dev = C.c_int(device)
        self._live = False
        _check(self.L.create_longwave(C.byref(self.s), num_levels, C.byref(grid), C.byref(dev)),
               "create_longwave")
        self._live = True

    def fluxes(self, optics: OpticsObject, t_surf, t_layers, t_levels, emis, out=None):
        """calculate_lw_fluxes; out = (up, down) preallocated float64[levels, n] arrays."""
        up, dn = out if out is not None else (np.empty((self.num_levels, self.n), dtype=DTYPE),
                                              np.empty((self.num_levels, self.n), dtype=DTYPE))
        _check(self.L.calculate_lw_fluxes(C.byref(self.s), C.byref(optics.o), float(t_surf),
                                          _dp(_f64(t_layers)), _dp(_f64(t_levels)),
                                          _dp(_f64(emis)), _dp(up), _dp(dn)),
               "calculate_lw_fluxes")
        return up, dn

    def column(self, gas: OpticsObject, p_mb, t_surf, t_layers, t_levels, emis, window=None):
        """Broadband fluxes of one column; window = (first, last) grid indices restricts
        the integration (spectral split with halo points)."""
        up = np.empty(self.num_levels, dtype=DTYPE)
        dn = np.empty(self.num_levels, dtype=DTYPE)
        first, last = window if window is not None else (0, self.n - 1)
        _check(self.L.calculate_lw_column_window(
            C.byref(self.s), C.byref(gas.o), _dp(_f64(p_mb)), float(t_surf), _dp(_f64(t_layers)),
            _dp(_f64(t_levels)), _dp(_f64(emis)) if emis is not None else None, first, last,
            _dp(up), _dp(dn)), "calculate_lw_column")
        return up, dn

    def close(self):
        if self._live:
            _check(self.L.destroy_longwave(C.byref(self.s)), "destroy_longwave")
            self._live = False

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class ShortwaveSolver:
    def __init__(self, num_levels: int, grid: SpectralGrid, device: int = 0):
        self.L = load()
        self.s = Shortwave()
        self.num_levels = num_levels
        self.n = int(grid.n)
        dev = C.c_int(device)
        self._live = False
        _check(self.L.create_shortwave(C.byref(self.s), num_levels, C.byref(grid), C.byref(dev)),
               "create_shortwave")
        self._live = True

    def fluxes(self, optics: OpticsObject, mu_dir, mu_dif, alb_dir, alb_dif, tsi, solar,
               out=None):
        """calculate_sw_fluxes; out = (up, down) preallocated float64[levels, n] arrays."""
        up, dn = out if out is not None else (np.empty((self.num_levels, self.n), dtype=DTYPE),
                                              np.empty((self.num_levels, self.n), dtype=DTYPE))
        _check(self.L.calculate_sw_fluxes(C.byref(self.s), C.byref(optics.o), float(mu_dir),
                                          float(mu_dif), _dp(_f64(alb_dir)), _dp(_f64(alb_dif)),
                                          float(tsi), _dp(_f64(solar)), _dp(up), _dp(dn)),
               "calculate_sw_fluxes")
        return up, dn

    def column(self, gas: OpticsObject, p_mb, mu_dir, mu_dif, alb_dir, alb_dif, tsi, solar,
               window=None):
        up = np.empty(self.num_levels, dtype=DTYPE)
        dn = np.empty(self.num_levels, dtype=DTYPE)
        if alb_dir is None:
            p1 = p2 = p3 = None
        else:
            a1 = _f64(alb_dir)
            a2 = a1 if alb_dif is alb_dir else _f64(alb_dif)
            p1, p2, p3 = _dp(a1), _dp(a2), _dp(_f64(solar))
        first, last = window if window is not None else (0, self.n - 1)
        _check(self.L.calculate_sw_column_window(
            C.byref(self.s), C.byref(gas.o), _dp(_f64(p_mb)), float(mu_dir), float(mu_dif), p1, p2,
            float(tsi), p3, first, last, _dp(up), _dp(dn)), "calculate_sw_column")
        return up, dn

    def close(self):
        if self._live:
            _check(self.L.destroy_shortwave(C.byref(self.s)), "destroy_shortwave")
            self._live = False

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def solar_flux(grid: SpectralGrid, path: str) -> np.ndarray:
    L = load()
    s = SolarFlux()
    _check(L.create_solar_flux(C.byref(s), C.byref(grid), path.encode()), "create_solar_flux")
    out = np.ctypeslib.as_array(s.incident_flux, shape=(int(s.n),)).copy()
    _check(L.destroy_solar_flux(C.byref(s)), "destroy_solar_flux")
    return out


def table_on_grid_host(path: str, cols: int, grid: SpectralGrid) -> np.ndarray:
    """The HOST resampling of a CSV table to the grid (host/parse_csv.c: grt_table_on_grid,
    the reference's interpolate_to_grid rule) -- what the device resampler must reproduce."""
    L = load()
    p = _dbl_p()
    _check(L.grt_table_on_grid(path.encode(), cols, grid.w0, grid.dw, int(grid.n), C.byref(p)),
           "grt_table_on_grid")
    out = np.ctypeslib.as_array(p, shape=(int(grid.n),)).astype(np.float64)
    libc = C.CDLL(None)
    libc.free.argtypes = [C.c_void_p]
    libc.free(C.cast(p, C.c_void_p))
    return out


def integrate_fluxes(device: int, num_levels: int, grid: SpectralGrid, flux_dev_ptr: int):
    out = np.empty(num_levels, dtype=DTYPE)
    _check(load().integrate_fluxes(device, num_levels, C.byref(grid), C.c_void_p(flux_dev_ptr),
                                   _dp(out)), "integrate_fluxes")
    return out


def heating_rates(p_mb, up, down) -> np.ndarray:
    p = _f64(p_mb)
    rate = np.empty(len(p) - 1, dtype=DTYPE)
    _check(load().heating_rates(len(p), _dp(p), _dp(_f64(up)), _dp(_f64(down)), _dp(rate)),
           "heating_rates")
    return rate
