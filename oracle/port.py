"""TEST INFRASTRUCTURE (oracle) -- the CPU restatement of the hot path.

``oracle/libgrt_oracle.so`` is built from ``oracle/grt_oracle.c`` (plain C99);
file parsing and orchestration are restated here in numpy:

* HITRAN records   -> parse_hitran      (gas-optics/src/parse_HITRAN_file.c:224-413)
* CSV tables       -> read_table        (utilities/src/parse_csv.c:55-166)
* launch()         -> gas_optical_depth (gas-optics/src/launch.c:40-226, all three
                      optical_depth_methods: kernels.c:135-465, 514-581)
* column solve     -> column_fluxes     (framework/src/driver.c:360-424)

Parity status: pinned against the compiled reference (tests/test_oracle.py) and
the reference's known-answer vectors (tests/golden/reference_pins.json).
Only tests/, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline legs may
import this module.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "libgrt_oracle.so")

# amu masses as float literals and isotopologue counts (gas-optics/src/molecules.c:36-306)
MOLECULES = {1: (np.float32(18.010565), 9), 2: (np.float32(43.98983), 13),
             3: (np.float32(47.984745), 18), 4: (np.float32(44.001062), 5),
             5: (np.float32(27.994915), 9), 6: (np.float32(16.0313), 4),
             7: (np.float32(31.98983), 6)}
MBTOATM = np.float64(np.float32(0.000986923))   # gas_optics.c:445

_lib = None
_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)


def available() -> bool:
    return os.path.exists(LIB)


def lib():
    global _lib
    if _lib is None:
        if not available():
            raise FileNotFoundError(f"{LIB} missing: run `make -C oracle port`")
        L = C.CDLL(LIB, mode=C.RTLD_LOCAL)
        L.ora_tips_q.restype = C.c_double
        L.ora_tips_q.argtypes = [C.c_int, C.c_double, C.c_int]
        L.ora_voigt.argtypes = [C.c_double, C.c_uint64, C.c_double, C.c_double, C.c_double,
                                C.c_double, _dp]
        L.ora_layer_quantities.argtypes = [C.c_int, _dp, _dp, _dp, _dp, _dp]
        L.ora_partial_pressure.argtypes = [C.c_int, _dp, _dp, _dp, _dp, _dp]
        L.ora_line_params.argtypes = [C.c_int, C.c_double, C.c_int, C.c_uint64, C.c_int, _ip,
                                      _dp, _dp, _dp, _dp, _dp, _dp, _dp, _dp, _dp, _dp, _dp, _dp,
                                      _dp, _dp]
        L.ora_line_sample.argtypes = [C.c_uint64, C.c_int, _dp, _dp, _dp, _dp, _dp, C.c_double,
                                      C.c_double, C.c_uint64, _dp]
        _up = C.POINTER(C.c_uint64)
        L.ora_bins_shape.argtypes = [C.c_uint64, C.c_double, C.c_double, _up]
        L.ora_sort_lines.argtypes = [C.c_uint64, C.c_int, _dp, _dp, _dp, _dp]
        for fn in (L.ora_bin_sweep, L.ora_line_sweep):
            fn.argtypes = [C.c_uint64, C.c_int, _dp, _dp, _dp, _dp, _dp, C.c_uint64, C.c_double,
                           C.c_double, C.c_double, _dp, _dp]
        L.ora_bins_interpolate.argtypes = [C.c_int, C.c_uint64, C.c_double, C.c_double, C.c_double,
                                           _dp, _dp]
        L.ora_h2o_continuum.argtypes = [C.c_uint64, C.c_int, _dp, _dp, _dp, _dp, _dp, _dp, _dp,
                                        _dp, _dp]
        L.ora_o3_continuum.argtypes = [C.c_uint64, C.c_int, _dp, _dp, _dp]
        L.ora_cfc.argtypes = [C.c_uint64, C.c_int, _dp, _dp, _dp, _dp]
        L.ora_cia.argtypes = [C.c_uint64, C.c_int, _dp, _dp, _dp, _dp, _dp, _dp]
        L.ora_rayleigh.argtypes = [C.c_int, C.c_double, C.c_double, C.c_uint64, _dp, _dp, _dp, _dp]
        L.ora_add_optics.argtypes = [C.c_uint64, C.c_int, _dp, _dp, _dp, _dp, _dp, _dp]
        L.ora_lw_fluxes.argtypes = [C.c_int, C.c_double, C.c_double, C.c_uint64, C.c_double, _dp,
                                    _dp, _dp, _dp, _dp, _dp, _dp]
        L.ora_sw_fluxes.argtypes = [C.c_int, C.c_uint64, _dp, _dp, _dp, C.c_double, C.c_double,
                                    _dp, _dp, C.c_double, _dp, _dp, _dp]
        L.ora_broadband.argtypes = [C.c_int, C.c_uint64, C.c_double, _dp, _dp]
        L.ora_table_to_grid.argtypes = [_dp, _dp, C.c_uint64, C.c_double, C.c_double, C.c_uint64,
                                        _dp]
        L.ora_scale_strengths.argtypes = [C.c_int, C.c_uint64, _ip, _dp, _dp, _dp]
        _lib = L
    return _lib


def _d(a):
    a = np.ascontiguousarray(a, dtype=np.float64)
    return a, a.ctypes.data_as(_dp)


def grid_size(w0: float, wn: float, dw: float) -> int:
    """utilities/src/spectral_grid.c:61"""
    return int(np.ceil((wn - w0)/dw) + 1.)


def tips_q(mol: int, t: float, iso: int) -> float:
    return float(lib().ora_tips_q(mol, t, iso))


def voigt(w, n, wres, center, gamma, alpha) -> np.ndarray:
    k = np.zeros(n)
    lib().ora_voigt(w, n, wres, center, gamma, alpha, k.ctypes.data_as(_dp))
    return k


def parse_hitran(path: str, mol: int, w0: float, wn: float) -> dict:
    """Fixed-width HITRAN-2012 reader; yair/yself/en/n/d rounded through float32
    (parse_HITRAN_file.c:79-100, 197-213), strengths scaled (:372-384)."""
    raw = np.fromfile(path, dtype=np.uint8)
    reclen = 161
    assert raw.size % reclen == 0, "records must be 160 characters + newline"
    rec = raw.reshape(-1, reclen)
    molfield = np.char.strip(rec[:, 0:2].copy().view("S2").ravel().astype(str)).astype(int)
    rec = rec[molfield == mol]

    def col(a, b, dtype=np.float64):
        return rec[:, a:b].copy().view(f"S{b - a}").ravel().astype(str).astype(dtype)

    iso_c = rec[:, 2].copy().view("S1").ravel().astype(str)
    iso = np.array([10 if c == "0" else (ord(c) - ord("A") + 11 if c.isalpha() else int(c))
                    for c in iso_c], dtype=np.int32)
    v = col(3, 15)
    keep = (v >= w0) & (v <= wn)
    f32 = lambda x: x.astype(np.float32).astype(np.float64)
    out = dict(iso=iso[keep], v=v[keep], s=col(15, 25)[keep], yair=f32(col(35, 40))[keep],
               yself=f32(col(40, 45))[keep], en=f32(col(45, 55))[keep],
               n=f32(col(55, 59))[keep], d=f32(col(59, 67))[keep])
    out = {k: np.ascontiguousarray(a) for k, a in out.items()}
    s, sp = _d(out["s"])
    en, enp = _d(out["en"])
    vv, vp = _d(out["v"])
    iso_a = np.ascontiguousarray(out["iso"], dtype=np.int32)
    lib().ora_scale_strengths(mol, len(vv), iso_a.ctypes.data_as(_ip), enp, vp, sp)
    out["s"] = s
    out["iso"] = iso_a
    return out


def read_table(path: str, w0: float, dw: float, nw: int, cols: int = 2) -> np.ndarray:
    """CSV (one header line) -> column 1 resampled to the grid, zero outside."""
    data = np.loadtxt(path, delimiter=",", skiprows=1, ndmin=2)
    assert data.shape[1] == cols, f"{path}: expected {cols} columns"
    x, xp = _d(data[:, 0])
    y, yp = _d(data[:, 1])
    out = np.zeros(nw)
    lib().ora_table_to_grid(xp, yp, len(x), w0, dw, nw, out.ctypes.data_as(_dp))
    return out


def solar_flux(path: str, w0: float, dw: float, nw: int) -> np.ndarray:
    """shortwave/src/solar_flux.c:27-90: resample + sequential trapezoid normalisation."""
    c = read_table(path, w0, dw, nw)
    w = w0 + np.arange(nw)*dw
    total = 0.
    for a in 0.5*(c[:-1] + c[1:])*(w[1:] - w[:-1]):
        total += a
    return c/total


WAVENUMBER_SWEEP, LINE_SWEEP, LINE_SAMPLE = 0, 1, 2      # gas_optics.h:88-93
BIN_WIDTH = 1.                                           # gas_optics.c:73


class GasOptics:
    """launch() of gas-optics/src/launch.c; method as enum OpticalDepthMethod (the drivers
    use line_sample, the library default is wavenumber_sweep)."""

    def __init__(self, num_levels, w0, wn, dw, hitran, h2o_ctm_dir=None, o3_ctm=None,
                 method=LINE_SAMPLE):
        self.method = method
        self.levels = num_levels
        self.w0, self.wn, self.dw = w0, wn, dw
        self.nw = grid_size(w0, wn, dw)
        self.hitran = hitran
        self.mols = []          # (id, lines) in add order
        self.h2o = None
        self.o3 = None
        self.h2o_dir, self.o3_file = h2o_ctm_dir, o3_ctm
        self.cfcs = []          # (id, xs)
        self.cias = []          # (id0, id1, xs)
        self.x = {}
        self.x_cfc = {}
        self.x_cia = {}

    def add_molecule(self, mol):
        self.mols.append((mol, parse_hitran(self.hitran, mol, self.w0, self.wn)))
        if mol == 1 and self.h2o_dir:
            t = lambda name, c: read_table(os.path.join(self.h2o_dir, name), self.w0, self.dw,
                                           self.nw, c)
            self.h2o = dict(S296=t("296MTCKD25_S.csv", 2), F296=t("296MTCKD25_F.csv", 2),
                            CKDS=t("CKDS.csv", 4), CKDF=t("CKDF.csv", 4))
        if mol == 3 and self.o3_file:
            self.o3 = read_table(self.o3_file, self.w0, self.dw, self.nw)

    def add_cfc(self, cfc, path):
        self.cfcs.append((cfc, read_table(path, self.w0, self.dw, self.nw)))

    def add_cia(self, a, b, path):
        self.cias.append((a, b, read_table(path, self.w0, self.dw, self.nw)))

    def set_ppmv(self, kind, ident, ppmv):
        {"mol": self.x, "cfc": self.x_cfc, "cia": self.x_cia}[kind][ident] = \
            np.asarray(ppmv, dtype=np.float64)*1.e-6

    def optical_depth(self, p_mb, t) -> np.ndarray:
        L = lib()
        nlay = self.levels - 1
        p, pp = _d(np.asarray(p_mb, dtype=np.float64)*MBTOATM)
        t, tp = _d(t)
        n, pavg, tavg = np.zeros(nlay), np.zeros(nlay), np.zeros(nlay)
        L.ora_layer_quantities(nlay, pp, tp, n.ctypes.data_as(_dp), pavg.ctypes.data_as(_dp),
                               tavg.ctypes.data_as(_dp))
        tau = np.zeros((nlay, self.nw))
        taup = tau.ctypes.data_as(_dp)
        shape = (C.c_uint64*5)()
        L.ora_bins_shape(self.nw, self.dw, BIN_WIDTH, shape)
        bins_tau = np.zeros((nlay, int(shape[2]), 3))       # launch.c:61-64: both zeroed
        for mol, ln in self.mols:
            x, xp = _d(self.x.get(mol, np.zeros(self.levels)))
            ps, ns = np.zeros(nlay), np.zeros(nlay)
            L.ora_partial_pressure(nlay, pp, xp, n.ctypes.data_as(_dp), ps.ctypes.data_as(_dp),
                                   ns.ctypes.data_as(_dp))
            nl = len(ln["v"])
            if nl > 0:
                amu, niso = MOLECULES[mol]
                mass = np.float64(amu)/6.023e23
                arr = [np.zeros((nlay, nl)) for _ in range(4)]
                L.ora_line_params(mol, mass, niso, nl, nlay, ln["iso"].ctypes.data_as(_ip),
                                  _d(ln["v"])[1], _d(ln["s"])[1], _d(ln["yair"])[1],
                                  _d(ln["yself"])[1], _d(ln["en"])[1], _d(ln["n"])[1],
                                  _d(ln["d"])[1], pavg.ctypes.data_as(_dp),
                                  tavg.ctypes.data_as(_dp), ps.ctypes.data_as(_dp),
                                  *[a.ctypes.data_as(_dp) for a in arr])
                ptrs = [a.ctypes.data_as(_dp) for a in arr]
                if self.method == LINE_SAMPLE:
                    L.ora_line_sample(nl, nlay, *ptrs, ns.ctypes.data_as(_dp), self.w0, self.dw,
                                      self.nw, taup)
                else:
                    if self.method == WAVENUMBER_SWEEP:
                        L.ora_sort_lines(nl, nlay, *ptrs)                   # launch.c:138
                    sweep = L.ora_bin_sweep if self.method == WAVENUMBER_SWEEP else L.ora_line_sweep
                    sweep(nl, nlay, *ptrs, ns.ctypes.data_as(_dp), self.nw, self.w0, self.dw,
                          BIN_WIDTH, taup, bins_tau.ctypes.data_as(_dp))
            if mol == 1 and self.h2o is not None:
                h = self.h2o
                L.ora_h2o_continuum(self.nw, nlay, taup, _d(h["S296"])[1],
                                    tavg.ctypes.data_as(_dp), ps.ctypes.data_as(_dp),
                                    ns.ctypes.data_as(_dp), _d(h["CKDS"])[1], _d(h["F296"])[1],
                                    pavg.ctypes.data_as(_dp), _d(h["CKDF"])[1])
            elif mol == 3 and self.o3 is not None:
                L.ora_o3_continuum(self.nw, nlay, _d(self.o3)[1], ns.ctypes.data_as(_dp), taup)
        for cfc, xs in self.cfcs:
            x, xp = _d(self.x_cfc.get(cfc, np.zeros(self.levels)))
            L.ora_cfc(self.nw, nlay, n.ctypes.data_as(_dp), xp, _d(xs)[1], taup)
        for a, b, xs in self.cias:
            x1, x1p = _d(self.x_cia.get(a, np.zeros(self.levels)))
            x2, x2p = _d(self.x_cia.get(b, np.zeros(self.levels)))
            L.ora_cia(self.nw, nlay, pp, tavg.ctypes.data_as(_dp), x1p, x2p, _d(xs)[1], taup)
        if self.method != LINE_SAMPLE:
            # launch.c:212-218: the remote sums are interpolated onto the grid last
            L.ora_bins_interpolate(nlay, self.nw, self.w0, self.dw, BIN_WIDTH,
                                   bins_tau.ctypes.data_as(_dp), taup)
        return tau


def rayleigh(p_mb, w0, dw, nw):
    nlay = len(p_mb) - 1
    p, pp = _d(np.asarray(p_mb, dtype=np.float64)*MBTOATM)
    t, tp = _d(np.zeros(len(p_mb)))
    n, a, b = np.zeros(nlay), np.zeros(nlay), np.zeros(nlay)
    lib().ora_layer_quantities(nlay, pp, tp, n.ctypes.data_as(_dp), a.ctypes.data_as(_dp),
                               b.ctypes.data_as(_dp))
    omega, g, tau = (np.zeros((nlay, nw)) for _ in range(3))
    lib().ora_rayleigh(nlay, w0, dw, nw, n.ctypes.data_as(_dp), omega.ctypes.data_as(_dp),
                       g.ctypes.data_as(_dp), tau.ctypes.data_as(_dp))
    return omega, g, tau


def add_optics(parts):
    """parts: list of (g, omega, tau) arrays of equal shape."""
    shape = parts[0][2].shape
    n = int(np.prod(shape))
    g_in, gp = _d(np.stack([p[0] for p in parts]))
    o_in, op = _d(np.stack([p[1] for p in parts]))
    t_in, tp = _d(np.stack([p[2] for p in parts]))
    g, o, t = (np.zeros(shape) for _ in range(3))
    lib().ora_add_optics(n, len(parts), gp, op, tp, g.ctypes.data_as(_dp),
                         o.ctypes.data_as(_dp), t.ctypes.data_as(_dp))
    return g, o, t


def lw_fluxes(w0, dw, tau, omega, t_surf, t_layers, t_levels, emis):
    nlay, nw = tau.shape
    up, dn = np.zeros((nlay + 1, nw)), np.zeros((nlay + 1, nw))
    lib().ora_lw_fluxes(nlay + 1, w0, dw, nw, float(t_surf), _d(t_layers)[1], _d(t_levels)[1],
                        _d(tau)[1], _d(omega)[1], _d(emis)[1], up.ctypes.data_as(_dp),
                        dn.ctypes.data_as(_dp))
    return up, dn


def sw_fluxes(omega, g, tau, mu_dir, mu_dif, alb_dir, alb_dif, tsi, solar):
    nlay, nw = tau.shape
    up, dn = np.zeros((nlay + 1, nw)), np.zeros((nlay + 1, nw))
    lib().ora_sw_fluxes(nlay + 1, nw, _d(omega)[1], _d(g)[1], _d(tau)[1], float(mu_dir),
                        float(mu_dif), _d(alb_dir)[1], _d(alb_dif)[1], float(tsi), _d(solar)[1],
                        up.ctypes.data_as(_dp), dn.ctypes.data_as(_dp))
    return up, dn


def broadband(flux, dw):
    nlev, nw = flux.shape
    out = np.zeros(nlev)
    lib().ora_broadband(nlev, nw, dw, _d(flux)[1], out.ctypes.data_as(_dp))
    return out
