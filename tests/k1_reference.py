"""TEST INFRASTRUCTURE: the per-(line, layer) quantities of the line preparation, from the
compiled reference's own kernels (oracle/_ref, unmodified sources):

    calc_pressures_and_temperatures, calc_partial_pressures_and_number_densities
        utilities/src/curtis_godson.c:59,92
    calc_line_centers, calc_lorentz_hw, calc_doppler_hw      gas-optics/src/kernels.c:34,93,114

and, from them, the two floats rfm_voigt_line_shape derives first (RFM_voigt.c:94-95):
REPWID = (float)(SQRLN2/alpha), Y = (float)(REPWID*gamma).  The CUDA path's K1 records must
carry exactly these bits (tests/test_gpu_k1.py); tools/make_near_ties.py uses the same
functions to construct columns whose Y sits on a float rounding boundary.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

MBTOATM = np.float64(np.float32(0.000986923))          # gas_optics.c:445
SQRLN2 = np.float64(np.float32(0.832554611))           # RFM_voigt.c:79
# molecules.c:40-70: float literal amu, then mass /= 6.023e23
MASS_AMU = {1: 18.010565, 2: 43.98983, 3: 47.984745, 4: 44.001062, 5: 27.994915, 6: 16.0313,
            7: 31.98983}

_dp = C.POINTER(C.c_double)


def _p(a):
    assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(_dp)


def mass_grams(mol: int) -> float:
    return float(np.float64(np.float32(MASS_AMU[mol]))/6.023e23)


def declare(ref):
    L = ref.lib
    L.calc_pressures_and_temperatures.argtypes = [C.c_int, _dp, _dp, _dp, _dp]
    L.calc_partial_pressures_and_number_densities.argtypes = [C.c_int, _dp, _dp, _dp, _dp, _dp]
    L.calc_number_densities.argtypes = [C.c_int, _dp, _dp]
    L.calc_line_centers.argtypes = [C.c_uint64, C.c_int, _dp, _dp, _dp, _dp]
    L.calc_lorentz_hw.argtypes = [C.c_uint64, C.c_int, _dp, _dp, _dp, _dp, _dp, _dp, _dp]
    L.calc_doppler_hw.argtypes = [C.c_uint64, C.c_int, C.c_double, _dp, _dp, _dp]
    return L


def layer_quantities(ref, p_mb, t, ppmv):
    """(pavg, tavg, ps) of one molecule on every layer, as calculate_optical_depth + launch()
    form them (gas_optics.c:445-451, launch.c:68-104)."""
    L = declare(ref)
    p = np.ascontiguousarray(np.asarray(p_mb, dtype=np.float64)*MBTOATM)
    t = np.ascontiguousarray(t, dtype=np.float64)
    x = np.ascontiguousarray(np.asarray(ppmv, dtype=np.float64)*1.e-6)   # gas_optics.c:311
    layers = len(p) - 1
    pavg, tavg, n = np.zeros(layers), np.zeros(layers), np.zeros(layers)
    ps, ns = np.zeros(layers), np.zeros(layers)
    ref.check(L.calc_number_densities(layers, _p(p), _p(n)))
    ref.check(L.calc_pressures_and_temperatures(layers, _p(p), _p(t), _p(pavg), _p(tavg)))
    ref.check(L.calc_partial_pressures_and_number_densities(layers, _p(p), _p(x), _p(n), _p(ps),
                                                            _p(ns)))
    return pavg, tavg, ps


def widths(ref, mol: int, v0, delta, nair, yair, yself, pavg, tavg, ps):
    """(vnn, gamma, alpha), each [layers, lines], from the reference's kernels."""
    L = declare(ref)
    c = lambda a: np.ascontiguousarray(a, dtype=np.float64)
    v0, delta, nair, yair, yself = c(v0), c(delta), c(nair), c(yair), c(yself)
    pavg, tavg, ps = c(pavg), c(tavg), c(ps)
    lines, layers = len(v0), len(pavg)
    vnn = np.zeros((layers, lines))
    gamma = np.zeros((layers, lines))
    alpha = np.zeros((layers, lines))
    ref.check(L.calc_line_centers(lines, layers, _p(v0), _p(delta), _p(pavg), _p(vnn)))
    ref.check(L.calc_lorentz_hw(lines, layers, _p(nair), _p(yair), _p(yself), _p(tavg), _p(pavg),
                                _p(ps), _p(gamma)))
    ref.check(L.calc_doppler_hw(lines, layers, mass_grams(mol), _p(vnn), _p(tavg), _p(alpha)))
    return vnn, gamma, alpha


def repwid_y(gamma, alpha):
    """REPWID and Y as float32 arrays, plus the double product REPWID*gamma (RFM_voigt.c:94-95:
    float const REPWID = SQRLN2/DOPADJ; float const Y = REPWID*WIDADJ)."""
    repwid = (SQRLN2/alpha).astype(np.float32)
    yd = repwid.astype(np.float64)*gamma
    return repwid, yd.astype(np.float32), yd


def ulps_from_float_tie(yd):
    """Distance, in double ulps, of each value from the nearest float rounding boundary."""
    bits = np.ascontiguousarray(yd, dtype=np.float64).view(np.uint64)
    lo = (bits & np.uint64(0x1fffffff)).astype(np.int64)
    return np.abs(lo - 0x10000000)
