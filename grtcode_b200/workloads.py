"""The BASELINE.json configurations as seeded synthetic workloads.

Config 2 (the one the headline metric is quoted on): CIRC case 1, clear sky,
longwave 1-3250 cm-1 and shortwave 1-50000 cm-1 at 0.01 cm-1, seven HITRAN
molecules + H2O/O3 continua + three CIA pairs + the three CFCs of the case.
The column itself is the real CIRC-1 profile (grtcode_b200/data/circ1_profile.json,
copied from testing_harness/src/circ1.h by tools/make_golden.py); line lists
and tables are synthetic (no HITRAN data in the mount).
"""
from __future__ import annotations

import json
import os
from dataclasses import dataclass

import numpy as np

from . import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CIRC1 = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data", "circ1_profile.json")

LW_RANGE = (1., 3250.)
SW_RANGE = (1., 50000.)
# Lines per molecule, HITRAN2016 order of magnitude; every entry
# stays below the reference's 600000-lines-per-molecule limit (gas_optics.c:46).
LW_LINES = dict(synth.DEFAULT_LW_LINES)
SW_LINES = {synth.H2O: 250_000, synth.CO2: 400_000, synth.O3: 300_000, synth.N2O: 50_000,
            synth.CH4: 400_000, synth.CO: 2_000, synth.O2: 10_000}
MOLECULES = (synth.H2O, synth.CO2, synth.O3, synth.N2O, synth.CO, synth.CH4, synth.O2)
CIAS = ((0, 0), (1, 0), (1, 1))
CFCS = ("CFC11", "CFC12", "CCl4")


def circ1_column() -> synth.Column:
    """CIRC case 1: levels TOA -> surface; layer abundances are carried to levels by
    linear interpolation in pressure, as utilities/test/test_curtis_godson.c:102-110 does."""
    c = json.load(open(CIRC1))
    pl = np.array(c["level_pressure"])
    play = np.array(c["layer_pressure"])
    nlev = len(pl)

    def to_levels(layer_vals):
        a = np.array(layer_vals)
        x = np.empty(nlev)
        x[0], x[-1] = a[0], a[-1]
        for i in range(1, nlev - 1):
            x[i] = a[i-1] + (a[i] - a[i-1])*(pl[i] - play[i-1])/(play[i] - play[i-1])
        return x*1.e6     # mole fraction -> ppmv

    names = {synth.H2O: "H2O", synth.CO2: "CO2", synth.O3: "O3", synth.N2O: "N2O",
             synth.CO: "CO", synth.CH4: "CH4", synth.O2: "O2"}
    ppmv = {m: to_levels(c[f"{n}_abundance"]) for m, n in names.items()}
    cfc = {synth.CFC_IDS["CFC11"]: to_levels(c["CFC11_abundance"]),
           synth.CFC_IDS["CFC12"]: to_levels(c["CFC12_abundance"]),
           synth.CFC_IDS["CCl4"]: to_levels(c["CCl4_abundance"])}
    o2 = ppmv[synth.O2]
    cia = {synth.CIA_N2: np.full(nlev, 781000.), synth.CIA_O2: o2}
    mu0 = float(np.cos(np.deg2rad(c["solar_zenith_angle"])))
    return synth.Column(p=pl, t=np.array(c["level_temperature"]),
                        t_layer=np.array(c["layer_temperature"]),
                        t_surf=float(c["surface_temperature"]), ppmv=ppmv, cfc_ppmv=cfc,
                        cia_ppmv=cia, emissivity=1.0, albedo=0.2, mu0=mu0,
                        tsi=float(c["toa_solar_irradiance"])/mu0)


def perturbed(col: synth.Column, rank: int, step: int) -> synth.Column:
    """A distinct but equally sized column per (rank, step): temperature offset and
    CO2 scaling in the spirit of the RFMIP experiments."""
    rng = np.random.default_rng(1000*rank + step)
    dt = float(rng.uniform(-4., 4.))
    scale = float(rng.choice([0.5, 1., 2., 4.]))
    ppmv = dict(col.ppmv)
    ppmv[synth.CO2] = col.ppmv[synth.CO2]*scale
    return synth.Column(p=col.p, t=col.t + dt, t_layer=col.t_layer + dt, t_surf=col.t_surf + dt,
                        ppmv=ppmv, cfc_ppmv=col.cfc_ppmv, cia_ppmv=col.cia_ppmv,
                        emissivity=col.emissivity, albedo=col.albedo, mu0=col.mu0, tsi=col.tsi)


@dataclass
class Band:
    name: str
    w0: float
    wn: float
    dw: float
    lines: dict
    dataset: synth.Dataset | None = None

    @property
    def n(self) -> int:
        return int(np.ceil((self.wn - self.w0)/self.dw) + 1.)

    @property
    def num_lines(self) -> int:
        return int(sum(self.lines.values()))


def circ1_bands(dw: float = 0.01, scale: float = 1.) -> tuple[Band, Band]:
    """The two spectral bands of config 2; scale < 1 shrinks the line counts (tests)."""
    lw = Band("lw", *LW_RANGE, dw, {m: max(1, int(n*scale)) for m, n in LW_LINES.items()})
    sw = Band("sw", *SW_RANGE, dw, {m: max(1, int(n*scale)) for m, n in SW_LINES.items()})
    return lw, sw


def materialise(band: Band, root: str, seed: int) -> Band:
    # the CFC tables cover 500-1500 cm-1 and are zero elsewhere (no-extrapolation rule,
    # utilities.c:173-246); both bands carry them (BASELINE config 4 names them for the
    # full 1-50000 cm-1 range)
    cfcs = CFCS
    band.dataset = synth.make_dataset(os.path.join(root, band.name), band.w0, band.wn, band.lines,
                                      seed=seed, cfcs=cfcs, cias=CIAS)
    return band


def cpu_sample_band(band: Band, width: float, centre: float) -> Band:
    """A window of `width` <= 25 cm-1 of the band for the bounded CPU-baseline sample.

    The reference only ever uses lines inside the grid (parse_HITRAN_file.c:349,
    kernels.c:433), and in a window narrower than the 25 cm-1 cut-off every line is
    sampled at every grid point.  Choosing N = N_band*50/span lines therefore gives the
    same number of Voigt samples PER GRID POINT (N) as the full band has on average
    (line density x 50 cm-1), which is what the pts x layers / s metric costs."""
    assert width <= 25.
    w0 = max(band.w0, centre - width/2.)
    wn = min(band.wn, w0 + width)
    frac = 50./(band.wn - band.w0)
    lines = {m: max(1, int(round(n*frac))) for m, n in band.lines.items()}
    return Band(band.name + "_sample", w0, wn, band.dw, lines)
