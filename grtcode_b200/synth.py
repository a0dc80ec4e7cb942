"""Seeded synthetic inputs in the reference's own file formats.

No HITRAN data, continuum tables or RFMIP/CIRC files exist in the reference
mount, so every workload is generated here:

* HITRAN-2012 160-character ``.par`` records -- the format read by
  gas-optics/src/parse_HITRAN_file.c:79-100 (field widths) and :224-413;
* CSV tables with one header line -- utilities/src/parse_csv.c:55-166 -- for the
  H2O continuum (water_vapor_continuum.c:49-64), O3 continuum, CFC and CIA
  cross sections (cfcs.c:113, collision_induced_absorption.c:58) and the solar
  spectrum (shortwave/src/solar_flux.c:37-47);
* atmospheric profiles of the shapes BASELINE.json names.
"""
from __future__ import annotations

import os
from dataclasses import dataclass, field

import numpy as np

# HITRAN ids (gas-optics/src/molecules.h:33-87)
H2O, CO2, O3, N2O, CO, CH4, O2 = 1, 2, 3, 4, 5, 6, 7
MOL_NAMES = {1: "H2O", 2: "CO2", 3: "O3", 4: "N2O", 5: "CO", 6: "CH4", 7: "O2"}
# CIA ids (collision_induced_absorption.h:37-42)
CIA_N2, CIA_O2 = 0, 1

# Default longwave line counts per molecule for 1-3250 cm-1.
DEFAULT_LW_LINES = {H2O: 100_000, CO2: 300_000, O3: 250_000, N2O: 30_000,
                    CH4: 200_000, CO: 1_000, O2: 2_000}


def _fmt_leading(v: float, width: int, prec: int) -> str:
    """Fortran-style fixed field that drops the leading zero when needed
    (HITRAN writes '.0700' in F5.4 and '-.012000' in F8.6)."""
    s = f"{v:.{prec}f}"
    if len(s) > width:
        if s.startswith("0."):
            s = s[1:]
        elif s.startswith("-0."):
            s = "-" + s[2:]
    return s.rjust(width)[:width]


def make_lines(mol: int, n: int, w0: float, wn: float, rng: np.random.Generator):
    """Random line parameters for one molecule, sorted by wavenumber."""
    nb = max(1, min(8, n//50))
    centers = rng.uniform(w0, wn, nb)
    widths = rng.uniform(20., 120., nb)
    which = rng.integers(0, nb, n)
    band = centers[which] + widths[which]*rng.standard_normal(n)
    uniform = rng.uniform(w0, wn, n)
    v = np.where(rng.random(n) < 0.3, uniform, band)
    bad = (v < w0 + 1e-3) | (v > wn - 1e-3)
    v[bad] = rng.uniform(w0 + 1e-3, wn - 1e-3, int(bad.sum()))
    v = np.sort(np.round(v, 6))
    return dict(
        mol=np.full(n, mol, dtype=np.int32),
        iso=rng.choice([1, 2, 3], size=n, p=[0.9, 0.07, 0.03]).astype(np.int32),
        v=v,
        s=10.**rng.uniform(-30., -19., n),
        yair=rng.uniform(0.03, 0.11, n),
        yself=rng.uniform(0.1, 0.5, n),
        elo=rng.uniform(0., 4500., n),
        n=rng.uniform(0.4, 0.8, n),
        d=rng.uniform(-0.02, 0.015, n),
    )


def write_hitran(path: str, line_sets: list[dict]) -> int:
    """Write line sets as one HITRAN-2012 file ordered by wavenumber."""
    cat = {k: np.concatenate([ls[k] for ls in line_sets]) for k in line_sets[0]}
    order = np.argsort(cat["v"], kind="stable")
    tail = " "*60 + " "*6 + " "*12 + " " + "    1.0" + "    1.0"
    assert len(tail) == 15*4 + 6 + 12 + 1 + 7 + 7
    mol, iso, v, s = cat["mol"], cat["iso"], cat["v"], cat["s"]
    yair, yself, elo, nn, d = cat["yair"], cat["yself"], cat["elo"], cat["n"], cat["d"]
    with open(path, "w") as f:
        chunk = []
        for i in order:
            rec = (f"{mol[i]:2d}{iso[i]%10:1d}{v[i]:12.6f}{s[i]:10.3E}{0.0:10.3E}"
                   f"{_fmt_leading(yair[i], 5, 4)}{_fmt_leading(yself[i], 5, 3)}"
                   f"{elo[i]:10.4f}{_fmt_leading(nn[i], 4, 2)}{_fmt_leading(d[i], 8, 6)}"
                   f"{tail}")
            chunk.append(rec)
            if len(chunk) == 65536:
                f.write("\n".join(chunk) + "\n")
                chunk = []
        if chunk:
            f.write("\n".join(chunk) + "\n")
    return len(order)


def write_csv(path: str, x: np.ndarray, cols: list[np.ndarray], header: str = "x,y"):
    with open(path, "w") as f:
        f.write(header + "\n")
        for i in range(len(x)):
            f.write(",".join([f"{x[i]:.4f}"] + [f"{c[i]:.8e}" for c in cols]) + "\n")


CFC_IDS = {"CFC11": 0, "CFC12": 1, "CFC113": 2, "HCFC22": 5, "CCl4": 15, "SF6": 20}


@dataclass
class Dataset:
    """Paths of one generated input set."""
    root: str
    hitran: str
    h2o_ctm_dir: str
    o3_ctm: str
    cfc: dict = field(default_factory=dict)      # cfc id -> csv path
    cia: dict = field(default_factory=dict)      # (id0, id1) -> csv path
    solar: str = ""
    num_lines: dict = field(default_factory=dict)


def make_dataset(root: str, w0: float, wn: float, lines: dict[int, int], seed: int = 1,
                 cfcs: tuple = (), cias: tuple = ()) -> Dataset:
    """Generate line list + all tables under ``root`` (idempotent per seed)."""
    os.makedirs(root, exist_ok=True)
    rng = np.random.default_rng(seed)
    hitran = os.path.join(root, "synthetic.par")
    sets = [make_lines(m, n, w0, wn, rng) for m, n in lines.items() if n > 0]
    ds = Dataset(root=root, hitran=hitran, h2o_ctm_dir=os.path.join(root, "h2o_ctm"),
                 o3_ctm=os.path.join(root, "o3_ctm.csv"),
                 num_lines={m: n for m, n in lines.items()})
    stamp = os.path.join(root, ".stamp")
    key = f"{w0} {wn} {sorted(lines.items())} {seed} {cfcs} {cias}"
    fresh = not (os.path.exists(stamp) and open(stamp).read() == key)
    if fresh:
        write_hitran(hitran, sets)
    # H2O continuum: 10 cm-1 spacing, smooth positive (MT-CKD-like magnitudes).
    os.makedirs(ds.h2o_ctm_dir, exist_ok=True)
    x = np.arange(0., wn + 20., 10.)
    selfc = 1e-20*np.exp(-x/400.)*(1. + 0.3*np.sin(x/150.)) + 1e-26
    forc = 2e-22*np.exp(-x/500.)*(1. + 0.3*np.cos(x/170.)) + 1e-28
    tself = 0.02*np.exp(-x/3000.) + 0.005
    tfor = 0.003*np.exp(-x/2500.) + 0.0005
    zero = np.zeros_like(x)
    if fresh:
        write_csv(os.path.join(ds.h2o_ctm_dir, "296MTCKD25_S.csv"), x, [selfc])
        write_csv(os.path.join(ds.h2o_ctm_dir, "296MTCKD25_F.csv"), x, [forc])
        write_csv(os.path.join(ds.h2o_ctm_dir, "CKDS.csv"), x, [tself, zero, zero], "x,a,b,c")
        write_csv(os.path.join(ds.h2o_ctm_dir, "CKDF.csv"), x, [tfor, zero, zero], "x,a,b,c")
        xo = np.arange(max(w0, 400.), min(wn, 45000.) + 5., 5.)
        if len(xo) < 2:      # window outside the table's range: two points, as above
            xo = np.array([w0, wn + 5.])
        o3 = 1e-21*np.exp(-((xo - 1040.)/60.)**2) + 4e-21*np.exp(-((xo - 16600.)/2500.)**2) \
            + 1e-18*np.exp(-((xo - 39000.)/3000.)**2)
        write_csv(ds.o3_ctm, xo, [o3])
    for k, name in enumerate(cfcs):
        p = os.path.join(root, f"{name}.csv")
        ds.cfc[CFC_IDS[name]] = p
        if fresh:
            xc = np.arange(500., 1500.5, 0.5)
            c0 = 700. + 90.*k
            xs = 3e-18*np.exp(-((xc - c0)/12.)**2) + 1e-18*np.exp(-((xc - c0 - 250.)/20.)**2)
            write_csv(p, xc, [xs])
    for pair in cias:
        p = os.path.join(root, f"cia_{pair[0]}_{pair[1]}.csv")
        ds.cia[pair] = p
        if fresh:
            xc = np.arange(1., min(wn, 12000.) + 1., 2.)
            c0 = {(0, 0): 100., (1, 0): 1600., (1, 1): 1550.}.get(pair, 1000.)
            xs = 1e-46*np.exp(-((xc - c0)/150.)**2) + 2e-47*np.exp(-((xc - 2400.)/200.)**2)
            write_csv(p, xc, [xs])
    ds.solar = os.path.join(root, "solar.csv")
    if fresh:
        xs_ = np.arange(max(1., w0 - 1.), wn + 20., 10.)
        t_sun = 6000.
        planck = xs_**3/(np.exp(np.minimum(1.4387773538277202*xs_/t_sun, 600.)) - 1.)
        write_csv(ds.solar, xs_, [planck])
        with open(stamp, "w") as f:
            f.write(key)
    return ds


@dataclass
class Column:
    """One atmospheric column (levels ordered TOA -> surface, as in circ1.h)."""
    p: np.ndarray            # level pressure [mb]
    t: np.ndarray            # level temperature [K]
    t_layer: np.ndarray      # layer temperature [K]
    t_surf: float
    ppmv: dict               # molecule id -> [levels]
    cfc_ppmv: dict           # cfc id -> [levels]
    cia_ppmv: dict           # cia species id -> [levels]
    emissivity: float
    albedo: float
    mu0: float
    tsi: float = 1361.


def make_column(num_levels: int, rng: np.random.Generator, molecules=(H2O, CO2, O3),
                cfcs=(), co2_scale: float = 1., dt: float = 0.) -> Column:
    """Hydrostatic log-spaced pressure grid, tropical-like lapse rate +
    isothermal stratosphere +- N(0, 3 K); SURVEY.md 8(d)."""
    p = np.exp(np.linspace(np.log(0.01), np.log(1013.), num_levels))
    z = 7.3*np.log(1013./p)                      # km, scale-height estimate
    t = np.maximum(300. - 6.5*z, 210.) + np.where(z > 25., 1.8*(z - 25.), 0.)
    t = np.clip(t + rng.normal(0., 3., num_levels) + dt, 150., 340.)
    ppmv = {}
    for m in molecules:
        if m == H2O:
            ppmv[m] = np.maximum(2.5e4*np.exp(-z/2.), 4.)
        elif m == CO2:
            ppmv[m] = np.full(num_levels, 400.*co2_scale)
        elif m == O3:
            ppmv[m] = 0.03 + 8.*np.exp(-(np.log(p/10.)/1.2)**2)
        elif m == N2O:
            ppmv[m] = np.full(num_levels, 0.33)
        elif m == CO:
            ppmv[m] = np.full(num_levels, 0.1)
        elif m == CH4:
            ppmv[m] = np.full(num_levels, 1.8)
        elif m == O2:
            ppmv[m] = np.full(num_levels, 209460.)
        else:
            ppmv[m] = np.full(num_levels, 1e-3)
    return Column(
        p=p, t=t, t_layer=0.5*(t[:-1] + t[1:]), t_surf=float(t[-1] + rng.normal(0., 1.)),
        ppmv=ppmv, cfc_ppmv={c: np.full(num_levels, 2.5e-4) for c in cfcs},
        cia_ppmv={CIA_N2: np.full(num_levels, 781000.), CIA_O2: np.full(num_levels, 209460.)},
        emissivity=0.98, albedo=float(rng.uniform(0.05, 0.4)),
        mu0=float(rng.uniform(0.1, 1.)))
