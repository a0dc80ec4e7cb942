"""Python mirror of the headless drivers' file formats (include/grtcode_colset.h):
column sets in ("GRTCOLS1"), broadband flux profiles out ("GRTFLUX1").

The arrays carry the semantics of the reference's Atmosphere_t
(framework/src/driver.h:114-166) as its netCDF readers fill it
(rfmip-irf/src/rfmip-irf.c:162-468): pressure in mb, abundances in ppmv on levels, the
cosine of the solar zenith angle, grey surface albedo / emissivity per column.
"""
from __future__ import annotations

import struct
from dataclasses import dataclass

import numpy as np

from . import synth

PATH_LEN = 1024
COLS_MAGIC = b"GRTCOLS1"
FLUX_MAGIC = b"GRTFLUX1"


def _path(p: str | None) -> bytes:
    b = (p or "none").encode()
    if len(b) >= PATH_LEN:
        raise ValueError(f"path longer than {PATH_LEN - 1} bytes: {p}")
    return b.ljust(PATH_LEN, b"\0")


def write_column_set(path: str, columns, dataset: synth.Dataset, molecules, use_ctm: bool = True):
    """columns: [times][columns] of synth.Column (all with the same number of levels).
    The CFC / CIA species and their cross-section files come from `dataset`."""
    T, C = len(columns), len(columns[0])
    levels = len(columns[0][0].p)
    cfcs = list(dataset.cfc.keys())
    cias = list(dataset.cia.keys())
    species = [synth.CIA_N2, synth.CIA_O2] if cias else []
    with open(path, "wb") as f:
        f.write(COLS_MAGIC)
        f.write(struct.pack("<7i", T, C, levels, len(molecules), len(cfcs), len(species), len(cias)))
        f.write(np.asarray(molecules, dtype="<i4").tobytes())
        f.write(np.asarray(cfcs, dtype="<i4").tobytes())
        f.write(np.asarray(species, dtype="<i4").tobytes())
        f.write(np.asarray(cias, dtype="<i4").reshape(-1).tobytes())
        f.write(_path(dataset.h2o_ctm_dir if use_ctm else None))
        f.write(_path(dataset.o3_ctm if use_ctm else None))
        for c in cfcs:
            f.write(_path(dataset.cfc[c]))
        for c in cias:
            f.write(_path(dataset.cia[c]))

        def put(get, shape):
            a = np.empty((T, C) + shape, dtype="<f8")
            for t in range(T):
                for c in range(C):
                    a[t, c] = get(columns[t][c])
            f.write(a.tobytes())

        put(lambda col: col.p, (levels,))
        put(lambda col: col.t, (levels,))
        put(lambda col: col.t_layer, (levels - 1,))
        put(lambda col: col.t_surf, ())
        put(lambda col: col.mu0, ())
        put(lambda col: col.tsi, ())
        put(lambda col: col.albedo, ())
        put(lambda col: col.emissivity, ())
        for m in molecules:
            put(lambda col, m=m: col.ppmv[m], (levels,))
        for c in cfcs:
            put(lambda col, c=c: col.cfc_ppmv[c], (levels,))
        for s in species:
            put(lambda col, s=s: col.cia_ppmv[s], (levels,))


@dataclass
class FluxSet:
    rlu: np.ndarray      # (times, columns, levels) W m-2: the reference's rlucsaf on every level
    rld: np.ndarray
    rsu: np.ndarray
    rsd: np.ndarray
    hr_lw: np.ndarray    # (times, columns, layers) K day-1
    hr_sw: np.ndarray


def read_flux_set(path: str) -> FluxSet:
    with open(path, "rb") as f:
        if f.read(8) != FLUX_MAGIC:
            raise ValueError(f"{path} is not a GRTFLUX1 file")
        T, C, levels, _ = struct.unpack("<4i", f.read(16))
        out = []
        for k in range(6):
            n = levels if k < 4 else levels - 1
            a = np.frombuffer(f.read(8*T*C*n), dtype="<f8").reshape(T, C, n).copy()
            out.append(a)
    return FluxSet(*out)


def rfmip_like_columns(num_sites: int, num_experiments: int, num_levels: int = 61,
                       molecules=(1, 2, 3, 4, 5, 6, 7), cfcs=()):
    """BASELINE config 3 shape: `num_experiments` perturbation experiments (the "time"
    axis of the reference's RFMIP reader, one experiment per run) x `num_sites` columns."""
    out = []
    for e in range(num_experiments):
        scale = (1., 4., 0.5, 2., 3., 8.)[e % 6]
        row = []
        for s in range(num_sites):
            rng = np.random.default_rng(7000 + s)
            row.append(synth.make_column(num_levels, rng, molecules=tuple(molecules), cfcs=tuple(cfcs),
                                         co2_scale=scale, dt=4.*(e % 3 - 1)))
        out.append(row)
    return out
