"""One GPU's worth of the hot path behind the reference-facing C API.

A ``ColumnEngine`` owns, for every spectral band it is given, a ``GasOptics_t``, an
``Optics_t`` for the gas optical depths and the band's solver (``Longwave_t`` for a band
whose name starts with "lw", ``Shortwave_t`` otherwise) and runs columns through them in
the order the reference's ``column_calculation`` does (framework/src/driver.c:360-600):
set abundances -> calculate_optical_depth -> Rayleigh + add_optics + fluxes + broadband
integration (the last three fused in ``calculate_{lw,sw}_column``).

Everything goes through grtcode_b200.api -> libgrtcode.so -> libgrt_sm100.so; there is
no CPU path.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import api as _api


def pinned(n: int, fill: float) -> np.ndarray:
    """fp_t[n] (float64, or float32 in the single-precision build) in page-locked host memory
    (cudaMallocHost through the C ABI)."""
    ptr = C.c_void_p()
    if _api.cuda_layer().grt_host_alloc(C.byref(ptr), C.sizeof(_api.REAL)*n) != 0:
        raise RuntimeError("grt_host_alloc failed")
    arr = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(_api.REAL)), shape=(n,))
    arr[:] = fill
    return arr


class ColumnEngine:
    """bands: objects with name, w0, wn, dw and a materialised ``dataset`` (synth.Dataset).
    line_window: optional {band name: (min, max)} passed to add_molecule, the reference's
    own way of restricting the lines a GasOptics_t loads (gas_optics.h:117-119).
    windows: optional {band name: (first, last)} integration windows (spectral split).
    solar_override: optional {band name: float64[n]} incident solar spectrum to use
    instead of create_solar_flux on the band's own grid.
    method: optical_depth_method of create_gas_optics (gas_optics.h:45-50); the reference's
    drivers pass line_sample (driver.c:618-624), None is the library default."""

    def __init__(self, bands, column, device: int, molecules, line_window=None, windows=None,
                 solar_override=None, method: int | None = _api.LINE_SAMPLE):
        api = _api
        api.load()
        self.device = device
        self.real_bytes = C.sizeof(api.REAL)
        self.levels = len(column.p)
        self.parts = []
        for band in bands:
            ds = band.dataset
            grid = api.spectral_grid(band.w0, band.wn, band.dw)
            go = api.GasOptics(self.levels, grid, ds.hitran, ds.h2o_ctm_dir, ds.o3_ctm,
                               method=method, device=device)
            win = (line_window or {}).get(band.name, (None, None))
            for m in molecules:
                go.add_molecule(m, *win)
            for c, path in ds.cfc.items():
                go.add_cfc(c, path)
            for (a, b), path in ds.cia.items():
                go.add_cia(a, b, path)
            gas = api.OpticsObject(self.levels - 1, grid, device)
            n = int(grid.n)
            part = dict(band=band, grid=grid, go=go, gas=gas, n=n,
                        window=(windows or {}).get(band.name))
            if band.name.startswith("lw"):
                part["solver"] = api.LongwaveSolver(self.levels, grid, device)
                part["emis"] = pinned(n, column.emissivity)
            else:
                part["solver"] = api.ShortwaveSolver(self.levels, grid, device)
                part["alb"] = pinned(n, column.albedo)
                sol = (solar_override or {}).get(band.name)
                if sol is None:
                    sol = api.solar_flux(grid, ds.solar)
                part["solar"] = pinned(n, 0.)
                part["solar"][:] = sol
            self.parts.append(part)
        self.launches = 0
        self.h2d = 0
        self.d2h = 0
        self.last = {}

    def _set_abundances(self, part, col):
        go = part["go"]
        for m, v in col.ppmv.items():
            go.set_ppmv("mol", m, v)
        for c, v in col.cfc_ppmv.items():
            if c in part["band"].dataset.cfc:
                go.set_ppmv("cfc", c, v)
        for c, v in col.cia_ppmv.items():
            go.set_ppmv("cia", c, v)

    def step_reference_api(self, col):
        """The same column through the reference's OWN entry points only
        (column_calculation, framework/src/driver.c:360-424): calculate_optical_depth,
        rayleigh_scattering, add_optics (which allocates its result, optics.c:84-148),
        calculate_lw_fluxes / calculate_sw_fluxes returning the spectral fluxes
        [levels x n] to host arrays (longwave.h:56-68, shortwave.h:56-69).  The host arrays
        are page-locked and allocated once.  Returns {"lw"/"sw": (up, down)} spectral."""
        api = _api
        out = {}
        h2d = d2h = 0
        for part in self.parts:
            self._set_abundances(part, col)
            part["go"].optical_depth(col.p, col.t, part["gas"])
            h2d += self.real_bytes*self.levels*(2 + 53 + 21 + 2)
            n = part["n"]
            if "ray" not in part:
                part["ray"] = api.OpticsObject(self.levels - 1, part["grid"], self.device)
                part["flux_up"] = pinned(self.levels*n, 0.).reshape(self.levels, n)
                part["flux_dn"] = pinned(self.levels*n, 0.).reshape(self.levels, n)
            api.rayleigh_scattering(part["ray"], col.p)
            tot = api.add_optics([part["gas"], part["ray"]])
            dst = (part["flux_up"], part["flux_dn"])
            if part["band"].name.startswith("lw"):
                part["emis"][:] = col.emissivity
                part["solver"].fluxes(tot, col.t_surf, col.t_layer, col.t, part["emis"], out=dst)
                h2d += self.real_bytes*n
            else:
                part["alb"][:] = col.albedo
                part["solver"].fluxes(tot, col.mu0, 0.5, part["alb"], part["alb"], col.tsi,
                                      part["solar"], out=dst)
                h2d += 3*self.real_bytes*n
            d2h += 2*self.real_bytes*self.levels*n
            tot.close()
            out[part["band"].name[:2]] = dst
        self.h2d_ref, self.d2h_ref = h2d, d2h
        return out

    def step(self, col, upload_surface: bool, only=None):
        """One column.  upload_surface=True is the end-to-end form: every host array the
        API takes is passed (and copied to the GPU) again.  only: band names to run (default
        all).  Returns {"lw"/"sw": (flux_up[levels], flux_down[levels], kernel stats)} in
        W m-2."""
        out = {}
        launches = 0
        h2d = 0
        for part in self.parts:
            if only is not None and part["band"].name[:2] not in only:
                continue
            go = part["go"]
            self._set_abundances(part, col)
            go.optical_depth(col.p, col.t, part["gas"])
            h2d += self.real_bytes*self.levels*(2 + 53 + 21 + 2)
            name = part["band"].name[:2]
            if name == "lw":
                send = upload_surface
                if part["emis"][0] != col.emissivity:      # grey surface of this column
                    part["emis"][:] = col.emissivity
                    send = True
                up, dn = part["solver"].column(part["gas"], col.p, col.t_surf, col.t_layer,
                                               col.t, part["emis"] if send else None,
                                               window=part["window"])
                h2d += self.real_bytes*part["n"] if send else 0
            else:
                send = upload_surface
                if part["alb"][0] != col.albedo:
                    part["alb"][:] = col.albedo
                    send = True
                alb = part["alb"] if send else None
                up, dn = part["solver"].column(part["gas"], col.p, col.mu0, 0.5, alb, alb,
                                               col.tsi, part["solar"] if send else None,
                                               window=part["window"])
                h2d += 2*self.real_bytes*part["n"] if send else 0
            # the sweep kernels were only enqueued by calculate_optical_depth: their event
            # times are read after the solver call, which synchronised the stream
            st = go.stats()
            launches += st["launches"] + 1
            out[name] = (up, dn, st)
        self.launches = launches
        self.h2d = h2d
        self.d2h = 2*2*self.real_bytes*self.levels + 2*32
        self.last = out
        return out

    def close(self):
        for part in self.parts:
            part["solver"].close()
            part["gas"].close()
            if "ray" in part:
                part["ray"].close()
                for key in ("flux_up", "flux_dn"):
                    _api.cuda_layer().grt_host_free(C.c_void_p(part[key].ctypes.data))
            part["go"].close()
        self.parts = []
