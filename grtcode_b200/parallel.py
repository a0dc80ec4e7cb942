"""Multi-GPU decomposition of the hot path: one process per GPU.

Two independent axes, both static:

* **Columns** (BASELINE configs 3 and 5).  Every (column, band) is independent; line
  lists and tables are replicated per GPU.  ``shard_columns`` is the block partition
  ``cols[r*C/G : (r+1)*C/G)``; there is NO data-path collective -- results are gathered
  once at the end (``gather_columns``), a few hundred bytes per column.

* **Spectral range** (config 4).  The global grid index range is cut into G contiguous
  chunks that share their end points (``split_spectral_range``), so that the trapezoid
  sums of the chunks add up to the global trapezoid sum (framework/src/driver.c:302-319)
  interval by interval.  Rank r builds its ``GasOptics_t`` on the chunk plus a halo of
  ceil(25/dw) points on either side (lines centred outside a grid do not contribute,
  kernels.c:431-433) and integrates only the points it owns
  (``calculate_{lw,sw}_column_window``).  The one exchange step is a SUM all-reduce of
  the per-rank partial broadband integrals ``[levels x {up, down}]``
  (``reduce_broadband``): ~1 KB per column, latency only.

The process group is torch.distributed: ``nccl`` with CUDA tensors on the GPU box,
``gloo`` with CPU tensors in the CPU tests (tests/test_distributed.py, world size 2).
"""
from __future__ import annotations

import math
import time
from dataclasses import dataclass

import numpy as np

LINE_CUTOFF = 25.     # cm-1, the reference's default wcutoff (gas_optics.c:88)


def shard_columns(num_columns: int, world: int, rank: int) -> range:
    """Block partition of SURVEY.md 8e.1: cols[r*C/G : (r+1)*C/G)."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world of {world}")
    lo = (rank*num_columns)//world
    hi = ((rank + 1)*num_columns)//world
    return range(lo, hi)


def grid_size(w0: float, wn: float, dw: float) -> int:
    """n = ceil((wn - w0)/dw) + 1 (utilities/src/spectral_grid.c:61)."""
    return int(math.ceil((wn - w0)/dw) + 1.)


@dataclass(frozen=True)
class SpectralChunk:
    """Rank `rank` OWNS global grid points [first, last] (its share of the trapezoid
    intervals; `last` is the next chunk's `first`) and COMPUTES on [lo, hi] = the owned
    points plus a halo of ceil(25/dw) + 1 + ceil(max|delta| p/dw) points on either side, clipped
    to the global grid:
    the reference only lets lines centred inside the grid contribute (kernels.c:431-433),
    so lines centred within the cut-off of an owned point must lie on the rank's grid."""
    rank: int
    first: int
    last: int
    lo: int
    hi: int
    w_lo: float         # wavenumber of `lo` on the global grid, w0 + lo*dw
    w_hi: float
    dw: float

    @property
    def n(self) -> int:
        """Grid points the rank computes on."""
        return self.hi - self.lo + 1

    @property
    def window(self) -> tuple[int, int]:
        """Owned points as indices of the rank's own grid (the integration window)."""
        return self.first - self.lo, self.last - self.lo

    @property
    def wn_arg(self) -> float:
        """Upper bound to pass to create_spectral_grid so that n = ceil((wn - w0)/dw) + 1
        (spectral_grid.c:61) comes out as hi - lo + 1 whatever the rounding of w_hi."""
        return self.w_hi - 0.25*self.dw


def split_spectral_range(w0: float, wn: float, dw: float, world: int,
                         cost: np.ndarray | None = None, shift: float = 0.) -> list[SpectralChunk]:
    """Cut the grid [w0, wn] @ dw into `world` chunks.  Neighbouring chunks share one grid
    point; every chunk keeps the global spacing and starts on a global grid point,
    w(i) = w0 + i*dw (spectral_grid.c:94).  Without `cost` the chunks hold (almost) equal
    numbers of trapezoid intervals; with `cost` (float[n-1], the estimated work of each
    interval, e.g. ``interval_cost``) the cuts equalise the summed cost instead -- lines
    cluster in bands, so equal point counts are not equal work."""
    n = grid_size(w0, wn, dw)
    intervals = n - 1
    if world < 1 or intervals < world:
        raise ValueError(f"cannot split {intervals} intervals over {world} ranks")
    # F of kernels.c:417, plus one point: lines in the half interval beyond a grid end are
    # dropped by the parser (parse_HITRAN_file.c:349), so their reach must end short of
    # the owned points
    # `shift` [cm-1] = max |delta|*p over the line list and column: a rank's parser keeps lines
    # by UNSHIFTED centre while the kernel windows a line around its pressure-SHIFTED centre
    # (kernels.c:44,431), so the halo has to cover the cut-off plus the largest shift
    halo = int(math.ceil(float(np.float32(LINE_CUTOFF))/dw)) + 1 + int(math.ceil(abs(shift)/dw))
    if cost is None:
        cuts = [(r*intervals)//world for r in range(world + 1)]
    else:
        c = np.asarray(cost, dtype=np.float64)
        if c.shape != (intervals,) or not np.all(np.isfinite(c)) or np.any(c < 0.):
            raise ValueError("cost must be a finite, non-negative float[n-1]")
        cum = np.concatenate([[0.], np.cumsum(c)])
        if cum[-1] <= 0.:
            cuts = [(r*intervals)//world for r in range(world + 1)]
        else:
            cuts = [0]
            for r in range(1, world):
                k = int(np.searchsorted(cum, cum[-1]*r/world, side="left"))
                # every chunk keeps at least one interval
                k = min(max(k, cuts[-1] + 1), intervals - (world - r))
                cuts.append(k)
            cuts.append(intervals)
    chunks = []
    for r in range(world):
        a, b = cuts[r], cuts[r + 1]
        lo = max(a - halo, 0)
        hi = min(b + halo, n - 1)
        chunks.append(SpectralChunk(r, a, b, lo, hi, w0 + lo*dw, w0 + hi*dw, dw))
    return chunks


def max_pressure_shift(hitran_path: str, molecules=None) -> float:
    """Largest |delta| [cm-1 atm-1] of a HITRAN .par file (field 9 = columns 60-67,
    parse_HITRAN_file.c:79-100), optionally over the given molecule ids only."""
    raw = np.fromfile(hitran_path, dtype=np.uint8)
    if raw.size == 0:
        return 0.
    nl = int(np.argmax(raw == 10))
    rec = nl + 1
    if raw[nl] != 10 or nl < 160 or raw.size % rec:
        raise ValueError(f"{hitran_path}: not fixed-length HITRAN records")
    tab = raw.reshape(-1, rec)
    d = np.ascontiguousarray(tab[:, 59:67]).view("S8").ravel().astype(np.float64)
    if molecules is not None:
        mol = np.ascontiguousarray(tab[:, 0:2]).view("S2").ravel().astype(np.int64)
        d = d[np.isin(mol, np.asarray(list(molecules), dtype=np.int64))]
    return float(np.abs(d).max()) if d.size else 0.


def line_centres(hitran_path: str, molecules=None) -> np.ndarray:
    """Transition wavenumbers of a HITRAN .par file (160-character records, field 2 =
    columns 4-15, parse_HITRAN_file.c:79-100), optionally of the given molecule ids only."""
    raw = np.fromfile(hitran_path, dtype=np.uint8)
    if raw.size == 0:
        return np.zeros(0)
    nl = int(np.argmax(raw == 10))
    rec = nl + 1
    if raw[nl] != 10 or nl < 160 or raw.size % rec:
        raise ValueError(f"{hitran_path}: not fixed-length HITRAN records")
    tab = raw.reshape(-1, rec)
    v = np.ascontiguousarray(tab[:, 3:15]).view("S12").ravel().astype(np.float64)
    if molecules is not None:
        mol = np.ascontiguousarray(tab[:, 0:2]).view("S2").ravel().astype(np.int64)
        v = v[np.isin(mol, np.asarray(list(molecules), dtype=np.int64))]
    return v


def interval_cost(w0: float, wn: float, dw: float, centres: np.ndarray,
                  point_cost: float | None = None, doppler_slope: float = 7.3e-8) -> np.ndarray:
    """Estimated work of each trapezoid interval of the grid, in units of one far-wing line:
    a line spreads its cost evenly over the 2*ceil(25/dw)+1 points of its window
    (kernels.c:417,435-437); a line at wavenumber v costs 1 + doppler_slope*v/dw (its
    Doppler core, and with it the number of samples outside the cheap far-wing branch,
    widens with v: +17 % measured between the two halves of config 4 at 0.01 cm-1); and
    every grid point costs `point_cost` lines for the continua, the solvers and the
    integration (default 2*dw: about 4 ns per point against layers*window/1e12 s per line,
    measured on B200).  Used to balance ``split_spectral_range``; only the relative sizes
    matter."""
    n = grid_size(w0, wn, dw)
    F = int(math.ceil(float(np.float32(LINE_CUTOFF))/dw))
    v = np.asarray(centres, dtype=np.float64)
    fc = np.floor((v - w0)/dw + 0.5).astype(np.int64)
    keep = (fc >= 0) & (fc < n)
    weight = 1. + doppler_slope*np.abs(v[keep])/dw
    hist = np.bincount(fc[keep], weights=weight, minlength=n).astype(np.float64)
    cum = np.concatenate([[0.], np.cumsum(hist)])
    idx = np.arange(n)
    dens = (cum[np.minimum(idx + F, n - 1) + 1] - cum[np.maximum(idx - F, 0)])/(2*F + 1)
    dens += (2.*dw if point_cost is None else point_cost)
    return 0.5*(dens[:-1] + dens[1:])


def trapezoid(flux: np.ndarray, dw: float) -> np.ndarray:
    """sum_{i<n-1} 0.5*(f_i + f_{i+1})*dw along the last axis (driver.c:305-318)."""
    f = np.asarray(flux, dtype=np.float64)
    return (0.5*(f[..., :-1] + f[..., 1:])*dw).sum(axis=-1)


def _backend_device(dist):
    import torch
    if dist.get_backend() == "nccl":
        return torch.device("cuda", torch.cuda.current_device())
    return torch.device("cpu")


def reduce_broadband(partial_up, partial_down, group=None):
    """SUM all-reduce of per-rank partial broadband integrals [.., levels] (W m-2).
    Returns (up, down) as numpy arrays, identical on every rank."""
    import torch
    import torch.distributed as dist
    up = np.ascontiguousarray(partial_up, dtype=np.float64)
    dn = np.ascontiguousarray(partial_down, dtype=np.float64)
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return up.copy(), dn.copy()
    buf = torch.from_numpy(np.stack([up, dn])).to(_backend_device(dist))
    dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=group)
    out = buf.cpu().numpy()
    return out[0], out[1]


def gather_columns(local: np.ndarray, num_columns: int, group=None) -> np.ndarray | None:
    """Collect per-column results [local_columns, ...] of the block partition on rank 0
    (None elsewhere).  Not on the data path: called once after the last column."""
    import torch
    import torch.distributed as dist
    local = np.ascontiguousarray(local, dtype=np.float64)
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return local.copy()
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    tail = local.shape[1:]
    sizes = [len(shard_columns(num_columns, world, r)) for r in range(world)]
    dev = _backend_device(dist)
    pad = max(sizes)
    send = torch.zeros((pad,) + tail, dtype=torch.float64, device=dev)
    send[:local.shape[0]] = torch.from_numpy(local).to(dev)
    recv = [torch.empty_like(send) for _ in range(world)]
    dist.all_gather(recv, send, group=group)
    if rank != 0:
        return None
    return np.concatenate([recv[r][:sizes[r]].cpu().numpy() for r in range(world)], axis=0)


def measured_costs(cost: np.ndarray | None, chunks: list, seconds_per_rank) -> np.ndarray:
    """Per-interval costs rescaled so that every chunk's summed cost equals the compute time
    its rank MEASURED with the current cuts: inside a chunk the model's shape is kept, between
    chunks the measurement decides.  Cutting by the result (``split_spectral_range``) moves
    points from slow ranks to fast ones; the estimate includes each rank's halo and fixed
    per-launch work, which is what one more pass corrects."""
    n_int = chunks[-1].last
    c = np.ones(n_int) if cost is None else np.asarray(cost, dtype=np.float64).copy()
    for ch, secs in zip(chunks, seconds_per_rank):
        seg = slice(ch.first, ch.last)
        total = float(c[seg].sum())
        if total > 0.:
            c[seg] *= float(secs)/total
    return c


class SpectralSplitColumn:
    """One column whose spectral range is split over the ranks of the process group
    (config 4).  Every rank builds the engine for its own chunk (+ halo) of every band;
    ``run`` returns the all-reduced broadband fluxes
    {"lw"/"sw": (up[levels], down[levels])}.  The solar spectrum is normalised on the
    GLOBAL grid (solar_flux.c:73-84) and sliced.  balance=True places the cuts by estimated
    work (``interval_cost`` of the band's line list) instead of by point count."""

    def __init__(self, bands, column, device: int, molecules, world: int, rank: int,
                 balance: bool = True, cost_override: dict | None = None):
        """cost_override: {band name: float[n-1]} per-interval costs to cut by instead of the
        line-density model (see ``measured_costs``)."""
        from . import api
        from . import workloads as W
        from .column_engine import ColumnEngine
        self.chunks = {}
        self.all_chunks = {}
        self.costs = {}
        sub = []
        windows = {}
        solar = {}
        for band in bands:
            cost = None
            if balance and world > 1:
                # every rank derives the same cuts from the same line list
                cost = interval_cost(band.w0, band.wn, band.dw,
                                     line_centres(band.dataset.hitran, molecules))
            if cost_override is not None and band.name in cost_override:
                cost = np.asarray(cost_override[band.name], dtype=np.float64)
            # largest pressure shift of any line in this column [cm-1]: mb -> atm as the
            # library does (gas_optics.c:445)
            shift = max_pressure_shift(band.dataset.hitran, molecules) * \
                float(np.max(np.abs(column.p)))*0.000986923
            every = split_spectral_range(band.w0, band.wn, band.dw, world, cost, shift=shift)
            ch = every[rank]
            self.chunks[band.name] = ch
            self.all_chunks[band.name] = every
            self.costs[band.name] = cost
            sub.append(W.Band(band.name, ch.w_lo, ch.wn_arg, band.dw, band.lines, band.dataset))
            windows[band.name] = ch.window
            if not band.name.startswith("lw"):
                full = api.solar_flux(api.spectral_grid(band.w0, band.wn, band.dw),
                                      band.dataset.solar)
                solar[band.name] = full[ch.lo:ch.hi + 1].copy()
        self.engine = ColumnEngine(sub, column, device, molecules, windows=windows,
                                   solar_override=solar)
        for part in self.engine.parts:
            ch = self.chunks[part["band"].name]
            if part["n"] != ch.n:
                raise RuntimeError(f"chunk grid has {part['n']} points, expected {ch.n}")

    def run(self, col, group=None):
        t0 = time.perf_counter()
        out = self.engine.step(col, upload_surface=True)
        # this rank's own compute before the exchange (host clock; the calls are synchronous)
        self.compute_seconds = time.perf_counter() - t0
        names = sorted(out)
        up = np.stack([out[k][0] for k in names])
        dn = np.stack([out[k][1] for k in names])
        t1 = time.perf_counter()
        up, dn = reduce_broadband(up, dn, group)
        self.reduce_seconds = time.perf_counter() - t1
        return {k: (up[i], dn[i]) for i, k in enumerate(names)}

    def close(self):
        self.engine.close()
