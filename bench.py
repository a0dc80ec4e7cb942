"""Benchmark of the line-by-line hot path on B200 (contract: see DESIGN.md section 6).

    python bench.py --gpus N --steps K --warmup W            # this repository
    python bench.py --impl reference --gpus N --steps K ...   # the reference's CPU path

One step = one clear-sky column of BASELINE config 2 (CIRC case 1, 54 layers,
LW 1-3250 + SW 1-50000 cm-1 at 0.01 cm-1, 7 HITRAN molecules + continua + CIA +
3 CFCs): optical depths on both grids, Rayleigh mixing, the LW and SW solvers and
broadband integration.  metric = spectral points x layers / s of line-by-line tau.
N > 1 (torchrun): every rank runs its own columns on its own GPU, no data-path
collective (columns are independent: "weak" scaling); NCCL is used only for the
barrier and the max-over-ranks of the device time.

Further legs in the same JSON line:
  e2e_reference_api  the step through the reference's own signatures (spectral fluxes to host)
  rfmip              the north-star target: RFMIP-IRF-shaped columns at 0.1 cm-1, 225 distinct
                     columns per rank (8 ranks = the 1800-column job), with the reference's CPU
                     path timed on full columns of the same shape at N = 1 (rfmip.cpu_baseline)
  spectral_split     N > 1: one config-4 column split over the ranks + NCCL all-reduce (strong
                     scaling, per-rank compute and all-reduce seconds)
  cpu_baseline       the reference on a bounded sample of the step (per-column work only;
                     setup reported separately on both sides)
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "spectral pts x layers / s (line-by-line tau, LW+SW column incl. solvers)"
UNIT = "pts*layers/s"
# Algorithmic flops per far-wing Voigt sample.  Double build: 6 FP64 + 3 FP32.
# Single build (GRT_PRECISION=single, the reference's --enable-single-precision): the float
# chain t2 - v, *REPWID, x*x + Y*Y, reciprocal, c*K + acc = 7 FP32.
SINGLE = os.environ.get("GRT_PRECISION", "double") == "single"
FLOP_PER_SAMPLE = 7.0 if SINGLE else 9.0
DTYPE = "f32" if SINGLE else "f64"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=4)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--scale", type=float, default=float(os.environ.get("GRT_BENCH_SCALE", "1")),
                    help="shrink the line counts (development only; 1 = the BASELINE workload)")
    ap.add_argument("--dw", type=float, default=float(os.environ.get("GRT_BENCH_DW", "0.01")))
    ap.add_argument("--data", default=os.environ.get("GRT_BENCH_DATA", "/tmp/grt_bench"))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--method", default="line_sample",
                    choices=["line_sample", "wavenumber_sweep", "line_sweep"],
                    help="optical_depth_method; the reference drivers (and the headline number) "
                         "use line_sample (driver.c:618-624)")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region."""

    QUERY = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.QUERY}",
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, smax, reasons = [], [], set()
        for line in self.lines:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                smax.append(float(f[1]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown",
                                  "sw_power_cap"), f[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None,
                "sm_max_mhz": max(smax) if smax else None, "reasons": sorted(reasons),
                "samples": len(sm)}


def host_threads() -> int:
    """Every host core this process may run on (torchrun exports OMP_NUM_THREADS=1 to its
    workers, which would otherwise time a one-thread reference)."""
    return len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else os.cpu_count()


class ReferenceBand:
    """The reference's own CPU implementation (oracle/_ref: the unmodified sources,
    gcc -O3 -fopenmp) of one band, split the way the GPU arm is timed: `setup` = everything
    done once per run (create_gas_optics, HITRAN parse, add_molecule / add_cfc / add_cia --
    framework/src/driver.c:617-625), `column` = the per-column chain of column_calculation
    (driver.c:360-424): set_*_ppmv, calculate_optical_depth, rayleigh_scattering,
    add_optics, calculate_{lw,sw}_fluxes and the broadband trapezoid."""

    def __init__(self, band, levels: int, molecules, threads: int):
        from oracle import ref as R
        self.R = R
        self.band = band
        self.lw = band.name.startswith("lw")
        self.levels = levels
        self.molecules = molecules
        t0 = time.perf_counter()
        self.r = R.Ref(threads=threads)
        self.cores = self.r.max_threads()
        self.grid = self.r.grid(band.w0, band.wn, band.dw)
        ds = band.dataset
        self.go = R.RefGasOptics(self.r, levels, self.grid, ds.hitran, ds.h2o_ctm_dir, ds.o3_ctm,
                                 method=R.LINE_SAMPLE)
        for m in molecules:
            self.go.add_molecule(m)
        for c, path in ds.cfc.items():
            self.go.add_cfc(c, path)
        for (a, b), path in ds.cia.items():
            self.go.add_cia(a, b, path)
        self.solar = None if self.lw else R.solar_flux(self.r, self.grid, ds.solar)
        self.setup_s = time.perf_counter() - t0

    def column(self, col) -> tuple[float, np.ndarray, np.ndarray]:
        """Seconds of the per-column chain, broadband up / down fluxes [levels]."""
        R, r, ds = self.R, self.r, self.band.dataset
        from grtcode_b200 import synth
        t0 = time.perf_counter()
        for m in self.molecules:
            self.go.set_ppmv("mol", m, col.ppmv[m])
        for c in ds.cfc:
            self.go.set_ppmv("cfc", c, col.cfc_ppmv[c])
        if ds.cia:
            for c in (synth.CIA_N2, synth.CIA_O2):
                self.go.set_ppmv("cia", c, col.cia_ppmv[c])
        gas = R.RefOptics(r, self.levels - 1, self.grid)
        self.go.optical_depth(col.p, col.t, gas)
        ray = R.RefOptics(r, self.levels - 1, self.grid)
        R.rayleigh(r, ray, col.p)
        tot = R.add_optics(r, [gas, ray])
        n = int(self.grid.n)
        if self.lw:
            up, dn = R.lw_fluxes(r, tot, col.t_surf, col.t_layer, col.t, np.full(n, col.emissivity))
        else:
            alb = np.full(n, col.albedo)
            up, dn = R.sw_fluxes(r, tot, col.mu0, 0.5, alb, alb, col.tsi, self.solar)
        bu, bd = R.broadband(up, self.band.dw), R.broadband(dn, self.band.dw)
        for o in (tot, ray, gas):
            o.close()
        return time.perf_counter() - t0, bu, bd

    def close(self):
        t0 = time.perf_counter()
        self.go.close()
        self.setup_s += time.perf_counter() - t0


def sample_windows(band, nwin: int, width: float, step: int):
    """`nwin` windows of `width` >= 50 cm-1 spread over the band; `step` rotates their
    positions (golden-ratio offset) so that successive steps of the reference arm sample
    different parts of the band.  Lines are only loaded inside a window, so a grid point
    closer than 25 cm-1 to a window end sees fewer lines than in the full band: the window's
    line density is raised by 50/(50 - 625/width) so that the number of Voigt samples PER
    GRID POINT -- what the pts x layers / s metric costs -- equals the band's mean."""
    from grtcode_b200 import workloads as W
    span = band.wn - band.w0
    width = min(width, span)
    shift = (0.6180339887498949*step) % 1.
    corr = 50./(50. - 625./width) if width >= 50. else 2.5
    out = []
    for k in range(nwin):
        centre = band.w0 + ((k + shift) % nwin + 0.5)*span/nwin
        w0 = min(max(band.w0, centre - width/2.), band.wn - width)
        lines = {m: max(1, int(round(n*width/span*corr))) for m, n in band.lines.items()}
        out.append(W.Band(f"{band.name}_w{k}", w0, w0 + width, band.dw, lines))
    return out


def cpu_reference_sample(args, col, threads=None, step: int = 0):
    """One bounded sample of the config-2 step on the reference (oracle/_ref; the C
    restatement oracle/ when that is absent): windows of each band carrying the band's mean
    number of Voigt samples per grid point, per-column chain only (setup timed separately),
    scaled by grid points to the full LW+SW column.
    Returns (pts*layers/s, description dict, extrapolated seconds of a full step)."""
    from grtcode_b200 import workloads as W
    from oracle import ref as R
    if threads is None:
        threads = host_threads()
    lw, sw = W.circ1_bands(args.dw, args.scale)
    kind = "reference" if R.available() else "port"
    width = float(os.environ.get("GRT_CPU_WINDOW", "50"))
    nwin = {"lw": max(1, int(os.environ.get("GRT_CPU_WINDOWS_LW", os.environ.get("GRT_CPU_WINDOWS", "3")))),
            "sw": max(1, int(os.environ.get("GRT_CPU_WINDOWS_SW", os.environ.get("GRT_CPU_WINDOWS", "6"))))}
    layers = len(col.p) - 1
    secs = {"lw": 0., "sw": 0.}
    setup = {"lw": 0., "sw": 0.}
    pts = {"lw": 0, "sw": 0}
    nlines = {"lw": 0, "sw": 0}
    cores = None
    for band in (lw, sw):
        key = band.name
        for k, sb in enumerate(sample_windows(band, nwin[key], width, step)):
            W.materialise(sb, os.path.join(args.data, "cpu", f"{key}{k}"), seed=11 + 97*step + k)
            if kind == "reference":
                rb = ReferenceBand(sb, len(col.p), W.MOLECULES, threads)
                cores = rb.cores
                dt, _, _ = rb.column(col)
                rb.close()
                setup[key] += rb.setup_s
            else:
                from oracle import port as P
                t0 = time.perf_counter()
                go = P.GasOptics(len(col.p), sb.w0, sb.wn, sb.dw, sb.dataset.hitran,
                                 sb.dataset.h2o_ctm_dir, sb.dataset.o3_ctm)
                for m in W.MOLECULES:
                    go.add_molecule(m)
                    go.set_ppmv("mol", m, col.ppmv[m])
                setup[key] += time.perf_counter() - t0
                t0 = time.perf_counter()
                go.optical_depth(col.p, col.t)
                dt = time.perf_counter() - t0
                cores = os.cpu_count()
            secs[key] += dt
            pts[key] += sb.n
            nlines[key] = sb.num_lines
    full_n = {"lw": lw.n, "sw": sw.n}
    est_full = sum(secs[k]*full_n[k]/pts[k] for k in secs)
    value = (lw.n + sw.n)*layers/est_full
    detail = "; ".join(f"{k}: {nwin[k]} x {width:g} cm-1 @ {args.dw:g} ({pts[k]} pts, {nlines[k]} "
                       f"lines per window) {secs[k]:.2f} s" for k in secs)
    sample = ("per-column chain (set_ppmv .. broadband) of one CIRC-1 column on windows of each "
              "band holding the band's mean Voigt samples per grid point, scaled by grid points "
              "to the full LW+SW column (an extrapolation, see profiles/r02_reference_full_band.json "
              "for the un-windowed check); " + detail)
    info = {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample,
            "extrapolated": True, "sample_work_s": sum(secs.values()),
            "setup_s": sum(setup.values()),
            "setup_note": "create_gas_optics + HITRAN parse + add_* + destroy of the sample "
                          "windows (not part of value, as setup_s is not part of the GPU arm's)",
            "full_band_check": recorded_full_band_check()}
    return value, info, est_full


def recorded_full_band_check():
    """The un-windowed reference run of one whole band committed under profiles/ (tools/
    reference_full_band.py), beside the windowed estimate for the same band."""
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_reference_full_band.json")))
    if not files:
        return None
    try:
        return json.load(open(files[-1]))
    except (OSError, ValueError):
        return None


def run_reference(args):
    """--impl reference: the reference's CPU path on the host cores (rank 0 only)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from grtcode_b200 import workloads as W
    col = W.circ1_column()
    lw, sw = W.circ1_bands(args.dw, args.scale)
    layers = len(col.p) - 1
    est, walls = [], []
    info = None
    for it in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        _, info, est_full = cpu_reference_sample(args, W.perturbed(col, 0, it), step=it)
        if it >= args.warmup:
            est.append(est_full)
            walls.append(time.perf_counter() - t0)
    # whole-run figure: the timed steps' extrapolated full-column seconds, averaged
    mean_est = float(np.mean(est)) if est else float("nan")
    value = (lw.n + sw.n)*layers/mean_est
    info = dict(info, value=value, steps_averaged=len(est),
                sample_wall_ms_per_step=1e3*float(np.mean(walls)) if walls else None)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT,
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3*mean_est, "ms_per_step_note": "extrapolated full-column time",
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": DTYPE,
            "data": "synthetic HITRAN-format line lists + CIRC-1 profile",
            "config": workload_config(args, lw, sw, layers),
            "cpu_baseline": info,
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0,
                    "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def workload_config(args, lw, sw, layers):
    return {"workload": "BASELINE config 2: CIRC case 1 clear-sky LW+SW, 7 HITRAN molecules + "
                        "H2O/O3 continua + 3 CIA pairs + 3 CFCs, one column per step per GPU",
            "precision": "single (fp_t = float, --enable-single-precision)" if SINGLE else "double",
            "layers": layers, "dw_cm-1": args.dw, "lw_points": lw.n, "sw_points": sw.n,
            "lw_lines": lw.num_lines, "sw_lines": sw.num_lines, "line_scale": args.scale,
            "method": args.method, "columns_per_step_per_gpu": 1,
            "l2": "inputs larger than L2 (line records + tau > 2 GB per band)"}


def recorded_wing_traffic(args):
    """DRAM bytes (read + write) per wing_sweep_kernel launch, averaged over the LW and SW
    launches of a step, from the latest ncu capture committed under profiles/ (ncu cannot run
    inside a timed bench; the figure is only valid for the full-size line_sample workload it
    was captured on, else None)."""
    import glob
    if args.scale != 1. or args.method != "line_sample":
        return None
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_wing_dram.json")))
    if not files:
        return None
    rec = json.load(open(files[-1]))["launches"]
    return sum(r["dram_read_bytes"] + r["dram_write_bytes"] for r in rec)/len(rec)


def rfmip_column(gi: int, cfcs):
    """Column `gi` of the 1800-column RFMIP-IRF job (100 sites x 18 experiments; one distinct
    synthetic profile per site, perturbed per experiment in the spirit of the RFMIP
    experiments: CO2 scaling, +-4 K)."""
    from grtcode_b200 import synth
    from grtcode_b200 import workloads as W
    site, expt = gi % 100, gi // 100
    rng = np.random.default_rng(7000 + site)
    scale = (1., 4., 0.5, 2., 3., 8.)[expt % 6]
    return synth.make_column(61, rng, molecules=W.MOLECULES, cfcs=cfcs, co2_scale=scale,
                             dt=4.*(expt % 3 - 1))


def rfmip_leg(args, bands2, device, rank, world, barrier, stream, torch, dist):
    """North-star target, BASELINE config 3: RFMIP-IRF-shaped columns (61 levels, LW 1-3250 +
    SW 1-50000 cm-1 at 0.1 cm-1, the config-2 synthetic line lists, 7 molecules + continua +
    3 CIA pairs + 3 CFCs), end to end through the C API (host arrays in, broadband fluxes out),
    GRT_RFMIP_COLUMNS (default 225) distinct columns per rank -- so that 8 ranks run the
    whole 1800-column job -- sharded with no collective.  On one rank the reference's CPU path
    is timed on GRT_RFMIP_CPU_COLUMNS (default 1) FULL columns of the same shape."""
    from grtcode_b200 import workloads as W
    from grtcode_b200.column_engine import ColumnEngine
    ncols = int(os.environ.get("GRT_RFMIP_COLUMNS", "225"))
    if os.environ.get("GRT_BENCH_RFMIP", "1") == "0" or ncols < 1:
        return None
    lw3, sw3 = W.circ1_bands(0.1, args.scale)
    lw3.dataset, sw3.dataset = bands2[0].dataset, bands2[1].dataset
    cfcs = tuple(bands2[0].dataset.cfc.keys())
    t_setup = time.perf_counter()
    eng = ColumnEngine((lw3, sw3), rfmip_column(0, cfcs), device, W.MOLECULES)
    t_setup = time.perf_counter() - t_setup
    eng.step(rfmip_column(rank*ncols, cfcs), upload_surface=True)
    # the columns are built before the timed region (input generation is not the hot path)
    cols = [rfmip_column(rank*ncols + k, cfcs) for k in range(ncols)]
    e0 = torch.cuda.Event(enable_timing=True)
    e1 = torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record(stream)
    samples = 0
    kern = {}
    first_fluxes = None
    for k, c in enumerate(cols):
        out = eng.step(c, upload_surface=True)
        if k == 0:
            first_fluxes = out
        samples += sum(out[n][2]["samples"] for n in ("lw", "sw"))
        for n in ("lw", "sw"):
            for key in ("prep_ns", "wing_ns", "core_ns"):
                name = f"{n}_{key[:-3]}"
                kern[name] = kern.get(name, 0.) + out[n][2][key]*1e-6/ncols
    e1.record(stream)
    barrier()
    secs_local = e0.elapsed_time(e1)*1e-3
    secs = secs_local
    per_rank = [secs_local]
    if world > 1:
        t = torch.tensor([secs], device=f"cuda:{device}", dtype=torch.float64)
        allt = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(allt, t)
        per_rank = [float(x.item()) for x in allt]
        secs = max(per_rank)
    eng.close()
    res = {"columns_per_s": world*ncols/secs, "ms_per_column_per_gpu": 1e3*secs/ncols,
           "columns_timed_per_gpu": ncols, "columns_timed_total": world*ncols,
           "seconds_timed": secs, "seconds_per_rank": [round(x, 4) for x in per_rank],
           "setup_s": t_setup,
           "voigt_samples_per_column": samples//ncols,
           "kernel_ms_per_column": {k: round(v, 3) for k, v in kern.items()},
           "config": "BASELINE config 3 shape: 61 levels, LW 1-3250 (n=32491) + SW 1-50000 "
                     "(n=499991) at 0.1 cm-1, 7 molecules + continua + 3 CIA + 3 CFCs, distinct "
                     "columns (site, experiment), block-sharded over the GPUs, no collective",
           "seconds_for_1800_columns": 1800.*secs/(world*ncols),
           "seconds_for_1800_columns_note": (
               "measured: this run IS the 1800-column job" if world*ncols == 1800 else
               f"{world*ncols} columns timed, scaled linearly to 1800 (columns are independent)")}
    ncpu = int(os.environ.get("GRT_RFMIP_CPU_COLUMNS", "1"))
    if rank == 0 and world == 1 and ncpu > 0 and not args.no_cpu_baseline:
        res["cpu_baseline"] = rfmip_cpu_baseline(args, (lw3, sw3), cfcs, ncpu, first_fluxes)
        res["vs_cpu"] = res["columns_per_s"]/res["cpu_baseline"]["columns_per_s"]
        res["vs_cpu_note"] = ("one B200 against the reference's OpenMP path on this box's "
                              f"{res['cpu_baseline']['cores']} host cores, same column shape, "
                              "per-column work only on both sides (north_star target: >= 50)")
    return res


def rfmip_cpu_baseline(args, bands3, cfcs, ncols: int, gpu_first):
    """The reference's CPU path on `ncols` FULL config-3-shaped columns (no windows): the
    column loop of framework/src/driver.c:691-743 on oracle/_ref with every host core."""
    from grtcode_b200 import workloads as W
    from oracle import ref as R
    if not R.available():
        return {"unavailable": "oracle/_ref not built"}
    threads = host_threads()
    parts = [ReferenceBand(b, 61, W.MOLECULES, threads) for b in bands3]
    secs = {"lw": 0., "sw": 0.}
    err = 0.
    for k in range(ncols):
        col = rfmip_column(k, cfcs)
        for part in parts:
            dt, up, dn = part.column(col)
            secs[part.band.name] += dt
            if k == 0 and gpu_first is not None:
                g = gpu_first[part.band.name]
                err = max(err, float(np.abs(g[0] - up).max()), float(np.abs(g[1] - dn).max()))
    total = sum(secs.values())
    cores = parts[0].cores
    for part in parts:
        part.close()
    return {"columns_per_s": ncols/total, "seconds_per_column": total/ncols, "columns_timed": ncols,
            "cores": cores, "kind": "reference", "extrapolated": False,
            "seconds_per_column_by_band": {k: v/ncols for k, v in secs.items()},
            "setup_s": sum(p.setup_s for p in parts),
            "max_abs_flux_difference_vs_gpu_W_m2": err,
            "sample": f"{ncols} full column(s) of the RFMIP shape, LW+SW, line_sample, "
                      "per-column chain (set_ppmv .. broadband); setup (parse, add_*) separate"}


def spectral_split_leg(args, bands2, device, rank, world, barrier, torch, dist, sw_ms_one_gpu):
    """BASELINE config 4 at N > 1: ONE column's 1-50000 cm-1 range at 0.01 cm-1 (the SW band
    of config 2: same grid, line list and tables) split over the N ranks, each rank on its
    chunk + halo, the partial broadband integrals summed by one NCCL all-reduce.  Strong
    scaling: efficiency = (one-GPU time of the same band, measured in this run by every
    rank) / (N x the slowest rank's time)."""
    from grtcode_b200 import parallel as P
    from grtcode_b200 import workloads as W
    if world < 2 or os.environ.get("GRT_BENCH_SPLIT", "1") == "0":
        return None
    col = W.circ1_column()
    sw = bands2[1]

    def measure(part):
        """Per-rank [compute s, all-reduce s, total s, points, samples] of one split column."""
        part.run(W.perturbed(col, 0, 0))                   # warm-up (allocations, tile lists)
        comp, red, tot = [], [], []
        for k in range(3):
            barrier()
            t0 = time.perf_counter()
            part.run(W.perturbed(col, 0, 300 + k))
            torch.cuda.synchronize(device)
            t1 = time.perf_counter()
            comp.append(part.compute_seconds)
            red.append(part.reduce_seconds)
            tot.append(t1 - t0)
        mine = torch.tensor([min(comp), float(np.median(red)), min(tot),
                             float(part.chunks["sw"].n),
                             float(part.engine.parts[0]["go"].stats()["samples"])],
                            device=f"cuda:{device}", dtype=torch.float64)
        allv = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allv, mine)
        return [[float(x) for x in v.tolist()] for v in allv]

    t_setup = time.perf_counter()
    part = P.SpectralSplitColumn([sw], col, device, W.MOLECULES, world, rank)
    t_setup = time.perf_counter() - t_setup
    cur = measure(part)
    rows = cur                                             # the best cuts' measurements
    passes = [{"cuts": "line-density model (parallel.interval_cost)",
               "seconds_split": max(r[2] for r in cur),
               "compute_seconds_per_rank": [round(r[0], 5) for r in cur]}]
    # cuts moved by the measured per-rank times (parallel.measured_costs); every rank takes
    # the same decisions because all of them see the same gathered numbers
    for it in range(int(os.environ.get("GRT_SPLIT_REBALANCE", "2"))):
        cost = P.measured_costs(part.costs["sw"], part.all_chunks["sw"], [r[0] for r in cur])
        part.close()
        part = P.SpectralSplitColumn([sw], col, device, W.MOLECULES, world, rank,
                                     cost_override={"sw": cost})
        cur = measure(part)
        passes.append({"cuts": f"measured per-rank times, pass {it + 1}",
                       "seconds_split": max(r[2] for r in cur),
                       "compute_seconds_per_rank": [round(r[0], 5) for r in cur]})
        if max(r[2] for r in cur) < max(r[2] for r in rows):
            rows = cur
    part.close()
    slowest = max(r[2] for r in rows)
    one = sw_ms_one_gpu*1e-3
    return {"workload": "BASELINE config 4: one CIRC-1 column, 1-50000 cm-1 at 0.01 cm-1 "
                        "(n=4999901), 7 molecules + continua + 3 CIA pairs, spectral range "
                        "split over the ranks, NCCL all-reduce of the broadband fluxes",
            "scaling": "strong", "n_gpus": world,
            "seconds_one_gpu": one, "seconds_split": slowest,
            "speedup": one/slowest, "efficiency": one/(world*slowest),
            "compute_seconds_per_rank": [round(r[0], 5) for r in rows],
            "allreduce_seconds_per_rank": [round(r[1], 6) for r in rows],
            "points_per_rank": [int(r[3]) for r in rows],
            "voigt_samples_per_rank": [int(r[4]) for r in rows],
            "allreduce_bytes": 2*8*len(col.p), "setup_s": t_setup, "passes": passes,
            "limiter": "the slowest rank's compute time (chunk cost imbalance + the halo and "
                       "fixed per-launch costs every rank repeats); the all-reduce is "
                       f"{1e6*float(np.median([r[1] for r in rows])):.0f} us"}


def reference_api_leg(args, eng, col0, rank, barrier, stream, torch, dist, world, device,
                      pts_layers):
    """e2e_reference_api: the config-2 step through the reference's own signatures only
    (ColumnEngine.step_reference_api): spectral fluxes [levels x n] of both bands are copied
    to (page-locked) host arrays every step -- 4.7 GB per column at config 2 -- instead of the
    1.8 KB of broadband values the fused column calls return."""
    from grtcode_b200 import workloads as W
    if args.method != "line_sample":
        return None
    steps = max(1, min(args.steps, 3))
    eng.step_reference_api(W.perturbed(col0, rank, 400))          # allocations, first touch
    e0 = torch.cuda.Event(enable_timing=True)
    e1 = torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record(stream)
    for k in range(steps):
        out = eng.step_reference_api(W.perturbed(col0, rank, 401 + k))
    e1.record(stream)
    barrier()
    secs = e0.elapsed_time(e1)*1e-3
    if world > 1:
        t = torch.tensor([secs], device=f"cuda:{device}", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        secs = float(t.item())
    dw = args.dw
    bb = {k: (float(np.sum(0.5*(v[0][0, :-1] + v[0][0, 1:])*dw)),
              float(np.sum(0.5*(v[1][-1, :-1] + v[1][-1, 1:])*dw))) for k, v in out.items()}
    return {"value": world*pts_layers*steps/secs, "unit": UNIT, "ms_per_step": 1e3*secs/steps,
            "steps": steps, "h2d_bytes_per_step": eng.h2d_ref, "d2h_bytes_per_step": eng.d2h_ref,
            "path": "set_*_ppmv, calculate_optical_depth, rayleigh_scattering, add_optics, "
                    "calculate_lw_fluxes / calculate_sw_fluxes with host flux arrays "
                    "(longwave.h:56-68, shortwave.h:56-69), page-locked",
            "broadband_of_returned_spectra_W_m2": {k: {"up_toa": v[0], "down_sfc": v[1]}
                                                   for k, v in bb.items()}}


def run_ours(args):
    import torch
    import torch.distributed as dist
    from grtcode_b200 import api
    from grtcode_b200 import workloads as W

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not api.libraries_present():
        raise RuntimeError("libgrt_sm100.so / libgrtcode.so are not built; there is no fallback")
    api.load()
    if api.num_gpus() < 1:
        raise RuntimeError("no CUDA device visible: this library has no CPU path")
    device = local
    torch.cuda.set_device(device)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", device))

    col0 = W.circ1_column()
    lw, sw = W.circ1_bands(args.dw, args.scale)
    # Synthetic line lists and tables: rank 0 of the node writes them, everyone reads.
    if local == 0:
        W.materialise(lw, args.data, seed=1)
        W.materialise(sw, args.data, seed=2)
    if world > 1:
        dist.barrier()
    if local != 0:
        W.materialise(lw, args.data, seed=1)
        W.materialise(sw, args.data, seed=2)
    t_setup = time.perf_counter()
    from grtcode_b200.column_engine import ColumnEngine
    method = {"wavenumber_sweep": 0, "line_sweep": 1, "line_sample": 2}[args.method]
    eng = ColumnEngine((lw, sw), col0, device, W.MOLECULES, method=method)
    t_setup = time.perf_counter() - t_setup
    layers = eng.levels - 1
    pts_layers = (lw.n + sw.n)*layers

    stream = torch.cuda.ExternalStream(api.stream_handle(device), device=device)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(device)

    def timed(upload: bool, first_step: int):
        """K steps bracketed by barrier + synchronize; device time from CUDA events on the
        library's stream; returns (max-over-ranks seconds, per-kernel stats of the run)."""
        e0 = torch.cuda.Event(enable_timing=True)
        e1 = torch.cuda.Event(enable_timing=True)
        wing_ns = 0
        wing_flop = 0.
        kern = {"prep_ns": 0, "wing_ns": 0, "core_ns": 0}
        band_ms = {}
        barrier()
        e0.record(stream)
        for k in range(args.steps):
            out = eng.step(W.perturbed(col0, rank, first_step + k), upload)
            for name in ("lw", "sw"):
                st = out[name][2]
                wing_ns += st["wing_ns"]
                wing_flop += FLOP_PER_SAMPLE*(st["samples"] - st["core_samples"])
                for key in kern:
                    kern[key] += st[key]
                    band_ms[f"{name}_{key[:-3]}"] = band_ms.get(f"{name}_{key[:-3]}", 0.) + \
                        st[key]*1e-6/args.steps
        e1.record(stream)
        barrier()
        secs = e0.elapsed_time(e1)*1e-3
        if world > 1:
            t = torch.tensor([secs], device=f"cuda:{device}", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            secs = float(t.item())
        kern = dict(kern, band_ms=band_ms)
        return secs, wing_ns, wing_flop, kern, out

    # warm-up (the first step also uploads the surface arrays the device-resident run reuses)
    for k in range(max(args.warmup, 1)):
        eng.step(W.perturbed(col0, rank, k), upload_surface=True)
    barrier()

    clocks = ClockSampler(device)
    clocks.start()
    secs, wing_ns, wing_flop, kern, out = timed(False, 100)
    clk = clocks.stop()
    launches = eng.launches*args.steps
    secs_e2e, _, _, _, out_e2e = timed(True, 200)
    h2d, d2h = eng.h2d, eng.d2h

    # the same step through the reference's own signatures: Rayleigh + add_optics
    # materialised, calculate_{lw,sw}_fluxes returning [levels x n] spectral fluxes to the host
    ref_api = None
    if os.environ.get("GRT_BENCH_REFAPI", "1") != "0":
        ref_api = reference_api_leg(args, eng, col0, rank, barrier, stream, torch, dist, world,
                                    device, pts_layers)
    # one-GPU time of the 1-50000 cm-1 band alone (the config-4 column): the denominator of
    # the spectral-split leg's strong-scaling efficiency
    sw_ms = None
    if world > 1:
        e0 = torch.cuda.Event(enable_timing=True)
        e1 = torch.cuda.Event(enable_timing=True)
        eng.step(W.perturbed(col0, 0, 0), True, only=("sw",))
        barrier()
        e0.record(stream)
        for k in range(3):
            eng.step(W.perturbed(col0, 0, 300 + k), True, only=("sw",))
        e1.record(stream)
        barrier()
        sw_ms = e0.elapsed_time(e1)/3.
    eng.close()
    rfmip = rfmip_leg(args, (lw, sw), device, rank, world, barrier, stream, torch, dist)
    split = spectral_split_leg(args, (lw, sw), device, rank, world, barrier, torch, dist, sw_ms)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    value = world*pts_layers*args.steps/secs
    e2e = world*pts_layers*args.steps/secs_e2e
    fp64_peak, fp32_peak = api.measure_fma_peak(device)
    launches_wing = 2*args.steps
    achieved = wing_flop/(wing_ns*1e-9)/1e12 if wing_ns else None
    samples_step = sum(out[n][2]["samples"] for n in ("lw", "sw"))
    roofline = {
        "bound": "fp32" if SINGLE else "fp64",
        "kernel": "wing_sweep_kernel" if args.method == "line_sample" else "bin_gather_kernel",
        "achieved": achieved, "peak": fp32_peak if SINGLE else fp64_peak, "unit": "TFLOP/s",
        "frac": achieved/(fp32_peak if SINGLE else fp64_peak) if achieved else None,
        "traffic": recorded_wing_traffic(args),
        "peak_source": "measured here: dependent-free " + ("FFMA" if SINGLE else "DFMA") +
                       " chains, all SMs (grt_measure_fma_peak); MEASURED_PEAKS.json has no "
                       "non-tensor FP64 / FP32 figure",
        "flop_per_sample": FLOP_PER_SAMPLE,
        "avg_launch_ms": wing_ns*1e-6/launches_wing if wing_ns else None,
        "samples_per_s": (wing_flop/FLOP_PER_SAMPLE)/(wing_ns*1e-9) if wing_ns else None,
        "fp32_peak": fp32_peak,
        "kernel_share_of_step": {k: v*1e-9/secs for k, v in kern.items() if k != "band_ms"},
        "kernel_ms_per_step": {k: round(v, 3) for k, v in kern["band_ms"].items()},
    }
    # The unit that actually binds the exact-rounding far-wing chain is the XU pipe (one
    # F2F.F32.F64 + one MUFU.RCP per sample at 16 lanes/clk/SM, profiles/*ubench_pipes.txt);
    # reported beside the FP64 figure SURVEY.md 8(d) asks for.  A lower bound: region-1 and
    # line-centre samples issue more XU instructions than two.
    sm_mhz = clk.get("sm_mhz") or clk.get("sm_max_mhz")
    if roofline["samples_per_s"] and sm_mhz:
        xu_peak = torch.cuda.get_device_properties(device).multi_processor_count*16.*sm_mhz*1e6
        # single build: no F2F, and half of the point pairs take the MUFU-free reciprocal
        xu_ops = 0.5 if SINGLE else 2.
        roofline["xu_pipe"] = {"ops_per_sample": xu_ops,
                               "achieved_lane_ops_per_s": xu_ops*roofline["samples_per_s"],
                               "peak_lane_ops_per_s": xu_peak,
                               "frac": xu_ops*roofline["samples_per_s"]/xu_peak}
    if args.method != "line_sample":
        # the flop bookkeeping above is the wing sweep's: no roofline line for the binned methods
        roofline = None
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3*secs/args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": DTYPE,
            "data": "synthetic HITRAN-format line lists + CIRC-1 profile",
            "config": workload_config(args, lw, sw, layers),
            "clocks": clk,
            "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "ms_per_step": 1e3*secs_e2e/args.steps},
            "gpu_launches": launches,
            "roofline": roofline,
            "voigt_samples_per_step": samples_step,
            "setup_s": t_setup,
            "rfmip": rfmip,
            "e2e_reference_api": ref_api,
            "spectral_split": split,
            "fluxes_W_m2": {"lw_up_toa": float(out["lw"][0][0]), "lw_dn_sfc": float(out["lw"][1][-1]),
                            "sw_up_toa": float(out["sw"][0][0]), "sw_dn_sfc": float(out["sw"][1][-1])}}
    if world == 1 and not args.no_cpu_baseline:
        _, info, _ = cpu_reference_sample(args, col0)
        line["cpu_baseline"] = info
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
