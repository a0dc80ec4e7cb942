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
FLOP_PER_SAMPLE = 9.0        # 6 FP64 + 3 FP32 per far-wing Voigt sample


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
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic HITRAN-format line lists + CIRC-1 profile",
            "config": workload_config(args, lw, sw, layers),
            "cpu_baseline": info,
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0,
                    "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def workload_config(args, lw, sw, layers):
    return {"workload": "BASELINE config 2: CIRC case 1 clear-sky LW+SW, 7 HITRAN molecules + "
                        "H2O/O3 continua + 3 CIA pairs + 3 CFCs, one column per step per GPU",
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


def rfmip_probe(args, bands2, device, rank, world, barrier, stream, torch, dist, ncols: int = 4):
    """Secondary figure, BASELINE config 3 shape: RFMIP-IRF-like columns (61 levels, LW 1-3250
    + SW 1-50000 cm-1 at 0.1 cm-1, the same synthetic line lists as config 2), `ncols` columns
    per rank, end to end through the C API (host arrays in, broadband fluxes out).  Columns
    are sharded over the ranks with no collective; columns/s is the whole-job figure."""
    from grtcode_b200 import synth
    from grtcode_b200 import workloads as W
    from grtcode_b200.column_engine import ColumnEngine
    if os.environ.get("GRT_BENCH_RFMIP", "1") == "0":
        return None
    lw3, sw3 = W.circ1_bands(0.1, args.scale)
    lw3.dataset, sw3.dataset = bands2[0].dataset, bands2[1].dataset
    cfcs = tuple(bands2[0].dataset.cfc.keys())

    def column(site: int, expt: int):
        rng = np.random.default_rng(7000 + site)
        scale = (1., 4., 0.5, 2., 3., 8.)[expt % 6]
        return synth.make_column(61, rng, molecules=W.MOLECULES, cfcs=cfcs, co2_scale=scale,
                                 dt=4.*(expt % 3 - 1))

    eng = ColumnEngine((lw3, sw3), column(0, 0), device, W.MOLECULES)
    eng.step(column(rank, 0), upload_surface=True)
    e0 = torch.cuda.Event(enable_timing=True)
    e1 = torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record(stream)
    samples = 0
    kern = {}
    for k in range(ncols):
        out = eng.step(column(100*rank + k, k), upload_surface=True)
        samples += sum(out[n][2]["samples"] for n in ("lw", "sw"))
        for n in ("lw", "sw"):
            for key in ("prep_ns", "wing_ns", "core_ns"):
                name = f"{n}_{key[:-3]}"
                kern[name] = kern.get(name, 0.) + out[n][2][key]*1e-6/ncols
    e1.record(stream)
    barrier()
    secs = e0.elapsed_time(e1)*1e-3
    if world > 1:
        t = torch.tensor([secs], device=f"cuda:{device}", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        secs = float(t.item())
    eng.close()
    return {"columns_per_s": world*ncols/secs, "ms_per_column_per_gpu": 1e3*secs/ncols,
            "columns_timed_per_gpu": ncols, "voigt_samples_per_column": samples//ncols,
            "kernel_ms_per_column": {k: round(v, 3) for k, v in kern.items()},
            "config": "BASELINE config 3 shape: 61 levels, LW 1-3250 (n=32491) + SW 1-50000 "
                      "(n=499991) at 0.1 cm-1, 7 molecules + continua + 3 CIA + 3 CFCs, "
                      "columns sharded over the GPUs, no collective",
            "seconds_for_1800_columns": 1800.*secs/(world*ncols)}


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

    rfmip = rfmip_probe(args, (lw, sw), device, rank, world, barrier, stream, torch, dist)

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
        "bound": "fp64",
        "kernel": "wing_sweep_kernel" if args.method == "line_sample" else "bin_gather_kernel",
        "achieved": achieved, "peak": fp64_peak, "unit": "TFLOP/s",
        "frac": achieved/fp64_peak if achieved else None,
        "traffic": recorded_wing_traffic(args),
        "peak_source": "measured here: dependent-free DFMA chains, all SMs (grt_measure_fma_peak); "
                       "MEASURED_PEAKS.json has no FP64 figure",
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
        roofline["xu_pipe"] = {"ops_per_sample": 2, "achieved_lane_ops_per_s": 2.*roofline["samples_per_s"],
                               "peak_lane_ops_per_s": xu_peak,
                               "frac": 2.*roofline["samples_per_s"]/xu_peak}
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3*secs/args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64",
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
            "fluxes_W_m2": {"lw_up_toa": float(out["lw"][0][0]), "lw_dn_sfc": float(out["lw"][1][-1]),
                            "sw_up_toa": float(out["sw"][0][0]), "sw_dn_sfc": float(out["sw"][1][-1])}}
    if world == 1 and not args.no_cpu_baseline:
        _, info = cpu_reference_sample(args, col0)
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
