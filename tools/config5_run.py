"""BASELINE config 5 shape on one GPU: ERA5-style slab columns (37 levels, clear sky, LW 1-3250
+ SW 1-50000 cm-1 at 0.1 cm-1, 7 molecules + continua + 3 CIA pairs + 3 CFCs, the config-2
synthetic line lists), GRT_C5_COLUMNS distinct columns (default 512) end to end through the C
API.  Prints one JSON line: columns/s, ms per column, the sum of the kernel times per column
and the share of the elapsed time that is not kernel time (host gaps, copies, syncs) --
the figure that says whether a batched entry point would buy anything.

    python tools/config5_run.py > gpurun_out/config5.json
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch                                              # noqa: E402

from grtcode_b200 import api, synth                       # noqa: E402
from grtcode_b200 import workloads as W                   # noqa: E402
from grtcode_b200.column_engine import ColumnEngine       # noqa: E402


def slab_column(k: int, cfcs):
    """Column k of the slab: a latitude-like spread of surface temperature and humidity."""
    rng = np.random.default_rng(90000 + k)
    return synth.make_column(37, rng, molecules=W.MOLECULES, cfcs=cfcs,
                             dt=float(rng.uniform(-25., 8.)))


def main():
    ncols = int(os.environ.get("GRT_C5_COLUMNS", "512"))
    data = os.environ.get("GRT_BENCH_DATA", "/tmp/grt_bench")
    api.load()
    lw2, sw2 = W.circ1_bands(0.01, 1.)
    W.materialise(lw2, data, seed=1)
    W.materialise(sw2, data, seed=2)
    lw, sw = W.circ1_bands(0.1, 1.)
    lw.dataset, sw.dataset = lw2.dataset, sw2.dataset
    cfcs = tuple(lw2.dataset.cfc.keys())
    t0 = time.perf_counter()
    eng = ColumnEngine((lw, sw), slab_column(0, cfcs), 0, W.MOLECULES)
    setup = time.perf_counter() - t0
    stream = torch.cuda.ExternalStream(api.stream_handle(0), device=0)
    cols = [slab_column(k, cfcs) for k in range(ncols)]
    eng.step(cols[0], upload_surface=True)
    e0 = torch.cuda.Event(enable_timing=True)
    e1 = torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(0)
    e0.record(stream)
    kern = {}
    samples = 0
    wall0 = time.perf_counter()
    for c in cols:
        out = eng.step(c, upload_surface=True)
        for n in ("lw", "sw"):
            samples += out[n][2]["samples"]
            for key in ("prep_ns", "wing_ns", "core_ns"):
                kern[f"{n}_{key[:-3]}"] = kern.get(f"{n}_{key[:-3]}", 0.) + out[n][2][key]*1e-6/ncols
    e1.record(stream)
    torch.cuda.synchronize(0)
    wall = time.perf_counter() - wall0
    secs = e0.elapsed_time(e1)*1e-3
    eng.close()
    ksum = sum(kern.values())
    line = {"workload": "BASELINE config 5 shape: ERA5-style slab columns, 37 levels, LW 1-3250 + SW "
                        "1-50000 cm-1 at 0.1 cm-1, one B200, columns independent (8 GPUs shard the "
                        "~65k columns of a slab with no collective)",
            "columns": ncols, "seconds": secs, "wall_seconds": wall, "columns_per_s": ncols/secs,
            "ms_per_column": 1e3*secs/ncols, "setup_s": setup,
            "kernel_ms_per_column": {k: round(v, 3) for k, v in kern.items()},
            "gas_kernel_ms_per_column": ksum,
            "not_gas_kernel_share": 1. - ksum/(1e3*secs/ncols),
            "not_gas_kernel_note": "elapsed minus K0/K1/wing/core event times: the two fused solver "
                                   "kernels (~1.5 ms at 0.1 cm-1), copies, host gaps, syncs",
            "voigt_samples_per_column": samples//ncols,
            "hours_for_65536_columns_on_8_gpus": 65536/(8*ncols/secs)/3600.,
            "last_fluxes_W_m2": {"lw_up_toa": float(out["lw"][0][0]), "sw_dn_sfc": float(out["sw"][1][-1])}}
    print(json.dumps(line))


if __name__ == "__main__":
    main()
