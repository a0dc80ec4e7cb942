"""Where the wing sweep's SM cycles go, per line class (development build only).

    make -B -C grtcode_b200/csrc EXTRA=-DGRT_DBG_TIMERS=1 && make -C grtcode_b200/host
    python tools/class_timers.py [--dw 0.01] [--scale 1]

Runs one CIRC-1 LW+SW column (BASELINE config 2 at --dw 0.01, the config-3 grid at 0.1) and
prints, per class of (line, tile) visit, the share of warp cycles and the cycles per visit.
The timers cost time themselves (two clock reads and two atomics per class per batch): the
shares are what matters, not the absolute cycles.
"""
import argparse
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from grtcode_b200 import api
from grtcode_b200 import workloads as W
from grtcode_b200.column_engine import ColumnEngine

NAMES = ["classify", "fast", "r1", "far_edge", "near", "lorentz_full", "lorentz_edge", "tma_wait",
         "epilogue", "item_total"]

ap = argparse.ArgumentParser()
ap.add_argument("--dw", type=float, default=0.01)
ap.add_argument("--scale", type=float, default=1.)
ap.add_argument("--data", default="/tmp/grt_bench")
args = ap.parse_args()
api.load()
lib = api.cuda_layer()
if not hasattr(lib, "grt_debug_timers"):
    sys.exit("libgrt_sm100.so was not built with -DGRT_DBG_TIMERS=1")
lw2, sw2 = W.circ1_bands(0.01, args.scale)
W.materialise(lw2, args.data, seed=1)
W.materialise(sw2, args.data, seed=2)
lw, sw = W.circ1_bands(args.dw, args.scale)
lw.dataset, sw.dataset = lw2.dataset, sw2.dataset
col = W.circ1_column()
n = len(NAMES)
cyc = (C.c_ulonglong*n)()
vis = (C.c_ulonglong*n)()
for band in (lw, sw):
    eng = ColumnEngine((band,), col, 0, W.MOLECULES)
    eng.step(col, upload_surface=True)
    lib.grt_debug_timers(cyc, vis, n)            # reset after the warm-up
    out = eng.step(col, upload_surface=True)
    lib.grt_debug_timers(cyc, vis, n)
    st = out[band.name[:2]][2]
    total = float(cyc[NAMES.index("item_total")])
    print(f"{band.name} @ {args.dw}: wing {st['wing_ns']*1e-6:.1f} ms (with timers), "
          f"{st['samples']:.3e} samples")
    for k, name in enumerate(NAMES):
        per = cyc[k]/vis[k] if vis[k] else 0.
        print(f"  {name:13s} {100.*cyc[k]/total:6.2f} % of item cycles   {vis[k]:14d} visits "
              f"{per:10.1f} cycles/visit")
    eng.close()
