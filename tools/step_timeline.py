"""Host-side timeline of one config-2 step: wall time of every C-API call (they are
synchronous) beside the kernel times the library reports, to see where the GPU idles."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from grtcode_b200 import api
from grtcode_b200 import workloads as W
from grtcode_b200.column_engine import ColumnEngine

api.load()
lw, sw = W.circ1_bands(0.01, float(os.environ.get("GRT_BENCH_SCALE", "1")))
W.materialise(lw, "/tmp/grt_bench", seed=1)
W.materialise(sw, "/tmp/grt_bench", seed=2)
col = W.circ1_column()
eng = ColumnEngine((lw, sw), col, 0, W.MOLECULES)
for it in range(3):
    eng.step(W.perturbed(col, 0, it), upload_surface=False)
c = W.perturbed(col, 0, 7)
T = {}
t_all = time.perf_counter()
for part in eng.parts:
    name = part["band"].name
    go = part["go"]
    t0 = time.perf_counter()
    for m, v in c.ppmv.items():
        go.set_ppmv("mol", m, v)
    for k, v in c.cfc_ppmv.items():
        if k in part["band"].dataset.cfc:
            go.set_ppmv("cfc", k, v)
    for k, v in c.cia_ppmv.items():
        go.set_ppmv("cia", k, v)
    t1 = time.perf_counter()
    go.optical_depth(c.p, c.t, part["gas"])
    t2 = time.perf_counter()
    st = go.stats()
    t3 = time.perf_counter()
    if name.startswith("lw"):
        part["solver"].column(part["gas"], c.p, c.t_surf, c.t_layer, c.t, None)
    else:
        part["solver"].column(part["gas"], c.p, c.mu0, 0.5, None, None, c.tsi, None)
    t4 = time.perf_counter()
    kern = (st["prep_ns"] + st["wing_ns"] + st["core_ns"])*1e-6
    print(f"{name}: set_ppmv {1e3*(t1-t0):.3f} ms  optical_depth {1e3*(t2-t1):.3f} ms "
          f"(kernels {kern:.3f}: prep {st['prep_ns']*1e-6:.3f} wing {st['wing_ns']*1e-6:.3f} "
          f"core {st['core_ns']*1e-6:.3f})  stats {1e3*(t3-t2):.3f} ms  solver {1e3*(t4-t3):.3f} ms")
print(f"step wall {1e3*(time.perf_counter()-t_all):.3f} ms")
eng.close()
