"""Setup cost of one band of BASELINE config 2, stage by stage (host clock)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from grtcode_b200 import api
from grtcode_b200 import workloads as W

api.load()
lw, sw = W.circ1_bands(0.01, 1.)
t0 = time.perf_counter()
W.materialise(lw, "/tmp/grt_bench", seed=1)
W.materialise(sw, "/tmp/grt_bench", seed=2)
print("materialise (synthetic files) %.2f s" % (time.perf_counter() - t0))
col = W.circ1_column()
for band in (lw, sw):
    ds = band.dataset
    t = [time.perf_counter()]
    mark = lambda: t.append(time.perf_counter())
    grid = api.spectral_grid(band.w0, band.wn, band.dw)
    go = api.GasOptics(len(col.p), grid, ds.hitran, ds.h2o_ctm_dir, ds.o3_ctm, device=0)
    mark()
    per = []
    for m in W.MOLECULES:
        t1 = time.perf_counter()
        go.add_molecule(m)
        per.append(time.perf_counter() - t1)
    mark()
    for c, path in ds.cfc.items():
        go.add_cfc(c, path)
    for (a, b), path in ds.cia.items():
        go.add_cia(a, b, path)
    mark()
    gas = api.OpticsObject(len(col.p) - 1, grid, 0)
    for m, v in col.ppmv.items():
        go.set_ppmv("mol", m, v)
    go.optical_depth(col.p, col.t, gas)
    mark()
    go.optical_depth(col.p, col.t, gas)
    mark()
    if band.name.startswith("sw"):
        api.solar_flux(grid, ds.solar)
    mark()
    d = [b - a for a, b in zip(t[:-1], t[1:])]
    print(f"{band.name}: create (continua tables) {d[0]:.2f}  add_molecule x7 {d[1]:.2f}  "
          f"cfc+cia tables {d[2]:.2f}  first optical_depth (line upload) {d[3]:.2f}  "
          f"second {d[4]:.2f}  solar {d[5]:.2f}  [s]")
    print("   add_molecule one by one: " + " ".join(f"{x:.3f}" for x in per))

# the HITRAN conversion alone: parse_hitran_file for one molecule with the table cache dropped,
# on the device and by the host threads
import ctypes as C
L = api.load()
L.grt_hitran_use_device.argtypes = [C.c_int]
L.parse_hitran_file.argtypes = [C.c_void_p, C.c_char_p, C.c_int, C.c_double, C.c_double]
for band in (lw, sw):
    for where in ("device", "host", "device", "host"):
        L.grt_hitran_cache_clear()
        if where == "host":
            os.environ["GRT_PARSE"] = "host"
        else:
            os.environ.pop("GRT_PARSE", None)
        L.grt_hitran_use_device(0)
        buf = C.create_string_buffer(256)
        t1 = time.perf_counter()
        rc = L.parse_hitran_file(buf, band.dataset.hitran.encode(), 2, band.w0, band.wn)
        print(f"{band.name}: parse_hitran_file (file read + conversion + one molecule's filter) "
              f"{where}: {time.perf_counter() - t1:.3f} s  rc {rc}")
