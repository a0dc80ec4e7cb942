"""One BASELINE config-1 sized column (59 layers, 1-3000 cm-1 @ 0.1, 6.5e5 lines)
for ncu: setup, two warm-up calls, one profiled-size call."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from grtcode_b200 import api, synth

scale = float(os.environ.get("GRT_LINES_SCALE", "1"))
lines = {1: int(100000*scale), 2: int(300000*scale), 3: int(250000*scale)}
ds = synth.make_dataset("/tmp/c1_%g" % scale, 1., 3000., lines, seed=1)
g = api.spectral_grid(1., 3000., 0.1)
col = synth.make_column(60, np.random.default_rng(1))
method = int(os.environ.get("GRT_METHOD", "2"))
go = api.GasOptics(60, g, ds.hitran, ds.h2o_ctm_dir, ds.o3_ctm, method=method)
for m in (1, 2, 3):
    go.add_molecule(m)
    go.set_ppmv("mol", m, col.ppmv[m])
gas = api.OpticsObject(59, g)
for it in range(int(os.environ.get("GRT_REPS", "3"))):
    t0 = time.time()
    go.optical_depth(col.p, col.t, gas)
    dt = time.time() - t0
    st = go.stats()
    print("tau secs %.4f samples/s %.3e" % (dt, st["samples"]/dt), st)
