"""Developer check on the GPU box: parity numbers + a timing of one bigger column."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from grtcode_b200 import api, synth
from tests import parity

api.load()
print("gpus", api.num_gpus())
print("fma peak (fp64, fp32) TF/s", api.measure_fma_peak(0))
for args in [
    dict(root="/tmp/p1", w0=500., wn=700., dw=0.1, num_levels=20, lines={1: 300, 2: 400, 3: 200}, seed=3),
    dict(root="/tmp/p2", w0=1., wn=3000., dw=0.1, num_levels=60, lines={1: 2000, 2: 3000, 3: 2000}, seed=1,
         cfcs=("CFC11", "CFC12"), cias=((0, 0), (1, 0), (1, 1))),
    dict(root="/tmp/p3", w0=1000., wn=1030., dw=0.01, num_levels=12, lines={1: 150, 2: 150, 6: 100}, seed=7),
]:
    t0 = time.time()
    res = parity.run_column_pair(**args)
    rel = parity.relative_error(res["got"]["tau"], res["want"]["tau"])
    print({k: v for k, v in res.items() if k not in ("got", "want")}, "secs", time.time() - t0)
    bad = np.argwhere(rel > 1e-9)
    print("bad points", len(bad), bad[:10].tolist())
    for (l, f) in bad[:5]:
        print("   ", l, f, res["got"]["tau"][l, f], res["want"]["tau"][l, f], rel[l, f])

# timing: config-1 sized column (6.5e5 lines)
import torch
ds = synth.make_dataset("/tmp/c1", 1., 3000., {1: 100000, 2: 300000, 3: 250000}, seed=1)
g = api.spectral_grid(1., 3000., 0.1)
rng = np.random.default_rng(1)
col = synth.make_column(60, rng)
t0 = time.time()
go = api.GasOptics(60, g, ds.hitran, ds.h2o_ctm_dir, ds.o3_ctm)
for m in (1, 2, 3):
    go.add_molecule(m)
    go.set_ppmv("mol", m, col.ppmv[m])
print("setup secs", time.time() - t0)
gas = api.OpticsObject(59, g)
for it in range(3):
    t0 = time.time()
    go.optical_depth(col.p, col.t, gas)
    dt = time.time() - t0
    st = go.stats()
    print("tau secs", dt, st, "samples/s %.3e" % (st["samples"]/dt), "pts*layers/s %.3e" % (59*g.n/dt))
