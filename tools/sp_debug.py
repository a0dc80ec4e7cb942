"""Development aid: per-level shortwave fluxes of both sides at one spectral point of the
single-precision config-1-shape case (GRT_PRECISION=single)."""
import sys, tempfile
import numpy as np
sys.path.insert(0, ".")
from grtcode_b200 import api, synth
from tests import parity
name, w0, wn, dw, levels, lines, seed, cfcs = ("config1_shape_0.1", 1., 3000., 0.1, 60,
    {synth.H2O: 2000, synth.CO2: 3000, synth.O3: 2000}, 1, ("CFC11", "CFC12"))
res = parity.run_column_pair(tempfile.mkdtemp(), w0, wn, dw, levels, lines, seed, cfcs=cfcs)
got, want = res["got"], res["want"]
absd = np.abs(got["sw_dn"].astype(np.float64) - want["sw_dn"].astype(np.float64))
lev, pt = np.unravel_index(int(absd.argmax()), absd.shape)
print("worst", lev, pt, "solar", got["solar"][pt], want["solar"][pt])
for l in range(levels):
    k = max(l - 1, 0)
    print(l, "dn", got["sw_dn"][l, pt], want["sw_dn"][l, pt], "up", got["sw_up"][l, pt], want["sw_up"][l, pt],
          "| layer above: tau", got["tau_total"][k, pt], want["tau_total"][k, pt],
          "omega", got["omega_total"][k, pt], want["omega_total"][k, pt])
# how many points differ by more than 1e-3 relative at the surface
rel = parity.relative_error(got["sw_dn"][-1].astype(np.float64), want["sw_dn"][-1].astype(np.float64))
big = np.nonzero((rel > 1e-3) & (want["sw_dn"][-1] > 1e-6))[0]
print("surface dn: points off by > 1e-3:", len(big), big[:20], rel[big[:20]])
