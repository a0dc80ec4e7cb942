"""Validation of bench.py's windowed reference estimate: the compiled reference (oracle/_ref,
all host cores) on ONE WHOLE band of BASELINE config 2, un-windowed, beside the windowed
estimate for the same band and column.  Writes profiles/r02_reference_full_band.json, which
bench.py attaches to its cpu_baseline as `full_band_check`.

    python tools/reference_full_band.py [lw|sw]      # lw: ~2-3 min on 16-32 cores
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np            # noqa: E402

import bench                  # noqa: E402
from grtcode_b200 import workloads as W   # noqa: E402


def main():
    which = sys.argv[1] if len(sys.argv) > 1 else "lw"
    data = os.environ.get("GRT_BENCH_DATA", "/tmp/grt_bench")
    lw, sw = W.circ1_bands(0.01, 1.)
    band = lw if which == "lw" else sw
    W.materialise(band, data, seed=1 if which == "lw" else 2)
    col = W.circ1_column()
    threads = bench.host_threads()
    rb = bench.ReferenceBand(band, len(col.p), W.MOLECULES, threads)
    full_s, up, dn = rb.column(col)
    rb.close()
    layers = len(col.p) - 1
    # the windowed estimate bench.py would make for this band (several window sets)
    width = float(os.environ.get("GRT_CPU_WINDOW", "50"))
    nwin = 3 if which == "lw" else 6
    est = []
    for step in range(4):
        secs = pts = 0
        for k, sb in enumerate(bench.sample_windows(band, nwin, width, step)):
            W.materialise(sb, os.path.join(data, "cpu_check", f"{which}{k}"), seed=11 + 97*step + k)
            wb = bench.ReferenceBand(sb, len(col.p), W.MOLECULES, threads)
            dt, _, _ = wb.column(col)
            wb.close()
            secs += dt
            pts += sb.n
        est.append(secs*band.n/pts)
    out = {"band": which, "points": band.n, "layers": layers, "lines": band.num_lines,
           "cores": rb.cores, "full_band_seconds": full_s, "full_band_setup_seconds": rb.setup_s,
           "full_band_pts_layers_per_s": band.n*layers/full_s,
           "windowed_estimates_seconds": est, "windowed_mean_seconds": float(np.mean(est)),
           "windows": f"{nwin} x {width:g} cm-1 per estimate, 4 window sets",
           "windowed_over_full": float(np.mean(est))/full_s,
           "broadband_W_m2": {"up_toa": float(up[0]), "down_sfc": float(dn[-1])},
           "when": time.strftime("%Y-%m-%dT%H:%M:%SZ", time.gmtime())}
    path = os.path.join(ROOT, "gpurun_out" if os.path.isdir(os.path.join(ROOT, "gpurun_out")) else
                        "profiles", "r02_reference_full_band.json")
    json.dump(out, open(path, "w"), indent=1)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
