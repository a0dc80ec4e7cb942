"""A/B of kernel variants on the GPU box: for each variant directory under
grtcode_b200/variants/ named on the command line ("base" = the in-tree build) run a short
bench (config 2 at 0.01 cm-1 + 4 RFMIP-shaped columns at 0.1 cm-1) and print the kernel times.

    python tools/ab_run.py base merge1 base:GRT_SWEEP_LANES=8,GRT_SWEEP_PTS=8 ...
"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for spec in sys.argv[1:]:
    # NAME or NAME:VAR=value,VAR=value (library variant + environment knobs)
    name, _, knobs = spec.partition(":")
    env = dict(os.environ, GRT_RFMIP_COLUMNS="6", GRT_RFMIP_CPU_COLUMNS="0", GRT_BENCH_REFAPI="0")
    for kv in filter(None, knobs.split(",")):
        k, _, v = kv.partition("=")
        env[k] = v
    if name != "base":
        env["GRT_LIB_DIR"] = os.path.join(ROOT, "grtcode_b200", "variants", name)
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "2", "--warmup",
                          "2", "--no-cpu-baseline"] + os.environ.get("GRT_AB_ARGS", "").split(),
                         env=env, capture_output=True, text=True)
    lines = [l for l in out.stdout.splitlines() if l.startswith("{")]
    if not lines:
        print(spec, "FAILED", out.stderr[-1500:])
        continue
    d = json.loads(lines[0])
    if d.get("roofline") is None:          # binned methods: the step time only
        print(f"{spec:12s} config2 {d['ms_per_step']:7.1f} ms/step ({d['config']['method']})", flush=True)
        continue
    k2 = d["roofline"]["kernel_ms_per_step"]
    k3 = d["rfmip"]["kernel_ms_per_column"]
    print(f"{spec:12s} config2 {d['ms_per_step']:7.1f} ms/step  lw_wing {k2['lw_wing']:6.1f} sw_wing "
          f"{k2['sw_wing']:6.1f} sw_core {k2['sw_core']:5.1f} | rfmip column "
          f"{d['rfmip']['ms_per_column_per_gpu']:6.2f} ms  lw_wing {k3['lw_wing']:5.2f} sw_wing "
          f"{k3['sw_wing']:5.2f} sw_core {k3['sw_core']:4.2f} prep {k3['lw_prep'] + k3['sw_prep']:4.2f} | "
          f"fluxes {d['fluxes_W_m2']['lw_up_toa']:.9f} {d['fluxes_W_m2']['sw_dn_sfc']:.9f}", flush=True)
