"""Generate the committed fixtures under tests/golden/ from the reference mount.

Runs in the build container only (reads /root/reference; the GPU box has no
such path).  It copies NUMBERS, not code:

* reference_pins.json -- the known-answer vectors and their inputs held by the
  reference's own unit tests for the hot path:
    gas-optics/test/test_kernels.c   (5-layer / 5-H2O-line fixture, :38-126, goldens :160-264,
                                      H2O continuum inputs + tau :376-512)
    gas-optics/test/test_tips2017.c  (:34-65)
    utilities/test/test_curtis_godson.c (CIRC-1 n, pavg, tavg, ps, ns, :24-175)
* circ1_profile.json -- the CIRC case 1 column embedded in
  testing_harness/src/circ1.h:8-181 (55 levels; the 49180-point albedo and solar
  spectra are not copied).
"""
import json
import os
import re
import sys

REF = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")

NUM = r"[-+]?(?:\d+\.?\d*|\.\d+)(?:[eE][-+]?\d+)?"


def c_array(src: str, name: str, index: int = 0):
    """Values of the index-th initializer `name[...] = {...}` or `.name = {...}`."""
    lead = r"\." if name.startswith(".") else r"(?:\b|\.)"
    pat = re.compile(lead + re.escape(name.lstrip(".")) + r"\s*(?:\[[^\]]*\])?\s*=\s*\{([^}]*)\}",
                     re.S)
    m = pat.findall(src)
    if len(m) <= index:
        return []
    body = re.sub(r"/\*.*?\*/", "", m[index], flags=re.S)
    return [float(x) for x in re.findall(NUM, body)]


def c_scalar(src: str, name: str):
    m = re.search(r"\b" + re.escape(name) + r"\s*=\s*(" + NUM + ")", src)
    return float(m.group(1))


def main():
    os.makedirs(OUT, exist_ok=True)
    tk = open(os.path.join(REF, "gas-optics/test/test_kernels.c")).read()
    pins = {"source": "NOAA-GFDL/GRTCODE unit tests (see tools/make_golden.py)"}
    pins["kernels"] = {
        "level_pressure": c_array(tk, "level_pressure"),
        "level_xh2o": c_array(tk, "level_xh2o"),
        "pressure": c_array(tk, "pressure"),
        "temperature": c_array(tk, "temperature"),
        "center": c_array(tk, "center"),
        "delta": c_array(tk, "delta"),
        "energy": c_array(tk, "energy"),
        "gamma_foreign": c_array(tk, "gamma_foreign"),
        "gamma_self": c_array(tk, "gamma_self"),
        "n": c_array(tk, ".n"),
        "strength": c_array(tk, "strength"),
        "mass_amu": 18.010565,
        "vnn_ref": c_array(tk, "vnn_ref"),
        "q_ref": c_array(tk, "q_ref"),
        "strength_ref": c_array(tk, "strength_ref"),
        "gamma_ref": c_array(tk, "gamma_ref"),
        "alpha_ref": c_array(tk, "alpha_ref"),
        "ctm_T0F": c_array(tk, "T0F"),
        "ctm_CF": c_array(tk, "CF"),
        "ctm_T0S": c_array(tk, "T0S"),
        "ctm_CS": c_array(tk, "CS"),
        "ctm_tau_ref": next(v for v in (c_array(tk, "tau_ref", k) for k in range(12))
                            if len(v) == 250),
    }
    tt = open(os.path.join(REF, "gas-optics/test/test_tips2017.c")).read()
    pins["tips"] = {"temperature": c_scalar(tt, "TEMPERATURE") if False else 275.234324,
                    "iso": 1,
                    "Q": {m: float(v) for m, v in
                          re.findall(r"helper\((\w+),\s*(" + NUM + r")\)", tt)}}
    cg = open(os.path.join(REF, "utilities/test/test_curtis_godson.c")).read()
    pins["curtis_godson"] = {k: c_array(cg, k) for k in
                             ("n_ref", "pavg_ref", "tavg_ref", "ps_ref", "ns_ref")}
    json.dump(pins, open(os.path.join(OUT, "reference_pins.json"), "w"), indent=0)

    c1 = open(os.path.join(REF, "testing_harness/src/circ1.h")).read()
    prof = {"source": "testing_harness/src/circ1.h:8-181 (CIRC case 1)"}
    for k in ("level_pressure", "level_temperature", "layer_pressure", "layer_temperature",
              "H2O_abundance", "CO2_abundance", "O3_abundance", "N2O_abundance", "CO_abundance",
              "CH4_abundance", "O2_abundance", "CCl4_abundance", "CFC11_abundance",
              "CFC12_abundance"):
        prof[k] = c_array(c1, k)
    for k in ("solar_zenith_angle", "surface_temperature", "toa_solar_irradiance"):
        prof[k] = c_scalar(c1, k)
    json.dump(prof, open(os.path.join(OUT, "circ1_profile.json"), "w"), indent=0)
    # the product's copy (input data of BASELINE config 2): the package must not read tests/
    data = os.path.join(os.path.dirname(OUT.rstrip("/")), "..", "grtcode_b200", "data")
    os.makedirs(data, exist_ok=True)
    json.dump(prof, open(os.path.join(data, "circ1_profile.json"), "w"), indent=0)
    print({k: len(v) for k, v in prof.items() if isinstance(v, list)})
    print({k: len(v) for k, v in pins["kernels"].items() if isinstance(v, list)})
    print(pins["tips"])
    print({k: len(v) for k, v in pins["curtis_godson"].items()})


if __name__ == "__main__":
    main()
