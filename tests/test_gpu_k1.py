"""K1 (line preparation) against the compiled reference, bit for bit.

rfm_voigt_line_shape computes its line shape in `float` from two per-line floats,
REPWID = (float)(SQRLN2/alpha) and Y = (float)(REPWID*gamma) (RFM_voigt.c:94-95).  alpha is
IEEE arithmetic only; gamma contains POW(296/T, n) (kernels.c:105), which CUDA's libdevice and
the host libm round differently in the last bit now and then.  A flipped Y moves every sample
of its line by 6e-8 relative -- sixty times the parity tolerance -- so K1 hands every record
whose REPWID*gamma lies within 64 double ulps of a float rounding boundary to the host, which
re-evaluates it with the libm the reference calls (csrc/gas.cu: patch_near_ties).

The tests compare the float bits of REPWID and Y of every K1 record with the values formed
from the reference's own calc_line_centers / calc_lorentz_hw / calc_doppler_hw
(tests/k1_reference.py): on the bench's config-2 line lists (1.2e8 records) and on constructed
near-tie columns (tests/golden/near_ties.json) that do flip without the patch.
"""
import ctypes as C
import hashlib
import json
import os

import numpy as np
import pytest

from grtcode_b200 import synth
from tests import k1_reference as K

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def _reference():
    from oracle import ref as R
    if not R.available():
        pytest.skip("oracle/_ref not built")
    return R.Ref()


def _upload_order_params(api, hitran, mols, w0, wn):
    """Line parameters in upload order (add_molecule by add_molecule, file order inside),
    parsed by the host library's parse_hitran_file (bit-identical to the reference's parser:
    tests/test_host.py::test_hitran_parser_matches_oracle)."""
    from tests.test_host import _LineParams
    L = api.load()
    L.parse_hitran_file.argtypes = [C.POINTER(_LineParams), C.c_char_p, C.c_int, C.c_double,
                                    C.c_double]
    out = []
    for m in mols:
        lp = _LineParams()
        assert L.parse_hitran_file(C.byref(lp), hitran.encode(), m, w0, wn) == 0
        n = int(lp.num_lines)
        out.append({f: np.ctypeslib.as_array(getattr(lp, f), shape=(n,)).copy()
                    for f in ("vnn", "yair", "yself", "n", "d")} | {"mol": m, "count": n})
        L.free_line_params(C.byref(lp))
    return out


def _compare_records(api, ref, hitran, mols, w0, wn, dw, col, label):
    """One calculate_optical_depth; returns (records, REPWID mismatches, Y mismatches,
    records within 64 ulps of a tie, records patched)."""
    g = api.spectral_grid(w0, wn, dw)
    levels = len(col.p)
    go = api.GasOptics(levels, g, hitran, None, None, method=api.LINE_SAMPLE)
    for m in mols:
        go.add_molecule(m)
        go.set_ppmv("mol", m, col.ppmv[m])
    gas = api.OpticsObject(levels - 1, g)
    go.optical_depth(col.p, col.t, gas)
    patched = go.patch_counts()[0]
    params = _upload_order_params(api, hitran, mols, w0, wn)
    total = sum(p["count"] for p in params)
    order = go.debug_line_order(total).astype(np.int64)       # device position -> upload index
    want_rep = np.empty((levels - 1, total), dtype=np.float32)
    want_y = np.empty((levels - 1, total), dtype=np.float32)
    near = 0
    off = 0
    for p in params:
        pavg, tavg, ps = K.layer_quantities(ref, col.p, col.t, col.ppmv[p["mol"]])
        _, gamma, alpha = K.widths(ref, p["mol"], p["vnn"], p["d"], p["n"], p["yair"], p["yself"],
                                   pavg, tavg, ps)
        rep, y, yd = K.repwid_y(gamma, alpha)
        near += int((K.ulps_from_float_tie(yd) <= 64).sum())
        want_rep[:, off:off + p["count"]] = rep
        want_y[:, off:off + p["count"]] = y
        off += p["count"]
    bad_rep = bad_y = 0
    for layer in range(levels - 1):
        rep, y, _ = go.debug_records(layer, 0, total)
        bad_rep += int((rep.view(np.uint32) != want_rep[layer, order].view(np.uint32)).sum())
        bad_y += int((y.view(np.uint32) != want_y[layer, order].view(np.uint32)).sum())
    gas.close()
    go.close()
    n = total*(levels - 1)
    print(f"{label}: {n} records, REPWID mismatches {bad_rep}, Y mismatches {bad_y}, "
          f"{near} within 64 ulps of a float tie, {patched} patched on the host")
    return n, bad_rep, bad_y, near, patched


def test_k1_record_bits_match_the_reference_on_the_bench_line_lists(built, tmp_path):
    """Both bands of the headline workload (config 2: 8.83e5 + 1.412e6 lines x 54 layers =
    1.24e8 records): zero REPWID / Y mismatches."""
    from grtcode_b200 import workloads as W
    ref = _reference()
    col = W.circ1_column()
    lw, sw = W.circ1_bands(0.01, 1.)
    # K1 does not depend on the grid resolution: a coarse grid keeps tau small
    total = 0
    for band, seed in ((lw, 1), (sw, 2)):
        W.materialise(band, str(tmp_path), seed=seed)
        n, bad_rep, bad_y, near, patched = _compare_records(
            built, ref, band.dataset.hitran, W.MOLECULES, band.w0, band.wn, 1.0, col, band.name)
        assert bad_rep == 0 and bad_y == 0
        assert patched >= near          # every near-tie record inside the grid went to the host
        total += n
    assert total >= 100_000_000


def _near_tie_case():
    fx = json.load(open(os.path.join(HERE, "golden", "near_ties.json")))
    return fx


def test_constructed_near_ties_flip_without_the_patch_and_not_with_it(built, tmp_path):
    """Columns whose level temperatures put one line per layer within ~1 double ulp of a
    float tie of Y (tools/make_near_ties.py).  With GRT_K1_PATCH=0 some of them round the
    other way than the reference (libdevice pow vs the host libm); with the patch none."""
    ref = _reference()
    fx = _near_tie_case()
    c = fx["case"]
    lines = {int(k): v for k, v in c["lines"].items()}
    mols = tuple(lines)
    ds = synth.make_dataset(str(tmp_path), c["w0"], c["wn"], lines, seed=c["seed"])
    assert hashlib.sha256(open(ds.hitran, "rb").read()).hexdigest() == fx["hitran_sha256"]
    flips_unpatched = 0
    near_total = 0
    for entry in fx["columns"]:
        rng = np.random.default_rng(entry["seed"])
        col = synth.make_column(c["levels"], rng, molecules=mols)
        col.t = np.array([float.fromhex(h) for h in entry["t_hex"]])
        os.environ["GRT_K1_PATCH"] = "0"
        try:
            _, bad_rep, bad_y, near, patched = _compare_records(
                built, ref, ds.hitran, mols, c["w0"], c["wn"], c["dw"], col, "unpatched")
        finally:
            del os.environ["GRT_K1_PATCH"]
        assert bad_rep == 0 and patched == 0
        flips_unpatched += bad_y
        near_total += near
        _, bad_rep, bad_y, near, patched = _compare_records(
            built, ref, ds.hitran, mols, c["w0"], c["wn"], c["dw"], col, "patched")
        assert bad_rep == 0 and bad_y == 0
        assert patched >= near >= 50
    print(f"{near_total} near-tie records, {flips_unpatched} flipped without the patch")
    assert flips_unpatched >= 1


def test_a_flipped_y_is_visible_in_tau_and_the_patch_removes_it(built, tmp_path):
    """End to end on the first constructed column: optical depths against the reference with
    and without the patch (the 1e-9 tolerance north_star states)."""
    from tests import parity
    fx = _near_tie_case()
    c = fx["case"]
    lines = {int(k): v for k, v in c["lines"].items()}
    mols = tuple(lines)
    ds = synth.make_dataset(str(tmp_path), c["w0"], c["wn"], lines, seed=c["seed"])
    worst = {}
    for entry in fx["columns"][:3]:
        rng = np.random.default_rng(entry["seed"])
        col = synth.make_column(c["levels"], rng, molecules=mols)
        col.t = np.array([float.fromhex(h) for h in entry["t_hex"]])
        want = parity.oracle_column(ds, c["w0"], c["wn"], c["dw"], col, mols, use_ctm=False,
                                    solvers=False)
        for patch in ("0", "1"):
            os.environ["GRT_K1_PATCH"] = patch
            try:
                got = parity.gpu_column(ds, c["w0"], c["wn"], c["dw"], col, mols, use_ctm=False,
                                        solvers=False)
            finally:
                del os.environ["GRT_K1_PATCH"]
            rel = parity.relative_error(got["tau"], want["tau"])
            worst[patch] = max(worst.get(patch, 0.), float(rel.max()))
    print(f"tau max rel: unpatched {worst['0']:.3e}, patched {worst['1']:.3e}")
    assert worst["1"] <= 1e-9
