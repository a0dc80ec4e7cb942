"""A C99 caller written against the reference's public headers, compiled with this
repository's include/ and linked with libgrtcode.so (tests/c/test_api_dropin.c): the
source-level drop-in.  The host section (HOST_ONLY refusals, grids, error text) runs
anywhere; the GPU section replays the reference's own unit-test call sequences."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "c", "test_api_dropin.c")


@pytest.fixture(scope="module")
def binary(built, tmp_path_factory):
    out = str(tmp_path_factory.mktemp("cdrop")/"test_api_dropin")
    lib = os.path.join(ROOT, "grtcode_b200")
    cmd = ["gcc", "-std=gnu99", "-O1", "-Wall", "-Wextra", "-Werror", "-I", os.path.join(ROOT, "include"),
           "-o", out, SRC, "-L", lib, "-lgrtcode", f"-Wl,-rpath,{lib}", "-lm"]
    proc = subprocess.run(cmd, capture_output=True, text=True)
    assert proc.returncode == 0, proc.stderr
    return out


def test_c_caller_compiles_and_host_section_passes(binary):
    proc = subprocess.run([binary, "host"], capture_output=True, text=True, timeout=120)
    assert proc.returncode == 0, proc.stdout + proc.stderr
    assert "0 failed expectations" in proc.stdout


@pytest.mark.gpu
def test_c_caller_gpu_sections(binary, tmp_path):
    from grtcode_b200 import synth
    ds = synth.make_dataset(str(tmp_path), 500., 700., {synth.H2O: 400, synth.CO2: 500}, seed=5,
                            cfcs=("CFC11",), cias=((0, 0),))
    files = [ds.hitran, ds.h2o_ctm_dir, ds.o3_ctm, ds.cfc[synth.CFC_IDS["CFC11"]],
             ds.cia[(0, 0)], ds.solar]
    proc = subprocess.run([binary, "gpu", *files], capture_output=True, text=True, timeout=600)
    assert proc.returncode == 0, proc.stdout + proc.stderr
    assert "0 failed expectations" in proc.stdout
