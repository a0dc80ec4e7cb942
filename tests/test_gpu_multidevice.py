"""Two GPUs driven from ONE process (api.py and create_device allow any device id): kernel
attributes such as the opt-in dynamic shared memory are per device, so the second device's
first launch must configure its own (csrc/gas.cu: launch_wing_impl).  Skipped on a one-GPU
box; gpurun --gpus 2 runs it."""
import numpy as np
import pytest

from grtcode_b200 import api, synth
from tests import parity

pytestmark = pytest.mark.gpu


def test_second_device_in_the_same_process(built, tmp_path):
    api.load()
    if api.num_gpus() < 2:
        pytest.skip("needs two GPUs in one process")
    lines = {synth.H2O: 200, synth.CO2: 300}
    w0, wn, levels = 600., 700., 12
    ds = parity.build_dataset(str(tmp_path), w0, wn, lines, 5)
    mols = tuple(lines)
    col = synth.make_column(levels, np.random.default_rng(5), molecules=mols)
    for dw in (0.1, 0.01):              # both wing-kernel variants (with the abscissa table)
        taus = []
        for device in (0, 1, 0):
            got = parity.gpu_column(ds, w0, wn, dw, col, mols, device=device)
            taus.append(got["tau"])
        rel = parity.relative_error(taus[1], taus[0])
        assert rel.max() <= 1e-12, (dw, rel.max())
        assert np.array_equal(taus[2] > 0, taus[0] > 0)
    want = parity.oracle_column(ds, w0, wn, 0.1, col, mols, solvers=False)
    got = parity.gpu_column(ds, w0, wn, 0.1, col, mols, solvers=False, device=1)
    assert parity.relative_error(got["tau"], want["tau"]).max() <= 1e-9
