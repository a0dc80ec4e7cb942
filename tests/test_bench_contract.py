"""The reference arm of bench.py (runs on the host cores, no GPU needed): JSON contract and
thread count."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(env_extra, *args):
    env = dict(os.environ)
    env.update(env_extra)
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference",
                          *args], env=env, capture_output=True, text=True, timeout=900)
    assert out.returncode == 0, out.stderr[-2000:]
    return out.stdout


def test_reference_arm_line_and_threads(built, tmp_path):
    """One bounded sample per band.  torchrun exports OMP_NUM_THREADS=1 to its workers; the
    arm must still time the reference on every host core and say how many it used."""
    from oracle import ref as R
    if not R.available():
        pytest.skip("compiled reference not built (no /root/reference on this machine)")
    text = _run({"OMP_NUM_THREADS": "1", "GRT_CPU_WINDOWS": "1",
                 "GRT_BENCH_DATA": str(tmp_path)}, "--steps", "1", "--warmup", "0")
    lines = [l for l in text.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["higher_is_better"] is True
    assert d["unit"] == "pts*layers/s" and d["value"] > 0.
    assert d["config"]["workload"].startswith("BASELINE config 2")
    assert d["cpu_baseline"]["kind"] == "reference"
    assert d["cpu_baseline"]["cores"] == len(os.sched_getaffinity(0))
    assert d["cpu_baseline"]["value"] == d["value"] == d["e2e"]["value"]
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0


def test_reference_arm_other_ranks_do_nothing(built):
    """Under torchrun only rank 0 runs the reference; the others exit 0 without output."""
    text = _run({"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"}, "--gpus", "2")
    assert [l for l in text.splitlines() if l.startswith("{")] == []


def test_product_arm_fails_loudly_without_a_gpu(built):
    """There is no CPU path in the product: without a CUDA device bench.py must stop with an
    error, not fall back to anything (skipped where a GPU is visible)."""
    from grtcode_b200 import api
    api.load()
    if api.num_gpus() > 0:
        pytest.skip("a GPU is visible here")
    proc = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "1",
                           "--warmup", "0"], capture_output=True, text=True, timeout=600)
    assert proc.returncode != 0
    assert "no CUDA device visible" in proc.stderr
    assert not [l for l in proc.stdout.splitlines() if l.startswith("{")]


def test_reference_arm_under_torchrun(built, tmp_path):
    """Launched exactly as the driver launches it for N > 1 (torchrun, one process per GPU):
    rank 0 alone prints the line, timed on every host core although torchrun exports
    OMP_NUM_THREADS=1; the other rank exits 0 without output."""
    from oracle import ref as R
    if not R.available():
        pytest.skip("compiled reference not built (no /root/reference on this machine)")
    env = dict(os.environ)
    env.update({"GRT_CPU_WINDOWS": "1", "GRT_BENCH_DATA": str(tmp_path)})
    env.pop("OMP_NUM_THREADS", None)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", "29541", os.path.join(ROOT, "bench.py"),
           "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0"]
    out = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=900)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["n_gpus"] == 2
    assert d["cpu_baseline"]["cores"] == len(os.sched_getaffinity(0))
