"""CPU: the example smoke runs of the reference's CI (.github/workflows/test.yaml:73-77)."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("cmd", [
    ["examples/ray_ddp_example.py", "--smoke-test"],
    ["examples/ray_ddp_example.py", "--smoke-test", "--tune"],
    ["examples/ray_horovod_example.py", "--smoke-test"],
    ["examples/ray_ddp_tune.py", "--smoke-test"],
])
def test_example_smoke(cmd):
    p = subprocess.run([sys.executable] + cmd, cwd=ROOT, capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-4000:]


def test_bench_reference_arm_contract():
    """bench.py --impl reference prints ONE JSON line with the contract's keys (tiny model here)."""
    import json
    p = subprocess.run([sys.executable, "bench.py", "--impl", "reference", "--model", "boring", "--ref-batch", "4",
                        "--steps", "2", "--warmup", "1"], cwd=ROOT, capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stderr[-3000:]
    line = json.loads(p.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["unit"] == "samples/s" and line["value"] > 0
    for k in ("metric", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "dtype", "data",
              "config", "cpu_baseline", "e2e"):
        assert k in line, k
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["cpu_baseline"]["kind"] == "reference"


def test_bench_reference_arm_under_torchrun():
    """The driver launches the reference arm like the GPU arm: under torch.distributed.run for N > 1.
    Rank 0 alone works (its gloo workers hold their own rendezvous, the elastic agent's environment must
    not leak into them); the other ranks exit 0; one JSON line, well inside a minute."""
    import json
    import time
    t0 = time.time()
    p = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29731", "bench.py", "--impl", "reference",
                        "--gpus", "2", "--model", "mnist_mlp", "--steps", "2", "--warmup", "1"],
                       cwd=ROOT, capture_output=True, text=True, timeout=120)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-4000:]
    lines = [ln for ln in p.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1, p.stdout
    line = json.loads(lines[0])
    assert line["impl"] == "reference" and line["n_gpus"] == 2 and line["value"] > 0
    assert "batch 128/worker" in line["config"]["workload"]
    assert time.time() - t0 < 60


def test_bench_gpu_arm_fails_loudly_without_cuda():
    import torch
    if torch.cuda.is_available():
        pytest.skip("this check is for the CPU box")
    p = subprocess.run([sys.executable, "bench.py", "--steps", "1", "--warmup", "1"], cwd=ROOT, capture_output=True,
                       text=True, timeout=300)
    assert p.returncode != 0 and "no CPU fallback" in (p.stderr + p.stdout)
