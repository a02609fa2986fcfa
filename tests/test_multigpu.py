"""Cross-process path over real NVLink: one process per GPU (torchrun), CUDA-IPC peer mapping,
parity against NCCL on the same tensors.  Needs >= 2 GPUs; on the one-GPU box the same kernels
and protocol are covered by tests/test_kernels_gpu.py through the local group."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(world, hide):
    env = dict(os.environ, B2C_TEST_HIDE_DEVICES="1" if hide else "0")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(29600 + world + (10 if hide else 0)),
           os.path.join(ROOT, "tests", "_mp_worker.py")]
    p = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0 and "MP_WORKER_OK" in p.stdout, p.stdout[-3000:] + p.stderr[-6000:]


def _need(n):
    if torch.cuda.device_count() < n:
        pytest.skip(f"needs {n} GPUs, have {torch.cuda.device_count()}")


@pytest.mark.parametrize("world", [2, 4, 8])
def test_ipc_path_matches_nccl(world):
    _need(world)
    _run(world, hide=False)


def test_ipc_path_with_ray_style_hidden_devices():
    """Each worker sees ONE GPU (CUDA_VISIBLE_DEVICES=<k>, device index 0) as a Ray actor
    does; peers are mapped through CUDA IPC although they cannot be enumerated."""
    _need(2)
    _run(2, hide=True)
