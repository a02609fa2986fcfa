"""GPU end-to-end through the drop-in classes: real worker processes, one GPU each, gradients through
libb2c.so.  Restates ray_lightning/tests/test_ddp_gpu.py (train :32-47, device contract :83-103) and
test_horovod.py's GPU cases (:87-136); plus BASELINE config #5's shape: independent 2-GPU groups running
at the same time (4 GPUs)."""
import os
import threading

import pytest
import torch

from ray_lightning_accelerators_b200 import HorovodRayAccelerator, RayAccelerator
from ray_lightning_accelerators_b200 import _ray as ray
from ray_lightning_accelerators_b200.mini_pl import Callback

from .utils import BoringModel, get_trainer, train_test

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]


def _need(n):
    if torch.cuda.device_count() < n:
        pytest.skip(f"needs {n} GPUs, have {torch.cuda.device_count()}")


@pytest.fixture(autouse=True)
def _clean():
    yield
    ray.shutdown()


class _DeviceContract(Callback):
    """Inside every worker: the actor sees ONE GPU, uses index 0, parameters live on it, and the
    gradient path is the b2c library (its communicator exists and has launched kernels by the end)."""

    def on_epoch_end(self, trainer, pl_module):
        assert trainer.root_gpu == 0
        assert int(os.environ["CUDA_VISIBLE_DEVICES"]) == trainer.local_rank
        assert torch.cuda.current_device() == 0
        assert all(p.is_cuda for p in pl_module.parameters())
        comm = trainer.accelerator_backend.b2c_comm
        assert comm is not None and comm.stats()["calls_allreduce"] > 0 and comm.stats()["calls_broadcast"] > 0
        comm.check_alive()


def test_ray_accelerator_two_gpus_train_and_device_contract(tmp_path):
    _need(2)
    trainer = get_trainer(tmp_path, RayAccelerator(num_workers=2, use_gpu=True), use_gpu=True,
                          callbacks=[_DeviceContract()])
    train_test(trainer, BoringModel())
    trained = BoringModel.load_from_checkpoint(trainer.checkpoint_callback.best_model_path)
    assert trained.val_epoch == 1
    assert ray.live_actor_count() == 0


def test_ray_accelerator_two_gpus_with_fused_optimizer(tmp_path, monkeypatch):
    """B2C_FUSED_OPT=1: optimizer.step() is folded into the bucket launches on every worker (fused.FusedOptimizer);
    the reference's train contract (weights change, come back to the driver, checkpoint loads) still holds and
    the replicas end up bit-identical."""
    _need(2)
    monkeypatch.setenv("B2C_FUSED_OPT", "1")   # actors inherit the driver's environment

    class Check(Callback):
        def on_epoch_end(self, trainer, pl_module):
            assert type(trainer.optimizers[0]).__name__ == "FusedOptimizer"
            comm = trainer.accelerator_backend.b2c_comm
            assert comm.stats()["fused_steps"] > 0
            comm.check_alive()
            import torch.distributed as dist
            flat = torch.cat([p.detach().reshape(-1) for p in pl_module.parameters()])
            ref = flat.clone()
            dist.broadcast(ref, src=0)
            assert torch.equal(ref, flat), "replicas differ after fused steps"

    trainer = get_trainer(tmp_path, RayAccelerator(num_workers=2, use_gpu=True), use_gpu=True, callbacks=[Check()])
    train_test(trainer, BoringModel())


def test_horovod_accelerator_two_gpus(tmp_path):
    _need(2)
    trainer = get_trainer(tmp_path, HorovodRayAccelerator(num_hosts=1, num_slots=2, use_gpu=True), use_gpu=True)
    train_test(trainer, BoringModel())


def test_horovod_accelerator_four_gpus_buckets_in_multicast_arena(tmp_path):
    """From 4 slots up the shim's gradient buckets are allocated inside the NVLink-multicast arena."""
    _need(4)

    class Check(Callback):
        def on_epoch_end(self, trainer, pl_module):
            from ray_lightning_accelerators_b200 import hvd_shim
            comm = hvd_shim._st["comm"]
            assert comm.nvls_enabled, comm.nvls_error
            assert all(comm.in_nvls_arena(p.grad) for p in pl_module.parameters())
            comm.check_alive()

    trainer = get_trainer(tmp_path, HorovodRayAccelerator(num_hosts=1, num_slots=4, use_gpu=True), use_gpu=True,
                          callbacks=[Check()])
    train_test(trainer, BoringModel())


def test_gpu_runs_match_cpu_gloo_reference(tmp_path):
    """Same seed, same data: the weights the driver gets back from the 2-GPU b2c run equal the ones
    from the reference's CPU configuration (gloo) to fp32 round-off (Linear(32,2), 10 SGD steps)."""
    _need(2)
    outs = {}
    for name, use_gpu in (("cpu", False), ("gpu", True)):
        os.environ["PL_GLOBAL_SEED"] = "0"
        torch.manual_seed(0)
        model = BoringModel()
        trainer = get_trainer(tmp_path / name, RayAccelerator(num_workers=2, use_gpu=use_gpu), use_gpu=use_gpu)
        trainer.fit(model)
        outs[name] = torch.cat([p.detach().reshape(-1) for p in model.parameters()])
    os.environ.pop("PL_GLOBAL_SEED", None)
    assert torch.allclose(outs["cpu"], outs["gpu"], rtol=1e-5, atol=1e-6), (outs["cpu"] - outs["gpu"]).abs().max()


def test_two_concurrent_two_gpu_groups(tmp_path):
    """Tune-style: independent trials, each with its own 2-GPU group (disjoint GPUs, separate
    communicators), training at the same time."""
    _need(4)
    errs = []

    def trial(i):
        try:
            trainer = get_trainer(tmp_path / f"t{i}", RayAccelerator(num_workers=2, use_gpu=True), use_gpu=True,
                                  callbacks=[_TrialDevices()])
            train_test(trainer, BoringModel())
        except BaseException as e:  # noqa: BLE001
            errs.append(e)

    ths = [threading.Thread(target=trial, args=(i,)) for i in range(2)]
    for t in ths:
        t.start()
    for t in ths:
        t.join()
    assert not errs, errs


class _TrialDevices(Callback):
    def on_epoch_end(self, trainer, pl_module):
        assert torch.cuda.device_count() == 1 and torch.cuda.current_device() == 0
        trainer.accelerator_backend.b2c_comm.check_alive()
