"""CPU, world_size 2 over gloo: the plumbing contract of ray_lightning/tests/test_ddp.py
(config #1: RayAccelerator(num_workers=2, use_gpu=False))."""
import pytest
from torch.utils.data import DistributedSampler

from ray_lightning_accelerators_b200 import RayAccelerator
from ray_lightning_accelerators_b200 import _ray as ray
from ray_lightning_accelerators_b200.mini_pl import Callback, EarlyStopping

from .utils import BoringModel, get_trainer, train_test


@pytest.fixture(autouse=True)
def _clean_actors():
    yield
    ray.shutdown()


def test_constructor_surface():
    acc = RayAccelerator()
    assert (acc.num_workers, acc.num_cpus_per_worker, acc.use_gpu, acc.init_hook) == (1, 1, False, None)
    assert acc.nickname == "ddp_ray" and acc.workers == [] and acc.require_distributed_sampler
    acc = RayAccelerator(num_workers=4, cpus_per_worker=3, use_gpu=False)  # README.md:26 spelling
    assert acc.num_cpus_per_worker == 3
    import pickle
    acc.workers = ["handle"]
    clone = pickle.loads(pickle.dumps(acc))
    assert clone.workers == [] and clone.num_workers == 4


class _SamplerCheck(Callback):
    """loaders get DistributedSampler(num_replicas=2, rank=global_rank); shuffle only for train
    (test_ddp.py:45-79).  Runs inside every worker."""

    def on_train_start(self, trainer, pl_module):
        tr = trainer.train_dataloader.sampler
        va = trainer.val_dataloaders[0].sampler
        assert isinstance(tr, DistributedSampler) and isinstance(va, DistributedSampler)
        assert tr.num_replicas == 2 and va.num_replicas == 2
        assert tr.rank == trainer.global_rank and va.rank == trainer.global_rank
        assert tr.shuffle and not va.shuffle


def test_two_workers_train_load_sampler_actors(tmp_path):
    """One fit, four contracts of the reference's test_ddp.py: exactly num_workers actors
    alive during training and none after (:29-42), distributed samplers (:45-79), the driver's
    weights moved (:82-96, utils.py:117-126), the returned checkpoint loads (:99-107)."""
    seen = []
    acc = RayAccelerator(num_workers=2)
    orig = acc.train

    def counting_train():
        seen.append(ray.live_actor_count())
        return orig()

    acc.train = counting_train
    trainer = get_trainer(tmp_path, acc, callbacks=[_SamplerCheck()])
    train_test(trainer, BoringModel())
    assert seen == [2]
    assert ray.live_actor_count() == 0
    trained = BoringModel.load_from_checkpoint(trainer.checkpoint_callback.best_model_path)
    assert trained.val_epoch == 1


def test_train_single_worker(tmp_path):
    train_test(get_trainer(tmp_path, RayAccelerator(num_workers=1)), BoringModel())


def test_predict(tmp_path):
    """predict_test of the reference (tests/utils.py:137-152, test_ddp.py:99-115): after one short epoch on
    2 workers the model the DRIVER holds classifies the test split with accuracy >= 0.5.  MNIST cannot be
    downloaded here; the synthetic, MNIST-shaped digits of examples/mnist_module.py stand in."""
    import torch
    from torch.utils.data import DataLoader

    from examples.mnist_module import MNISTClassifier, synthetic_mnist
    torch.manual_seed(0)
    model = MNISTClassifier({"layer_1": 32, "layer_2": 32, "lr": 1e-2, "batch_size": 32})
    trainer = get_trainer(tmp_path, RayAccelerator(num_workers=2), limit_train_batches=10, limit_val_batches=2)
    trainer.fit(model)
    correct = total = 0
    with torch.no_grad():
        for x, y in DataLoader(synthetic_mnist(512, 7), batch_size=64):
            correct += int((model(x).argmax(1) == y).sum())
            total += len(y)
    assert correct / total >= 0.5, f"accuracy {correct / total:.3f}"


def test_early_stop(tmp_path):
    """constant val_loss + patience 2 stops a 500-epoch run early, on every worker at the same
    epoch (the stop flag is all-reduced on the control group, C9).  The reference pins
    val_epoch == 2 (test_ddp.py:118-134), a count that comes out of PTL 1.1.7's own callback
    ordering [upstream]; this trainer stops after the first (best) validation plus exactly
    `patience` non-improving ones = 3."""
    acc = RayAccelerator(num_workers=2)
    trainer = get_trainer(tmp_path, acc, max_epochs=500, limit_train_batches=4,
                          callbacks=[EarlyStopping(monitor="val_loss", patience=2)])
    trainer.fit(BoringModel())
    trained = BoringModel.load_from_checkpoint(trainer.checkpoint_callback.best_model_path)
    assert trained.val_epoch == 3, trained.val_epoch


def test_worker_error_surfaces(tmp_path):
    class Boom(Callback):
        def on_train_start(self, trainer, pl_module):
            raise RuntimeError("boom in worker")

    trainer = get_trainer(tmp_path, RayAccelerator(num_workers=2), callbacks=[Boom()])
    with pytest.raises(Exception, match="boom in worker"):
        trainer.fit(BoringModel())
