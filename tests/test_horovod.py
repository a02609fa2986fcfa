"""CPU (gloo backend of the hvd shim): the contract of ray_lightning/tests/test_horovod.py:51-84 —
train / load for num_slots in {1, 2} — plus Horovod's Average semantics against the oracle."""
import numpy as np
import pytest
import torch

from oracle import oracle as O
from ray_lightning_accelerators_b200 import HorovodRayAccelerator
from ray_lightning_accelerators_b200 import _ray as ray
from ray_lightning_accelerators_b200.ray_horovod import HorovodExecutor

from .utils import BoringModel, get_trainer, train_test


@pytest.fixture(autouse=True)
def _clean():
    yield
    ray.shutdown()


def test_constructor_surface():
    acc = HorovodRayAccelerator()
    assert (acc.num_hosts, acc.num_slots, acc.use_gpu, acc.nickname) == (1, 1, False, "horovod_ray")
    import pickle
    clone = pickle.loads(pickle.dumps(HorovodRayAccelerator(num_slots=4, use_gpu=False)))
    assert clone.num_slots == 4 and clone.executor is None
    with pytest.raises(RuntimeError, match="one NVSwitch domain"):
        HorovodExecutor(HorovodExecutor.create_settings(30), num_hosts=2, num_slots=1)


@pytest.mark.parametrize("num_slots", [1, 2])
def test_train_and_load(tmp_path, num_slots):
    trainer = get_trainer(tmp_path, HorovodRayAccelerator(num_hosts=1, num_slots=num_slots, use_gpu=False))
    train_test(trainer, BoringModel())
    trained = BoringModel.load_from_checkpoint(trainer.checkpoint_callback.best_model_path)
    assert trained.val_epoch == 1


def _avg_worker(_w):
    import torch
    from oracle import oracle as O
    from ray_lightning_accelerators_b200 import hvd_shim as hvd
    hvd.init()
    x = torch.from_numpy(O.payload("randn", 1000, hvd.rank()))
    out = hvd.allreduce(x, op=hvd.Average)
    p = torch.nn.Parameter(torch.full((5,), float(hvd.rank() + 3)))
    hvd.broadcast_parameters({"p": p}, root_rank=0)
    # DistributedOptimizer: gradients averaged before step, LR untouched by the optimizer wrapper
    w = torch.nn.Parameter(torch.zeros(4))
    opt = hvd.DistributedOptimizer(torch.optim.SGD([w], lr=1.0))
    opt.zero_grad()
    (w * float(hvd.rank() + 1)).sum().backward()
    opt.synchronize()
    with opt.skip_synchronize():
        opt.step()
    n = hvd.join()
    return out.numpy(), p.detach().numpy(), w.detach().numpy(), n


def test_shim_average_broadcast_optimizer():
    ex = HorovodExecutor(HorovodExecutor.create_settings(30), num_hosts=1, num_slots=2, use_gpu=False)
    ex.start()
    try:
        outs = ex.execute(_avg_worker)
    finally:
        ex.shutdown()
    want = O.hvd_average_f32([O.payload("randn", 1000, r) for r in range(2)])
    for avg, p, w, n in outs:
        assert np.array_equal(avg.view(np.uint32), want.view(np.uint32))   # W=2: bit-exact
        assert (p == 3.0).all()                                            # rank 0's value everywhere
        assert np.allclose(w, -1.5)                                        # mean of grads 1 and 2, lr 1
        assert n == 1
