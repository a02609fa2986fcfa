"""CPU: session.py and tune.py keep the reference's surface and error behaviour
(ray_lightning/session.py:17-63, ray_lightning/tune.py:68-101,128-142,188-199)."""
import os

import pytest
import torch

from ray_lightning_accelerators_b200 import RayAccelerator, session, tune
from ray_lightning_accelerators_b200 import _ray as ray
from ray_lightning_accelerators_b200.tune import (TuneReportCallback, TuneReportCheckpointCallback,
                                                    is_session_enabled)

from .utils import BoringModel, get_trainer


@pytest.fixture(autouse=True)
def _reset():
    session._reset_session_for_tests()
    yield
    session._reset_session_for_tests()
    ray.shutdown()


def test_session_errors_and_queue():
    with pytest.raises(ValueError, match="outside an Pytorch Lightning run"):
        session.get_actor_rank()
    session.init_session(rank=3, queue=None)
    assert session.get_actor_rank() == 3
    with pytest.raises(ValueError, match="twice"):
        session.init_session(rank=0, queue=None)
    with pytest.raises(ValueError, match="queue was not initialized"):
        session.put_queue("x")

    class Q:
        def __init__(self):
            self.items = []

        def put(self, item):
            self.items.append(item)

    q = Q()
    session.set_session_queue(q)
    session.put_queue("hello")
    assert q.items == [(3, "hello")]


def test_callback_argument_validation():
    with pytest.raises(ValueError, match="Invalid trigger time"):
        TuneReportCallback(on="never")
    cb = TuneReportCallback("loss")
    assert cb._metrics == ["loss"] and cb._on == ["validation_end"]
    assert not is_session_enabled()


def test_report_dict_shapes():
    class T:
        running_sanity_check = False
        callback_metrics = {"val_loss": torch.tensor(0.5), "acc": torch.tensor(0.9)}

    assert TuneReportCallback()._get_report_dict(T, None) == {"val_loss": 0.5, "acc": pytest.approx(0.9)}
    assert TuneReportCallback(["acc"])._get_report_dict(T, None) == {"acc": pytest.approx(0.9)}
    assert TuneReportCallback({"loss": "val_loss"})._get_report_dict(T, None) == {"loss": 0.5}
    T.running_sanity_check = True
    assert TuneReportCallback()._get_report_dict(T, None) is None


def test_tune_report_and_checkpoint_through_the_queue(tmp_path):
    """Inside a trial session the accelerator creates the queue, rank 0's closures run in the
    DRIVER process: training_iteration == max_epochs (test_tune.py:33-58) and the checkpoint
    exists on disk (test_tune.py:61-88)."""
    max_epochs = 2
    with tune.local_tune.session(str(tmp_path / "trial")) as trial:
        assert is_session_enabled()
        cbs = [TuneReportCheckpointCallback(metrics={"loss": "val_loss"}, filename="ckpt", on="validation_end")]
        trainer = get_trainer(tmp_path, RayAccelerator(num_workers=2), max_epochs=max_epochs, callbacks=cbs)
        trainer.fit(BoringModel())
    assert [r["training_iteration"] for r in trial["reports"]] == list(range(1, max_epochs + 1))
    assert all(r["loss"] == 1.0 for r in trial["reports"])
    assert len(trial["checkpoints"]) == max_epochs
    ckpt = torch.load(os.path.join(trial["checkpoints"][-1], "ckpt"), weights_only=False)
    assert "state_dict" in ckpt and ckpt["val_epoch"] == max_epochs
