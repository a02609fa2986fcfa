"""Tune reporting callbacks for training that runs in remote workers.

Same public surface as ray_lightning/tune.py: TuneReportCallback(metrics=None,
on="validation_end") (tune.py:68-75), TuneReportCheckpointCallback(metrics=None,
filename="checkpoint", on="validation_end") (tune.py:188-195), TUNE_INSTALLED,
is_session_enabled (tune.py:10-22).  Rank 0 only; the closure is shipped through the session
queue and executed in the driver process (tune.py:97-101,128-142).  Control plane only: no
arithmetic, nothing here touches the GPU path.

ray.tune is absent from this image, so `tune` resolves to a minimal in-process stand-in
(`local_tune`) with report() / checkpoint_dir() / is_session_enabled(); with real Ray Tune
installed the import below picks it up instead.
"""
import contextlib
import functools
import os
from typing import Dict, List, Union

import torch

from .session import get_actor_rank, put_queue

try:  # pragma: no cover - ray is not installable here
    from ray import tune
    from ray.tune import is_session_enabled
    TUNE_INSTALLED = True
except ImportError:
    TUNE_INSTALLED = False

    class _LocalTune:
        """What a Tune trial offers to the function running inside it."""

        def __init__(self):
            self._session = None

        @contextlib.contextmanager
        def session(self, trial_dir):
            self._session = {"dir": trial_dir, "reports": [], "checkpoints": []}
            try:
                yield self._session
            finally:
                self._session = None

        def report(self, **metrics):
            if self._session is None:
                raise RuntimeError("tune.report called outside a trial session")
            it = len(self._session["reports"]) + 1
            self._session["reports"].append(dict(metrics, training_iteration=it))

        @contextlib.contextmanager
        def checkpoint_dir(self, step):
            if self._session is None:
                raise RuntimeError("tune.checkpoint_dir called outside a trial session")
            d = os.path.join(self._session["dir"], f"checkpoint_{step:06d}")
            os.makedirs(d, exist_ok=True)
            yield d
            self._session["checkpoints"].append(d)

        def is_session_enabled(self):
            return self._session is not None

    tune = local_tune = _LocalTune()

    def is_session_enabled():
        return tune.is_session_enabled()


def _tune_report(report):
    """Module-level on purpose: the queued item is pickled BY REFERENCE to this function, so it
    runs against the driver's `tune` (the live trial session), not a copy of the worker's."""
    tune.report(**report)


def _tune_checkpoint(checkpoint_dict, global_step, filename):
    with tune.checkpoint_dir(step=global_step) as checkpoint_dir:
        atomic_save(checkpoint_dict, os.path.join(checkpoint_dir, filename))


def atomic_save(checkpoint, filepath):
    tmp = filepath + ".part"
    torch.save(checkpoint, tmp)
    os.replace(tmp, filepath)


class TuneCallback:
    """Dispatches the trainer events named in `on` to _handle (what
    ray.tune.integration.pytorch_lightning.TuneCallback does upstream)."""
    _allowed = ["init_start", "init_end", "fit_start", "fit_end", "sanity_check_start", "sanity_check_end",
                "epoch_start", "epoch_end", "batch_start", "validation_batch_start", "validation_batch_end",
                "test_batch_start", "test_batch_end", "batch_end", "train_start", "train_end",
                "validation_start", "validation_end", "test_start", "test_end", "keyboard_interrupt"]

    def __init__(self, on: Union[str, List[str]] = "validation_end"):
        if not isinstance(on, list):
            on = [on]
        if any(w not in self._allowed for w in on):
            raise ValueError(f"Invalid trigger time selected: {on}. Must be one of {self._allowed}")
        self._on = on

    def _handle(self, trainer, pl_module):
        raise NotImplementedError

    def on_event(self, event, trainer, pl_module):
        if event in self._on:
            self._handle(trainer, pl_module)


class TuneReportCallback(TuneCallback):
    """Report metrics to Tune from rank 0 of a remote training run.

    metrics: None (everything in trainer.callback_metrics), a name, a list of names, or a
    {tune_name: trainer_name} dict."""

    def __init__(self, metrics: Union[None, str, List[str], Dict[str, str]] = None,
                 on: Union[str, List[str]] = "validation_end"):
        super().__init__(on)
        if isinstance(metrics, str):
            metrics = [metrics]
        self._metrics = metrics

    def _get_report_dict(self, trainer, pl_module):
        if getattr(trainer, "running_sanity_check", False):
            return None
        if not self._metrics:
            return {k: v.item() for k, v in trainer.callback_metrics.items()}
        report = {}
        for key in self._metrics:
            source = self._metrics[key] if isinstance(self._metrics, dict) else key
            report[key] = trainer.callback_metrics[source].item()
        return report

    def _handle(self, trainer, pl_module):
        if get_actor_rank() == 0:
            report = self._get_report_dict(trainer, pl_module)
            if report is not None:
                put_queue(functools.partial(_tune_report, report))


class _TuneCheckpointCallback(TuneCallback):
    """Save a checkpoint into the trial's checkpoint dir (executed on the driver)."""

    def __init__(self, filename: str = "checkpoint", on: Union[str, List[str]] = "validation_end"):
        super().__init__(on)
        self._filename = filename

    @staticmethod
    def _create_checkpoint(checkpoint_dict: dict, global_step: int, filename: str):
        _tune_checkpoint(checkpoint_dict, global_step, filename)

    def _handle(self, trainer, pl_module):
        if getattr(trainer, "running_sanity_check", False):
            return
        checkpoint_dict = trainer.checkpoint_connector.dump_checkpoint()
        global_step = trainer.global_step
        if get_actor_rank() == 0:
            put_queue(functools.partial(_tune_checkpoint, checkpoint_dict, global_step, self._filename))


class TuneReportCheckpointCallback(TuneCallback):
    """Checkpoint, then report (Tune registers a checkpoint only on the next report)."""

    def __init__(self, metrics: Union[None, str, List[str], Dict[str, str]] = None,
                 filename: str = "checkpoint", on: Union[str, List[str]] = "validation_end"):
        super().__init__(on)
        self._checkpoint = _TuneCheckpointCallback(filename, on)
        self._report = TuneReportCallback(metrics, on)

    def _handle(self, trainer, pl_module):
        self._checkpoint._handle(trainer, pl_module)
        self._report._handle(trainer, pl_module)
