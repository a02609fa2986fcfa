"""The slice of PyTorch-Lightning 1.1.7 the reference's accelerators and tests lean on
(ray_lightning/tests/utils.py:94-114 get_trainer, :24-91 BoringModel), restated as a small
plain-PyTorch trainer.  pytorch_lightning is not installable in this image (and 1.1.7 predates
torch 2.x), so `Trainer.fit` -> accelerator.setup / train / teardown -> worker-side step loop is
provided here.  Nothing in this file is on the GPU data path; it only decides WHEN the
accelerator's collectives run.
"""
import os

import torch
import torch.distributed as dist
from torch.utils.data import DataLoader, DistributedSampler


class LightningModule(torch.nn.Module):
    """Model protocol: training_step / validation_step / configure_optimizers / *_dataloader."""

    def __init__(self):
        super().__init__()
        self.trainer = None

    def log(self, name, value, **_):
        if self.trainer is not None:
            if not torch.is_tensor(value):
                value = torch.tensor(float(value))
            self.trainer.callback_metrics[name] = value.detach()

    # optional hooks
    def training_step_end(self, out):
        return out

    def training_epoch_end(self, outputs):
        pass

    def validation_epoch_end(self, outputs):
        pass

    def on_save_checkpoint(self, checkpoint):
        pass

    def on_load_checkpoint(self, checkpoint):
        pass

    def val_dataloader(self):
        return None

    @classmethod
    def load_from_checkpoint(cls, path, *args, **kwargs):
        ckpt = torch.load(path, map_location="cpu", weights_only=False)
        model = cls(*args, **kwargs)
        model.load_state_dict(ckpt["state_dict"])
        model.on_load_checkpoint(ckpt)
        return model


class StepModule(torch.nn.Module):
    """What DDP wraps on a worker: forward(batch, batch_idx) == module.training_step(...), so
    one DDP.forward per optimisation step, exactly like PTL's LightningDistributedDataParallel."""

    def __init__(self, module):
        super().__init__()
        self.module = module

    def forward(self, batch, batch_idx):
        return self.module.training_step(batch, batch_idx)


class Callback:
    """Base class; override on_<event>(trainer, pl_module)."""


class EarlyStopping(Callback):
    def __init__(self, monitor="val_loss", patience=3, min_delta=0.0, mode="min"):
        self.monitor, self.patience, self.min_delta, self.mode = monitor, patience, min_delta, mode
        self.best, self.wait = None, 0

    def on_validation_end(self, trainer, pl_module):
        if trainer.running_sanity_check or self.monitor not in trainer.callback_metrics:
            return
        cur = float(trainer.callback_metrics[self.monitor])
        better = self.best is None or (cur < self.best - self.min_delta if self.mode == "min" else cur > self.best + self.min_delta)
        if better:
            self.best, self.wait = cur, 0
        else:
            self.wait += 1
            if self.wait >= self.patience:
                trainer.should_stop = True


class ModelCheckpoint(Callback):
    def __init__(self, dirpath=None):
        self.dirpath = dirpath
        self.best_model_path = ""

    def on_validation_end(self, trainer, pl_module):
        if trainer.running_sanity_check:
            return
        if trainer.global_rank == 0:
            d = self.dirpath or os.path.join(trainer.default_root_dir, "checkpoints")
            os.makedirs(d, exist_ok=True)
            path = os.path.join(d, f"epoch={trainer.current_epoch}.ckpt")
            torch.save(trainer.checkpoint_connector.dump_checkpoint(), path)
            self.best_model_path = path


class _CheckpointConnector:
    def __init__(self, trainer):
        self.trainer = trainer

    def dump_checkpoint(self):
        t = self.trainer
        module = t.get_model()
        ckpt = {"epoch": t.current_epoch, "global_step": t.global_step,
                "state_dict": {k: v.detach().cpu().clone() for k, v in module.state_dict().items()}}
        module.on_save_checkpoint(ckpt)
        return ckpt


class LightningDataModule:
    def setup(self, stage=None):
        pass

    def train_dataloader(self):
        raise NotImplementedError

    def val_dataloader(self):
        return None

    def test_dataloader(self):
        return None


class Trainer:
    """Trainer(default_root_dir, gpus, max_epochs, limit_train_batches, limit_val_batches,
    callbacks, accelerator, checkpoint_callback, progress_bar_refresh_rate) — the kwargs of
    ray_lightning/tests/utils.py:94-114.  fit() runs on the driver; run_training() on a worker."""

    def __init__(self, default_root_dir=None, gpus=0, max_epochs=1, limit_train_batches=None,
                 limit_val_batches=None, progress_bar_refresh_rate=0, checkpoint_callback=True,
                 callbacks=None, accelerator=None, max_steps=None, **_unused):
        self.default_root_dir = default_root_dir or os.getcwd()
        self.gpus = gpus
        self.on_gpu = bool(gpus)
        self.max_epochs = max_epochs
        self.max_steps = max_steps
        self.limit_train_batches = limit_train_batches
        self.limit_val_batches = limit_val_batches
        self.callbacks = list(callbacks or [])
        self.checkpoint_callback = ModelCheckpoint() if checkpoint_callback is True else (checkpoint_callback or None)
        self.accelerator_backend = accelerator
        if accelerator is not None:
            accelerator.trainer = self
        self.checkpoint_connector = _CheckpointConnector(self)
        self.model = None
        self.datamodule = None
        self.use_ddp = False
        self.use_horovod = False
        self.local_rank = 0
        self.global_rank = 0
        self.world_size = 1
        self.root_gpu = 0 if gpus else None
        self.callback_metrics = {}
        self.running_sanity_check = False
        self.should_stop = False
        self.global_step = 0
        self.current_epoch = 0
        self.train_losses = []

    # ------------------------------------------------------------ driver side
    def fit(self, model, datamodule=None):
        acc = self.accelerator_backend
        if acc is None:
            raise ValueError("this mini trainer needs accelerator=RayAccelerator(...) / HorovodRayAccelerator(...)")
        self.datamodule = datamodule
        acc.trainer = self
        acc.setup(model)
        try:
            self.results = acc.train()
        finally:
            acc.teardown()
        return 1

    def get_model(self):
        m = self.model
        while isinstance(m, (torch.nn.parallel.DistributedDataParallel, StepModule)) or getattr(m, "_is_b2c_data_parallel", False):
            m = m.module
        return m

    # ------------------------------------------------------------ worker side
    def _all_callbacks(self):
        cbs = list(self.callbacks)
        if self.checkpoint_callback is not None:
            cbs.append(self.checkpoint_callback)
        return cbs

    def _fire(self, event, module):
        for cb in self._all_callbacks():
            if hasattr(cb, "on_event"):
                cb.on_event(event, self, module)
            fn = getattr(cb, "on_" + event, None)
            if fn is not None:
                fn(self, module)

    def _loader(self, loader, shuffle):
        acc = self.accelerator_backend
        if loader is None or not getattr(acc, "require_distributed_sampler", False):
            return loader
        if isinstance(loader.sampler, DistributedSampler):
            return loader
        sampler = DistributedSampler(loader.dataset, shuffle=shuffle, **acc.distributed_sampler_kwargs)
        return DataLoader(loader.dataset, batch_size=loader.batch_size, sampler=sampler,
                          num_workers=loader.num_workers, collate_fn=loader.collate_fn,
                          pin_memory=loader.pin_memory, drop_last=loader.drop_last)

    def _to_device(self, batch, device):
        if torch.is_tensor(batch):
            return batch.to(device, non_blocking=True)
        if isinstance(batch, (list, tuple)):
            return type(batch)(self._to_device(b, device) for b in batch)
        if isinstance(batch, dict):
            return {k: self._to_device(v, device) for k, v in batch.items()}
        return batch

    def _sync_should_stop(self):
        """Early-stop flag agreed across workers on the stock torch process group (C9: stays
        a control collective, never routed through the gradient kernels)."""
        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
            flag = torch.tensor([1.0 if self.should_stop else 0.0], device=dev)
            dist.all_reduce(flag, op=dist.ReduceOp.MAX)
            self.should_stop = bool(flag.item() > 0)

    def run_training(self, wrapped, module, optimizers, schedulers, device, backward_fn=None, step_fn=None):
        """The step loop.  `wrapped` is what forward is called through (DDP or the bare module),
        `module` the LightningModule.  backward_fn / step_fn let the Horovod path synchronise."""
        module.trainer = self
        self.model = wrapped
        self.optimizers = optimizers
        src = self.datamodule if self.datamodule is not None else module
        if self.datamodule is not None:
            self.datamodule.setup("fit")
        train_loader = self._loader(src.train_dataloader(), shuffle=True)
        val_loader = self._loader(src.val_dataloader(), shuffle=False) if hasattr(src, "val_dataloader") else None
        self.train_dataloader, self.val_dataloaders = train_loader, [val_loader] if val_loader is not None else []
        self._fire("fit_start", module)
        self._fire("train_start", module)
        results = 1
        for epoch in range(self.max_epochs):
            self.current_epoch = epoch
            if isinstance(getattr(train_loader, "sampler", None), DistributedSampler):
                train_loader.sampler.set_epoch(epoch)
            wrapped.train()
            self._fire("epoch_start", module)
            outputs = []
            for batch_idx, batch in enumerate(train_loader):
                if self.limit_train_batches is not None and batch_idx >= self.limit_train_batches:
                    break
                batch = self._to_device(batch, device)
                self._fire("batch_start", module)
                for opt in optimizers:
                    opt.zero_grad(set_to_none=True)
                out = self._training_step(wrapped, module, batch, batch_idx)
                loss = out["loss"] if isinstance(out, dict) else out
                (backward_fn or (lambda l: l.backward()))(loss)
                for opt in optimizers:
                    (step_fn or (lambda o: o.step()))(opt)
                self.global_step += 1
                self.train_losses.append(loss.detach())
                outputs.append({"loss": loss.detach()})
                self._fire("batch_end", module)
                if self.max_steps is not None and self.global_step >= self.max_steps:
                    self.should_stop = True
                    break
            module.training_epoch_end(outputs)
            for sch in schedulers:
                sch.step()
            if val_loader is not None:
                self._run_validation(wrapped, module, val_loader, device)
            else:
                self._fire("validation_end", module)
            self._fire("epoch_end", module)
            self._sync_should_stop()
            if self.should_stop:
                break
        self._fire("train_end", module)
        self._fire("fit_end", module)
        return results

    def _training_step(self, wrapped, module, batch, batch_idx):
        # through the DDP wrapper (over a StepModule) so its pre-forward and gradient hooks fire —
        # PTL routes training_step through LightningDistributedDataParallel.forward the same way
        if wrapped is not module:
            return module.training_step_end(wrapped(batch, batch_idx))
        return module.training_step_end(module.training_step(batch, batch_idx))

    @torch.no_grad()
    def _run_validation(self, wrapped, module, loader, device):
        wrapped.eval()
        self._fire("validation_start", module)
        outs = []
        for batch_idx, batch in enumerate(loader):
            if self.limit_val_batches is not None and batch_idx >= self.limit_val_batches:
                break
            outs.append(module.validation_step(self._to_device(batch, device), batch_idx))
        module.validation_epoch_end(outs)
        self._fire("validation_end", module)
        wrapped.train()
