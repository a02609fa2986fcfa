"""Test helpers restating the behavioural contract of ray_lightning/tests/utils.py:
BoringModel (:24-91), get_trainer (:94-114), train_test (:117-126), load_test (:129-134)."""
import torch
from torch.utils.data import DataLoader, Dataset

from ray_lightning_accelerators_b200.mini_pl import LightningModule, Trainer


class RandomDataset(Dataset):
    def __init__(self, size, length, seed=0):
        self.len = length
        self.data = torch.randn(length, size, generator=torch.Generator().manual_seed(seed))

    def __getitem__(self, index):
        return self.data[index]

    def __len__(self):
        return self.len


class BoringModel(LightningModule):
    def __init__(self):
        super().__init__()
        self.layer = torch.nn.Linear(32, 2)
        self.val_epoch = 0

    def forward(self, x):
        return self.layer(x)

    def loss(self, batch, prediction):
        return torch.nn.functional.mse_loss(prediction, torch.ones_like(prediction))

    def training_step(self, batch, batch_idx):
        return {"loss": self.loss(batch, self.layer(batch))}

    def validation_step(self, batch, batch_idx):
        self.layer(batch)
        loss = torch.tensor(1.0)
        self.log("val_loss", loss)
        return {"x": loss}

    def validation_epoch_end(self, outputs):
        torch.stack([x["x"] for x in outputs]).mean()
        self.val_epoch += 1

    def configure_optimizers(self):
        optimizer = torch.optim.SGD(self.layer.parameters(), lr=0.1)
        lr_scheduler = torch.optim.lr_scheduler.StepLR(optimizer, step_size=1)
        return [optimizer], [lr_scheduler]

    def train_dataloader(self):
        return DataLoader(RandomDataset(32, 64))

    def val_dataloader(self):
        return DataLoader(RandomDataset(32, 64, seed=1))

    def on_save_checkpoint(self, checkpoint):
        checkpoint["val_epoch"] = self.val_epoch

    def on_load_checkpoint(self, checkpoint):
        self.val_epoch = checkpoint["val_epoch"]


def get_trainer(dir, accelerator, use_gpu=False, max_epochs=1, limit_train_batches=10, limit_val_batches=10,
                progress_bar_refresh_rate=0, callbacks=None):
    return Trainer(default_root_dir=str(dir), gpus=1 if use_gpu else 0, max_epochs=max_epochs,
                   limit_train_batches=limit_train_batches, limit_val_batches=limit_val_batches,
                   progress_bar_refresh_rate=progress_bar_refresh_rate, checkpoint_callback=True,
                   callbacks=callbacks or [], accelerator=accelerator)


def train_test(trainer, model):
    """fit() == 1 and the DRIVER's copy of the weights moved by > 0.1: rank 0's weights really
    came back (ray_ddp.py:185-189)."""
    initial = torch.tensor([float(torch.sum(torch.abs(x.detach()))) for x in model.parameters()])
    result = trainer.fit(model)
    post = torch.tensor([float(torch.sum(torch.abs(x.detach()))) for x in model.parameters()])
    assert result == 1, "trainer failed"
    assert torch.norm(initial - post) > 0.1


def load_test(trainer, model):
    trainer.fit(model)
    trained = BoringModel.load_from_checkpoint(trainer.checkpoint_callback.best_model_path)
    assert trained is not None, "loading model failed"
    return trained
