"""MNISTClassifier for the examples: the model of examples/ray_ddp_example.py:18-58 in the reference
(on top of ray.tune.examples.mnist_ptl_mini.LightningMNISTClassifier [upstream]) with a synthetic,
MNIST-shaped dataset — there is no network here to download MNIST from."""
import torch
import torch.nn.functional as F
from torch.utils.data import DataLoader, TensorDataset

from ray_lightning_accelerators_b200.mini_pl import LightningModule


def synthetic_mnist(n, seed):
    """Separable synthetic digits: class k lights up rows 2k..2k+7, plus noise."""
    g = torch.Generator().manual_seed(seed)
    y = torch.randint(0, 10, (n,), generator=g)
    x = 0.3 * torch.rand(n, 1, 28, 28, generator=g)
    for k in range(10):
        x[y == k, :, 2 * k:2 * k + 8, :] += 0.7
    return TensorDataset(x, y)


class MNISTClassifier(LightningModule):
    def __init__(self, config, data_dir=None):
        super().__init__()
        self.data_dir = data_dir
        self.lr = config["lr"]
        self.batch_size = config["batch_size"]
        self.layer_1 = torch.nn.Linear(28 * 28, config["layer_1"])
        self.layer_2 = torch.nn.Linear(config["layer_1"], config["layer_2"])
        self.layer_3 = torch.nn.Linear(config["layer_2"], 10)

    def forward(self, x):
        x = x.view(x.size(0), -1)
        x = torch.relu(self.layer_1(x))
        x = torch.relu(self.layer_2(x))
        return F.log_softmax(self.layer_3(x), dim=1)

    def configure_optimizers(self):
        return torch.optim.Adam(self.parameters(), lr=self.lr)

    def training_step(self, batch, batch_idx):
        x, y = batch
        loss = F.nll_loss(self(x), y)
        self.log("ptl/train_loss", loss)
        return loss

    def validation_step(self, batch, batch_idx):
        x, y = batch
        logits = self(x)
        return {"val_loss": F.nll_loss(logits, y), "val_accuracy": (logits.argmax(1) == y).float().mean()}

    def validation_epoch_end(self, outputs):
        self.log("ptl/val_loss", torch.stack([o["val_loss"] for o in outputs]).mean())
        self.log("ptl/val_accuracy", torch.stack([o["val_accuracy"] for o in outputs]).mean())

    def train_dataloader(self):
        return DataLoader(synthetic_mnist(2048, 0), batch_size=self.batch_size, drop_last=True)

    def val_dataloader(self):
        return DataLoader(synthetic_mnist(512, 1), batch_size=self.batch_size, drop_last=True)
