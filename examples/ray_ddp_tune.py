"""DDP + Tune-style search + LightningDataModule — the shape of the reference's examples/ray_ddp_tune.py
(--num-workers, --use-gpu, --num-samples, --num-epochs, --smoke-test): every trial trains through its own
RayAccelerator group and reports / checkpoints through TuneReportCheckpointCallback."""
import argparse
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from torch.utils.data import DataLoader  # noqa: E402

from examples.mnist_module import MNISTClassifier, synthetic_mnist  # noqa: E402
from ray_lightning_accelerators_b200 import RayAccelerator, tune as rl_tune  # noqa: E402
from ray_lightning_accelerators_b200 import _ray as ray  # noqa: E402
from ray_lightning_accelerators_b200.mini_pl import LightningDataModule, Trainer  # noqa: E402
from ray_lightning_accelerators_b200.tune import TuneReportCheckpointCallback  # noqa: E402


class MNISTDataModule(LightningDataModule):
    def __init__(self, batch_size=32):
        self.batch_size = batch_size

    def setup(self, stage=None):
        self.train, self.val = synthetic_mnist(2048, 0), synthetic_mnist(512, 1)

    def train_dataloader(self):
        return DataLoader(self.train, batch_size=self.batch_size, drop_last=True)

    def val_dataloader(self):
        return DataLoader(self.val, batch_size=self.batch_size, drop_last=True)


def train_mnist(config, num_epochs, num_workers, use_gpu, trial_dir):
    model = MNISTClassifier(config)
    dm = MNISTDataModule(config["batch_size"])
    cb = TuneReportCheckpointCallback({"loss": "ptl/val_loss", "acc": "ptl/val_accuracy"}, filename="checkpoint",
                                      on="validation_end")
    trainer = Trainer(default_root_dir=trial_dir, max_epochs=num_epochs, gpus=int(use_gpu), callbacks=[cb],
                      accelerator=RayAccelerator(num_workers=num_workers, use_gpu=use_gpu))
    trainer.fit(model, datamodule=dm)


def tune_mnist(num_samples, num_epochs, num_workers, use_gpu, root):
    import random
    rng = random.Random(1)
    results = []
    for i in range(num_samples):
        config = {"layer_1": rng.choice([32, 64, 128]), "layer_2": rng.choice([64, 128, 256]),
                  "lr": 10 ** rng.uniform(-4, -1), "batch_size": rng.choice([32, 64, 128])}
        trial_dir = os.path.join(root, f"trial_{i}")
        with rl_tune.local_tune.session(trial_dir) as trial:
            train_mnist(config, num_epochs, num_workers, use_gpu, trial_dir)
        assert trial["reports"][-1]["training_iteration"] == num_epochs
        assert os.path.exists(os.path.join(trial["checkpoints"][-1], "checkpoint"))
        results.append((trial["reports"][-1]["loss"], config))
    print("Best hyperparameters found were: ", min(results, key=lambda r: r[0])[1])


if __name__ == "__main__":
    parser = argparse.ArgumentParser()
    parser.add_argument("--num-workers", type=int, default=1)
    parser.add_argument("--use-gpu", action="store_true")
    parser.add_argument("--num-samples", type=int, default=10)
    parser.add_argument("--num-epochs", type=int, default=10)
    parser.add_argument("--smoke-test", action="store_true")
    args, _ = parser.parse_known_args()
    try:
        if args.smoke_test:
            tune_mnist(1, 1, 1, False, tempfile.mkdtemp(prefix="b2c_tune_"))
        else:
            tune_mnist(args.num_samples, args.num_epochs, args.num_workers, args.use_gpu, tempfile.mkdtemp(prefix="b2c_tune_"))
    finally:
        ray.shutdown()
