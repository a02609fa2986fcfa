"""DDP training through RayAccelerator — same command line as the reference's
examples/ray_ddp_example.py:121-171 (--num-workers, --use-gpu, --tune, --num-samples, --num-epochs,
--smoke-test), synthetic MNIST-shaped data.

    python examples/ray_ddp_example.py --smoke-test
    python examples/ray_ddp_example.py --num-workers 8 --use-gpu        # gradients through libb2c.so
"""
import argparse
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from examples.mnist_module import MNISTClassifier  # noqa: E402
from ray_lightning_accelerators_b200 import RayAccelerator, tune as rl_tune  # noqa: E402
from ray_lightning_accelerators_b200 import _ray as ray  # noqa: E402
from ray_lightning_accelerators_b200.mini_pl import Trainer  # noqa: E402
from ray_lightning_accelerators_b200.tune import TuneReportCallback  # noqa: E402


def train_mnist(config, data_dir=None, num_epochs=10, num_workers=1, use_gpu=False, callbacks=None, root=None):
    model = MNISTClassifier(config, data_dir)
    trainer = Trainer(default_root_dir=root, max_epochs=num_epochs, gpus=int(use_gpu), callbacks=callbacks or [],
                      accelerator=RayAccelerator(num_workers=num_workers, use_gpu=use_gpu))
    trainer.fit(model)
    return trainer, model


def tune_mnist(data_dir, num_samples=10, num_epochs=10, num_workers=1, use_gpu=False, root=None):
    """The reference hands this loop to ray.tune.run (examples/ray_ddp_example.py:79-118); Ray Tune is
    not installed here, so the trials run one after the other inside the minimal trial session of
    ray_lightning_accelerators_b200.tune — the reporting path (worker -> queue -> driver) is the real one."""
    import random
    rng = random.Random(0)
    best = None
    for i in range(num_samples):
        config = {"layer_1": rng.choice([32, 64, 128]), "layer_2": rng.choice([64, 128, 256]),
                  "lr": 10 ** rng.uniform(-4, -1), "batch_size": rng.choice([32, 64, 128])}
        trial_dir = os.path.join(root or tempfile.gettempdir(), f"trial_{i}")
        with rl_tune.local_tune.session(trial_dir) as trial:
            metrics = {"loss": "ptl/val_loss", "acc": "ptl/val_accuracy"}
            train_mnist(config, data_dir, num_epochs, num_workers, use_gpu,
                        callbacks=[TuneReportCallback(metrics, on="validation_end")], root=trial_dir)
        last = trial["reports"][-1]
        print(f"trial {i}: {config} -> {last}")
        if best is None or last["loss"] < best[1]["loss"]:
            best = (config, last)
    print("Best hyperparameters found were: ", best[0])
    return best


if __name__ == "__main__":
    parser = argparse.ArgumentParser()
    parser.add_argument("--num-workers", type=int, default=1, help="Number of training workers to use.")
    parser.add_argument("--use-gpu", action="store_true", help="Use GPU for training.")
    parser.add_argument("--tune", action="store_true", help="Hyperparameter search over a few trials.")
    parser.add_argument("--num-samples", type=int, default=10, help="Number of samples to tune.")
    parser.add_argument("--num-epochs", type=int, default=10, help="Number of epochs to train for.")
    parser.add_argument("--smoke-test", action="store_true", help="Finish quickly for testing")
    parser.add_argument("--address", required=False, type=str, help="unused here (no Ray cluster)")
    args, _ = parser.parse_known_args()

    num_epochs = 1 if args.smoke_test else args.num_epochs
    num_workers = 1 if args.smoke_test else args.num_workers
    use_gpu = False if args.smoke_test else args.use_gpu
    num_samples = 1 if args.smoke_test else args.num_samples
    data_dir = os.path.join(tempfile.gettempdir(), "mnist_data_")
    root = tempfile.mkdtemp(prefix="b2c_example_")
    try:
        if args.tune:
            tune_mnist(data_dir, num_samples, num_epochs, num_workers, use_gpu, root=root)
        else:
            config = {"layer_1": 32, "layer_2": 64, "lr": 1e-1, "batch_size": 32}
            trainer, model = train_mnist(config, data_dir, num_epochs, num_workers, use_gpu, root=root)
            print("done; driver-side model updated from rank 0:", sum(p.numel() for p in model.parameters()), "parameters")
    finally:
        ray.shutdown()
