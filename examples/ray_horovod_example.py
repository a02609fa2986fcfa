"""Horovod-style training through HorovodRayAccelerator — command line of the reference's
examples/ray_horovod_example.py (--num-hosts, --num-slots, --use-gpu, --num-epochs, --smoke-test).

    python examples/ray_horovod_example.py --smoke-test
    python examples/ray_horovod_example.py --num-slots 8 --use-gpu
"""
import argparse
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from examples.mnist_module import MNISTClassifier  # noqa: E402
from ray_lightning_accelerators_b200 import HorovodRayAccelerator  # noqa: E402
from ray_lightning_accelerators_b200 import _ray as ray  # noqa: E402
from ray_lightning_accelerators_b200.mini_pl import Trainer  # noqa: E402


def train_mnist(config, num_epochs=10, num_hosts=1, num_slots=1, use_gpu=False, root=None):
    model = MNISTClassifier(config)
    trainer = Trainer(default_root_dir=root, max_epochs=num_epochs, gpus=int(use_gpu),
                      accelerator=HorovodRayAccelerator(num_hosts=num_hosts, num_slots=num_slots, use_gpu=use_gpu))
    trainer.fit(model)
    return trainer, model


if __name__ == "__main__":
    parser = argparse.ArgumentParser()
    parser.add_argument("--num-hosts", type=int, default=1)
    parser.add_argument("--num-slots", type=int, default=1)
    parser.add_argument("--use-gpu", action="store_true")
    parser.add_argument("--num-epochs", type=int, default=10)
    parser.add_argument("--smoke-test", action="store_true")
    args, _ = parser.parse_known_args()
    num_epochs = 1 if args.smoke_test else args.num_epochs
    num_slots = 1 if args.smoke_test else args.num_slots
    use_gpu = False if args.smoke_test else args.use_gpu
    try:
        config = {"layer_1": 32, "layer_2": 64, "lr": 1e-1, "batch_size": 32}
        train_mnist(config, num_epochs, args.num_hosts, num_slots, use_gpu, root=tempfile.mkdtemp(prefix="b2c_example_"))
        print("done")
    finally:
        ray.shutdown()
