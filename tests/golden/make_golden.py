"""Generates tests/golden/*.npz by running the REAL upstream path (torch.distributed gloo
all_reduce / broadcast / DDP — what RayAccelerator(use_gpu=False) resolves to,
ray_lightning/ray_ddp.py:227) in this container.  The reference package itself cannot be
imported (ray / pytorch_lightning / horovod are absent and not installable), and it
holds no golden vectors of its own, so these fixtures are the pin.

Run from the repo root:   python tests/golden/make_golden.py
Inputs are NOT stored: they are regenerated from seeds by oracle.oracle.payload().
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import gloo_ref  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def collective_jobs(world):
    jobs = []
    sizes = [1, 3, 17, 1000] + ([27882] if world in (2, 8) else [])
    for kind in ("int", "randn", "wide"):
        for n in sizes:
            jobs.append(dict(name=f"ar_f32_{kind}_{n}", op="allreduce", dtype="f32", payload=kind, n=n))
    for kind in ("int", "randn"):
        for n in (1, 9, 1000):
            jobs.append(dict(name=f"ar_bf16_{kind}_{n}", op="allreduce", dtype="bf16", payload=kind, n=n))
    jobs.append(dict(name="ar_f32_special_24", op="allreduce", dtype="f32", payload="special", n=24))
    for n in (1, 15, 16, 17, 4099):
        jobs.append(dict(name=f"bc_{n}_root{(n % world)}", op="broadcast", n=n, root=n % world))
    return jobs


def main():
    meta = {"torch": torch.__version__, "numpy": np.__version__, "backend": "gloo"}
    for world in (2, 3, 4, 8):
        out = gloo_ref.run_collectives(world, collective_jobs(world))
        np.savez_compressed(os.path.join(HERE, f"gloo_collectives_w{world}.npz"), **out)
        print("world", world, len(out), "vectors")
    # config #1: MNIST MLP, W=2, 100 steps, loss curve + final params
    r = gloo_ref.run_ddp(2, "mnist_mlp", steps=100, batch=32, return_grads=True)
    np.savez_compressed(os.path.join(HERE, "gloo_ddp_mnist_mlp_w2.npz"),
                        losses=np.array(r["losses"], np.float64), params=r["params"], last_grads=r["grads"])
    r = gloo_ref.run_ddp(2, "boring", steps=10, batch=1, return_grads=True)
    np.savez_compressed(os.path.join(HERE, "gloo_ddp_boring_w2.npz"),
                        losses=np.array(r["losses"], np.float64), params=r["params"], last_grads=r["grads"])
    with open(os.path.join(HERE, "META.json"), "w") as f:
        json.dump(meta, f, indent=1)


if __name__ == "__main__":
    main()
