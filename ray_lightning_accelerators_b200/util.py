"""Driver-side plumbing (queue + result pump) and the NEW GPU-pinning / IPC-handle helpers.

Mirrors ray_lightning/util.py:9-13 (Unavailable), :71-85 (Queue), :88-93 (_handle_queue) and
:96-109 (process_results).  The reference's util.py contains no GPU pinning; the helpers at
the bottom are new (BASELINE.json north_star: "util.py GPU-to-actor pinning").
"""
import os
from typing import Callable, Dict, Optional

from . import _ray as ray


class Unavailable:
    """Placeholder for optional classes; must never be instantiated."""

    def __init__(self, *args, **kwargs):
        raise RuntimeError("This class should never be instantiated.")


class Queue(ray.Queue):
    """Driver<->worker queue (the reference creates it with actor_options={"num_cpus": 0})."""

    def __init__(self, maxsize: int = 0, actor_options: Optional[Dict] = None) -> None:
        super().__init__(maxsize=maxsize, actor_options=actor_options or {})


def _handle_queue(queue) -> None:
    """Run every callable the workers have queued — in THIS (the driver / Tune trial) process,
    which is the only place tune.report works."""
    while not queue.empty():
        (actor_rank, item) = queue.get()
        if isinstance(item, Callable):
            item()


def process_results(training_result_futures, queue, poll_s: float = 0.01):
    """Drain the queue while the workers train, surface worker exceptions as soon as a future
    fails, return every worker's result.  The reference spins on ray.wait(timeout=0); a short
    poll keeps the semantics and frees the core for the data loaders."""
    not_ready = list(training_result_futures)
    while not_ready:
        if queue:
            _handle_queue(queue)
        ready, not_ready = ray.wait(not_ready, num_returns=len(not_ready), timeout=poll_s)
        ray.get(ready)  # re-raises a worker failure immediately
    if queue:
        _handle_queue(queue)
    return ray.get(list(training_result_futures))


# ------------------------------------------------------------------ new: GPU pinning + handle exchange
def pin_gpu(use_gpu: bool, local_rank: int = 0) -> Optional[int]:
    """Select this worker's CUDA device and return its index.

    Ray-style actors see one GPU (CUDA_VISIBLE_DEVICES=<k>) and use index 0 — the contract
    ray_lightning/tests/test_ddp_gpu.py:83-103 asserts (root_gpu == 0, current_device() == 0).
    Horovod-on-Ray workers see every node GPU and use their local rank
    (ray_lightning/ray_horovod.py:153-155).  Peers are mapped by CUDA IPC handle either way, so a
    peer GPU does not have to be enumerable in this process."""
    if not use_gpu:
        return None
    import torch
    n = torch.cuda.device_count()
    if n < 1:
        raise RuntimeError("use_gpu=True but this worker sees no CUDA device "
                           f"(CUDA_VISIBLE_DEVICES={os.environ.get('CUDA_VISIBLE_DEVICES')!r})")
    idx = 0 if n == 1 else int(local_rank) % n
    torch.cuda.set_device(idx)
    return idx
