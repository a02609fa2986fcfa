"""CPU: the launcher seam (_ray.py) behaves like the slice of Ray the reference relies on
(ray_ddp.py:17-31 actors, util.py:71-109 queue + result pump)."""
import os
import pickle
import time

import pytest

from ray_lightning_accelerators_b200 import _ray as ray
from ray_lightning_accelerators_b200.util import Queue, process_results


@pytest.fixture(autouse=True)
def _clean():
    yield
    ray.shutdown()


class Counter:
    def __init__(self, start=0):
        self.v = start

    def add(self, n):
        self.v += n
        return self.v

    def env(self, key):
        return os.environ.get(key)

    def boom(self):
        raise ValueError("worker-side failure")

    def pid(self):
        return os.getpid()

    def sleep(self, s):
        time.sleep(s)
        return s


def test_actor_calls_are_ordered_and_state_lives_in_the_worker():
    a = ray.remote(Counter).options(num_cpus=2).remote(10)
    futs = [a.add.remote(1) for _ in range(5)]
    assert ray.get(futs) == [11, 12, 13, 14, 15]
    assert ray.get(a.pid.remote()) != os.getpid()
    assert ray.get(a.env.remote("OMP_NUM_THREADS")) == "2"   # Ray exports num_cpus the same way
    assert ray.live_actor_count() == 1
    ray.kill(a)
    assert ray.live_actor_count() == 0


def test_worker_exception_travels_with_its_traceback():
    a = ray.remote(Counter).remote()
    with pytest.raises(ray.RayTaskError, match="worker-side failure"):
        ray.get(a.boom.remote())
    assert ray.get(a.add.remote(1)) == 1  # the actor survives a failed call


def test_constructor_failure_surfaces_on_first_use():
    class Bad:
        def __init__(self):
            raise RuntimeError("cannot start")

        def ping(self):
            return 1

    a = ray.remote(Bad).remote()
    with pytest.raises(ray.RayTaskError, match="cannot start"):
        ray.get(a.ping.remote())


def test_put_get_wait():
    ref = ray.put({"x": [1, 2, 3]})
    assert ray.get(ref) == {"x": [1, 2, 3]}
    a = ray.remote(Counter).remote()
    slow, fast = a.sleep.remote(0.3), None
    ready, rest = ray.wait([slow], timeout=0)
    assert ready == [] and rest == [slow]
    ready, rest = ray.wait([slow], timeout=5)
    assert ready == [slow] and rest == []


def test_handle_survives_pickling_and_reaches_the_same_actor():
    a = ray.remote(Counter).remote(5)
    clone = pickle.loads(pickle.dumps(a))
    assert ray.get(clone.add.remote(1)) == 6
    assert ray.get(a.add.remote(1)) == 7


def _producer(q, rank):
    q.put((rank, "payload"))
    q.put((rank, lambda: rank * 10))
    return rank


def test_queue_is_shared_between_driver_and_workers_and_results_are_pumped():
    from ray_lightning_accelerators_b200.ray_ddp import RayExecutor
    q = Queue(actor_options={"num_cpus": 0})
    workers = [ray.remote(RayExecutor).remote() for _ in range(2)]
    futs = [w.execute.remote(_producer, q, r) for r, w in enumerate(workers)]
    assert process_results(futs, None) == [0, 1]
    items = []
    while not q.empty():
        items.append(q.get())
    assert sorted(i[0] for i in items) == [0, 0, 1, 1]
    assert sorted(i[1]() for i in items if callable(i[1])) == [0, 10]
    q.shutdown()


def test_gpu_actor_needs_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("CPU-box check")
    with pytest.raises(ray.RayActorError, match="no free GPU"):
        ray.remote(Counter).options(num_gpus=1).remote()
