"""Process launcher seam.

The reference launches its workers as Ray actors (ray_lightning/ray_ddp.py:17-31,92-97) and
moves the trainer / results through the Ray object store.  Ray is not installed in this image
and cannot be, so this module provides the SUBSET of the `ray` API the reference touches,
implemented with spawned local processes: remote actor classes with .options(num_cpus,
num_gpus).remote(), method.remote() futures, get / wait / put / kill, and a queue actor.
This module is the single seam where the real `ray` would plug in; nothing else in the package
imports ray.

GPU actors get exactly one device, Ray-style: CUDA_VISIBLE_DEVICES=<k> is exported in the
child BEFORE torch is imported, so the worker's device index is 0 (the pinning contract of
ray_lightning/tests/test_ddp_gpu.py:83-103).
"""
import itertools
import multiprocessing as mp
import os
import queue as _queue
import threading
import traceback

import cloudpickle

_ctx = mp.get_context("spawn")
_ids = itertools.count()


class RayTaskError(RuntimeError):
    pass


class RayActorError(RuntimeError):
    pass


# ------------------------------------------------------------------ object store (driver-local)
class ObjectRef:
    def __init__(self, payload):
        self._payload = payload  # cloudpickled bytes: travels with the ref into the actors

    def _value(self):
        return cloudpickle.loads(self._payload)


def put(obj):
    return ObjectRef(cloudpickle.dumps(obj))


# ------------------------------------------------------------------ futures
class _Future:
    def __init__(self, actor):
        self._actor = actor
        self._ev = threading.Event()
        self._ok = None
        self._val = None

    def _set(self, ok, val):
        self._ok, self._val = ok, val
        self._ev.set()

    def done(self):
        return self._ev.is_set()

    def result(self, timeout=None):
        if not self._ev.wait(timeout):
            raise TimeoutError("remote call still running")
        if self._ok:
            return self._val
        raise self._val


def get(x, timeout=None):
    if isinstance(x, (list, tuple)):
        return [get(v, timeout) for v in x]
    if isinstance(x, ObjectRef):
        return x._value()
    if isinstance(x, _Future):
        return x.result(timeout)
    raise TypeError(f"ray.get: unsupported {type(x)}")


def wait(futures, num_returns=1, timeout=None):
    """Like ray.wait: (ready, not_ready); returns as soon as num_returns are ready or the
    timeout (seconds; 0 = poll) expires."""
    import time
    deadline = None if timeout is None else time.monotonic() + timeout
    while True:
        ready = [f for f in futures if f.done()]
        if len(ready) >= min(num_returns, len(futures)) or (deadline is not None and time.monotonic() >= deadline):
            ready = ready[:max(num_returns, 0)] if len(ready) > num_returns else ready
            rest = [f for f in futures if f not in ready]
            return ready, rest
        time.sleep(0.001)


# ------------------------------------------------------------------ GPU pool (one id per GPU actor)
_gpu_lock = threading.Lock()
_gpu_free = None


def _visible_gpus():
    env = os.environ.get("CUDA_VISIBLE_DEVICES")
    if env is not None and env.strip() != "":
        return [g.strip() for g in env.split(",") if g.strip() != ""]
    try:
        import torch
        return [str(i) for i in range(torch.cuda.device_count())]
    except Exception:
        return []


def _acquire_gpu():
    global _gpu_free
    with _gpu_lock:
        if _gpu_free is None:
            _gpu_free = _visible_gpus()
        if not _gpu_free:
            raise RayActorError("no free GPU for an actor requesting num_gpus=1")
        return _gpu_free.pop(0)


def _release_gpu(g):
    with _gpu_lock:
        if g is not None and _gpu_free is not None:
            _gpu_free.append(g)
            _gpu_free.sort()


# ------------------------------------------------------------------ actor process
def _actor_main(conn, cls_payload, init_payload, env):
    os.environ.update(env)
    try:
        cls = cloudpickle.loads(cls_payload)
        args, kwargs = cloudpickle.loads(init_payload)
        obj = cls(*args, **kwargs)
        conn.send((-1, True, None))
    except BaseException:
        conn.send((-1, False, traceback.format_exc()))
        return
    while True:
        try:
            msg = conn.recv()
        except EOFError:
            return
        if msg is None:
            return
        call_id, name, payload = msg
        try:
            args, kwargs = cloudpickle.loads(payload)
            out = getattr(obj, name)(*args, **kwargs)
            conn.send((call_id, True, cloudpickle.dumps(out)))
        except BaseException:
            conn.send((call_id, False, traceback.format_exc()))


class _Method:
    def __init__(self, handle, name):
        self._h, self._name = handle, name

    def remote(self, *args, **kwargs):
        return self._h._call(self._name, args, kwargs)


class ActorHandle:
    def __init__(self, cls, args, kwargs, options):
        self._options = dict(options)
        self._gpu = None
        env = {}
        if int(self._options.get("num_gpus", 0) or 0) >= 1:
            self._gpu = _acquire_gpu()
            env["CUDA_VISIBLE_DEVICES"] = self._gpu
        ncpu = self._options.get("num_cpus", None)
        if ncpu:  # Ray exports OMP_NUM_THREADS=num_cpus for actors that reserve CPUs
            env["OMP_NUM_THREADS"] = str(max(1, int(ncpu)))
        self._conn, child = _ctx.Pipe()
        self._proc = _ctx.Process(target=_actor_main, daemon=True,
                                  args=(child, cloudpickle.dumps(cls), cloudpickle.dumps((args, kwargs)), env))
        self._proc.start()
        child.close()
        self._pending = {}
        self._lock = threading.Lock()
        self._alive = True
        boot = _Future(self)
        self._pending[-1] = boot
        self._reader = threading.Thread(target=self._read_loop, daemon=True)
        self._reader.start()
        try:
            boot.result()
        except BaseException:
            self._shutdown()
            raise

    def _read_loop(self):
        while True:
            try:
                call_id, ok, payload = self._conn.recv()
            except (EOFError, OSError):
                break
            with self._lock:
                fut = self._pending.pop(call_id, None)
            if fut is None:
                continue
            if ok:
                fut._set(True, cloudpickle.loads(payload) if payload is not None else None)
            else:
                fut._set(False, RayTaskError(payload))
        self._alive = False
        with self._lock:
            pending, self._pending = self._pending, {}
        for fut in pending.values():
            fut._set(False, RayActorError("actor died before the call returned"))

    def _call(self, name, args, kwargs):
        if not self._alive:
            raise RayActorError("actor is dead")
        fut = _Future(self)
        call_id = next(_ids)
        with self._lock:
            self._pending[call_id] = fut
        self._conn.send((call_id, name, cloudpickle.dumps((args, kwargs))))
        return fut

    def __getattr__(self, name):
        if name.startswith("_"):
            raise AttributeError(name)
        return _Method(self, name)

    def _shutdown(self):
        if self._proc is None:
            return
        try:
            self._conn.send(None)
        except Exception:
            pass
        self._proc.join(timeout=5)
        if self._proc.is_alive():
            self._proc.terminate()
            self._proc.join(timeout=5)
        try:
            self._conn.close()
        except Exception:
            pass
        _release_gpu(self._gpu)
        self._gpu = None
        self._alive = False
        self._proc = None

    def _is_alive(self):
        return self._proc is not None and self._proc.is_alive()


class ActorClass:
    def __init__(self, cls, options=None):
        self._cls = cls
        self._options = dict(options or {})

    def options(self, **kw):
        return ActorClass(self._cls, {**self._options, **kw})

    def remote(self, *args, **kwargs):
        h = ActorHandle(self._cls, args, kwargs, self._options)
        _live_actors.append(h)
        return h


_live_actors = []


def remote(cls=None, **options):
    if cls is None:
        return lambda c: ActorClass(c, options)
    return ActorClass(cls, options)


def kill(actor, no_restart=True):
    actor._shutdown()
    if actor in _live_actors:
        _live_actors.remove(actor)


def live_actor_count():
    """How many actor processes are alive (the reference's test_ddp.py:29-42 asserts on this
    through ray.state.actors())."""
    return sum(1 for a in _live_actors if a._is_alive())


def get_node_ip_address():
    return "127.0.0.1"


# ------------------------------------------------------------------ queue actor
_manager = None


class Empty(Exception):
    pass


class Queue:
    """The part of ray.util.queue.Queue the reference uses (util.py:71-93): put / get / empty /
    shutdown, picklable so it can be shipped to the workers."""

    def __init__(self, maxsize=0, actor_options=None):
        global _manager
        if _manager is None:
            _manager = _ctx.Manager()
        self.maxsize = maxsize
        self._q = _manager.Queue(maxsize)

    def put(self, item, block=True, timeout=None):
        self._q.put(cloudpickle.dumps(item), block, timeout)

    def get(self, block=True, timeout=None):
        try:
            return cloudpickle.loads(self._q.get(block, timeout))
        except _queue.Empty:
            raise Empty

    def empty(self):
        return self._q.empty()

    def qsize(self):
        return self._q.qsize()

    def shutdown(self):
        self._q = None


def shutdown():
    global _manager
    for a in list(_live_actors):
        kill(a)
    if _manager is not None:
        _manager.shutdown()
        _manager = None


def is_real_ray():
    return False
