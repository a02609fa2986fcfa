"""Process launcher seam.

The reference launches its workers as Ray actors (ray_lightning/ray_ddp.py:17-31,92-97) and
moves the trainer / results through the Ray object store.  Ray is not installed in this image
and cannot be, so this module provides the SUBSET of the `ray` API the reference touches,
implemented with local worker processes (plain `python -m ..._actor_main` children, so the
user's script is never re-imported and needs no __main__ guard): remote actor classes with .options(num_cpus,
num_gpus).remote(), method.remote() futures, get / wait / put / kill, and a queue actor.
This module is the single seam where the real `ray` would plug in; nothing else in the package
imports ray.

GPU actors get exactly one device, Ray-style: CUDA_VISIBLE_DEVICES=<k> is exported in the
child BEFORE torch is imported, so the worker's device index is 0 (the pinning contract of
ray_lightning/tests/test_ddp_gpu.py:83-103).
"""
import itertools
import os
import secrets
import subprocess
import sys
import tempfile
import threading
import time
from multiprocessing.connection import Client

import cloudpickle

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


# ------------------------------------------------------------------ actor handles
class _Method:
    def __init__(self, handle, name):
        self._h, self._name = handle, name

    def remote(self, *args, **kwargs):
        return self._h._call(self._name, args, kwargs)


def _connect(address, authkey, proc=None, timeout=120.0):
    deadline = time.monotonic() + timeout
    while True:
        try:
            return Client(address, family="AF_UNIX", authkey=authkey)
        except (FileNotFoundError, ConnectionRefusedError):
            if proc is not None and proc.poll() is not None:
                raise RayActorError(f"actor process exited with code {proc.returncode} before it was reachable")
            if time.monotonic() > deadline:
                raise RayActorError("timed out connecting to the actor process")
            time.sleep(0.01)


class ActorHandle:
    """Client side of one actor.  The creating handle owns the process; a pickled copy that
    travels to another process reconnects to the same actor (Ray handle semantics)."""

    def __init__(self, cls=None, args=(), kwargs=None, options=None, _attach=None):
        self._options = dict(options or {})
        self._gpu = None
        self._proc = None
        self._boot = None
        self._pending = {}
        self._lock = threading.Lock()
        if _attach is not None:
            self._address, self._authkey = _attach
            self._conn = _connect(self._address, self._authkey)
            self._start_reader()
            return
        env = dict(os.environ)
        if int(self._options.get("num_gpus", 0) or 0) >= 1:
            self._gpu = _acquire_gpu()
            env["CUDA_VISIBLE_DEVICES"] = self._gpu
        ncpu = self._options.get("num_cpus", None)
        if ncpu:  # Ray exports OMP_NUM_THREADS=num_cpus for actors that reserve CPUs
            env["OMP_NUM_THREADS"] = str(max(1, int(ncpu)))
        self._authkey = secrets.token_bytes(16)
        self._address = os.path.join(tempfile.gettempdir(), f"b2c-actor-{os.getpid()}-{next(_ids)}-{secrets.token_hex(4)}.sock")
        pkg_root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
        env["PYTHONPATH"] = pkg_root + os.pathsep + env.get("PYTHONPATH", "")
        self._proc = subprocess.Popen([sys.executable, "-m", "ray_lightning_accelerators_b200._actor_main",
                                       self._address, self._authkey.hex()], env=env)
        try:
            self._conn = _connect(self._address, self._authkey, self._proc)
            self._start_reader()
            # boot is asynchronous so that several actors start (import torch ...) in parallel;
            # the first method call waits for it and re-raises a constructor failure
            self._boot = self._send(("__init__", None, cloudpickle.dumps(cls), cloudpickle.dumps((args, kwargs or {})),
                                     list(sys.path)))
        except BaseException:
            self._shutdown()
            raise

    def __reduce__(self):
        return (_attach_handle, (self._address, self._authkey, self._options))

    def _start_reader(self):
        self._alive = True
        self._reader = threading.Thread(target=self._read_loop, daemon=True)
        self._reader.start()

    def _read_loop(self):
        while True:
            try:
                call_id, ok, payload = self._conn.recv()
            except (EOFError, OSError, TypeError):
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

    def _send(self, msg):
        """msg = (kind, None, ...): slot 1 is replaced by the call id."""
        if not self._alive:
            raise RayActorError("actor is dead")
        fut = _Future(self)
        call_id = next(_ids)
        with self._lock:
            self._pending[call_id] = fut
            self._conn.send((msg[0], call_id) + tuple(msg[2:]))
        return fut

    def _call(self, name, args, kwargs):
        boot = getattr(self, "_boot", None)
        if boot is not None:
            try:
                boot.result()
            except BaseException:
                self._shutdown()
                raise
            self._boot = None
        return self._send(("call", None, name, cloudpickle.dumps((args, kwargs))))

    def __getattr__(self, name):
        if name.startswith("_"):
            raise AttributeError(name)
        return _Method(self, name)

    def _shutdown(self):
        if self._proc is None:
            return
        try:
            self._send(("__exit__", None)).result(timeout=5)
        except Exception:
            pass
        try:
            self._proc.wait(timeout=5)
        except Exception:
            self._proc.kill()
            self._proc.wait(timeout=5)
        try:
            self._conn.close()
        except Exception:
            pass
        try:
            os.unlink(self._address)
        except OSError:
            pass
        _release_gpu(self._gpu)
        self._gpu = None
        self._alive = False
        self._proc = None

    def _is_alive(self):
        return self._proc is not None and self._proc.poll() is None


def _attach_handle(address, authkey, options):
    return ActorHandle(options=options, _attach=(address, authkey))


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
class Empty(Exception):
    pass


class _QueueActor:
    """Lives in its own actor process (the reference's Queue wraps such an actor,
    util.py:17-85); items are opaque cloudpickled bytes."""

    def __init__(self, maxsize):
        import collections
        self.maxsize = maxsize
        self.items = collections.deque()

    def qsize(self):
        return len(self.items)

    def put_nowait(self, item):
        if self.maxsize > 0 and len(self.items) >= self.maxsize:
            return False
        self.items.append(item)
        return True

    def get_nowait(self):
        if not self.items:
            return (False, None)
        return (True, self.items.popleft())


class Queue:
    """The part of ray.util.queue.Queue the reference uses (util.py:71-93): put / get / empty /
    qsize / shutdown.  Picklable: the copy shipped to a worker talks to the same actor."""

    def __init__(self, maxsize=0, actor_options=None):
        self.maxsize = maxsize
        self.actor = ActorClass(_QueueActor, actor_options or {}).remote(maxsize)

    def put(self, item, block=True, timeout=None):
        payload = cloudpickle.dumps(item)
        deadline = None if timeout is None else time.monotonic() + timeout
        while not get(self.actor.put_nowait.remote(payload)):
            if not block or (deadline is not None and time.monotonic() > deadline):
                raise RuntimeError("queue full")
            time.sleep(0.005)

    def get(self, block=True, timeout=None):
        deadline = None if timeout is None else time.monotonic() + timeout
        while True:
            ok, payload = get(self.actor.get_nowait.remote())
            if ok:
                return cloudpickle.loads(payload)
            if not block or (deadline is not None and time.monotonic() > deadline):
                raise Empty
            time.sleep(0.005)

    def empty(self):
        return get(self.actor.qsize.remote()) == 0

    def qsize(self):
        return get(self.actor.qsize.remote())

    def shutdown(self):
        if self.actor is not None and self.actor._proc is not None:
            kill(self.actor)
        self.actor = None


def shutdown():
    for a in list(_live_actors):
        kill(a)
