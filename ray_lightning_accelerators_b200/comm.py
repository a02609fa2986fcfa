"""Python handles over the b2c C-ABI: ``Comm`` (one rank per process, peers mapped
with CUDA IPC) and ``LocalGroup`` (W virtual ranks on one GPU, one cooperative
launch per collective).

PyTorch is used here for device memory and streams only; every byte of the data
path moves inside libb2c.so's sm_100a kernels.
"""
import ctypes
import os

import torch

from . import _lib
from ._lib import (ALGO_AUTO, ALGO_NVLS, ALGO_ONESHOT, ALGO_TWOSHOT, BF16, F32, U8, B2cError)  # noqa: F401

_DTYPES = {torch.float32: F32, torch.bfloat16: BF16}
_ELEM = {F32: 4, BF16: 2}

DEFAULT_ARENA_BYTES = int(os.environ.get("B2C_ARENA_BYTES", 256 << 20))

# env knobs (the accelerator constructor signatures are frozen)
_ENV_PARAMS = {
    "B2C_CROSSOVER_BYTES": _lib.PARAM_CROSSOVER_BYTES,
    "B2C_MAX_BLOCKS_ONESHOT": _lib.PARAM_MAX_BLOCKS_ONESHOT,
    "B2C_MAX_BLOCKS_TWOSHOT": _lib.PARAM_MAX_BLOCKS_TWOSHOT,
    "B2C_MAX_BLOCKS_BCAST": _lib.PARAM_MAX_BLOCKS_BCAST,
    "B2C_THREADS": _lib.PARAM_THREADS,
    "B2C_TIMEOUT_MS": _lib.PARAM_TIMEOUT_MS,
    "B2C_TWOSHOT_VARIANT": _lib.PARAM_TWOSHOT_VARIANT,
    "B2C_ZEROCOPY": _lib.PARAM_ZEROCOPY,
    "B2C_ZEROCOPY_MIN_BYTES": _lib.PARAM_ZEROCOPY_MIN_BYTES,
    "B2C_NVLS_MIN_BYTES": _lib.PARAM_NVLS_MIN_BYTES,
    "B2C_MAX_BLOCKS_NVLS": _lib.PARAM_MAX_BLOCKS_NVLS,
    "B2C_UNROLL_NVLS": _lib.PARAM_UNROLL_NVLS,
    "B2C_THREADS_NVLS": _lib.PARAM_THREADS_NVLS,
    "B2C_PDL": _lib.PARAM_PDL,
    "B2C_NVLS_LEAN": _lib.PARAM_NVLS_LEAN,
    "B2C_BCAST_MC": _lib.PARAM_BCAST_MC,
    "B2C_MC_ACQ_POLL": _lib.PARAM_MC_ACQ_POLL,
    "B2C_NVLS_STAGE_BYTES": _lib.PARAM_NVLS_STAGE_BYTES,
}


def _stream_ptr(stream):
    if stream is None:
        stream = torch.cuda.current_stream()
    return ctypes.c_void_p(stream.cuda_stream)


def _wire(t, wire_dtype):
    d = _DTYPES.get(t.dtype)
    if d is None:
        raise TypeError(f"b2c allreduce supports float32 / bfloat16 tensors, got {t.dtype}")
    if wire_dtype is None:
        return d, d
    w = _DTYPES.get(wire_dtype)
    if w is None:
        raise TypeError(f"unsupported wire dtype {wire_dtype}")
    return d, w


def _check_tensor(t):
    if not t.is_cuda:
        raise ValueError("b2c operates on CUDA tensors only (there is no CPU path)")
    if not t.is_contiguous():
        raise ValueError("b2c needs a contiguous tensor (DDP buckets are flat 1-D)")


def fused_shard(lib, count, world, rank):
    first, n = ctypes.c_size_t(), ctypes.c_size_t()
    _lib.check(lib.b2c_fused_shard(int(count), int(world), int(rank), ctypes.byref(first), ctypes.byref(n)))
    return first.value, n.value


class _Base:
    def _apply_env(self):
        for env, key in _ENV_PARAMS.items():
            if env in os.environ:
                self.set_param(key, int(os.environ[env]))

    def set_param(self, key, value):
        _lib.check(self._lib.b2c_set_param(self._first, key, int(value)))

    def get_param(self, key):
        v = ctypes.c_long()
        _lib.check(self._lib.b2c_get_param(self._first, key, ctypes.byref(v)))
        return v.value

    def stats(self):
        s = _lib.Stats()
        _lib.check(self._lib.b2c_get_stats(self._first, ctypes.byref(s)))
        return s.as_dict()

    def phase_us(self):
        """(to barrier A, A to B, B to end) of the last kernel's channel 0, in microseconds (syncs)."""
        out = (ctypes.c_uint64 * 4)()
        _lib.check(self._lib.b2c_get_phase_ns(self._first, ctypes.byref(out)))
        t = list(out)
        if t[2] < t[1]:  # kernels with a single barrier (one-shot, broadcast) leave slot 2 stale
            t[2] = t[3]
        return ((t[1] - t[0]) / 1e3, (t[2] - t[1]) / 1e3, (t[3] - t[2]) / 1e3)

    def phase_ring_us(self):
        """[(start->A, A->B, B->end, start_ns)] of this comm's last (up to 32) kernels, oldest first, in
        microseconds (synchronises).  start->A is the wait for the slowest rank, A->B the data phase."""
        n = _lib.PHASE_RING
        out = (ctypes.c_uint64 * (4 * n))()
        total = ctypes.c_uint32()
        got = self._lib.b2c_get_phase_ring(self._first, out, n, ctypes.byref(total))
        if got < 0:
            _lib.check(got)
        rows = []
        for i in range(got):
            t = [out[4 * i + k] for k in range(4)]
            if t[2] < t[1]:
                t[2] = t[3]
            rows.append(((t[1] - t[0]) / 1e3, (t[2] - t[1]) / 1e3, (t[3] - t[2]) / 1e3, t[0]))
        return rows, total.value

    def check_alive(self):
        """Raise if a kernel reported a flag-barrier timeout (sticky)."""
        st = self.stats()
        if st["device_status"]:
            raise B2cError(-int(st["device_status"]),
                           "flag barrier timed out (epoch {}, channel {}, waiting for rank {})".format(
                               *st["device_status_detail"]))


class Comm(_Base):
    """One rank of a cross-process group.  ``exchange`` is any callable that turns this
    rank's blob (bytes) into the list of all ranks' blobs in rank order — the Ray object
    store in the reference's world, a torch.distributed store / all_gather_object here."""

    def __init__(self, rank, world, device=None, arena_bytes=None, exchange=None):
        self._lib = _lib.load()
        if device is None:
            device = torch.cuda.current_device()
        self.rank, self.world, self.device = int(rank), int(world), int(device)
        self.arena_bytes = int(arena_bytes or DEFAULT_ARENA_BYTES)
        h = ctypes.c_void_p()
        _lib.check(self._lib.b2c_comm_create(self.rank, self.world, self.device, self.arena_bytes, ctypes.byref(h)))
        self._h = h
        self._first = h
        self._exchange = exchange
        self._registered = {}
        self._nvls = False
        self.nvls_error = None
        self._apply_env()
        if self.world > 1:
            if exchange is None:
                raise ValueError("world > 1 needs an `exchange` callable to swap IPC handles")
            blobs = exchange(self.export())
            self.import_(blobs)

    @classmethod
    def from_process_group(cls, group=None, device=None, arena_bytes=None):
        """Bootstrap over an initialised torch.distributed group (control plane only:
        W x 256 bytes, once).  ray_lightning/ray_ddp.py:232-237 creates that group."""
        import torch.distributed as dist
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        # Handle exchanges may happen later than set-up and from another thread (a DDP comm hook registers a
        # bucket the first time it sees it, on the autograd thread, while the main thread is free to issue
        # its own collectives).  They therefore never share a communicator with the training loop: when the
        # group is NCCL a private gloo group carries the blobs (created here, collectively, on the main thread).
        side = group
        if world > 1 and dist.get_backend(group) != "gloo":
            ranks = dist.get_process_group_ranks(group) if group is not None else None
            side = dist.new_group(ranks=ranks, backend="gloo")

        def exchange(blob):
            out = [None] * world
            dist.all_gather_object(out, blob, group=side)
            return out

        return cls(rank, world, device=device, arena_bytes=arena_bytes, exchange=exchange)

    @property
    def side_stream(self):
        """THE high-priority stream of this comm for reductions that overlap backward.  One per comm: calls on
        a comm must be stream-ordered, so every user (DDP reducer, hvd optimizers) shares it."""
        if getattr(self, "_side_stream", None) is None:
            self._side_stream = torch.cuda.Stream(device=self.device, priority=-1)
        return self._side_stream

    def export(self):
        buf = ctypes.create_string_buffer(_lib.EXPORT_BLOB_BYTES)
        n = ctypes.c_size_t(_lib.EXPORT_BLOB_BYTES)
        _lib.check(self._lib.b2c_comm_export(self._h, buf, ctypes.byref(n)))
        return buf.raw[:n.value]

    def import_(self, blobs):
        if len(blobs) != self.world:
            raise ValueError(f"need {self.world} blobs in rank order, got {len(blobs)}")
        joined = b"".join(bytes(b) for b in blobs)
        _lib.check(self._lib.b2c_comm_import(self._h, joined, len(joined) // self.world))

    def register(self, t):
        """Collective: map this tensor's memory into every peer so that allreduce can run in
        place on it with no staging pass (zero-copy two-shot).  Every rank must call it with
        its corresponding tensor (same size) in the same order.  Returns True when every rank
        could export its buffer (plain cudaMalloc memory), False otherwise — the staged path
        keeps working either way."""
        _check_tensor(t)
        key = (t.data_ptr(), t.numel() * t.element_size())
        if self.world == 1:
            self._registered[key] = False
            return False
        # ALWAYS collective: a rank-local cache hit must not skip the exchange (the same address can be a
        # fresh allocation on a peer).  The hit/miss flag travels with the blob; unless every rank has this
        # exact buffer registered already, everyone drops what it has and maps the current buffers.
        cached = key in self._registered
        buf = ctypes.create_string_buffer(_lib.REG_BLOB_BYTES)
        n = ctypes.c_size_t(_lib.REG_BLOB_BYTES)
        rc = self._lib.b2c_register_export(self._h, ctypes.c_void_p(key[0]), key[1], buf, ctypes.byref(n))
        blobs = self._exchange((cached, buf.raw if rc == 0 else None))
        if all(c for c, _ in blobs):
            return self._registered[key]
        if self._registered.pop(key, False):
            _lib.check(self._lib.b2c_deregister(self._h, ctypes.c_void_p(key[0])))
        ok = all(b is not None for _, b in blobs)
        if ok:
            joined = b"".join(bytes(b) for _, b in blobs)
            _lib.check(self._lib.b2c_register_import(self._h, ctypes.c_void_p(key[0]), key[1], joined,
                                                     _lib.REG_BLOB_BYTES))
        self._registered[key] = ok
        return ok

    def enable_nvls(self, arena_bytes=512 << 20):
        """Collective: set up the NVLink-multicast arena (in-switch reduction).  Returns True when
        every rank succeeded; on False nothing changes and the peer-to-peer paths keep working."""
        if self.world < 2 or self._nvls:
            return self._nvls
        blob = ctypes.create_string_buffer(_lib.NVLS_BLOB_BYTES)
        n = ctypes.c_size_t(_lib.NVLS_BLOB_BYTES)
        rc = self._lib.b2c_nvls_export(self._h, int(arena_bytes), blob, ctypes.byref(n))
        self.nvls_error = None if rc == 0 else _lib.last_error()
        blobs = self._exchange(blob.raw if rc == 0 else None)
        if any(b is None for b in blobs):
            return False
        joined = b"".join(bytes(b) for b in blobs)
        rc = self._lib.b2c_nvls_import(self._h, joined, _lib.NVLS_BLOB_BYTES)
        if rc != 0:
            self.nvls_error = _lib.last_error()
        oks = self._exchange(rc == 0)      # doubles as the barrier: every device added before any bind
        if not all(oks):
            return False
        rc = self._lib.b2c_nvls_bind(self._h)
        if rc != 0:
            self.nvls_error = _lib.last_error()
        oks = self._exchange(rc == 0)      # every rank bound before the first multimem access
        self._nvls = all(oks)
        if self._nvls:
            base, nbytes = ctypes.c_void_p(), ctypes.c_size_t()
            _lib.check(self._lib.b2c_nvls_arena(self._h, ctypes.byref(base), ctypes.byref(nbytes)))
            self._nvls_base, self._nvls_bytes, self._nvls_top = base.value, nbytes.value, 0
        return self._nvls

    def autotune(self, sizes=(32 << 10, 128 << 10, 512 << 10, 2 << 20, 8 << 20), iters=5, force=False):
        """Collective, once per comm: measure where each allreduce path starts to win at THIS world size on THIS
        box and set the AUTO thresholds from it (the crossover is measured, not assumed).
        Paths timed per size (fp32, CUDA events, median, max over ranks): the staged path AUTO would otherwise take
        (one-shot up to the crossover, two-shot above; column "oneshot"), zero-copy two-shot on
        a registered buffer, in-switch (NVLS) in place when the multicast arena is up (device time of back-to-back
        calls; the host's launch path is kept out of the measurement).  Rank 0's table decides
        for everyone (thresholds must be identical on all ranks).  B2C_AUTOTUNE=0 keeps the built-in defaults;
        explicit B2C_* threshold variables still win afterwards.  Returns the table (also kept in .tuning)."""
        if self.world == 1 or (getattr(self, "tuning", None) is not None and not force):
            return getattr(self, "tuning", None)
        if os.environ.get("B2C_AUTOTUNE", "1") == "0":
            self.tuning = {"skipped": "B2C_AUTOTUNE=0"}
            return self.tuning
        dev = torch.device("cuda", self.device)
        rows = []
        with torch.cuda.device(dev):
            for S in sizes:
                n = S // 4
                plain = torch.ones(n, device=dev)          # never registered: AUTO stages it (one-shot / two-shot)
                x = torch.ones(n, device=dev)
                reg = self.register(x)
                y = None
                if self._nvls:
                    try:
                        y = self.nvls_empty(n).fill_(1.0)
                    except MemoryError:
                        y = None
                cands = [("oneshot", plain, ALGO_AUTO)]     # "oneshot" = the staged path AUTO falls back to at this size
                if reg:
                    cands.append(("zerocopy", x, ALGO_TWOSHOT))
                if y is not None:
                    cands.append(("nvls", y, ALGO_NVLS))
                row = {"bytes": S}
                for name, t, algo in cands:
                    for _ in range(3):
                        self.allreduce(t, scale=1.0 / self.world, algo=algo)
                    ts = []
                    reps = 4
                    for _ in range(iters):
                        # device time only: the GPU is kept busy while the host enqueues, and `reps` calls run back
                        # to back between the two events (a single call after a sync would time the host's launch path)
                        self.barrier()
                        torch.cuda._sleep(300000)
                        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                        a.record()
                        for _ in range(reps):
                            self.allreduce(t, scale=1.0 / self.world, algo=algo)
                        b.record()
                        b.synchronize()
                        ts.append(a.elapsed_time(b) * 1e3 / reps)
                    ts.sort()
                    row[name] = ts[len(ts) // 2]
                if y is not None:
                    self.nvls_free(y)
                if reg:
                    self.deregister(x)
                rows.append(row)
        allrows = self._exchange(rows)
        merged = []
        for i, S in enumerate(sizes):   # per size and path: the slowest rank's median
            m = {"bytes": S}
            for k in ("oneshot", "zerocopy", "nvls"):
                vals = [r[i][k] for r in allrows if k in r[i]]
                if len(vals) == self.world:
                    m[k] = max(vals)
            merged.append(m)

        def first_win(path, margin=0.97):
            """Smallest measured size from which `path` beats the staged one-shot at every larger measured size."""
            thr = None
            for m in reversed(merged):
                if path in m and m[path] <= margin * m["oneshot"]:
                    thr = m["bytes"]
                else:
                    break
            return thr

        big = 1 << 40
        zc, nv = first_win("zerocopy"), first_win("nvls")
        decided = {"zerocopy_min_bytes": zc if zc is not None else (sizes[-1] * 2 if any("zerocopy" in m for m in merged) else big),
                   "nvls_min_bytes": nv if nv is not None else (sizes[-1] * 2 if any("nvls" in m for m in merged) else 0)}
        decided = self._exchange(decided)[0]          # rank 0 decides
        self.set_param(_lib.PARAM_ZEROCOPY_MIN_BYTES, decided["zerocopy_min_bytes"])
        if decided["nvls_min_bytes"]:
            self.set_param(_lib.PARAM_NVLS_MIN_BYTES, decided["nvls_min_bytes"])
        self._apply_env()                              # explicit environment settings keep the last word
        self.tuning = {"world": self.world, "table_us": merged, **decided}
        return self.tuning

    def nvls_empty(self, numel, dtype=torch.float32):
        """A tensor INSIDE the multicast arena (first-fit allocation, 256-byte aligned; give it back with
        nvls_free): allreduce on it runs in place in the switch with no staging.  Every rank must issue
        the same sequence of allocations and frees so that offsets match."""
        if not self._nvls:
            raise RuntimeError("multicast arena is not set up (enable_nvls)")
        es = torch.empty(0, dtype=dtype).element_size()
        nbytes = (numel * es + 255) // 256 * 256
        p = ctypes.c_void_p()
        rc = self._lib.b2c_nvls_alloc(self._h, nbytes, ctypes.byref(p))
        if rc != 0:
            raise MemoryError(_lib.last_error())
        ptr = p.value

        class _Raw:  # __cuda_array_interface__ v3: zero-copy view of raw device memory
            pass

        raw = _Raw()
        raw.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3,
                                        "strides": None}
        flat = torch.as_tensor(raw, device=torch.device("cuda", self.device))
        return flat[:numel * es].view(dtype)

    def nvls_free(self, t):
        """Return a tensor obtained from nvls_empty to the arena (same order on every rank)."""
        if self._nvls and getattr(self, "_h", None):
            _lib.check(self._lib.b2c_nvls_free(self._h, ctypes.c_void_p(t.data_ptr())))

    def torch_pool(self):
        """A torch.cuda.MemPool whose segments come out of the multicast arena: tensors allocated
        under `torch.cuda.use_mem_pool(pool)` (DDP's gradient buckets, when DDP is constructed
        there) are reduced in place by the NVLS kernel.  One pool per process."""
        if not self._nvls:
            raise RuntimeError("multicast arena is not set up (enable_nvls)")
        if getattr(self, "_pool", None) is None:
            _lib.check(self._lib.b2c_nvls_pool_bind(self._h))
            self._pool_alloc = torch.cuda.memory.CUDAPluggableAllocator(_lib.LIB_PATH, "b2c_pool_alloc", "b2c_pool_free")
            self._pool = torch.cuda.MemPool(self._pool_alloc.allocator())
        return self._pool

    def in_nvls_arena(self, t):
        return self._nvls and self._nvls_base <= t.data_ptr() < self._nvls_base + self._nvls_bytes

    @property
    def nvls_enabled(self):
        return self._nvls

    def deregister(self, t):
        self.deregister_ptr(t.data_ptr(), t.numel() * t.element_size())

    def deregister_ptr(self, ptr, nbytes):
        """Local: forget a registration (the memory is about to be freed or was replaced)."""
        if self._registered.pop((ptr, nbytes), False):
            _lib.check(self._lib.b2c_deregister(self._h, ctypes.c_void_p(ptr)))

    def allreduce(self, t, scale=None, wire_dtype=None, algo=ALGO_AUTO, stream=None):
        """In place: t = sum_r t_r * scale (default 1/world), fp32 accumulate in rank order."""
        _check_tensor(t)
        d, w = _wire(t, wire_dtype)
        if scale is None:
            scale = 1.0 / self.world
        _lib.check(self._lib.b2c_allreduce(self._h, ctypes.c_void_p(t.data_ptr()), t.numel(), d, w,
                                           ctypes.c_float(scale), algo, _stream_ptr(stream)))
        return t

    def broadcast(self, t, root=0, stream=None):
        _check_tensor(t)
        _lib.check(self._lib.b2c_broadcast(self._h, ctypes.c_void_p(t.data_ptr()), t.numel() * t.element_size(),
                                           int(root), _stream_ptr(stream)))
        return t

    def fused_shard(self, count):
        """(first element, number of elements) of the shard of a `count`-element bucket this rank owns in
        fused_step: the optimizer state buffers cover exactly that range."""
        return fused_shard(self._lib, count, self.world, self.rank)

    def fused_step(self, grad, param, opt, state1, state2, scale=None, bcast=None, bcast_root=0, stream=None):
        """Gradient average + optimizer update in one launch (b2c_fused_step): `param` (same layout as `grad`) is
        updated on every rank from the averaged `grad`; `opt` is a _lib.Opt; state1/2 are this rank's fp32 shard
        state tensors (or None).  `bcast`: optional uint8 tensor inside the multicast arena replicated from
        `bcast_root` in the same launch."""
        _check_tensor(grad)
        _check_tensor(param)
        if scale is None:
            scale = 1.0 / self.world
        bp, bn = (ctypes.c_void_p(bcast.data_ptr()), bcast.numel() * bcast.element_size()) if bcast is not None else (None, 0)
        _lib.check(self._lib.b2c_fused_step(self._h, ctypes.c_void_p(grad.data_ptr()), ctypes.c_void_p(param.data_ptr()),
                                            grad.numel(), _DTYPES[grad.dtype], ctypes.c_float(scale), ctypes.byref(opt),
                                            ctypes.c_void_p(state1.data_ptr()) if state1 is not None else None,
                                            ctypes.c_void_p(state2.data_ptr()) if state2 is not None else None,
                                            bp, bn, int(bcast_root), _stream_ptr(stream)))

    def barrier(self, stream=None):
        _lib.check(self._lib.b2c_barrier(self._h, _stream_ptr(stream)))

    def close(self):
        if getattr(self, "_h", None):
            h, self._h = self._h, None
            _lib.check(self._lib.b2c_comm_destroy(h))

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class LocalGroup(_Base):
    """W virtual ranks on one GPU.  Same kernels, same flag protocol, one cooperative
    launch per collective (grid.y = W) so the ranks' CTAs are guaranteed co-resident."""

    def __init__(self, world, device=None, arena_bytes=None):
        self._lib = _lib.load()
        if device is None:
            device = torch.cuda.current_device()
        self.world, self.device = int(world), int(device)
        self.arena_bytes = int(arena_bytes or DEFAULT_ARENA_BYTES)
        self._arr = (ctypes.c_void_p * self.world)()
        _lib.check(self._lib.b2c_group_create_local(self.world, self.device, self.arena_bytes, self._arr))
        self._first = ctypes.c_void_p(self._arr[0])
        self._open = True
        self._apply_env()

    def _ptrs(self, tensors):
        if len(tensors) != self.world:
            raise ValueError(f"need {self.world} tensors (one per virtual rank)")
        for t in tensors:
            _check_tensor(t)
        return (ctypes.c_void_p * self.world)(*[t.data_ptr() for t in tensors])

    def allreduce(self, tensors, scale=None, wire_dtype=None, algo=ALGO_AUTO, stream=None):
        d, w = _wire(tensors[0], wire_dtype)
        n = tensors[0].numel()
        if any(t.numel() != n or t.dtype != tensors[0].dtype for t in tensors):
            raise ValueError("all ranks must pass the same count and dtype")
        if scale is None:
            scale = 1.0 / self.world
        _lib.check(self._lib.b2c_group_allreduce(self._arr, self._ptrs(tensors), n, d, w, ctypes.c_float(scale),
                                                 algo, _stream_ptr(stream)))
        return tensors

    def broadcast(self, tensors, root=0, stream=None):
        nbytes = tensors[0].numel() * tensors[0].element_size()
        if any(t.numel() * t.element_size() != nbytes for t in tensors):
            raise ValueError("all ranks must pass the same byte count")
        _lib.check(self._lib.b2c_group_broadcast(self._arr, self._ptrs(tensors), nbytes, int(root),
                                                 _stream_ptr(stream)))
        return tensors

    def fused_step(self, grads, params, opt, state1, state2, scale=None, stream=None):
        """b2c_group_fused_step: per virtual rank gradient / parameter buckets and shard state tensors."""
        if scale is None:
            scale = 1.0 / self.world
        n = grads[0].numel()

        def arr(ts):
            return (ctypes.c_void_p * self.world)(*[(t.data_ptr() if t is not None else None) for t in ts])

        _lib.check(self._lib.b2c_group_fused_step(self._arr, self._ptrs(grads), self._ptrs(params), n, _DTYPES[grads[0].dtype],
                                                  ctypes.c_float(scale), ctypes.byref(opt), arr(state1), arr(state2),
                                                  _stream_ptr(stream)))

    def barrier(self, stream=None):
        _lib.check(self._lib.b2c_group_barrier(self._arr, _stream_ptr(stream)))

    def close(self):
        if getattr(self, "_open", False):
            self._open = False
            _lib.check(self._lib.b2c_group_destroy_local(self._arr, self.world))

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
