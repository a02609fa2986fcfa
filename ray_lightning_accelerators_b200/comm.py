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

        def exchange(blob):
            out = [None] * world
            dist.all_gather_object(out, blob, group=group)
            return out

        return cls(rank, world, device=device, arena_bytes=arena_bytes, exchange=exchange)

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
        if key in self._registered:
            return self._registered[key]
        if self.world == 1:
            self._registered[key] = False
            return False
        buf = ctypes.create_string_buffer(_lib.REG_BLOB_BYTES)
        n = ctypes.c_size_t(_lib.REG_BLOB_BYTES)
        rc = self._lib.b2c_register_export(self._h, ctypes.c_void_p(key[0]), key[1], buf, ctypes.byref(n))
        blobs = self._exchange(buf.raw if rc == 0 else None)
        ok = all(b is not None for b in blobs)
        if ok:
            joined = b"".join(bytes(b) for b in blobs)
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

    def nvls_empty(self, numel, dtype=torch.float32):
        """A tensor INSIDE the multicast arena (bump allocation, 256-byte aligned, never freed):
        allreduce on it runs in place in the switch with no staging.  Every rank must allocate
        the same sequence of sizes so that offsets match."""
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
        key = (t.data_ptr(), t.numel() * t.element_size())
        if self._registered.pop(key, False):
            _lib.check(self._lib.b2c_deregister(self._h, ctypes.c_void_p(key[0])))

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
