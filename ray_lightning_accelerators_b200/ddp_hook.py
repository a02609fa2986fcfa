"""Where the b2c kernels plug into torch DDP — the worker-side half of the hot path.

Reference call sites replaced:
  C2  initial parameter/buffer broadcast inside the DDP constructor
      (torch:nn/parallel/distributed.py:879-890, reached from ray_ddp.py:218-219)
      -> ``broadcast_module_states`` after constructing DDP with init_sync=False
  C3  per-bucket gradient allreduce of the c10d Reducer
      -> ``allreduce_hook`` registered with DDP.register_comm_hook
  C4  per-forward buffer broadcast (_sync_buffers, distributed.py:2214-2257)
      -> ``BufferSync`` forward pre-hook on ONE flat buffer (one launch per step)

No NCCL on any of the three; the torch process group stays for control collectives (C9).
"""
import contextlib
import ctypes
import os

import torch
import torch.distributed as dist

from . import _lib
from .comm import ALGO_AUTO, Comm


class B2cHookState:
    """State object handed to DDP.register_comm_hook (the `state` of hook(state, bucket))."""

    INLINE_BYTES = int(os.environ.get("B2C_HOOK_INLINE_BYTES", 1 << 20))

    def __init__(self, comm, wire_dtype=None, algo=ALGO_AUTO, register_buckets=True):
        self.comm = comm
        # first time a bucket buffer is seen it is registered with every peer (collective, on the
        # autograd thread, control-plane bytes only) so later steps reduce it in place
        self.register_buckets = register_buckets and wire_dtype is None
        self.wire_dtype = wire_dtype
        self.algo = algo
        # high-priority side stream: the reduce of a large bucket overlaps the rest of backward
        self.stream = torch.cuda.Stream(device=comm.device, priority=-1)
        self.buckets_reduced = 0
        self.timing = None  # optional list of (start_event, end_event, nbytes) filled by bench.py
        self._prepared = {}  # bucket index -> (data_ptr, call arguments): the per-step path does no set-up
        self._call = comm._lib.b2c_allreduce
        self._scale = ctypes.c_float(1.0 / comm.world)
        self._side_used = False

    def enable_timing(self):
        if getattr(self, "cpp_hook", None) is not None:
            self.cpp_hook.enable_timing(self.cpp_hook_id, True)
        self.timing = []

    def collect_timing_us(self):
        """[(microseconds, bytes)] of every bucket allreduce since enable_timing(); synchronises."""
        if getattr(self, "cpp_hook", None) is not None:
            return [(us, nb) for us, nb in self.cpp_hook.collect_timing(self.cpp_hook_id)]
        torch.cuda.synchronize()
        out = [(e0.elapsed_time(e1) * 1e3, nb) for e0, e1, nb in (self.timing or [])]
        self.timing = [] if self.timing is not None else None
        return out

    def prepare(self, index, buf):
        comm = self.comm
        wire = self.wire_dtype if (self.wire_dtype is not None and buf.dtype == torch.float32) else None
        old = self._prepared.get(index)
        if old is not None and old[0] != buf.data_ptr():
            comm.deregister_ptr(old[0], old[6])  # torch rebuilt its buckets: the old buffer is gone
        if self.register_buckets and not comm.in_nvls_arena(buf):
            # collective over the comm's private side channel (never the training loop's communicator);
            # once per bucket buffer, and again on every rank when torch re-allocates its buckets
            comm.register(buf)
        from .comm import _check_tensor, _wire
        _check_tensor(buf)
        d, w = _wire(buf, wire)
        nbytes = buf.numel() * buf.element_size()
        ent = (buf.data_ptr(), ctypes.c_void_p(buf.data_ptr()), buf.numel(), d, w, nbytes <= self.INLINE_BYTES, nbytes)
        self._prepared[index] = ent
        return ent


def allreduce_hook(state, bucket):
    """DDP communication hook: hook(state, bucket) -> Future[Tensor]
    (torch:nn/parallel/distributed.py:1987-2067).  With a hook registered the bucket is NOT
    pre-divided by world size (distributed.py:2021-2023); the kernel owns the 1/W.

    Runs on the autograd engine thread and never blocks the host.  Small buckets (latency-bound,
    <= 1 MiB) are enqueued straight on the backward stream: no stream switch, no extra events, a
    handful of microseconds of host time.  Large buckets go to a high-priority side stream that
    waits for the backward stream on the device, so the reduction overlaps the rest of backward;
    the returned Future is completed while that stream is current so torch records the event the
    Reducer later waits on."""
    buf = bucket.buffer()
    index = bucket.index()
    ent = state._prepared.get(index)
    if ent is None or ent[0] != buf.data_ptr():
        ent = state.prepare(index, buf)
    _, ptr, numel, d, w, inline, nbytes = ent
    comm = state.comm
    if inline and state.timing is None:
        if state._side_used:
            # calls on one comm must stay stream-ordered: a large bucket of this backward pass may
            # still be running on the side stream
            torch.cuda.current_stream().wait_stream(state.stream)
        rc = state._call(comm._h, ptr, numel, d, w, state._scale, state.algo,
                         ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
        if rc != 0:
            _lib.check(rc)
        fut = torch.futures.Future(devices=[buf.device])
        fut.set_result(buf)
        state.buckets_reduced += 1
        return fut
    state._side_used = True
    state.stream.wait_stream(torch.cuda.current_stream(buf.device))
    with torch.cuda.stream(state.stream):
        if state.timing is not None:
            e0 = torch.cuda.Event(enable_timing=True)
            e1 = torch.cuda.Event(enable_timing=True)
            e0.record(state.stream)
        rc = state._call(comm._h, ptr, numel, d, w, state._scale, state.algo, ctypes.c_void_p(state.stream.cuda_stream))
        if rc != 0:
            _lib.check(rc)
        if state.timing is not None:
            e1.record(state.stream)
            state.timing.append((e0, e1, nbytes))
        fut = torch.futures.Future(devices=[buf.device])
        fut.set_result(buf)
    state.buckets_reduced += 1
    return fut


def _flat_groups(tensors, cap_bytes):
    """Group tensors by dtype into runs of at most cap_bytes (like _broadcast_coalesced's
    250 MiB buffers, torch:distributed/utils.py:289-328)."""
    by_dtype = {}
    for t in tensors:
        by_dtype.setdefault(t.dtype, []).append(t)
    for dtype, ts in by_dtype.items():
        run, size = [], 0
        for t in ts:
            nb = t.numel() * t.element_size()
            if run and size + nb > cap_bytes:
                yield run
                run, size = [], 0
            run.append(t)
            size += nb
        if run:
            yield run


@torch.no_grad()
def broadcast_module_states(comm, module, root=0, cap_bytes=64 << 20, stream=None):
    """C2: copy rank `root`'s parameters and buffers to every rank, bit for bit.
    Coalesced: flatten -> one b2c_broadcast per <= cap_bytes run -> copy back."""
    tensors = [p.data for p in module.parameters()] + [b.data for b in module.buffers()]
    tensors = [t for t in tensors if t.is_cuda and t.numel() > 0]
    nbytes = 0
    for run in _flat_groups(tensors, cap_bytes):
        flat = torch.cat([t.reshape(-1) for t in run])
        comm.broadcast(flat, root=root, stream=stream)
        off = 0
        for t in run:
            n = t.numel()
            t.copy_(flat[off:off + n].view_as(t))
            off += n
        nbytes += flat.numel() * flat.element_size()
    return nbytes


class BufferSync:
    """C4: the per-forward broadcast of module buffers (BatchNorm running stats,
    num_batches_tracked) from rank 0.  At set-up every buffer is re-pointed at a slice of ONE
    flat byte tensor (16-byte aligned slices), so the per-step work is a single kernel launch
    with no flatten / unflatten copies; in-place updates by the modules keep the views valid."""

    def __init__(self, comm, module, root=0):
        self.comm, self.root = comm, root
        bufs = [b for b in module.buffers() if b.is_cuda and b.numel() > 0]
        self.count = len(bufs)
        self.nbytes = 0
        self.flat = None
        if not bufs or comm.world == 1:
            return
        offs, total = [], 0
        for b in bufs:
            offs.append(total)
            total += (b.numel() * b.element_size() + 15) // 16 * 16
        # inside the multicast arena when it is up: the root then multicasts the bytes in place (one launch,
        # nothing copied on the receiving side)
        self.flat = None
        if getattr(comm, "nvls_enabled", False):
            try:
                self.flat = comm.nvls_empty(total, torch.uint8).zero_()
            except MemoryError:
                self.flat = None
        if self.flat is None:
            self.flat = torch.zeros(total, dtype=torch.uint8, device=bufs[0].device)
        with torch.no_grad():
            for b, off in zip(bufs, offs):
                nb = b.numel() * b.element_size()
                view = self.flat[off:off + nb].view(b.dtype).view(b.shape)
                view.copy_(b.data)
                b.data = view
        self.nbytes = total

    def __call__(self, module=None, inputs=None):
        if self.flat is not None:
            self.comm.broadcast(self.flat, root=self.root)


def _want_nvls(comm):
    """In-switch reduction pays off from ~4 ranks up (wire bytes (1 + 1/W) S against 2 (W-1)/W S for
    peer-to-peer); B2C_NVLS=0/1 overrides."""
    env = os.environ.get("B2C_NVLS")
    if env is not None:
        return env == "1"
    return comm.world >= 4


class B2cDataParallel(torch.nn.Module):
    """Data-parallel wrapper with b2c's own gradient reducer (reducer.B2cReducer) instead of torch's c10d
    Reducer: what the reference's workers get from DistributedDataParallel(model, device_ids,
    find_unused_parameters=True) (PTL 1.1.7 configure_ddp, reached from ray_ddp.py:218-219), with the same
    observable behaviour — parameters and buffers start as rank 0's (C2), buffers follow rank 0 before every
    forward (C4), after backward every ``.grad`` holds the average over ranks (C3), ``no_sync()`` skips the
    reduction, unused parameters are tolerated — and none of its per-step overheads: gradients accumulate
    straight into the communication buckets (no copy-in / copy-out), no NCCL collective of any kind runs
    inside a step (torch's Reducer allreduces a used-parameter bitmap per step), and nothing blocks the host.

    Difference to torch DDP that a user can see: the gradient of a parameter that took no part in a backward
    pass is a tensor of zeros (torch leaves it ``None`` when it is unused on every rank)."""

    _is_b2c_data_parallel = True   # unwrapped like DistributedDataParallel by the trainer (mini_pl.Trainer.get_model)

    def __init__(self, module, comm, bucket_cap_mb=25, root=0, sync_states=True):
        super().__init__()
        from .reducer import B2cReducer
        self.module = module
        self.comm = comm
        if sync_states:
            broadcast_module_states(comm, module, root=root)       # C2
        if comm.world > 1 and _want_nvls(comm) and not comm.nvls_enabled:
            grad_bytes = sum(p.numel() * p.element_size() for p in module.parameters() if p.requires_grad)
            comm.enable_nvls(int(os.environ.get("B2C_NVLS_ARENA_BYTES", 2 * grad_bytes + (96 << 20))))
        if comm.world > 1:
            comm.autotune()                                         # measured crossover table, once per comm
        self.bufsync = BufferSync(comm, module, root=root)          # C4 (buffers re-pointed into one flat tensor)
        self.reducer = B2cReducer(module.parameters(), comm, op_average=True, bucket_cap_bytes=int(bucket_cap_mb) << 20)
        self.require_backward_grad_sync = True
        self._c4_folded = False        # set by fuse_optimizer: the buffer broadcast rides in the last bucket's launch
        self._buffers_synced = False

    def fuse_optimizer(self, optimizer):
        """Fold `optimizer`'s step into the bucket launches (fused.FusedOptimizer); returns the object to use in
        its place.  Falls back to the optimizer itself when it is of an unsupported kind."""
        from . import fused
        if not fused.supported(optimizer) or self._c4_folded or self.reducer.launch_fn is not None:
            return optimizer
        for b in self.reducer.buckets:   # checked BEFORE anything is re-pointed: the fused step is fp32, one param_group per bucket
            groups = {id(g) for g in optimizer.param_groups for p in g["params"] if any(p is q for q in b.params)}
            owned = sum(1 for q in b.params if any(q is p for g in optimizer.param_groups for p in g["params"]))
            if b.flat.dtype != torch.float32 or len(groups) != 1 or owned != len(b.params):
                import warnings
                warnings.warn("b2c: fused optimizer step not applicable (needs fp32 buckets whose parameters all belong to "
                              "one param_group of this optimizer); keeping optimizer.step()")
                return optimizer
        return fused.FusedOptimizer(optimizer, self)

    def forward(self, *args, **kwargs):
        if self.training:
            if self.bufsync.flat is not None:
                if self._c4_folded and self._buffers_synced:
                    self._buffers_synced = False   # already replicated by the last fused step of the previous backward
                else:
                    self.bufsync()
            if torch.is_grad_enabled():
                self.reducer.rearm()
        return self.module(*args, **kwargs)

    @contextlib.contextmanager
    def no_sync(self):
        """Accumulate gradients locally; the next backward outside the context reduces the sum (DDP.no_sync)."""
        old = self.reducer._enabled
        self.reducer._enabled = False
        try:
            yield
        finally:
            self.reducer._enabled = old

    def zero_grad(self, set_to_none=False):
        self.reducer.zero_grad()

    # bench / tests: per-bucket device times
    @property
    def _b2c_state(self):
        return self.reducer

    def close(self):
        self.reducer.close()


def wrap_ddp(model, comm, device_ids=None, wire_dtype=None, find_unused_parameters=True,
             bucket_cap_mb=25, gradient_as_bucket_view=False, process_group=None, reducer=None):
    """Build the data-parallel wrapper the way the reference's workers do (PTL 1.1.7 configure_ddp:
    DistributedDataParallel(model, device_ids, find_unused_parameters=True), reached from
    ray_ddp.py:218-219) with the three collectives re-routed to b2c.

    reducer="b2c" (default; B2C_REDUCER overrides): B2cDataParallel, b2c's own bucket reducer.
    reducer="torch": torch's DistributedDataParallel with a comm hook per bucket (kept for the bf16-wire mode and
    as the like-for-like comparison against stock DDP)."""
    reducer = reducer or os.environ.get("B2C_REDUCER", "b2c")
    if reducer == "b2c" and wire_dtype is None:
        return B2cDataParallel(model, comm, bucket_cap_mb=bucket_cap_mb)
    from torch.nn.parallel import DistributedDataParallel as DDP
    broadcast_module_states(comm, model, root=0)                 # C2 (instead of init_sync)
    bufsync = BufferSync(comm, model, root=0)                    # C4 set-up (re-points buffers)
    pool_ctx = contextlib.nullcontext()
    if wire_dtype is None and comm.world > 1 and _want_nvls(comm):
        # NVLink multicast arena sized for the gradients (x2 for allocator rounding); when it comes up
        # DDP's buckets are allocated INSIDE it and every bucket allreduce runs in the switch, in place
        grad_bytes = sum(p.numel() * p.element_size() for p in model.parameters() if p.requires_grad)
        arena = int(os.environ.get("B2C_NVLS_ARENA_BYTES", 2 * grad_bytes + (64 << 20)))
        if comm.enable_nvls(arena):
            pool_ctx = torch.cuda.use_mem_pool(comm.torch_pool())
    with pool_ctx:
        ddp = DDP(model, device_ids=device_ids, find_unused_parameters=find_unused_parameters,
                  bucket_cap_mb=bucket_cap_mb, gradient_as_bucket_view=gradient_as_bucket_view,
                  broadcast_buffers=False, init_sync=False, process_group=process_group)
    state = B2cHookState(comm, wire_dtype=wire_dtype)
    cpp = _lib.load_torch_hook() if (os.environ.get("B2C_CPP_HOOK") == "1" and wire_dtype is None) else None
    if cpp is not None:
        # opt-in: the same hook logic as a c10d::CommHookInterface, no Python frame per bucket
        state.cpp_hook_id = cpp.register_hook(ddp.reducer, _lib.fn_address("b2c_allreduce"),
                                              _lib.fn_address("b2c_last_error"), comm._h.value, 1.0 / comm.world,
                                              ALGO_AUTO, B2cHookState.INLINE_BYTES)
        state.cpp_hook = cpp
    else:
        ddp.register_comm_hook(state, allreduce_hook)            # C3
    if bufsync.flat is not None:
        ddp.register_forward_pre_hook(bufsync)                   # C4 per step
    ddp._b2c_state = state
    ddp._b2c_bufsync = bufsync
    return ddp


def comm_from_default_group(arena_bytes=None):
    """Create the b2c communicator over the already-initialised torch process group
    (ray_ddp.py:232-237 creates it); the group only carries the 256-byte IPC blobs."""
    if not dist.is_initialized():
        raise RuntimeError("torch.distributed is not initialised; call init_ddp_connection first")
    return Comm.from_process_group(arena_bytes=arena_bytes)
