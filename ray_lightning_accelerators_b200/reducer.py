"""B2cReducer — the gradient-bucket reducer shared by the DDP path and the Horovod shim.

What it replaces: torch's c10d Reducer behind ray_lightning/ray_ddp.py:218-219 and
Horovod's fusion buffer behind ray_horovod.py:158-159.  Both copy every gradient into a communication
buffer and back; the c10d Reducer additionally runs one NCCL allreduce of its "locally used parameters"
bitmap per step when find_unused_parameters=True (what PTL 1.1.7 passes).  Here

  * every parameter's ``.grad`` IS a slice of a static flat bucket (16-byte aligned slices), so autograd
    accumulates straight into communication memory: no copy-in, no copy-out;
  * buckets live inside the NVLink-multicast arena when it is up (in-switch reduction, in place), else in
    ordinary torch memory registered with every peer once (zero-copy peer-to-peer two-shot);
  * ``register_post_accumulate_grad_hook`` counts arrivals; a bucket is launched on the comm's side stream
    as soon as its last gradient has been accumulated AND every earlier bucket has been launched (fixed
    launch order on every rank, whatever order gradients arrive in); an autograd-engine callback queued at the
    start of each backward pass launches whatever is left (unused parameters contribute their zeros) and makes
    the training stream wait for the side stream;
  * nothing on this path touches NCCL or the host: no bitmap allreduce, no synchronisation.

Buckets are formed like DDP forms them (reverse parameter order, a small first bucket so the reduction starts
early, then ``bucket_cap_mb``), per (dtype, device).
"""
import torch

from .comm import ALGO_AUTO

FIRST_BUCKET_BYTES = 1 << 20      # torch.distributed._DEFAULT_FIRST_BUCKET_BYTES


class Bucket:
    __slots__ = ("flat", "params", "offsets", "ready", "launched", "in_arena", "index", "param_flat", "opt_state")

    def __init__(self, flat, params, offsets, in_arena, index):
        self.flat, self.params, self.offsets = flat, params, offsets
        self.ready, self.launched = 0, False
        self.in_arena, self.index = in_arena, index
        self.param_flat = None      # fused optimizer: the parameters themselves as views of a flat bucket
        self.opt_state = None

    @property
    def nbytes(self):
        return self.flat.numel() * self.flat.element_size()


def plan_buckets(params, cap_bytes, first_bytes=FIRST_BUCKET_BYTES):
    """Reverse parameter order (gradients arrive roughly back to front), runs of one (dtype, device), a small
    first bucket, then `cap_bytes`.  Returns a list of parameter lists."""
    runs, run, size, key = [], [], 0, None
    cap = first_bytes
    for p in reversed(list(params)):
        k = (p.dtype, p.device)
        nb = p.numel() * p.element_size()
        if run and (k != key or size + nb > cap):
            runs.append(run)
            run, size = [], 0
            cap = cap_bytes
        key = k
        run.append(p)
        size += nb
    if run:
        runs.append(run)
    return runs


class B2cReducer:
    """Owns the gradient buckets of `params` and reduces them across `comm` during backward.

    on_bucket_reduced(bucket) may be set to run (on the side stream, stream-ordered after the allreduce) once
    per bucket per backward pass.  `launch_fn(bucket, stream)` may replace the plain allreduce — the fused
    optimizer step plugs in there."""

    def __init__(self, params, comm, op_average=True, bucket_cap_bytes=25 << 20, first_bucket_bytes=FIRST_BUCKET_BYTES,
                 algo=ALGO_AUTO, fallback_allreduce=None, world=None):
        self.comm = comm
        self.world = int(world if world is not None else (comm.world if comm is not None else 1))
        self.scale = (1.0 / self.world) if op_average else 1.0
        self.algo = algo
        self.params = [p for p in params if p.requires_grad]
        self.fallback_allreduce = fallback_allreduce      # CPU plumbing tests: gloo all_reduce on the flat bucket
        self.buckets = []
        self._bucket_of = {}
        self._handles = []
        self._next = 0                # buckets launch strictly in index order
        self._callback_queued = False
        self._enabled = True          # no_sync(): accumulate locally, reduce nothing
        self._grads_seen = False
        self._finalize_at_backward_end = True   # DDP semantics; the hvd shim finalises in synchronize()
        self.launch_fn = None
        self.timing = None            # optional [(start_event, end_event, nbytes)]
        self.passes = 0
        for run in plan_buckets(self.params, bucket_cap_bytes, first_bucket_bytes):
            self._make_bucket(run)
        self._check_symmetric()
        for p in self.params:
            self._handles.append(p.register_post_accumulate_grad_hook(self._on_grad))

    # ------------------------------------------------------------------ set-up
    def _make_bucket(self, run):
        offs, total = [], 0
        es = run[0].element_size()
        for p in run:
            offs.append(total)
            total += (p.numel() * es + 15) // 16 * 16 // es     # 16-byte aligned slices: legal vector targets
        comm = self.comm
        flat, in_arena = None, False
        if comm is not None and run[0].is_cuda and getattr(comm, "nvls_enabled", False):
            try:
                flat = comm.nvls_empty(total, run[0].dtype).zero_()
                in_arena = True
            except MemoryError:
                flat = None
        if flat is None:
            flat = torch.zeros(total, dtype=run[0].dtype, device=run[0].device)
        b = Bucket(flat, list(run), offs, in_arena, len(self.buckets))
        for p, off in zip(run, offs):
            p.grad = flat[off:off + p.numel()].view_as(p)
            p._b2c_grad_view = p.grad
            self._bucket_of[p] = b
        if comm is not None and flat.is_cuda and not in_arena and self.world > 1:
            comm.register(flat)      # collective, once: zero-copy peer-to-peer from the first step on
        self.buckets.append(b)

    def _check_symmetric(self):
        """The in-switch kernels address a bucket by its OFFSET in the multicast arena: every rank must hold every
        bucket at the same offset with the same size.  The allocator is deterministic, so this only fails when the
        ranks issued different allocation sequences — then nothing may be reduced in place."""
        comm = self.comm
        if comm is None or self.world == 1 or not any(b.in_arena for b in self.buckets):
            return
        mine = [(b.flat.data_ptr() - comm._nvls_base if b.in_arena else -1, b.nbytes) for b in self.buckets]
        everyone = comm._exchange(mine)
        if any(o != everyone[0] for o in everyone):
            raise RuntimeError(f"gradient buckets sit at different offsets of the multicast arena on different ranks: {everyone}")

    @property
    def stream(self):
        return self.comm.side_stream if (self.comm is not None and self.buckets and self.buckets[0].flat.is_cuda) else None

    def enable_timing(self):
        self.timing = []

    def collect_timing_us(self):
        torch.cuda.synchronize()
        out = [(e0.elapsed_time(e1) * 1e3, nb) for e0, e1, nb in (self.timing or [])]
        self.timing = [] if self.timing is not None else None
        return out

    # ------------------------------------------------------------------ backward
    def _on_grad(self, p):
        view = p._b2c_grad_view
        if p.grad is not view:       # the gradient tensor was replaced (zero_grad(set_to_none=True)): adopt it
            view.copy_(p.grad)
            p.grad = view
        if not self._enabled:
            return
        if self._finalize_at_backward_end and not self._callback_queued:
            # runs once the whole backward pass is done (the hook point the c10d Reducer uses as well)
            torch.autograd.Variable._execution_engine.queue_callback(self._finalize)
            self._callback_queued = True
        b = self._bucket_of[p]
        if b.launched:
            # Horovod raises here too: a second backward pass arrived before synchronize() / step() consumed
            # the first one, and this bucket's reduction is already in flight
            raise AssertionError("gradients were computed more than once before the reduction of the previous backward "
                                 "pass was consumed: call optimizer.synchronize() / step() after every backward, or "
                                 "accumulate under no_sync()")
        b.ready += 1
        self._grads_seen = True
        while self._next < len(self.buckets) and self.buckets[self._next].ready >= len(self.buckets[self._next].params):
            self._launch(self.buckets[self._next])
            self._next += 1

    def _launch(self, b):
        if b.launched:
            return
        b.launched = True
        if self.world == 1 and self.launch_fn is None:
            return
        flat = b.flat
        if self.comm is not None and flat.is_cuda:
            st = self.comm.side_stream
            st.wait_stream(torch.cuda.current_stream(flat.device))
            if self.timing is not None:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(st)
            if self.launch_fn is not None:
                self.launch_fn(b, st)
            else:
                self.comm.allreduce(flat, scale=self.scale, algo=self.algo, stream=st)
            if self.timing is not None:
                e1.record(st)
                self.timing.append((e0, e1, b.nbytes))
        elif self.fallback_allreduce is not None:
            self.fallback_allreduce(flat, self.scale)

    def _finalize(self):
        """End of a backward pass: launch the buckets whose gradients never all arrived (unused parameters hold
        zeros), in index order, then make the training stream wait for the reductions."""
        self._callback_queued = False
        if not self._enabled:
            return
        for b in self.buckets[self._next:]:
            self._launch(b)
        self.wait()
        for b in self.buckets:
            b.ready, b.launched = 0, False
        self._next = 0
        self._grads_seen = False
        self.passes += 1

    def flush(self):
        """Explicit end-of-backward for callers outside an autograd pass (hvd.synchronize())."""
        if self._grads_seen or any(b.launched for b in self.buckets):
            self._finalize()

    def wait(self):
        st = self.stream
        if st is not None:
            torch.cuda.current_stream().wait_stream(st)

    def rearm(self):
        """Before a forward pass: a gradient that was set to None (optimizer.zero_grad(set_to_none=True), the
        default of current torch) would make autograd allocate a fresh tensor and cost one copy into the bucket
        per parameter.  Point it back at its zeroed slice instead — one memset per bucket."""
        for b in self.buckets:
            none = [p for p in b.params if p.grad is None]
            if not none:
                continue
            if len(none) == len(b.params):
                b.flat.zero_()
            else:
                for p in none:
                    p._b2c_grad_view.zero_()
            for p in none:
                p.grad = p._b2c_grad_view

    def zero_grad(self):
        """One memset per bucket; gradients stay views of the buckets."""
        for b in self.buckets:
            b.flat.zero_()
            for p in b.params:
                if p.grad is not p._b2c_grad_view:
                    p.grad = p._b2c_grad_view

    def close(self):
        for h in self._handles:
            h.remove()
        self._handles = []
        if self.comm is not None:
            for b in self.buckets:
                if b.in_arena:
                    try:
                        self.comm.nvls_free(b.flat)
                    except Exception:
                        pass
                elif b.flat.is_cuda and self.world > 1:
                    self.comm.deregister(b.flat)
        self.buckets = []
