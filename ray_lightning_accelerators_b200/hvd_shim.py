"""`horovod.torch` surface used behind HorovodRayAccelerator (ray_lightning/ray_horovod.py:149-162
and PTL 1.1.7's HorovodAccelerator [upstream]): init, rank, size, local_rank,
broadcast_parameters, broadcast_optimizer_state, DistributedOptimizer (+ synchronize /
skip_synchronize), allreduce, join, shutdown.

Horovod itself is absent from this image.  On GPUs the backend is the b2c library: gradients
live in static flat buckets (registered with every peer once, so the per-step allreduce runs
in place over NVLink — no fusion-buffer pack/unpack, no background negotiation thread), and
Average = sum / size with fp32 accumulation and one rounding.  On CPUs it is gloo, which is what
the reference's CI builds Horovod with (.github/workflows/test.yaml:39).
"""
import contextlib
import os

import torch
import torch.distributed as dist

Average, Sum = "average", "sum"
_BUCKET_BYTES = int(os.environ.get("B2C_HVD_BUCKET_BYTES", 64 << 20))  # Horovod's fusion threshold

_st = {"init": False, "rank": 0, "size": 1, "local_rank": 0, "comm": None, "gpu": False, "own_pg": False}


def init(rank=None, size=None, local_rank=None, address=None, use_gpu=None):
    """Rendezvous.  horovod.ray.RayExecutor exports the coordinates as env vars on each worker;
    so does ray_horovod.HorovodExecutor here."""
    if _st["init"]:
        return
    rank = int(os.environ.get("HOROVOD_RANK", 0)) if rank is None else rank
    size = int(os.environ.get("HOROVOD_SIZE", 1)) if size is None else size
    local_rank = int(os.environ.get("HOROVOD_LOCAL_RANK", rank)) if local_rank is None else local_rank
    address = address or os.environ.get("B2C_HVD_ADDRESS")
    use_gpu = (os.environ.get("B2C_HVD_USE_GPU", "0") == "1") if use_gpu is None else use_gpu
    _st.update(rank=rank, size=size, local_rank=local_rank, gpu=bool(use_gpu))
    if use_gpu:
        n = torch.cuda.device_count()
        torch.cuda.set_device(0 if n == 1 else local_rank % n)
    if not dist.is_initialized():
        if address is None:
            raise RuntimeError("hvd.init(): no rendezvous address (B2C_HVD_ADDRESS)")
        kw = {"device_id": torch.device("cuda", torch.cuda.current_device())} if use_gpu else {}
        dist.init_process_group("nccl" if use_gpu else "gloo", init_method=address, rank=rank, world_size=size, **kw)
        _st["own_pg"] = True
    if use_gpu:
        from .comm import Comm  # fails loudly if libb2c.so is missing: there is no fallback
        _st["comm"] = Comm.from_process_group()
    _st["init"] = True


def _need():
    if not _st["init"]:
        raise ValueError("Horovod has not been initialized; use hvd.init().")


def rank():
    _need()
    return _st["rank"]


def size():
    _need()
    return _st["size"]


def local_rank():
    _need()
    return _st["local_rank"]


def is_initialized():
    return _st["init"]


def shutdown():
    if _st["comm"] is not None:
        torch.cuda.synchronize()
        dist.barrier()
        _st["comm"].close()
        _st["comm"] = None
    if _st["own_pg"] and dist.is_initialized():
        dist.destroy_process_group()
    _st.update(init=False, own_pg=False)


def join():
    """Every rank reached the end of training (ray_horovod.py:159 -> PTL calls hvd.join())."""
    _need()
    if _st["comm"] is not None:
        _st["comm"].barrier()
        torch.cuda.current_stream().synchronize()
    elif _st["size"] > 1:
        dist.barrier()
    return _st["size"] - 1


def allreduce_(tensor, op=Average, name=None):
    _need()
    scale = 1.0 / _st["size"] if op == Average else 1.0
    if _st["comm"] is not None and tensor.is_cuda:
        if tensor.dtype not in (torch.float32, torch.bfloat16):
            # no silent detour through NCCL: the GPU path is libb2c.so or nothing
            raise TypeError(f"hvd.allreduce on GPU supports float32 / bfloat16 tensors, got {tensor.dtype}; "
                            f"cast first (metrics: .float())")
        _st["comm"].allreduce(tensor, scale=scale)
    elif _st["size"] > 1:
        dist.all_reduce(tensor)
        if op == Average:
            tensor.div_(_st["size"])
    return tensor


def allreduce(tensor, op=Average, name=None):
    return allreduce_(tensor.clone(), op=op, name=name)


def _bcast_tensor(t, root):
    if _st["size"] == 1:
        return
    if _st["comm"] is not None and t.is_cuda:
        c = t if t.is_contiguous() else t.contiguous()
        _st["comm"].broadcast(c, root=root)
        if c is not t:
            t.copy_(c)
    else:
        dist.broadcast(t, src=root)


@torch.no_grad()
def broadcast_parameters(params, root_rank=0):
    """In-place copy of rank `root_rank`'s tensors (state_dict or named_parameters) to all."""
    _need()
    items = sorted(params.items()) if isinstance(params, dict) else sorted(list(params), key=lambda kv: kv[0])
    tensors = [p.data if isinstance(p, torch.nn.Parameter) else p for _, p in items if torch.is_tensor(p)]
    if _st["comm"] is not None:
        # coalesce like the DDP path: one launch per dtype run instead of one per tensor
        from .ddp_hook import _flat_groups
        cuda = [t for t in tensors if t.is_cuda and t.numel() > 0]
        for run in _flat_groups(cuda, 64 << 20):
            flat = torch.cat([t.reshape(-1) for t in run])
            _st["comm"].broadcast(flat, root=root_rank)
            off = 0
            for t in run:
                t.copy_(flat[off:off + t.numel()].view_as(t))
                off += t.numel()
        tensors = [t for t in tensors if not t.is_cuda]
    for t in tensors:
        _bcast_tensor(t, root_rank)


def broadcast_optimizer_state(optimizer, root_rank=0):
    """Tensor state is broadcast in place; scalar hyper-parameters follow rank 0 via the control
    group.  Empty at step 0 for SGD/Adam (state is created lazily by the first step)."""
    _need()
    if isinstance(optimizer, DistributedOptimizer):
        optimizer = optimizer._opt
    if _st["size"] == 1:
        return
    groups = [{k: v for k, v in g.items() if k != "params"} for g in optimizer.param_groups]
    box = [groups]
    dist.broadcast_object_list(box, src=root_rank)
    for g, src in zip(optimizer.param_groups, box[0]):
        g.update(src)
    for p in (p for g in optimizer.param_groups for p in g["params"]):
        for k, v in sorted(optimizer.state.get(p, {}).items()):
            if torch.is_tensor(v):
                _bcast_tensor(v, root_rank)


class DistributedOptimizer:
    """hvd.DistributedOptimizer(optimizer, named_parameters=...): gradients are averaged across
    ranks before step().  Built on reducer.B2cReducer — the same bucket reducer the DDP path uses:
    gradients are views into static flat buckets (inside the NVLink-multicast arena from 4 ranks up,
    registered with every peer below that); a bucket is reduced on the comm's side stream as soon as the
    last of its gradients has been accumulated and every earlier bucket has been launched (fixed order on
    every rank), overlapping the rest of backward; synchronize() launches what is left and makes the
    current stream wait."""

    def __init__(self, optimizer, named_parameters=None, op=Average, bucket_bytes=None, backward_passes_per_step=1):
        _need()
        from .reducer import B2cReducer
        self._opt = optimizer
        self._op = op
        self._skip_sync = False
        self._passes_per_step = int(backward_passes_per_step)
        params = [p for g in optimizer.param_groups for p in g["params"] if p.requires_grad]
        self._params = params
        comm = _st["comm"]
        if comm is not None and params:
            from .ddp_hook import _want_nvls
            if _want_nvls(comm) and not comm.nvls_enabled:
                # from ~4 ranks up the gradient buckets live in the NVLink-multicast arena and are
                # reduced in the switch, in place (collective set-up, once)
                total = sum(p.numel() * p.element_size() for p in params)
                comm.enable_nvls(int(os.environ.get("B2C_NVLS_ARENA_BYTES", 2 * total + (96 << 20))))
        if comm is not None and comm.world > 1:
            comm.autotune()      # measured crossover table, once per comm
        cap = bucket_bytes or _BUCKET_BYTES

        def gloo_allreduce(flat, scale):
            if _st["size"] > 1:
                dist.all_reduce(flat)
                if scale != 1.0:
                    flat.mul_(scale)

        self._red = B2cReducer(params, comm, op_average=(op == Average), bucket_cap_bytes=cap, first_bucket_bytes=cap,
                               fallback_allreduce=gloo_allreduce, world=_st["size"])
        # Horovod semantics: the reduction is completed by synchronize() / step(), not by the end of backward
        self._red._finalize_at_backward_end = False
        self._buckets = [(b.flat, b.params) for b in self._red.buckets]

    @property
    def _synced(self):
        return not (self._red._grads_seen or any(b.launched for b in self._red.buckets))

    def synchronize(self):
        """Wait for every gradient allreduce.  Buckets whose gradients never all arrived (unused
        parameters) are reduced now, with whatever they hold (zeros), like Horovod does."""
        self._red.flush()

    @contextlib.contextmanager
    def skip_synchronize(self):
        self._skip_sync = True
        try:
            yield
        finally:
            self._skip_sync = False

    def step(self, closure=None):
        if not self._skip_sync and not self._synced:
            self.synchronize()
        return self._opt.step(closure) if closure is not None else self._opt.step()

    def zero_grad(self, set_to_none=False):
        self._red.zero_grad()  # one memset per bucket; gradients stay views

    @property
    def param_groups(self):
        return self._opt.param_groups

    @property
    def state(self):
        return self._opt.state

    def state_dict(self):
        return self._opt.state_dict()

    def load_state_dict(self, sd):
        return self._opt.load_state_dict(sd)

    def __getattr__(self, name):
        return getattr(self._opt, name)
