"""RayAccelerator — DDP training on worker actors, gradients synchronised by the b2c kernels.

Drop-in for ray_lightning/ray_ddp.py: same constructor (ray_ddp.py:79-83), attributes and
method names; what changes underneath (BASELINE.json north_star) is worker set-up and
process-group init: with use_gpu=True every worker additionally creates a b2c communicator
(CUDA-IPC handles swapped through the process group), DDP is built with its three collectives
(C2 initial broadcast, C3 bucket allreduce, C4 buffer broadcast) re-routed to libb2c.so, and
the communicator is closed before the actors are killed.  use_gpu=False is the reference's own
CPU configuration — stock gloo DDP (ray_ddp.py:227) — and is kept as is.

The method sequence on a worker follows what PTL 1.1.7's DDPSpawnAccelerator.ddp_train runs
(entered at ray_ddp.py:218-219): set_world_ranks -> init_ddp_connection -> init_device ->
model_to_device -> optimizers -> get_device_ids -> configure_ddp -> train loop ->
transfer_distrib_spawn_state_on_fit_end.
"""
import os
import socket
from collections import defaultdict
from typing import Callable, List, Optional

import torch

from . import _ray as ray
from .mini_pl import StepModule
from .session import init_session
from .tune import is_session_enabled
from .util import Queue, pin_gpu, process_results


def setup_address():
    """tcp://ip:port of a free port on this worker (the reference borrows the helper from
    ray.util.sgd.torch.utils, ray_ddp.py:10,162-163)."""
    s = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return f"tcp://127.0.0.1:{port}"


class RayExecutor:
    """Runs arbitrary functions inside a worker process (ray_ddp.py:17-31)."""

    def set_env_var(self, key: str, value: str):
        os.environ[key] = value

    def get_node_ip(self):
        return ray.get_node_ip_address()

    def execute(self, fn: Callable, *args, **kwargs):
        return fn(*args, **kwargs)


_RemoteExecutor = ray.remote(RayExecutor)


class RayAccelerator:
    """Args:
        num_workers: number of training workers (one GPU each when use_gpu).
        num_cpus_per_worker: CPUs reserved per worker (README.md:26 documents the same knob as
            ``cpus_per_worker``; both spellings are accepted).
        use_gpu: one GPU per worker; gradients then go through the b2c CUDA library (required —
            there is no fallback).
        init_hook: function run once on every worker right after it starts.
    """

    def __init__(self, num_workers: int = 1, num_cpus_per_worker: int = 1, use_gpu: bool = False,
                 init_hook: Optional[Callable] = None, cpus_per_worker: Optional[int] = None):
        if cpus_per_worker is not None:
            num_cpus_per_worker = cpus_per_worker
        self.trainer = None
        self.nickname = "ddp_ray"
        self.num_workers = num_workers
        self.num_cpus_per_worker = num_cpus_per_worker
        self.use_gpu = use_gpu
        self.workers: List = []
        self.init_hook = init_hook
        self.ddp_plugin = None
        self.b2c_comm = None

    # ------------------------------------------------------------------ driver side
    def _create_worker(self):
        return _RemoteExecutor.options(num_cpus=self.num_cpus_per_worker, num_gpus=int(self.use_gpu)).remote()

    def setup(self, model):
        assert hasattr(self, "trainer") and self.trainer is not None
        if self.use_gpu:
            from . import _lib
            _lib.load()  # fail on the driver, loudly, before any actor starts
        self.trainer.use_ddp = True
        self.trainer.model = model
        self.workers = [self._create_worker() for _ in range(self.num_workers)]
        if self.init_hook:
            ray.get([w.execute.remote(self.init_hook) for w in self.workers])

    def teardown(self):
        def shutdown_remote():
            from ray_lightning_accelerators_b200.ray_ddp import _close_worker_comm
            _close_worker_comm()  # unmap peers + free the arena BEFORE the process group / actor goes away
            if torch.distributed.is_initialized():
                torch.distributed.destroy_process_group()
            if torch.cuda.is_available():
                torch.cuda.empty_cache()

        # Graceful only when training ended normally: after a worker failure its peers may still sit in a control
        # collective and would never reach shutdown_remote — kill them instead of waiting for a watchdog, so the
        # original exception is what the user sees.
        try:
            if not getattr(self, "_failed", False):
                ray.get([w.execute.remote(shutdown_remote) for w in self.workers], timeout=60)
        except Exception:  # noqa: BLE001
            pass
        finally:
            for w in self.workers:
                ray.kill(w, no_restart=True)
            self.workers = []

    def __getstate__(self):
        # actor handles never travel to the workers (intent of ray_ddp.py:123-130)
        d = self.__dict__.copy()
        d["workers"] = []
        d["b2c_comm"] = None
        return d

    def __setstate__(self, d):
        d["workers"] = []
        self.__dict__.update(d)

    def get_local_ranks(self):
        node_ips = ray.get([w.get_node_ip.remote() for w in self.workers])
        seen = defaultdict(int)
        global_to_local = [None] * self.num_workers
        for global_rank, ip in enumerate(node_ips):
            global_to_local[global_rank] = seen[ip]
            seen[ip] += 1
        if self.use_gpu and len(set(node_ips)) > 1:
            raise RuntimeError("b2c maps peers with CUDA IPC inside one NVSwitch domain; workers on "
                               f"several nodes ({sorted(set(node_ips))}) are out of scope")
        return global_to_local

    def train(self):
        if "PL_GLOBAL_SEED" in os.environ:
            seed = os.environ["PL_GLOBAL_SEED"]
            ray.get([w.set_env_var.remote("PL_GLOBAL_SEED", seed) for w in self.workers])
        self.ddp_address = ray.get(self.workers[0].execute.remote(setup_address))
        self.global_to_local = self.get_local_ranks()

        trainer = self.trainer
        assert trainer is not None
        trainer_ref = ray.put(trainer)
        self.trainer = None  # the accelerator is pickled into every worker; not the trainer twice

        queue = None
        if is_session_enabled():
            queue = Queue(actor_options={"num_cpus": 0})

        futures = [self.workers[i].execute.remote(self.train_remote, trainer_ref, i, queue)
                   for i in range(self.num_workers)]
        self._failed = True
        try:
            results = process_results(futures, queue)
            self._failed = False
        finally:
            self.trainer = trainer   # also on failure: the caller's Trainer keeps its accelerator usable
        results, best_path, state_dict = results[0]
        self.trainer.model.load_state_dict(state_dict)
        if self.trainer.checkpoint_callback:
            self.trainer.checkpoint_callback.best_model_path = best_path
        if queue:
            queue.shutdown()
        return results

    # ------------------------------------------------------------------ worker side
    def train_remote(self, trainer, global_rank: int, queue=None):
        assert isinstance(self, RayAccelerator)
        if isinstance(trainer, ray.ObjectRef):
            trainer = ray.get(trainer)
        self.trainer = trainer
        self.trainer.accelerator_backend = self
        self.global_rank = global_rank
        model = self.trainer.model
        if queue is not None:
            init_session(rank=global_rank, queue=queue)
        self.ddp_train(process_idx=global_rank, mp_queue=None, model=model)
        return self.results, self.best_model_path, self.model_state_dict

    def ddp_train(self, process_idx, mp_queue, model):
        seed = os.environ.get("PL_GLOBAL_SEED")
        if seed is not None:
            torch.manual_seed(int(seed))
        self.set_world_ranks(process_idx)
        self.init_ddp_connection(self.trainer.global_rank, self.trainer.world_size)
        self.init_device(process_idx, is_master=(self.trainer.global_rank == 0))
        self.model_to_device(model)
        optimizers, schedulers = _init_optimizers(model)
        device_ids = self.get_device_ids()
        wrapped = self.configure_ddp(model, device_ids)
        if self.use_gpu and os.environ.get("B2C_FUSED_OPT", "0") == "1" and hasattr(wrapped, "fuse_optimizer"):
            # opt-in: optimizer.step() folded into the bucket launches (fused.FusedOptimizer)
            optimizers = [wrapped.fuse_optimizer(o) for o in optimizers]
        device = torch.device("cuda", self.trainer.root_gpu) if self.use_gpu else torch.device("cpu")
        results = self.trainer.run_training(wrapped, model, optimizers, schedulers, device)
        self.transfer_distrib_spawn_state_on_fit_end(model, mp_queue, results)

    def init_ddp_connection(self, global_rank: int, world_size: int, is_slurm_managing_tasks: bool = True) -> None:
        """Process-group creation (ray_ddp.py:222-237) plus, on GPUs, the b2c communicator:
        the torch group only carries control traffic from here on."""
        torch_backend = "nccl" if self.use_gpu else "gloo"
        if not torch.distributed.is_initialized():
            kwargs = {}
            if self.use_gpu:
                idx = pin_gpu(True, self.global_to_local[global_rank])
                kwargs["device_id"] = torch.device("cuda", idx)
            torch.distributed.init_process_group(backend=torch_backend, init_method=self.ddp_address,
                                                 rank=global_rank, world_size=world_size, **kwargs)
        if self.use_gpu and self.b2c_comm is None:
            from .ddp_hook import comm_from_default_group
            self.b2c_comm = comm_from_default_group()
            _set_worker_comm(self.b2c_comm)

    def set_world_ranks(self, process_idx: int):
        self.trainer.local_rank = self.global_to_local[self.global_rank]
        self.trainer.global_rank = self.global_rank
        self.trainer.world_size = self.num_workers

    def init_device(self, process_idx: int, is_master: bool):
        if self.use_gpu:
            # the launcher narrowed CUDA_VISIBLE_DEVICES to one GPU: index 0 (ray_ddp.py:245-253)
            self.trainer.root_gpu = pin_gpu(True, self.trainer.local_rank)

    def get_device_ids(self):
        return [self.trainer.root_gpu] if self.use_gpu else None

    def model_to_device(self, model):
        if self.use_gpu:
            model.cuda(self.trainer.root_gpu)
        else:
            model.cpu()

    def configure_ddp(self, model, device_ids):
        step_module = StepModule(model)
        if self.use_gpu:
            from .ddp_hook import wrap_ddp
            wire = torch.bfloat16 if os.environ.get("B2C_WIRE_DTYPE", "") == "bf16" else None
            return wrap_ddp(step_module, self.b2c_comm, device_ids=device_ids, wire_dtype=wire,
                            find_unused_parameters=True)
        return torch.nn.parallel.DistributedDataParallel(step_module, device_ids=None, find_unused_parameters=True)

    def transfer_distrib_spawn_state_on_fit_end(self, model, mp_queue, results):
        self.results = results
        self.model_state_dict = {k: v.detach().cpu() for k, v in model.state_dict().items()}
        best_model_path = None
        if self.trainer.checkpoint_callback is not None:
            best_model_path = self.trainer.checkpoint_callback.best_model_path
        self.best_model_path = best_model_path

    @property
    def distributed_sampler_kwargs(self):
        kwargs = dict(num_replicas=self.num_workers, rank=self.global_rank)
        if self.ddp_plugin is not None:
            kwargs = self.ddp_plugin.distributed_sampler_kwargs(kwargs)
        return kwargs

    @property
    def require_distributed_sampler(self):
        return True


# one communicator per worker process; closed from teardown's shutdown_remote
_worker_comm = None


def _set_worker_comm(comm):
    global _worker_comm
    _worker_comm = comm


def _close_worker_comm():
    global _worker_comm
    if _worker_comm is not None:
        if torch.cuda.is_available():
            torch.cuda.synchronize()
        if torch.distributed.is_initialized():
            torch.distributed.barrier()  # nobody unmaps while a peer may still be reading
        _worker_comm.close()
        _worker_comm = None


def _init_optimizers(model):
    out = model.configure_optimizers()
    if isinstance(out, torch.optim.Optimizer):
        return [out], []
    if isinstance(out, (list, tuple)) and len(out) == 2 and isinstance(out[0], (list, tuple)):
        return list(out[0]), list(out[1])
    if isinstance(out, (list, tuple)):
        return list(out), []
    if isinstance(out, dict):
        sch = out.get("lr_scheduler")
        return [out["optimizer"]], ([sch] if sch is not None else [])
    raise TypeError(f"configure_optimizers returned {type(out)}")
