"""oracle/gloo_ref.py — TEST INFRASTRUCTURE, NOT PRODUCT.

Runs the REAL upstream implementation the reference's CPU path resolves to:
RayAccelerator(use_gpu=False) => torch.distributed "gloo" process group over tcp://
(ray_lightning/ray_ddp.py:227,232-237) + stock torch DistributedDataParallel
(entered through ray_ddp.py:218-219).  Ray is absent from this image, so the W actors
(ray_ddp.py:92-97, num_cpus=1 each => one torch thread per worker) are W spawned
processes here; everything below the launcher is the upstream code, unmodified.

Used by: tests/golden/make_golden.py (golden vectors), tests (oracle pinning, CPU),
bench.py (cpu_baseline leg and --impl reference).  Never imported by the product package.
"""
import os
import socket
import sys
import time

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _init(rank, world, port, threads):
    torch.set_num_threads(threads)
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)


# ---------------------------------------------------------------- raw collectives
def _collectives_worker(rank, world, port, jobs, q):
    from oracle import oracle as O
    _init(rank, world, port, 1)
    out = {}
    for job in jobs:
        kind = job["op"]
        if kind == "allreduce":
            x = O.payload(job["payload"], job["n"], rank, job.get("seed", 1234))
            if job["dtype"] == "bf16":
                t = torch.from_numpy(x).to(torch.bfloat16)
            else:
                t = torch.from_numpy(x.copy())
            # torch default hook semantics (default_hooks.py:18-32): div_ by W, then SUM
            t.div_(world)
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
            res = t.view(torch.uint16).numpy().copy() if job["dtype"] == "bf16" else t.numpy().copy()
        elif kind == "broadcast":
            rng = np.random.default_rng(job.get("seed", 77) + rank)
            t = torch.from_numpy(rng.integers(0, 256, size=job["n"], dtype=np.uint8))
            dist.broadcast(t, src=job["root"])
            res = t.numpy().copy()
        else:
            raise ValueError(kind)
        out[job["name"]] = res
    # replica consistency: every rank must hold the same bits
    for name, res in out.items():
        t = torch.from_numpy(res.view(np.uint8).copy())
        ref = t.clone()
        dist.broadcast(ref, src=0)
        assert torch.equal(t, ref), f"gloo replicas differ on {name}"
    if rank == 0:
        q.put(out)
    dist.barrier()
    dist.destroy_process_group()


def run_collectives(world, jobs):
    """jobs: list of dicts {name, op, ...}.  Returns rank 0's {name: ndarray}."""
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_collectives_worker, args=(r, world, port, jobs, q)) for r in range(world)]
    for p in procs:
        p.start()
    out = q.get()
    for p in procs:
        p.join()
        assert p.exitcode == 0
    return out


# ---------------------------------------------------------------- DDP training (loss curve / throughput)
def _ddp_worker(rank, world, port, cfg, q):
    from examples import workloads
    _init(rank, world, port, cfg.get("threads", 1))
    name, steps, warmup, batch = cfg["model"], cfg["steps"], cfg.get("warmup", 0), cfg["batch"]
    model = workloads.build_model(name, seed=cfg.get("seed", 0))
    ddp = torch.nn.parallel.DistributedDataParallel(model, find_unused_parameters=cfg.get("find_unused", True))
    opt = model.configure_optimizers()
    losses = []
    t0 = None
    n_distinct = cfg.get("distinct_batches", steps + warmup)
    cache = {}
    for step in range(warmup + steps):
        if step == warmup:
            dist.barrier()
            t0 = time.perf_counter()
        k = step % n_distinct
        if k not in cache:
            cache[k] = workloads.synthetic_batch(name, batch, rank, k)
        x, y = cache[k]
        opt.zero_grad(set_to_none=True)
        loss = model.loss_fn(ddp(x), y)
        loss.backward()
        opt.step()
        losses.append(float(loss.item()))
    dist.barrier()
    dt = time.perf_counter() - t0
    grads = None
    if cfg.get("return_grads"):
        grads = torch.cat([p.grad.reshape(-1) for p in model.parameters()]).numpy().copy()
    params = torch.cat([p.detach().reshape(-1) for p in model.parameters()]).numpy().copy()
    allq = [None] * world
    dist.all_gather_object(allq, {"losses": losses, "dt": dt})
    if rank == 0:
        q.put({"per_rank": allq, "params": params, "grads": grads})
    dist.barrier()
    dist.destroy_process_group()


def run_ddp(world, model, steps, batch, warmup=0, threads=1, seed=0, return_grads=False, distinct_batches=None,
            find_unused=True):
    """Stock torch DDP over gloo with `world` one-thread workers.  Returns dict with
    per-rank loss curves, the max-over-ranks timed seconds, samples/s and rank-0 params."""
    cfg = dict(model=model, steps=steps, warmup=warmup, batch=batch, threads=threads, seed=seed,
               return_grads=return_grads, find_unused=find_unused)
    if distinct_batches:
        cfg["distinct_batches"] = distinct_batches
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_ddp_worker, args=(r, world, port, cfg, q)) for r in range(world)]
    for p in procs:
        p.start()
    out = q.get()
    for p in procs:
        p.join()
        assert p.exitcode == 0
    dt = max(r["dt"] for r in out["per_rank"])
    out["seconds"] = dt
    out["samples_per_sec"] = world * batch * steps / dt
    out["losses"] = [r["losses"] for r in out["per_rank"]]
    return out


# ---------------------------------------------------------------- gloo allreduce timing (busbw)
def _sweep_worker(rank, world, port, sizes, iters, warmup, q):
    _init(rank, world, port, 1)
    res = []
    for nbytes in sizes:
        t = torch.ones(nbytes // 4, dtype=torch.float32)
        for _ in range(warmup):
            dist.all_reduce(t)
        dist.barrier()
        t0 = time.perf_counter()
        for _ in range(iters):
            dist.all_reduce(t)
        dist.barrier()
        dt = (time.perf_counter() - t0) / iters
        tt = torch.tensor([dt], dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        res.append((nbytes, float(tt.item())))
    if rank == 0:
        q.put(res)
    dist.barrier()
    dist.destroy_process_group()


def run_allreduce_sweep(world, sizes, iters=10, warmup=2):
    """Returns [(bytes, seconds_per_allreduce)], seconds = max over ranks."""
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_sweep_worker, args=(r, world, port, list(sizes), iters, warmup, q)) for r in range(world)]
    for p in procs:
        p.start()
    out = q.get()
    for p in procs:
        p.join()
        assert p.exitcode == 0
    return out
