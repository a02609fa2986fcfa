#!/usr/bin/env python
"""bench.py — BASELINE.json's metric on BASELINE.json's config.

metric  : train samples/sec (ResNet-50, synthetic 224x224, batch 256 per GPU, fp32 DDP with 25 MB
          buckets through RayAccelerator's worker-side plugin methods), plus the gradient-allreduce
          bus bandwidth of the b2c kernels on that model's real bucket sizes.
launch  : python bench.py --gpus 1 ...            (single process)
          python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
                 --master-port P bench.py --gpus N --steps K --warmup W      (one rank per GPU)
arms    : default          = this repo (libb2c.so kernels on the gradient path; fails loudly without it)
          --impl reference = the reference's own CPU path: RayAccelerator(use_gpu=False) => stock torch
                             DDP over gloo on the host cores (oracle/gloo_ref.py), bounded sample.
One JSON line on stdout from rank 0.  Timing: CUDA events, barrier + synchronize on both sides, max
over ranks; every timed step touches far more than the 126 MB L2 (activations of a 256-image batch).
"""
import argparse
import json
import os
import socket
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

NVLINK_NOMINAL_GBPS = 900.0      # per direction per GPU (BASELINE.json)
NVLINK_MEASURED_COPY_GBPS = 770.0  # peer copy on this pool (B200_PROFILING.md)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0}, "fallback"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled DURING the timed region (recipe's clocks line)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu, self.rows, self.proc = gpu_index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.gpu)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                pass
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------- reference arm (CPU)
def run_reference(args):
    """The reference's own CPU path for this workload: W = --gpus workers, gloo, stock DDP, every
    host thread it can use (ray_ddp.py:92-97 gives each actor num_cpus_per_worker CPUs)."""
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return  # under torchrun only rank 0 works; the other ranks exit 0
    # Under torch.distributed.run the environment carries the elastic agent's rendezvous
    # (TORCHELASTIC_USE_AGENT_STORE=True makes a tcp:// init_process_group connect as a CLIENT to the
    # agent's store, torch/distributed/rendezvous.py:188-196).  The gloo workers spawned below hold
    # their own rendezvous on a fresh port, so every inherited coordinate has to go first.
    for k in list(os.environ):
        if k.startswith(("TORCHELASTIC_", "TORCH_NCCL_", "ROLE_", "GROUP_")) or k in (
                "MASTER_ADDR", "MASTER_PORT", "RANK", "WORLD_SIZE", "LOCAL_RANK", "LOCAL_WORLD_SIZE",
                "GROUP_RANK", "GROUP_WORLD_SIZE", "OMP_NUM_THREADS"):
            os.environ.pop(k)
    from oracle import gloo_ref
    cores = os.cpu_count() or 1
    world = max(1, args.gpus)
    threads = max(1, cores // world)
    batch = _ref_batch(args, world)
    steps = max(1, min(args.steps, args.ref_max_steps))
    warmup = max(1, min(args.warmup, 1))
    r = gloo_ref.run_ddp(world, args.model, steps=steps, batch=batch, warmup=warmup, threads=threads,
                         distinct_batches=2, find_unused=True)
    sps = r["samples_per_sec"]
    sample = (f"{args.model} fp32, stock torch DDP over gloo, {world} worker(s) x {threads} thread(s), "
              f"batch {batch}/worker (bounded sample: the GPU arm's 256/GPU workload at {batch * world} samples per step), "
              f"{steps} timed steps after {warmup} warm-up")
    line = {
        "impl": "reference", "metric": "train_samples_per_sec", "value": sps, "unit": "samples/s",
        "n_gpus": args.gpus, "steps": steps, "warmup": warmup, "ms_per_step": r["seconds"] / steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{_workload_name(args.model)}, torch DDP 25 MB buckets; "
                               f"GPU arm batch {args.batch}/GPU, CPU reference arm batch {batch}/worker",
                   "global_batch": batch * world, "parallelism": f"dp{world}", "backend": "gloo"},
        "cpu_baseline": {"value": sps, "unit": "samples/s", "cores": threads * world, "kind": "reference",
                         "sample": sample, "host_cores": cores},
        "e2e": {"value": sps, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------- b2c arm (GPU)
def _workload_name(model):
    return {"resnet50": "resnet50 synthetic 224x224 fp32 (BASELINE configs[2])",
            "mnist_mlp": "MNISTClassifier 784-32-64-10 MLP, synthetic 28x28 (BASELINE configs[1])",
            "bert_base": "BERT-base MLM seq 128 (BASELINE configs[3])"}.get(model, model)


def _ref_batch(args, world):
    """Per-worker batch of the CPU arm: the GPU arm's 256 images when one worker has the whole host,
    256 // W with W workers, so that a step costs about the same host time at every N."""
    return args.ref_batch if args.ref_batch > 0 else max(1, args.batch // max(1, world))


def _build_worker(args, rank, world, local):
    """Worker set-up through the reference-facing plugin: the same RayAccelerator methods a Ray actor
    runs inside train_remote (ray_ddp.py:199-267), in the order PTL's ddp_train calls them."""
    import torch
    from examples import workloads
    from ray_lightning_accelerators_b200 import RayAccelerator
    from ray_lightning_accelerators_b200.mini_pl import Trainer

    acc = RayAccelerator(num_workers=world, num_cpus_per_worker=1, use_gpu=True)
    trainer = Trainer(gpus=1, accelerator=acc, checkpoint_callback=False)
    acc.trainer = trainer
    acc.global_rank = rank
    acc.global_to_local = list(range(world)) if torch.cuda.device_count() > 1 else [0] * world
    addr = os.environ.get("MASTER_ADDR", "127.0.0.1")
    port = os.environ.get("MASTER_PORT") or str(_free_port())
    acc.ddp_address = f"tcp://{addr}:{port}"
    acc.set_world_ranks(rank)
    acc.init_ddp_connection(rank, world)          # process group + b2c communicator (IPC handles)
    acc.init_device(rank, is_master=(rank == 0))
    model = workloads.build_model(args.model, seed=0)
    acc.model_to_device(model)
    opt = model.configure_optimizers()
    if args.impl == "nccl":   # on-box comparator: what RayAccelerator(use_gpu=True) resolves to upstream (ray_ddp.py:227)
        ddp = torch.nn.parallel.DistributedDataParallel(model, device_ids=acc.get_device_ids(), find_unused_parameters=True)
    else:
        ddp = _wrap(acc, model, args)
    return acc, model, ddp, opt


def _wrap(acc, model, args):
    """configure_ddp, but wrapping the bare module: the bench calls forward(x) directly."""
    import torch
    from ray_lightning_accelerators_b200.ddp_hook import wrap_ddp
    wire = torch.bfloat16 if args.wire == "bf16" else None
    return wrap_ddp(model, acc.b2c_comm, device_ids=acc.get_device_ids(), wire_dtype=wire,
                    find_unused_parameters=True, reducer=args.reducer)


def _bucket_stats(state):
    """Per-launch device times recorded by the reducer / hook (CUDA events on the comm stream) around every b2c
    launch of the timed region, averaged per bucket size."""
    by_size = {}
    for us, nbytes in state.collect_timing_us():
        by_size.setdefault(nbytes, []).append(us)
    return {nb: sum(v) / len(v) for nb, v in by_size.items()}, sum(len(v) for v in by_size.values())


def _phase_split(comm, seq_bytes):
    """Attribute the in-step time of each bucket launch with the kernels' own phase clock (channel 0,
    %globaltimer): the ring keeps the last 32 kernels of the comm; `seq_bytes` is the per-step launch sequence
    (bucket bytes, or None for a launch that is not a bucket).  Returns {bytes: (wait_for_peers_us, data_us, kernel_us)}."""
    rows, total = comm.phase_ring_us()
    L = len(seq_bytes)
    if L == 0 or len(rows) < L:
        return {}
    steps = len(rows) // L
    rows = rows[len(rows) - steps * L:]
    acc = {}
    for i, (to_a, a_to_b, b_to_end, _) in enumerate(rows):
        nb = seq_bytes[i % L]
        if nb is None:
            continue
        acc.setdefault(nb, []).append((to_a, a_to_b, to_a + a_to_b + b_to_end))
    return {nb: tuple(sum(x[k] for x in v) / len(v) for k in range(3)) for nb, v in acc.items()}


PARITY_SIZES = [12415000, 31500000, 29664000, 27021000, 1625000, 4 * 27883]  # ResNet-50 buckets + one ragged size


def _parity_block(comm, rank, world, dev):
    """Cross-GPU correctness where the driver can see it: outside the timed region, on the real bucket sizes.
    b2c_allreduce(ALGO_AUTO) against torch.distributed all_reduce on a clone pre-divided by W (what the reference's
    default hook computes, default_hooks.py:18-32), payload families int / randn / wide / special, fp32 and bf16;
    b2c_broadcast against dist.broadcast for every root.  Bars are north_star's."""
    import torch
    import torch.distributed as dist
    from ray_lightning_accelerators_b200.comm import ALGO_AUTO
    names = {1: "oneshot", 2: "twoshot", 3: "nvls"}
    rows, ok = [], True

    def gen(kind, n, dtype):
        g = torch.Generator(device=dev).manual_seed(4242 + 17 * rank + n % 1000)
        if kind == "int":
            x = torch.randint(-8, 9, (n,), generator=g, device=dev).float()
        elif kind == "randn":
            x = torch.randn(n, generator=g, device=dev)
        elif kind == "wide":
            x = torch.randn(n, generator=g, device=dev) * torch.exp2(torch.randint(-20, 20, (n,), generator=g, device=dev).float())
        else:  # special values: zeros of both signs, infinities, NaN, denormals, large finite
            base = torch.tensor([0.0, -0.0, float("inf"), float("-inf"), float("nan"), 1e-42, -1e-42, 3.0e37,
                                 1.0, -1.0, 65504.0, 1.1754944e-38], device=dev)
            x = base.repeat(n // base.numel() + 1)[:n].roll(rank % 2)
        return x.to(dtype)

    for dtype, tag in ((torch.float32, "f32"), (torch.bfloat16, "bf16")):
        for S in PARITY_SIZES:
            n = S // (4 if dtype == torch.float32 else 2)
            for kind in ("int", "randn", "wide", "special"):
                src = gen(kind, n, dtype)
                ref = src.clone().div_(world)
                dist.all_reduce(ref)
                mag = src.float().abs() / world
                dist.all_reduce(mag)                       # sum_r |g_r| / W: the error denominator
                exact = src.double() / world
                dist.all_reduce(exact)                     # the mean in fp64: judges both arms in bf16
                t = None
                if comm.nvls_enabled:
                    try:
                        t = comm.nvls_empty(n, dtype)
                    except MemoryError:
                        t = None
                arena = t is not None
                if t is None:
                    t = torch.empty(n, dtype=dtype, device=dev)
                    comm.register(t)
                t.copy_(src)
                comm.allreduce(t, algo=ALGO_AUTO)
                torch.cuda.synchronize()
                comm.check_alive()
                path = names[comm.stats()["last_algo"]]
                got, want = t.float(), ref.float()
                ibits = torch.int16 if dtype == torch.bfloat16 else torch.int32
                fin = torch.isfinite(want) & torch.isfinite(got) & (mag < 3.0e38)
                den = mag[fin].clamp_min(1.1754944e-38)

                def worst(a, b):
                    return float(((a - b).abs()[fin] / den).max().item()) if bool(fin.any()) else 0.0

                err = worst(got, want)                                  # b2c against the reference path (NCCL)
                err_exact = worst(got.double(), exact)                  # b2c against the fp64 mean
                ref_err_exact = worst(want.double(), exact)             # NCCL against the fp64 mean
                nan_ok = bool((torch.isnan(got) == torch.isnan(want)).all())
                inf_ok = bool((got[torch.isinf(want)] == want[torch.isinf(want)]).all())
                bit = bool(torch.equal(t.view(ibits), ref.view(ibits)))
                rep = t.clone()
                dist.broadcast(rep, src=0)
                same = bool(torch.equal(rep.view(ibits), t.view(ibits)))
                if kind == "int":
                    num_ok = bit
                elif dtype == torch.float32:
                    num_ok = err <= 1e-6
                else:
                    # bf16: within one bf16 ulp (2^-7 of the magnitude) of the exact mean, and no less accurate than
                    # the reference path beyond half an ulp.  NCCL's ring / tree kernels accumulate bf16 hop by hop at
                    # small sizes, so b2c-vs-NCCL itself can exceed one ulp while b2c is the closer of the two.
                    num_ok = err_exact <= 2.0 ** -7 and err_exact <= ref_err_exact + 2.0 ** -9
                good = same and nan_ok and inf_ok and num_ok
                flags = torch.tensor([0 if good else 1], device=dev)
                dist.all_reduce(flags)
                good = int(flags.item()) == 0
                ok = ok and good
                rows.append({"op": "allreduce", "dtype": tag, "bytes": S, "payload": kind, "path": path, "in_arena": arena,
                             "max_rel_err": err, "max_rel_err_vs_fp64_mean": err_exact, "nccl_max_rel_err_vs_fp64_mean": ref_err_exact,
                             "bit_exact_vs_nccl": bit, "replicas_identical": same, "ok": good})
                if arena:
                    comm.nvls_free(t)
                else:
                    comm.deregister(t)
    for S in (212904, 102441032):
        for root in range(world):
            g = torch.Generator(device=dev).manual_seed(9000 + rank)
            src = torch.randint(0, 256, (S,), generator=g, device=dev, dtype=torch.uint8)
            a, b = src.clone(), src.clone()
            dist.broadcast(a, src=root)
            m0 = comm.stats()["bcast_mc_launches"]
            comm.broadcast(b, root=root)
            torch.cuda.synchronize()
            good = bool(torch.equal(a, b))
            flags = torch.tensor([0 if good else 1], device=dev)
            dist.all_reduce(flags)
            good = int(flags.item()) == 0
            ok = ok and good
            rows.append({"op": "broadcast", "bytes": S, "root": root, "bytes_equal_vs_nccl": good,
                         "path": "multicast" if comm.stats()["bcast_mc_launches"] > m0 else "pull", "ok": good})
    summary = {}
    for r in rows:
        if r["op"] == "allreduce":
            k = f'{r["dtype"]}/{r["path"]}'
            if r["payload"] in ("randn", "wide"):
                e = summary.setdefault(k, {"vs_nccl": 0.0, "vs_fp64_mean": 0.0, "nccl_vs_fp64_mean": 0.0})
                e["vs_nccl"] = max(e["vs_nccl"], r["max_rel_err"])
                e["vs_fp64_mean"] = max(e["vs_fp64_mean"], r["max_rel_err_vs_fp64_mean"])
                e["nccl_vs_fp64_mean"] = max(e["nccl_vs_fp64_mean"], r["nccl_max_rel_err_vs_fp64_mean"])
    return {"ok": ok, "checks": len(rows), "failed": [r for r in rows if not r["ok"]][:8],
            "paths": sorted({r["path"] for r in rows}), "max_rel_err_by_dtype_and_path": summary,
            "bars": "int payloads bit-exact vs NCCL; fp32 <= 1e-6 of sum|g|/W vs NCCL; bf16 within one bf16 ulp (2^-7 of sum|g|/W) of "
                    "the fp64 mean and no less accurate than NCCL's own result (+ half an ulp); NaN/inf classes equal; replicas "
                    "bitwise identical; broadcast byte-equal vs NCCL for every root",
            "rows": [r for r in rows if r["bytes"] in (31500000, 102441032) and r.get("payload", "randn") == "randn"]}


def run_b2c(args):
    import torch
    import torch.distributed as dist
    from examples import workloads

    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch N>1 with torch.distributed.run")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py (b2c arm) needs a CUDA device; there is no CPU fallback. "
                         "Use --impl reference for the reference's CPU path.")
    torch.backends.cudnn.benchmark = True
    # room in the multicast arena for the parity / allreduce sections' buffers as well
    os.environ.setdefault("B2C_NVLS_ARENA_BYTES", str(2 << 30))
    acc, model, ddp, opt = _build_worker(args, rank, world, local)
    stock = args.impl == "nccl"
    comm, state = acc.b2c_comm, (None if stock else ddp._b2c_state)
    own = (not stock) and hasattr(ddp, "fuse_optimizer")
    fused = False
    if own and args.fused_opt:
        opt = ddp.fuse_optimizer(opt)
        fused = type(opt).__name__ == "FusedOptimizer"
    dev = torch.device("cuda", torch.cuda.current_device())
    batch = args.batch
    K, W = args.steps, max(3, args.warmup)

    # ---- leg 1 ("value"): inputs already resident in HBM when the timed region starts
    x_dev, y_dev = [t.to(dev) for t in workloads.synthetic_batch(args.model, batch, rank, 0)]

    def step(x, y):
        opt.zero_grad(set_to_none=True)
        loss = model.loss_fn(ddp(x), y)
        loss.backward()
        opt.step()
        return loss

    for _ in range(W):
        step(x_dev, y_dev)
    torch.cuda.synchronize()
    dist.barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    if not stock and not args.no_bucket_timing:
        state.enable_timing()
    st0 = comm.stats()
    launches0 = st0["launches"]
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(K):
        step(x_dev, y_dev)
    e1.record()
    torch.cuda.synchronize()
    dist.barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    clocks = sampler.stop() if rank == 0 else None
    st1 = comm.stats()
    launches = st1["launches"] - launches0
    opt_name = type(getattr(opt, "_opt", opt)).__name__
    what = f"gradient average + {opt_name} step in one launch per bucket (weights travel back)" if fused else "gradient allreduce"
    if st1["nvls_launches"] > st0["nvls_launches"]:
        grad_kernel = "k_fused_step_nvls<SGD>" if fused else "k_nvls_inplace<f32>"
        grad_path = f"libb2c.so: {what}; in-switch reduction (multimem), buckets inside the multicast arena, in place"
    elif st1["zerocopy_launches"] > st0["zerocopy_launches"]:
        grad_kernel = f"k_fused_step_zc<SGD,W={world}>" if fused else f"k_allreduce_twoshot_zc<f32,W={world}>"
        grad_path = f"libb2c.so: {what}; zero-copy two-shot over peer-mapped buckets"
    elif world > 1:
        grad_kernel, grad_path = f"k_allreduce_twoshot<f32,W={world}>", "libb2c.so: staged two-shot / one-shot through the arena"
    elif fused:
        grad_kernel, grad_path = "k_fused_step_local<SGD>", f"libb2c.so: {opt_name} step over the gradient buckets (W=1: no peer to reduce with)"
    else:
        grad_kernel, grad_path = None, "W=1 and no fused step: nothing for libb2c.so to do inside a step"
    if stock:
        grad_kernel, grad_path = "ncclDevKernel_AllReduce_*", "stock torch DDP: c10d Reducer + NCCL all_reduce (libb2c.so not on the gradient path)"
    per_bucket_us, n_timed = _bucket_stats(state) if (not stock and state.timing is not None) else ({}, 0)
    if not stock:
        state.timing = None
    # device-side attribution of the in-step launches (own reducer: the per-step launch sequence is known)
    phases = {}
    if own and world > 1:
        seq = ([None] if (ddp.bufsync.flat is not None and not getattr(ddp, "_c4_folded", False)) else []) + \
              [b.nbytes for b in ddp.reducer.buckets]
        try:
            phases = _phase_split(comm, seq)
        except Exception as e:  # noqa: BLE001
            phases = {"error": str(e)[:120]}
    ms_total = float(ms.item())
    value = world * batch * K / (ms_total / 1e3)

    # ---- leg 2 ("e2e"): host buffers; H2D of the step's batch from pinned memory and a D2H read of
    # the loss inside the timed region, every step
    host = [workloads.synthetic_batch(args.model, batch, rank, s) for s in range(2)]
    host = [(x.pin_memory(), y.pin_memory()) for x, y in host]
    h2d = host[0][0].numel() * host[0][0].element_size() + host[0][1].numel() * host[0][1].element_size()

    copy_stream = torch.cuda.Stream()

    def prefetch(s):
        """H2D of step s's batch from pinned host memory on a copy stream (double buffering, what a
        pinned-memory data loader does); the step waits for it on the device."""
        xh, yh = host[s % 2]
        with torch.cuda.stream(copy_stream):
            x = xh.to(dev, non_blocking=True)
            y = yh.to(dev, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(copy_stream)
        return x, y, ev

    def e2e_run(n):
        nxt = prefetch(0)
        last_ = None
        for s in range(n):
            x, y, ev = nxt
            cur = torch.cuda.current_stream()
            cur.wait_event(ev)
            x.record_stream(cur)
            y.record_stream(cur)
            if s + 1 < n:
                nxt = prefetch(s + 1)          # overlaps with this step's compute
            last_ = float(step(x, y).item())   # device -> host read of the result, every step
        return last_

    e2e_run(2)
    torch.cuda.synchronize()
    dist.barrier()
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record()
    last = e2e_run(K)                          # K host->device batch copies inside the timed region
    f1.record()
    torch.cuda.synchronize()
    dist.barrier()
    ms2 = torch.tensor([f0.elapsed_time(f1)], device=dev, dtype=torch.float64)
    dist.all_reduce(ms2, op=dist.ReduceOp.MAX)
    e2e_value = world * batch * K / (float(ms2.item()) / 1e3)

    # ---- roofline of the dominant b2c kernel of the timed region (largest share of our device time)
    peaks, peak_kind = _peaks()
    # One kernel carries every bucket, so "the dominant kernel's average launch" is the mean over ALL its launches of
    # the timed region: achieved = sum of algorithmic bytes / sum of CUDA-event durations (the 4 KB bucket included).
    roofline = None
    if per_bucket_us:
        sizes_ = sorted(per_bucket_us)
        n_l = len(sizes_)
        us = sum(per_bucket_us[nb] for nb in sizes_) / n_l
        dom = max(sizes_)
        if world == 1:
            # fused SGD+momentum step, local: read g, w, m and write w, m = 20 bytes per fp32 parameter
            alg = sum(5.0 * nb for nb in sizes_) / n_l
            roofline = {"bound": "hbm", "kernel": grad_kernel, "achieved": alg / us / 1e3,
                        "peak": peaks["hbm_gbs"], "peak_kind": f"{peak_kind} HBM copy", "unit": "GB/s",
                        "frac": alg / us / 1e3 / peaks["hbm_gbs"], "traffic": None,
                        "bytes_per_launch": alg, "us_per_launch": us, "launches_per_step": n_l,
                        "largest_bucket": {"bytes": dom, "us": per_bucket_us[dom],
                                           "frac": 5.0 * dom / per_bucket_us[dom] / 1e3 / peaks["hbm_gbs"]},
                        "note": "in-step: the kernel shares the GPU with the tail of backward; mean over all bucket launches of a step"}
        else:
            f = 2.0 * (world - 1) / world   # NVLink bytes per GPU per direction per bucket byte (busbw numerator)
            alg = sum(f * nb for nb in sizes_) / n_l
            roofline = {"bound": "nvlink", "kernel": grad_kernel,
                        "achieved": alg / us / 1e3, "peak": NVLINK_NOMINAL_GBPS,
                        "peak_kind": "nominal NVLink 5 per direction (measured peer copy 770 GB/s)", "unit": "GB/s",
                        "frac": alg / us / 1e3 / NVLINK_NOMINAL_GBPS,
                        "frac_of_measured_copy": alg / us / 1e3 / NVLINK_MEASURED_COPY_GBPS, "traffic": None,
                        "bytes_per_launch": alg, "us_per_launch": us, "launches_per_step": n_l,
                        "largest_bucket": {"bytes": dom, "us": per_bucket_us[dom],
                                           "frac": f * dom / per_bucket_us[dom] / 1e3 / NVLINK_NOMINAL_GBPS},
                        "note": "in-step CUDA-event time, mean over all bucket launches of a step: it includes the wait for the "
                                "slowest rank at the entry barrier (grad_buckets[*].wait_for_peers_us) — rank skew, not kernel work"}
            ph = [phases[nb] for nb in sizes_ if nb in phases and isinstance(phases[nb], tuple)]
            if len(ph) == n_l:
                data_us = sum(p_[1] for p_ in ph) / n_l
                roofline["data_phase_us"] = data_us
                roofline["frac_data_phase"] = alg / data_us / 1e3 / NVLINK_NOMINAL_GBPS if data_us > 0 else None
                roofline["wait_for_peers_us"] = sum(p_[0] for p_ in ph) / n_l
    if roofline is not None:
        try:  # per-launch DRAM bytes of this kernel from the committed ncu capture
            with open(os.path.join(ROOT, "profiles", "r2_ncu_traffic.json")) as f:
                cap = json.load(f).get(roofline["kernel"])
            if cap:
                roofline["traffic"] = cap["dram_bytes_read"] + cap["dram_bytes_write"]
                roofline["traffic_note"] = cap.get("note")
            elif world > 1:
                roofline["traffic_note"] = ("null: the cross-GPU kernels cannot run under ncu on this pool (flag barriers with live peers; "
                                            "multi-rank commands under ncu are ruled out) and the NVLink counters are NOT_SUPPORTED "
                                            "(profiles/r2_nvlink_bytes_w8.json); DRAM traffic of the same kernel family on one GPU: "
                                            "profiles/r2_ncu_fused_summary.md")
        except Exception:
            pass
    buckets = []
    for nb in sorted(per_bucket_us):
        row = {"bytes": nb, "us": per_bucket_us[nb],
               "busbw_GBps": (2.0 * (world - 1) / world * nb / per_bucket_us[nb] / 1e3) if world > 1 else None}
        if nb in phases and isinstance(phases[nb], tuple):
            wait, data, kern = phases[nb]
            row.update({"wait_for_peers_us": wait, "data_phase_us": data, "kernel_us": kern,
                        "queue_and_launch_us": per_bucket_us[nb] - kern,
                        "busbw_data_phase_GBps": 2.0 * (world - 1) / world * nb / data / 1e3 if data > 0 else None})
        buckets.append(row)

    parity = _parity_block(comm, rank, world, dev) if (world > 1 and not stock and not args.no_parity) else None
    ar_sizes = None
    if own:
        # the buckets this reducer really launches (16-byte multiples) plus torch DDP's two largest for continuity
        ar_sizes = sorted({b.nbytes for b in ddp.reducer.buckets if b.nbytes >= (1 << 20)} | {31500000, 29664000})
    allreduce = _allreduce_section(args, comm, rank, world, dev, ar_sizes) if (not stock and not args.no_allreduce_section) else None
    cpu_baseline = _cpu_baseline(args) if (rank == 0 and world == 1 and not args.no_cpu_baseline) else None

    if rank == 0:
        line = {
            "impl": "b2c" if not stock else "nccl (stock torch DDP, on-box comparator; not the reference arm)",
            "metric": "train_samples_per_sec", "value": value, "unit": "samples/s", "n_gpus": world, "steps": K,
            "warmup": W, "ms_per_step": ms_total / K, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32" if args.wire != "bf16" else "f32 grads on a bf16 wire",
            "data": "synthetic",
            "config": {"workload": f"{_workload_name(args.model)}, batch {batch}/GPU, {opt_name}, 25 MB gradient "
                                   f"buckets via RayAccelerator(use_gpu=True) plugin methods; "
                                   f"the CPU reference arm (--impl reference) runs batch {_ref_batch(args, world)}/worker",
                       "global_batch": batch * world, "parallelism": f"dp{world}",
                       "l2": "inputs larger than L2 (activations of a 256-image batch >> 126 MB); no flush needed",
                       "reducer": "torch c10d Reducer + NCCL" if stock else ("b2c (own bucket reducer)" if own else "torch c10d Reducer + b2c comm hook"),
                       "fused_optimizer": fused, "grad_path": grad_path},
            "e2e": {"value": e2e_value, "unit": "samples/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
                    "ms_per_step": float(ms2.item()) / K, "last_loss": last},
            "gpu_launches": int(launches), "b2c_launches_timed": int(n_timed),
            "clocks": clocks, "roofline": roofline, "grad_buckets": buckets, "parity": parity, "allreduce": allreduce,
            "crossover_table": getattr(comm, "tuning", None),
        }
        if cpu_baseline is not None:
            line["cpu_baseline"] = cpu_baseline
        print(json.dumps(line), flush=True)
    dist.barrier()
    from ray_lightning_accelerators_b200.ray_ddp import _close_worker_comm
    _close_worker_comm()
    dist.destroy_process_group()
    if parity is not None and not parity["ok"]:
        raise SystemExit("parity block failed: see the 'parity' object of the JSON line")


# ---------------------------------------------------------------------------- Horovod-style arm
def run_hvd(args):
    """BASELINE configs[3]: the HorovodRayAccelerator worker path (ray_horovod.py:146-169) on the hvd shim —
    hvd.init, LR x size, broadcast_parameters / optimizer state, DistributedOptimizer, backward +
    synchronize() + step — with the model held in --param-dtype (bf16 for BERT: bf16 gradients, fp32
    accumulate in the kernel, one rounding).  --hvd-backend nccl runs the same loop with a per-bucket NCCL
    all_reduce instead of libb2c.so, as the on-box comparator."""
    import torch
    import torch.distributed as dist
    from examples import workloads
    from ray_lightning_accelerators_b200 import HorovodRayAccelerator, hvd_shim as hvd
    from ray_lightning_accelerators_b200.mini_pl import Trainer

    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    use_gpu = torch.cuda.is_available()
    if not use_gpu and not args.allow_cpu:
        raise SystemExit("bench.py --impl hvd needs a CUDA device (or --allow-cpu for the gloo plumbing check)")
    addr = os.environ.get("MASTER_ADDR", "127.0.0.1")
    port = os.environ.get("MASTER_PORT") or str(_free_port())
    hvd.init(rank=rank, size=world, local_rank=local, address=f"tcp://{addr}:{port}", use_gpu=use_gpu)
    stock = args.hvd_backend == "nccl"
    comm = hvd._st["comm"]
    if stock and comm is not None:   # comparator: drop the b2c communicator, the shim then uses the process group
        hvd._st["comm"] = None
    acc = HorovodRayAccelerator(num_hosts=1, num_slots=world, use_gpu=use_gpu)
    acc.trainer = Trainer(gpus=int(use_gpu), accelerator=acc, checkpoint_callback=False)
    model = workloads.build_model(args.model, seed=0)
    if args.param_dtype == "bf16":
        model = model.to(torch.bfloat16)
    optimizers, _, dev = acc._horovod_setup(model)
    opt = optimizers[0]
    batch, K, W = args.batch, args.steps, max(3, args.warmup)
    x, y = [t.to(dev) if t is not None else None for t in workloads.synthetic_batch(args.model, batch, rank, 0)]

    def step():
        opt.zero_grad()
        loss = model.loss_fn(model(x), y)
        loss.backward()
        opt.synchronize()
        with opt.skip_synchronize():
            opt.step()
        return loss

    def sync():
        if use_gpu:
            torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    for _ in range(W):
        step()
    sync()
    st0 = comm.stats() if (comm is not None and not stock) else None
    if use_gpu:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
    t0 = time.perf_counter()
    for _ in range(K):
        loss = step()
    if use_gpu:
        e1.record()
    sync()
    ms = e0.elapsed_time(e1) if use_gpu else (time.perf_counter() - t0) * 1e3
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    st1 = comm.stats() if st0 is not None else None
    grad_bytes = sum(f.numel() * f.element_size() for f, _ in opt._buckets)
    if rank == 0:
        line = {"impl": "b2c-hvd" if not stock else "hvd loop on NCCL all_reduce (on-box comparator)",
                "metric": "train_samples_per_sec", "value": world * batch * K / (ms / 1e3), "unit": "samples/s",
                "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms / K, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": args.param_dtype, "data": "synthetic",
                "config": {"workload": f"{args.model} seq 128 batch {batch}/GPU via HorovodRayAccelerator worker path "
                                       f"(BASELINE configs[3])", "global_batch": batch * world, "parallelism": f"dp{world}",
                           "grad_buckets": [f.numel() * f.element_size() for f, _ in opt._buckets],
                           "grad_bytes_per_step": grad_bytes},
                "last_loss": float(loss.item()),
                "gpu_launches": int(st1["launches"] - st0["launches"]) if st1 else 0,
                "paths": {k: int(st1[k] - st0[k]) for k in ("nvls_launches", "zerocopy_launches", "oneshot_launches",
                                                            "twoshot_launches")} if st1 else None}
        print(json.dumps(line), flush=True)
    if stock and comm is not None:
        hvd._st["comm"] = comm
    hvd.shutdown()


RESNET50_BUCKETS = [12415000, 31500000, 29664000, 27021000, 1625000]  # bytes, launch order


def _allreduce_section(args, comm, rank, world, dev, sizes=None):
    """The second half of BASELINE.json's metric: gradient-allreduce bus GB/s on the real bucket sizes.
    N>1: over NVLink, b2c vs stock NCCL in the same processes.  N=1: there is no link; the same kernels
    run as 8 virtual ranks on the one GPU (HBM-bound), reported against the measured HBM peak."""
    import torch
    import torch.distributed as dist
    from ray_lightning_accelerators_b200 import _lib
    from ray_lightning_accelerators_b200.comm import ALGO_AUTO, ALGO_TWOSHOT, LocalGroup
    iters = args.ar_iters
    sizes = list(sizes or RESNET50_BUCKETS)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    out = []
    peaks, peak_kind = _peaks()

    def timed(fn, barrier):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        ts = []
        for _ in range(iters):
            flush.fill_(1)               # L2 flush between timed launches
            torch.cuda._sleep(200000)    # keep the GPU busy so the host runs ahead of the stream
            barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); fn(); b.record(); b.synchronize()
            ts.append(a.elapsed_time(b) * 1e3)
        t = torch.tensor(ts, device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.sort().values[len(ts) // 2])

    def pipelined(make, fn, S, free=None):
        """Back-to-back launches over distinct buffers whose total footprint exceeds L2 (no flush
        needed): launch latency and barrier skew overlap with the previous kernel, as they do for the
        buckets of a real backward pass.  One event pair around 3 rounds; per-allreduce time."""
        k = max(2, (192 << 20) // S + 1)
        bufs = [make() for _ in range(k)]
        for b in bufs:
            fn(b)
        torch.cuda.synchronize()
        if world > 1:
            comm.barrier()
        a, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(3):
            for b in bufs:
                fn(b)
        b_.record()
        b_.synchronize()
        t = torch.tensor([a.elapsed_time(b_) * 1e3 / (3 * k)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if free is not None:
            for b in reversed(bufs):
                free(b)
        return float(t.item())

    if world > 1:
        for S in sizes:
            if comm.nvls_enabled:
                try:
                    x = comm.nvls_empty(S // 4)   # lives in the multicast arena: in-switch reduction, in place
                    x.normal_()
                except MemoryError:
                    x = torch.randn(S // 4, device=dev)
            else:
                x = torch.randn(S // 4, device=dev)
            if not comm.in_nvls_arena(x):
                comm.register(x)
            us = timed(lambda: comm.allreduce(x, algo=ALGO_AUTO), comm.barrier)
            path = {1: "oneshot", 2: "twoshot", 3: "nvls"}[comm.stats()["last_algo"]]
            y = torch.randn(S // 4, device=dev)
            us_nccl = timed(lambda: dist.all_reduce(y), comm.barrier)
            bus = 2.0 * (world - 1) / world * S
            row = {"bytes": S, "path": path, "us": us, "busbw_GBps": bus / us / 1e3, "frac_of_900": bus / us / 1e3 / NVLINK_NOMINAL_GBPS,
                   "nccl_us": us_nccl, "nccl_busbw_GBps": bus / us_nccl / 1e3}
            if S >= (8 << 20):
                def make():
                    if comm.nvls_enabled:
                        try:
                            t_ = comm.nvls_empty(S // 4)
                            t_.normal_()
                            return t_
                        except MemoryError:
                            pass
                    t_ = torch.randn(S // 4, device=dev)
                    comm.register(t_)
                    return t_
                def free(t_):
                    if comm.in_nvls_arena(t_):
                        comm.nvls_free(t_)
                    else:
                        comm.deregister(t_)
                up = pipelined(make, lambda t_: comm.allreduce(t_, algo=ALGO_AUTO), S, free)
                un = pipelined(lambda: torch.randn(S // 4, device=dev), lambda t_: dist.all_reduce(t_), S)
                row.update({"us_pipelined": up, "busbw_pipelined_GBps": bus / up / 1e3,
                            "frac_of_900_pipelined": bus / up / 1e3 / NVLINK_NOMINAL_GBPS,
                            "nccl_us_pipelined": un, "nccl_busbw_pipelined_GBps": bus / un / 1e3})
            out.append(row)
            if comm.in_nvls_arena(x):
                comm.nvls_free(x)
            else:
                comm.deregister(x)
        return {"mode": f"nvlink W={world}",
                "timing": "us: median of %d isolated launches, L2 flushed, ranks aligned, max over ranks; "
                          "us_pipelined: back-to-back launches over distinct buffers > L2, max over ranks" % iters,
                "buckets": out}
    g = LocalGroup(8, device=dev.index or 0, arena_bytes=8 << 20)
    g.set_param(_lib.PARAM_TIMEOUT_MS, 3000)
    for S in sizes:
        ts = [torch.randn(S // 4, device=dev) for _ in range(8)]
        us = timed(lambda: g.allreduce(ts, algo=ALGO_TWOSHOT), lambda: None)
        alg = 2.0 * 8 * S  # every virtual rank reads S and writes S (zero-copy two-shot, in place)
        out.append({"bytes": S, "us": us, "hbm_GBps": alg / us / 1e3, "frac_of_hbm_peak": alg / us / 1e3 / peaks["hbm_gbs"]})
    g.close()
    return {"mode": "one GPU, 8 virtual ranks (HBM-bound emulation, no NVLink at N=1)", "peak": peaks["hbm_gbs"],
            "peak_kind": f"{peak_kind} HBM copy GB/s", "timing": "median of %d, L2 flushed" % iters, "buckets": out}


def _cpu_baseline(args):
    """The reference's CPU path on this box's host cores, bounded sample (~10-30 s of CPU work):
    2 gloo workers (config #1's worker count) sharing all cores."""
    from oracle import gloo_ref
    cores = os.cpu_count() or 1
    world = 2
    threads = max(1, cores // world)
    rb = args.ref_batch if args.ref_batch > 0 else 16
    r = gloo_ref.run_ddp(world, args.model, steps=2, batch=rb, warmup=1, threads=threads, distinct_batches=1)
    return {"value": r["samples_per_sec"], "unit": "samples/s", "cores": world * threads, "host_cores": cores, "kind": "reference",
            "sample": f"{args.model} fp32, stock torch DDP over gloo (what RayAccelerator(use_gpu=False) runs, ray_ddp.py:227), "
                      f"{world} workers x {threads} threads, batch {rb}/worker, 2 timed steps after 1 warm-up"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b2c", choices=["b2c", "reference", "nccl", "hvd"])
    ap.add_argument("--param-dtype", default="f32", choices=["f32", "bf16"])
    ap.add_argument("--hvd-backend", default="b2c", choices=["b2c", "nccl"])
    ap.add_argument("--allow-cpu", action="store_true")
    ap.add_argument("--model", default="resnet50")
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--wire", default="f32", choices=["f32", "bf16"])
    ap.add_argument("--ar-iters", type=int, default=15)
    ap.add_argument("--ref-batch", type=int, default=0, help="CPU reference arm: batch per worker (0 = batch // workers)")
    ap.add_argument("--ref-max-steps", type=int, default=3)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--reducer", default="b2c", choices=["b2c", "torch"],
                    help="b2c: own bucket reducer (default); torch: c10d Reducer with the b2c comm hook")
    ap.add_argument("--fused-opt", dest="fused_opt", action="store_true", default=True,
                    help="fold optimizer.step() into the bucket launches (default with the own reducer)")
    ap.add_argument("--no-fused-opt", dest="fused_opt", action="store_false")
    ap.add_argument("--no-parity", action="store_true", help="skip the cross-GPU parity block (N > 1)")
    ap.add_argument("--no-allreduce-section", action="store_true", help="skip the bucket-size allreduce sweep after the training legs")
    ap.add_argument("--no-bucket-timing", action="store_true", help="no per-bucket CUDA events in the hook (host-bound tiny models)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    elif args.impl == "hvd":
        run_hvd(args)
    else:
        run_b2c(args)


if __name__ == "__main__":
    main()
