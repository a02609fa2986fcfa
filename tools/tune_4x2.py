"""BASELINE config #5: 4 concurrent Tune trials x 2 GPUs each (examples/ray_ddp_tune.py:67-79 of the reference:
every trial reserves its workers' GPUs with extra_gpu, so the trials hold DISJOINT GPU pairs and train at the same
time).  Each trial is an independent b2c communicator — own process group, own arena, own flags.

driver (no torchrun):  python tools/tune_4x2.py --out gpurun_out/r2_tune_4x2.json
  1. solo: trial 0 alone on GPUs {0,1}: MNIST-MLP steps/s (RayAccelerator worker path, own reducer + fused Adam)
     and a W=2 allreduce sweep;
  2. concurrent: 4 trials on {0,1},{2,3},{4,5},{6,7} at once, the same measurements inside each;
  3. report per-trial numbers and the concurrent / solo ratios (interference).
Every worker sees ONE GPU (CUDA_VISIBLE_DEVICES=<k>, device index 0), as a Ray actor does.
worker (spawned):      python tools/tune_4x2.py --worker --rank R --world 2 --gpu K --port P --out F
"""
import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

SWEEP = [4096, 111528, 1625000, 12415000, 31500000]


def worker(args):
    os.environ["CUDA_VISIBLE_DEVICES"] = str(args.gpu)
    import torch
    import torch.distributed as dist
    from examples import workloads
    from ray_lightning_accelerators_b200 import _lib
    from ray_lightning_accelerators_b200.comm import ALGO_AUTO, Comm
    from ray_lightning_accelerators_b200.ddp_hook import wrap_ddp
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    dist.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{args.port}", rank=args.rank, world_size=args.world, device_id=dev)
    comm = Comm.from_process_group(arena_bytes=96 << 20)
    comm.set_param(_lib.PARAM_TIMEOUT_MS, 10000)
    model = workloads.build_model("mnist_mlp", seed=0).to(dev)
    ddp = wrap_ddp(model, comm, device_ids=[0])
    opt = ddp.fuse_optimizer(model.configure_optimizers()) if hasattr(ddp, "fuse_optimizer") else model.configure_optimizers()
    x, y = [t.to(dev) for t in workloads.synthetic_batch("mnist_mlp", 32, args.rank, 0)]

    def step():
        opt.zero_grad()
        loss = model.loss_fn(ddp(x), y)
        loss.backward()
        opt.step()
        return loss

    for _ in range(50):
        step()
    torch.cuda.synchronize()
    dist.barrier()
    # start line shared by all trials: wall-clock time given by the driver
    while time.time() < args.start_at:
        time.sleep(0.0005)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        loss = step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    sweep = []
    flush = torch.empty(192 << 20, dtype=torch.uint8, device=dev)
    for S in SWEEP:
        t = torch.randn(S // 4, device=dev)
        comm.register(t)
        ts = []
        for i in range(13):
            flush.fill_(1)
            comm.barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); comm.allreduce(t, algo=ALGO_AUTO); b.record(); b.synchronize()
            if i >= 3:
                ts.append(a.elapsed_time(b) * 1e3)
        ts.sort()
        sweep.append({"bytes": S, "us": ts[len(ts) // 2]})
        comm.deregister(t)
    comm.check_alive()
    out = torch.tensor([ms] + [r["us"] for r in sweep], device=dev, dtype=torch.float64)
    dist.all_reduce(out, op=dist.ReduceOp.MAX)
    if args.rank == 0:
        res = {"gpus": args.group, "ms_per_step": float(out[0]) / args.steps, "samples_per_sec": 2 * 32 * args.steps / (float(out[0]) / 1e3),
               "last_loss": float(loss.item()), "sweep": [{"bytes": r["bytes"], "us": float(out[1 + i])} for i, r in enumerate(sweep)],
               "launches": comm.stats()["launches"]}
        with open(args.out, "w") as f:
            json.dump(res, f)
    dist.barrier()
    comm.close()
    dist.destroy_process_group()


def run_trials(groups, steps, tag, outdir):
    procs, outs = [], []
    start_at = time.time() + 30.0          # every trial finishes set-up and warm-up before the shared start line
    for gi, gpus in enumerate(groups):
        port = 29800 + 7 * gi + (100 if tag == "solo" else 0)
        out = os.path.join(outdir, f"_tune_{tag}_{gi}.json")
        outs.append(out)
        for r, k in enumerate(gpus):
            env = dict(os.environ)
            for v in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT"):
                env.pop(v, None)
            procs.append(subprocess.Popen([sys.executable, os.path.abspath(__file__), "--worker", "--rank", str(r), "--world", str(len(gpus)),
                                           "--gpu", str(k), "--port", str(port), "--out", out, "--steps", str(steps),
                                           "--start-at", repr(start_at), "--group", ",".join(map(str, gpus))], env=env))
    rcs = [p.wait(timeout=600) for p in procs]
    assert all(rc == 0 for rc in rcs), rcs
    res = []
    for o in outs:
        with open(o) as f:
            res.append(json.load(f))
        os.remove(o)
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--worker", action="store_true")
    ap.add_argument("--rank", type=int, default=0)
    ap.add_argument("--world", type=int, default=2)
    ap.add_argument("--gpu", type=int, default=0)
    ap.add_argument("--port", type=int, default=29800)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--start-at", type=float, default=0.0)
    ap.add_argument("--group", default="")
    ap.add_argument("--out", default="gpurun_out/r2_tune_4x2.json")
    ap.add_argument("--trials", type=int, default=4)
    args = ap.parse_args()
    if args.worker:
        return worker(args)
    outdir = os.path.dirname(os.path.abspath(args.out))
    os.makedirs(outdir, exist_ok=True)
    groups = [(2 * i, 2 * i + 1) for i in range(args.trials)]
    solo = run_trials(groups[:1], args.steps, "solo", outdir)[0]
    conc = run_trials(groups, args.steps, "conc", outdir)
    rep = {"config": "BASELINE configs[4]: 4 concurrent trials x 2 GPUs, MNIST-MLP batch 32/GPU, own reducer + fused Adam; W=2 allreduce sweep inside each trial",
           "solo": solo, "concurrent": conc,
           "interference": [{"gpus": c["gpus"], "samples_per_sec_ratio_vs_solo": c["samples_per_sec"] / solo["samples_per_sec"],
                             "sweep_us_ratio_vs_solo": [{"bytes": a["bytes"], "ratio": a["us"] / b["us"]} for a, b in zip(c["sweep"], solo["sweep"])]}
                            for c in conc]}
    with open(args.out, "w") as f:
        json.dump(rep, f, indent=1)
    print(json.dumps({"solo_sps": solo["samples_per_sec"], "concurrent_sps": [c["samples_per_sec"] for c in conc]}))
    print("TUNE_4X2_OK")


if __name__ == "__main__":
    main()
