"""Gradient-allreduce sweep over real NVLink (one process per GPU, launch with torchrun):

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
      --master-port 29511 tools/nvlink_sweep.py --out gpurun_out/sweep_wN.json

Per size / dtype / algorithm: L2 flush, GPU-side flag barrier to align the ranks, CUDA events
around ONE call on the launching stream, per-iteration time = max over ranks; reports median /
p99 and bus bandwidth busbw = 2(W-1)/W * S / t (nccl-tests definition).  The same loop times
torch.distributed NCCL all_reduce as the on-box comparator, and every configuration is
parity-checked against NCCL and the contract in the same run.
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ray_lightning_accelerators_b200 import _lib  # noqa: E402
from ray_lightning_accelerators_b200.comm import ALGO_NVLS, ALGO_ONESHOT, ALGO_TWOSHOT, Comm  # noqa: E402

REAL_BUCKETS = [111528, 1625000, 12415000, 27021000, 29664000, 31500000, 93760000]


def timed(fn, comm, flush, iters, warmup):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        flush.fill_(1)
        torch.cuda._sleep(200000)
        comm.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        e1.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)  # us
    t = torch.tensor(ts, device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)  # per-iteration max over ranks
    t = t.sort().values
    return float(t[len(t) // 2]), float(t[min(len(t) - 1, int(len(t) * 0.99))]), float(t.mean())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="gpurun_out/nvlink_sweep.json")
    ap.add_argument("--iters", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--max-bytes", type=int, default=1 << 30)
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--fast", action="store_true", help="only one-shot, zero-copy two-shot and NVLS")
    ap.add_argument("--nvls", action="store_true")
    ap.add_argument("--blocks", default="32,64,128")
    args = ap.parse_args()

    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    comm = Comm.from_process_group(arena_bytes=2 * args.max_bytes + (1 << 20))
    comm.set_param(_lib.PARAM_TIMEOUT_MS, 5000)
    comm.set_param(_lib.PARAM_NVLS_MIN_BYTES, 0)
    nvls = comm.enable_nvls(2 * args.max_bytes + (4 << 20)) if args.nvls else False
    if rank == 0:
        print("NVLS", nvls, comm.nvls_error, flush=True)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    sizes = sorted(set([1 << k for k in range(10, 31, 2 if args.quick else 1)] + REAL_BUCKETS))
    sizes = [s for s in sizes if s <= args.max_bytes]
    block_opts = [int(b) for b in args.blocks.split(",")]
    rows = []
    nvls_bufs = {}
    for dtype in (torch.float32, torch.bfloat16):
        for S in sizes:
            es = 4 if dtype == torch.float32 else 2
            n = S // es
            gen = torch.Generator(device="cuda").manual_seed(1234 + rank)
            x = torch.randn(n, device="cuda", generator=gen).to(dtype)
            # NCCL comparator: the reference path (ray_ddp.py:227 -> "nccl"), default-hook arithmetic
            ref = x.clone().div_(world)
            dist.all_reduce(ref)
            work = x.clone()

            def nccl_fn():
                dist.all_reduce(work)

            med, p99, mean = timed(nccl_fn, comm, flush, args.iters, args.warmup)
            busf = 2 * (world - 1) / world * S
            rows.append(dict(impl="nccl", W=world, dtype=str(dtype).split(".")[1], bytes=S, us_med=med, us_p99=p99,
                             us_mean=mean, busbw_GBps=busf / med / 1e3))
            if rank == 0:
                print(rows[-1], flush=True)
            configs = []
            algos = [(ALGO_ONESHOT, 0), (ALGO_TWOSHOT, 0), (ALGO_TWOSHOT, 1), (ALGO_TWOSHOT, 2)]
            if args.fast:
                algos = [(ALGO_ONESHOT, 0), (ALGO_TWOSHOT, 2)]
            if nvls:
                algos += [(ALGO_NVLS, 3), (ALGO_NVLS, 4)]  # 3 = staged through the multicast arena, 4 = in place in it
            for algo, variant in algos:
                if algo == ALGO_ONESHOT and S > (64 << 20):
                    continue
                for nb in block_opts:
                    if S <= (1 << 16) and nb != block_opts[0]:
                        continue  # geometry clamps tiny messages to a few channels anyway
                    configs.append((algo, variant, nb))
            for algo, variant, nb in configs:
                comm.set_param(_lib.PARAM_TWOSHOT_VARIANT, variant if variant < 2 else 0)
                comm.set_param(_lib.PARAM_ZEROCOPY, 1 if variant == 2 else 0)  # 2 = in place on registered buffers
                comm.set_param(_lib.PARAM_MAX_BLOCKS_ONESHOT, nb)
                comm.set_param(_lib.PARAM_MAX_BLOCKS_TWOSHOT, nb)
                comm.set_param(_lib.PARAM_MAX_BLOCKS_NVLS, nb)
                # parity first (same run): vs NCCL and bitwise across replicas
                if variant == 4:
                    if not nvls_bufs:
                        nvls_bufs["raw"] = (comm.nvls_empty(args.max_bytes, torch.uint8), comm.nvls_empty(args.max_bytes, torch.uint8))
                    nvls_bufs[dtype] = tuple(r.view(dtype) for r in nvls_bufs["raw"])
                    y = nvls_bufs[dtype][0][:n]
                    y.copy_(x)
                else:
                    y = x.clone()
                if variant == 2:
                    comm.register(y)
                comm.allreduce(y, algo=algo)
                torch.cuda.synchronize()
                comm.check_alive()
                denom = x.float().abs().div_(world)
                dist.all_reduce(denom)
                err = ((y.float() - ref.float()).abs() / denom.clamp_min(1e-30)).max().item()
                tol = 1e-6 if dtype == torch.float32 else world * 2.0 ** -8  # NCCL rounds W times in bf16
                y0 = y.clone()
                dist.broadcast(y0, src=0)
                same = bool(torch.equal(y0.view(torch.int16 if es == 2 else torch.int32),
                                        y.view(torch.int16 if es == 2 else torch.int32)))
                bitexact = bool(torch.equal(y.view(torch.int16 if es == 2 else torch.int32),
                                            ref.view(torch.int16 if es == 2 else torch.int32)))
                if variant == 4:
                    buf = nvls_bufs[dtype][1][:n]
                    buf.copy_(x)
                else:
                    buf = x.clone()
                if variant == 2:
                    comm.register(buf)

                def b2c_fn():
                    comm.allreduce(buf, algo=algo)

                med, p99, mean = timed(b2c_fn, comm, flush, args.iters, args.warmup)
                st = comm.stats()
                rows.append(dict(impl="b2c", W=world, dtype=str(dtype).split(".")[1], bytes=S,
                                 algo={1: "oneshot", 2: "twoshot", 3: "nvls"}[algo], variant=variant, zerocopy=(variant == 2),
                                 max_blocks=nb,
                                 blocks=st["last_blocks"], us_med=med, us_p99=p99, us_mean=mean,
                                 busbw_GBps=busf / med / 1e3, rel_err_vs_nccl=err, within_tol=bool(err <= tol),
                                 bitexact_vs_nccl=bitexact, replicas_identical=same))
                if rank == 0:
                    print(rows[-1], flush=True)
                assert same, "replicas differ"
                assert err <= tol, f"parity vs NCCL failed: {err}"
            del x, ref, work
    # broadcast: time + bit-exactness vs NCCL broadcast
    for S in (111528, 212904, 102441032):
        if S > args.max_bytes:
            continue
        src = torch.randint(0, 256, (S,), dtype=torch.uint8, device="cuda", generator=torch.Generator(device="cuda").manual_seed(rank))
        a, b = src.clone(), src.clone()
        dist.broadcast(a, src=0)
        comm.broadcast(b, root=0)
        torch.cuda.synchronize()
        assert torch.equal(a, b), "broadcast differs from NCCL"
        med, p99, mean = timed(lambda: comm.broadcast(b, root=0), comm, flush, args.iters, args.warmup)
        medn, p99n, _ = timed(lambda: dist.broadcast(a, src=0), comm, flush, args.iters, args.warmup)
        rows.append(dict(impl="b2c", op="broadcast", W=world, bytes=S, us_med=med, us_p99=p99, nccl_us_med=medn,
                         root_egress_GBps=(world - 1) * S / med / 1e3, bit_exact_vs_nccl=True))
        if rank == 0:
            print(rows[-1], flush=True)
    comm.check_alive()
    if rank == 0:
        os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
        with open(args.out, "w") as f:
            json.dump(rows, f)
    dist.barrier()
    comm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
