"""Launch-geometry tuning for the two large-bucket paths (zero-copy two-shot and NVLS in place) at the
current world size: sweeps channels x threads x unroll on the ResNet-50-sized buckets.  torchrun only."""
import argparse
import itertools
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ray_lightning_accelerators_b200 import _lib  # noqa: E402
from ray_lightning_accelerators_b200.comm import ALGO_NVLS, ALGO_TWOSHOT, Comm  # noqa: E402
from tools.nvlink_sweep import timed  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="gpurun_out/tune.json")
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--sizes", default="27021000,31500000,67108864,268435456")
    ap.add_argument("--nvls-blocks", default="8,16,24,32,48,64")
    ap.add_argument("--skip-zc", action="store_true")
    args = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    sizes = [int(s) for s in args.sizes.split(",")]
    nvls_blocks = [int(s) for s in args.nvls_blocks.split(",")]
    comm = Comm.from_process_group(arena_bytes=8 << 20)
    comm.set_param(_lib.PARAM_TIMEOUT_MS, 5000)
    nvls = comm.enable_nvls(max(sizes) + (4 << 20))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    rows = []
    ybig = comm.nvls_empty(max(sizes) // 4) if nvls else None
    for S in sizes:
        n = S // 4
        x = torch.randn(n, device="cuda")
        comm.register(x)
        busf = 2 * (world - 1) / world * S
        for nb, th, un in itertools.product((32, 64, 96, 128, 192), (256, 512), (1, 2, 4)):
            if args.skip_zc:
                break
            if world != 8 and un != 2:
                continue
            if th == 512 and un == 4:
                continue  # spills
            comm.set_param(_lib.PARAM_MAX_BLOCKS_TWOSHOT, nb)
            comm.set_param(_lib.PARAM_THREADS, th)
            comm.set_param(_lib.PARAM_UNROLL_ZC, un if world == 8 else 0)
            med, p99, _ = timed(lambda: comm.allreduce(x, algo=ALGO_TWOSHOT), comm, flush, args.iters, 3)
            rows.append(dict(path="zc", bytes=S, blocks=nb, threads=th, unroll=un, us=med, busbw=busf / med / 1e3))
            if rank == 0:
                print(rows[-1], flush=True)
        if nvls:
            y = ybig[:n]
            y.copy_(x)
            for nb, th, un in itertools.product(nvls_blocks, (256, 512), (8, 16)):
                comm.set_param(_lib.PARAM_MAX_BLOCKS_NVLS, nb)
                comm.set_param(_lib.PARAM_THREADS, th)
                comm.set_param(_lib.PARAM_UNROLL_NVLS, un)
                med, p99, _ = timed(lambda: comm.allreduce(y, algo=ALGO_NVLS), comm, flush, args.iters, 3)
                rows.append(dict(path="nvls", bytes=S, blocks=nb, threads=th, unroll=un, us=med, busbw=busf / med / 1e3))
                if rank == 0:
                    print(rows[-1], flush=True)
        del x
    comm.check_alive()
    if rank == 0:
        with open(args.out, "w") as f:
            json.dump(rows, f)
        for S in sizes:
            for path in ("zc", "nvls"):
                best = sorted([r for r in rows if r["bytes"] == S and r["path"] == path], key=lambda r: r["us"])[:4]
                print(S, path, [(r["blocks"], r["threads"], r["unroll"], round(r["us"], 1), round(r["busbw"])) for r in best])
    dist.barrier()
    comm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
