"""Kernel timing on ONE GPU: W virtual ranks (LocalGroup), CUDA events on the launching
stream.  HBM-bound here (all 'peers' are local); the NVLink numbers come from
tools/nvlink_sweep.py under torchrun.  Writes gpurun_out/local_sweep.json."""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ray_lightning_accelerators_b200 import _lib  # noqa: E402
from ray_lightning_accelerators_b200.comm import ALGO_ONESHOT, ALGO_TWOSHOT, LocalGroup  # noqa: E402


def hbm_bytes(algo, variant, W, S):
    """algorithmic HBM bytes of ONE launch over all W virtual ranks."""
    if W == 1:
        return 2 * S
    if algo == ALGO_ONESHOT:
        return W * (2 * S + W * S + S)  # stage r+w, read W copies, write result
    if variant == 2:
        return W * 2 * S  # zero-copy: each rank reads W shards (S) and writes W shards (S), in place
    per = 2 * S + S + S / W  # stage r+w, reduce-scatter read of W shards, own shard -> buf
    if variant == 0:
        per += (W - 1) * S / W + 2 * (W - 1) * S / W  # push to peers, gather local arena -> buf
    else:
        per += S / W + 2 * (W - 1) * S / W  # own arena write, pull + write
    return W * per


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--out", default="gpurun_out/local_sweep.json")
    args = ap.parse_args()
    torch.cuda.set_device(0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    rows = []
    sizes = [111528, 1 << 20, 1625000, 12415000, 27000000, 31500000, 64 << 20]
    for W in (1, 2, 4, 8):
        g = LocalGroup(W, device=0, arena_bytes=160 << 20)
        g.set_param(_lib.PARAM_TIMEOUT_MS, 3000)
        for dtype in (torch.float32, torch.bfloat16):
            for S in sizes:
                n = S // (4 if dtype == torch.float32 else 2)
                ts = [torch.randn(n, device="cuda").to(dtype) for _ in range(W)]
                for algo, variant in ((ALGO_ONESHOT, 0), (ALGO_TWOSHOT, 0), (ALGO_TWOSHOT, 1), (ALGO_TWOSHOT, 2)):
                    if W == 1 and algo != ALGO_ONESHOT:
                        continue
                    if algo == ALGO_ONESHOT and W * S > (512 << 20):
                        continue
                    g.set_param(_lib.PARAM_TWOSHOT_VARIANT, variant if variant < 2 else 0)
                    g.set_param(_lib.PARAM_ZEROCOPY, 1 if variant == 2 else 0)
                    for _ in range(3):
                        g.allreduce(ts, algo=algo)
                    torch.cuda.synchronize()
                    tot = 0.0
                    for _ in range(args.iters):
                        flush.fill_(1)  # L2 flush between timed launches
                        torch.cuda._sleep(300000)  # keep the GPU busy so the host runs ahead of the stream
                        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                        e0.record()
                        g.allreduce(ts, algo=algo)
                        e1.record()
                        e1.synchronize()
                        tot += e0.elapsed_time(e1)
                    us = tot / args.iters * 1e3
                    st = g.stats()
                    hb = hbm_bytes(algo, variant, W, n * ts[0].element_size())
                    rows.append(dict(W=W, dtype=str(dtype).split(".")[1], bytes=n * ts[0].element_size(),
                                     algo={1: "oneshot", 2: "twoshot"}[algo], variant=variant, us=round(us, 2),
                                     blocks=st["last_blocks"], hbm_GBps=round(hb / us / 1e3, 1)))
                    print(rows[-1], flush=True)
        g.check_alive()
        g.close()
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    with open(args.out, "w") as f:
        json.dump(rows, f, indent=0)


if __name__ == "__main__":
    main()
