"""Where does the fixed cost of an allreduce go?  Device-side phase clock (channel 0) for a few sizes / paths."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ray_lightning_accelerators_b200 import _lib  # noqa: E402
from ray_lightning_accelerators_b200.comm import ALGO_NVLS, ALGO_ONESHOT, ALGO_TWOSHOT, Comm  # noqa: E402

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
local = int(os.environ.get("LOCAL_RANK", rank))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
comm = Comm.from_process_group(arena_bytes=8 << 20)
comm.set_param(_lib.PARAM_TIMEOUT_MS, 5000)
nvls = comm.enable_nvls(80 << 20)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
big = comm.nvls_empty(67108864 // 4) if nvls else None
for S in (4096, 111528, 1625000, 12415000, 31500000, 67108864):
    n = S // 4
    x = torch.randn(n, device="cuda")
    comm.register(x)
    cases = [("oneshot", ALGO_ONESHOT, x), ("zc", ALGO_TWOSHOT, x)]
    if nvls:
        cases.append(("nvls", ALGO_NVLS, big[:n]))
    for name, algo, t in cases:
        if name == "oneshot" and S > (4 << 20):
            continue
        acc = [0.0, 0.0, 0.0]
        tot = 0.0
        for it in range(8):
            flush.fill_(1)
            torch.cuda._sleep(200000)
            comm.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            comm.allreduce(t, algo=algo)
            e1.record()
            e1.synchronize()
            ph = comm.phase_us()
            if it >= 3:
                acc = [a + p for a, p in zip(acc, ph)]
                tot += e0.elapsed_time(e1) * 1e3
        if rank == 0:
            print(f"S={S:9d} {name:8s} event={tot / 5:7.1f}us  kernel: start->A {acc[0] / 5:6.1f}  A->B {acc[1] / 5:7.1f}  B->end {acc[2] / 5:5.1f}  "
                  f"(event - kernel = {tot / 5 - sum(acc) / 5:5.1f})", flush=True)
dist.barrier()
comm.close()
dist.destroy_process_group()
