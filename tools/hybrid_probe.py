"""Probe: do the NVLS path and the peer-to-peer zero-copy path add up when they run at the same time
(two comms, two streams)?  If yes, one hybrid kernel can split every bucket between them.  torchrun only."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ray_lightning_accelerators_b200 import _lib  # noqa: E402
from ray_lightning_accelerators_b200.comm import ALGO_NVLS, ALGO_TWOSHOT, Comm  # noqa: E402

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
local = int(os.environ.get("LOCAL_RANK", rank))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
c1 = Comm.from_process_group(arena_bytes=8 << 20)
c2 = Comm.from_process_group(arena_bytes=8 << 20)
for c in (c1, c2):
    c.set_param(_lib.PARAM_TIMEOUT_MS, 5000)
assert c1.enable_nvls(300 << 20)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
abig = c1.nvls_empty(268435456 // 4)
abig.normal_()
for S in (31500000, 67108864, 268435456):
    for frac in (1.0, 0.9, 0.8, 0.7, 0.6):
        for p2p_blocks in (16, 32, 64):
            n1 = int(S * frac) // 4
            n2 = S // 4 - n1
            a = abig[:max(n1, 4)]
            b = torch.randn(max(n2, 4), device="cuda")
            c2.register(b)
            c2.set_param(_lib.PARAM_MAX_BLOCKS_TWOSHOT, p2p_blocks)
            ts = []
            for it in range(13):
                flush.fill_(1)
                torch.cuda._sleep(200000)
                c1.barrier()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                cur = torch.cuda.current_stream()
                s1.wait_stream(cur)
                s2.wait_stream(cur)
                c1.allreduce(a, algo=ALGO_NVLS, stream=s1)
                if n2 > 0:
                    c2.allreduce(b, algo=ALGO_TWOSHOT, stream=s2)
                cur.wait_stream(s1)
                cur.wait_stream(s2)
                e1.record()
                e1.synchronize()
                if it >= 3:
                    ts.append(e0.elapsed_time(e1) * 1e3)
            t = torch.tensor(ts, device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            med = float(t.sort().values[len(ts) // 2])
            if rank == 0:
                print(f"S={S} nvls_frac={frac} p2p_blocks={p2p_blocks} us={med:.1f} busbw={2 * (world - 1) / world * S / med / 1e3:.0f}", flush=True)
            if frac == 1.0:
                break
c1.check_alive(); c2.check_alive()
dist.barrier()
c1.close(); c2.close()
dist.destroy_process_group()
