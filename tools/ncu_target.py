"""Smallest program that launches each b2c kernel variant a few times on ONE GPU (local group of
8 virtual ranks + the W=1 path) — the target for `ncu --set full` captures (never a multi-rank
command under ncu)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ray_lightning_accelerators_b200 import _lib  # noqa: E402
from ray_lightning_accelerators_b200.comm import ALGO_ONESHOT, ALGO_TWOSHOT, LocalGroup  # noqa: E402

torch.cuda.set_device(0)
S = 31500000  # the 30.04 MiB ResNet-50 bucket
for W in (8, 1):
    g = LocalGroup(W, device=0, arena_bytes=(80 << 20) if W > 1 else (1 << 20))
    g.set_param(_lib.PARAM_TIMEOUT_MS, 3000)
    ts = [torch.randn(S // 4, device="cuda") for _ in range(W)]
    if W > 1:
        for zc in (1, 0):
            g.set_param(_lib.PARAM_ZEROCOPY, zc)
            for _ in range(3):
                g.allreduce(ts, algo=ALGO_TWOSHOT)
        small = [torch.randn(27882, device="cuda") for _ in range(W)]
        for _ in range(3):
            g.allreduce(small, algo=ALGO_ONESHOT)
        bs = [torch.empty(212904, dtype=torch.uint8, device="cuda") for _ in range(W)]
        for _ in range(3):
            g.broadcast(bs, root=0)
    else:
        for _ in range(3):
            g.allreduce(ts, algo=ALGO_ONESHOT)
    torch.cuda.synchronize()
    g.check_alive()
    g.close()
print("ncu_target ok")
