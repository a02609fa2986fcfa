"""Smallest program that launches ONE b2c kernel variant a few times on ONE GPU — the target for
`ncu --set full` captures (never a multi-rank command under ncu).  Keep the device footprint small:
ncu saves / restores every byte the kernel can reach between its replay passes.
usage: python tools/ncu_target.py {zc|staged|oneshot|w1|bcast}"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ray_lightning_accelerators_b200 import _lib  # noqa: E402
from ray_lightning_accelerators_b200.comm import ALGO_ONESHOT, ALGO_TWOSHOT, LocalGroup  # noqa: E402

mode = sys.argv[1] if len(sys.argv) > 1 else "zc"
torch.cuda.set_device(0)
S = 31500000  # the 30.04 MiB ResNet-50 bucket
W = 1 if mode == "w1" else 8
arena = (80 << 20) if mode == "staged" else (2 << 20)
g = LocalGroup(W, device=0, arena_bytes=arena)
g.set_param(_lib.PARAM_TIMEOUT_MS, 3000)
if mode in ("zc", "staged", "w1"):
    ts = [torch.randn(S // 4, device="cuda") for _ in range(W)]
    g.set_param(_lib.PARAM_ZEROCOPY, 0 if mode == "staged" else 1)
    for _ in range(3):
        g.allreduce(ts, algo=ALGO_ONESHOT if mode == "w1" else ALGO_TWOSHOT)
elif mode == "oneshot":
    ts = [torch.randn(27882, device="cuda") for _ in range(W)]
    for _ in range(3):
        g.allreduce(ts, algo=ALGO_ONESHOT)
else:
    ts = [torch.empty(212904, dtype=torch.uint8, device="cuda") for _ in range(W)]
    for _ in range(3):
        g.broadcast(ts, root=0)
torch.cuda.synchronize()
g.check_alive()
print("ncu_target ok", mode, g.stats()["last_blocks"])
g.close()
