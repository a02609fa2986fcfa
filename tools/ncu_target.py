"""Smallest program that launches ONE b2c kernel variant a few times on ONE GPU — the target for
`ncu --set full` captures (never a multi-rank command under ncu).  Keep the device footprint small:
ncu saves / restores every byte the kernel can reach between its replay passes.
usage: python tools/ncu_target.py {zc|staged|oneshot|w1|bcast|fused|fusedzc}"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ray_lightning_accelerators_b200 import _lib  # noqa: E402
from ray_lightning_accelerators_b200.comm import ALGO_ONESHOT, ALGO_TWOSHOT, LocalGroup  # noqa: E402

mode = sys.argv[1] if len(sys.argv) > 1 else "zc"
torch.cuda.set_device(0)
if mode in ("fused", "fusedzc"):
    # the fused optimizer step on the largest ResNet-50 bucket of the own reducer (26,058,752 bytes), SGD + momentum
    from ray_lightning_accelerators_b200.comm import Comm, fused_shard
    n = 26058752 // 4
    o = _lib.Opt()
    o.kind, o.lr, o.momentum = _lib.OPT_SGD, 0.1, 0.9
    if mode == "fused":
        comm = Comm(0, 1, device=0, arena_bytes=1 << 20)
        g_, w_, m_ = (torch.randn(n, device="cuda") for _ in range(3))
        flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
        for _ in range(4):
            flush.fill_(1)      # cold L2 for every profiled launch, as inside a training step
            comm.fused_step(g_, w_, o, m_, None)
        torch.cuda.synchronize()
        print("ncu_target ok fused", comm.stats()["last_blocks"])
        comm.close()
    else:
        W_ = 8
        grp = LocalGroup(W_, device=0, arena_bytes=2 << 20)
        grp.set_param(_lib.PARAM_TIMEOUT_MS, 3000)
        gs = [torch.randn(n, device="cuda") for _ in range(W_)]
        w0 = torch.randn(n, device="cuda")
        ws = [w0.clone() for _ in range(W_)]
        sh = [fused_shard(grp._lib, n, W_, r)[1] for r in range(W_)]
        ms = [torch.zeros(max(k, 4), device="cuda") for k in sh]
        for _ in range(3):
            grp.fused_step(gs, ws, o, ms, [None] * W_)
        torch.cuda.synchronize()
        grp.check_alive()
        print("ncu_target ok fusedzc", grp.stats()["last_blocks"])
        grp.close()
    sys.exit(0)
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
