"""Round-2 probe (torchrun, W >= 2): fixed cost of the NVLS allreduce and the broadcast paths.

For each size: the round-1 in-place NVLS kernel (unicast flags) against the lean kernel (multicast flag
barrier) with / without programmatic dependent launch and with 512 / 256-thread CTAs — isolated launches
(L2 flushed, ranks aligned, CUDA events around one call, max over ranks, median) and back-to-back trains over
distinct buffers larger than L2 — plus the device-side phase split of channel 0.  Broadcast: pull-from-root
against the multicast path (staged and in place) against NCCL.  Every configuration is parity-checked
against NCCL on the same tensors before it is timed.
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="gpurun_out/r2_probe.json")
    ap.add_argument("--iters", type=int, default=12)
    ap.add_argument("--sizes", default="4096,111528,1625000,12415000,27021000,29664000,31500000")
    ap.add_argument("--bcast", default="212992,8388608,102441040")
    args = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    comm = Comm.from_process_group(arena_bytes=64 << 20)
    comm.set_param(_lib.PARAM_TIMEOUT_MS, 5000)
    ok = comm.enable_nvls(1 << 30)
    if rank == 0:
        print("NVLS", ok, comm.nvls_error, flush=True)
    if not ok:
        raise SystemExit("no multicast on this box")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    rows = []

    def emit(r):
        rows.append(r)
        if rank == 0:
            print(json.dumps(r), flush=True)

    def isolated(fn, iters=args.iters, phases=False):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        ts, ph = [], [0.0, 0.0, 0.0]
        for _ in range(iters):
            flush.fill_(1)
            torch.cuda._sleep(200000)
            comm.barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); fn(); b.record(); b.synchronize()
            ts.append(a.elapsed_time(b) * 1e3)
            if phases:
                p = comm.phase_us()
                ph = [x + y / iters for x, y in zip(ph, p)]
        t = torch.tensor(ts, device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.sort().values[len(ts) // 2]), ph

    def train(bufs, fn, rounds=3):
        for b in bufs:
            fn(b)
        torch.cuda.synchronize()
        comm.barrier()
        a, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(rounds):
            for b in bufs:
                fn(b)
        b_.record(); b_.synchronize()
        t = torch.tensor([a.elapsed_time(b_) * 1e3 / (rounds * len(bufs))], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def bits(t):
        return t.view(torch.int32)

    # ---------------- parity of the lean kernel (int payload: bit-exact vs NCCL; randn: 1e-6 of sum|g|/W)
    for n in (4, 5, 1003, 27882, 406568, 3103747):
        for kind in ("int", "randn"):
            g = torch.Generator(device=dev).manual_seed(99 + rank)
            src = (torch.randint(-8, 9, (n,), generator=g, device=dev).float() if kind == "int"
                   else torch.randn(n, generator=g, device=dev))
            ref = src.clone().div_(world)
            dist.all_reduce(ref)
            denom = src.abs() / world
            dist.all_reduce(denom)
            for threads in (512, 256):
                comm.set_param(_lib.PARAM_THREADS_NVLS, threads)
                t = comm.nvls_empty(n)
                t.copy_(src)
                comm.allreduce(t, algo=ALGO_NVLS)
                torch.cuda.synchronize()
                comm.check_alive()
                if kind == "int":
                    assert torch.equal(bits(t), bits(ref)), ("lean int", n, threads)
                else:
                    err = float(((t - ref).abs().max() / denom.max()).item())
                    assert err <= 1e-6, ("lean randn", n, threads, err)
                t0 = t.clone()
                dist.broadcast(t0, src=0)
                assert torch.equal(bits(t0), bits(t)), "replicas differ"
                comm.nvls_free(t)
    comm.set_param(_lib.PARAM_THREADS_NVLS, 512)
    if rank == 0:
        print("lean parity ok", flush=True)

    # ---------------- allreduce fixed cost
    configs = [("r1", dict(lean=0, pdl=0, th=512, acq=1)), ("lean", dict(lean=1, pdl=0, th=512, acq=1)),
               ("lean+pdl", dict(lean=1, pdl=1, th=512, acq=1)), ("lean+pdl t256", dict(lean=1, pdl=1, th=256, acq=1)),
               ("lean+pdl relaxed-poll", dict(lean=1, pdl=1, th=512, acq=0))]
    for S in [int(x) for x in args.sizes.split(",")]:
        n = S // 4
        k = max(2, (160 << 20) // S + 1) if S >= (1 << 20) else 8
        bufs = [comm.nvls_empty(n) for _ in range(k)]
        for b in bufs:
            b.normal_()
        y = torch.randn(n, device=dev)
        bus = 2.0 * (world - 1) / world * S
        us_n, _ = isolated(lambda: dist.all_reduce(y))
        ys = [torch.randn(n, device=dev) for _ in range(k)]
        tr_n = train(ys, lambda t_: dist.all_reduce(t_))
        del ys
        emit(dict(op="allreduce", cfg="nccl", bytes=S, us=us_n, us_train=tr_n, busbw=bus / us_n / 1e3, busbw_train=bus / tr_n / 1e3))
        for name, cfg in configs:
            comm.set_param(_lib.PARAM_NVLS_LEAN, cfg["lean"])
            comm.set_param(_lib.PARAM_PDL, cfg["pdl"])
            comm.set_param(_lib.PARAM_THREADS_NVLS, cfg["th"])
            comm.set_param(_lib.PARAM_MC_ACQ_POLL, cfg["acq"])
            us, ph = isolated(lambda: comm.allreduce(bufs[0], algo=ALGO_NVLS), phases=True)
            tr = train(bufs, lambda t_: comm.allreduce(t_, algo=ALGO_NVLS))
            emit(dict(op="allreduce", cfg=name, bytes=S, us=us, us_train=tr, busbw=bus / us / 1e3, busbw_train=bus / tr / 1e3,
                      start_to_A=ph[0], A_to_B=ph[1], outside=us - sum(ph)))
        if S <= (2 << 20):
            x = torch.randn(n, device=dev)
            us, ph = isolated(lambda: comm.allreduce(x, algo=ALGO_ONESHOT), phases=True)
            emit(dict(op="allreduce", cfg="oneshot staged", bytes=S, us=us, busbw=bus / us / 1e3, start_to_A=ph[0]))
            comm.register(x)
            us, ph = isolated(lambda: comm.allreduce(x, algo=ALGO_TWOSHOT), phases=True)
            emit(dict(op="allreduce", cfg="zero-copy two-shot", bytes=S, us=us, busbw=bus / us / 1e3, start_to_A=ph[0], A_to_B=ph[1]))
        for b in reversed(bufs):
            comm.nvls_free(b)
    comm.set_param(_lib.PARAM_NVLS_LEAN, 1)
    comm.set_param(_lib.PARAM_PDL, 1)
    comm.set_param(_lib.PARAM_THREADS_NVLS, 512)
    comm.set_param(_lib.PARAM_MC_ACQ_POLL, 1)
    comm.check_alive()

    # ---------------- broadcast
    comm.set_param(_lib.PARAM_BCAST_INPLACE_MIN_BYTES, 0)   # "mc in place" rows really are in place at every size
    for S in [int(x) for x in args.bcast.split(",")]:
        g = torch.Generator(device=dev).manual_seed(7 + rank)
        src = torch.randint(0, 256, (S,), generator=g, device=dev, dtype=torch.uint8)
        for root in range(min(world, 2)):
            want = src.clone()
            dist.broadcast(want, src=root)
            for mc in (0, 1):
                comm.set_param(_lib.PARAM_BCAST_MC, mc)
                got = src.clone()
                comm.broadcast(got, root=root)
                torch.cuda.synchronize()
                assert torch.equal(got, want), ("bcast", S, root, mc)
            inp = comm.nvls_empty(S // 16 * 16, dtype=torch.uint8)
            inp.copy_(src[:inp.numel()])
            comm.broadcast(inp, root=root)
            torch.cuda.synchronize()
            assert torch.equal(inp, want[:inp.numel()]), ("bcast in place", S, root)
            comm.nvls_free(inp)
        a = src.clone()
        us_n, _ = isolated(lambda: dist.broadcast(a, src=0), iters=8)
        emit(dict(op="broadcast", cfg="nccl", bytes=S, us=us_n, GBps=S / us_n / 1e3))
        for mc in (0, 1):
            comm.set_param(_lib.PARAM_BCAST_MC, mc)
            us, _ = isolated(lambda: comm.broadcast(a, root=0), iters=8)
            emit(dict(op="broadcast", cfg="mc staged" if mc else "pull from root", bytes=S, us=us, GBps=S / us / 1e3))
        inp = comm.nvls_empty(S // 16 * 16, dtype=torch.uint8)
        us, _ = isolated(lambda: comm.broadcast(inp, root=0), iters=8)
        emit(dict(op="broadcast", cfg="mc in place", bytes=S, us=us, GBps=S / us / 1e3))
        comm.nvls_free(inp)
    comm.check_alive()
    if rank == 0:
        with open(args.out, "w") as f:
            json.dump({"world": world, "rows": rows}, f, indent=1)
        print("PROBE_OK", flush=True)
    dist.barrier()
    comm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
