"""Measured NVLink bytes per launch of every cross-GPU kernel variant, on real links (torchrun, W >= 2).

Evidence for SURVEY.md §8(d): achieved NVLink traffic per kernel against its algorithmic bytes.  A profiler is
the wrong tool for these kernels on this pool — they spin on flags another rank writes, ncu replays a kernel and
first re-runs the whole command without itself, and B200_PROFILING.md rules out multi-rank commands under ncu —
so the link traffic is read from the GPU's own NVLink throughput counters (NVML field values
NVLINK_THROUGHPUT_DATA_TX / _RX, summed over the 18 links, KiB) before and after N back-to-back launches of ONE
kernel variant on rank 0's GPU: bytes per launch = delta / N.  Duration comes from CUDA events around the same
N launches.  Counters include protocol-visible payload only (DATA, not RAW).
"""
import argparse
import json
import os
import subprocess
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ray_lightning_accelerators_b200 import _lib  # noqa: E402
from ray_lightning_accelerators_b200.comm import ALGO_NVLS, ALGO_ONESHOT, ALGO_TWOSHOT, Comm  # noqa: E402


class LinkCounters:
    """(tx_bytes, rx_bytes) summed over all links of one GPU, or None when the box exposes no link counters."""

    def __init__(self, index):
        self.index = index
        self.mode, self.how, self.errors = None, "unavailable", []
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
        except Exception as e:  # noqa: BLE001
            self.nv = None
            self.errors.append(f"nvml init: {type(e).__name__}: {e}")
        for mode in ("nvml_sum", "nvml_links", "smi"):
            try:
                self.mode = mode
                self.read()
                self.how = {"nvml_sum": "NVML field values NVLINK_THROUGHPUT_DATA_TX/RX (KiB), scope = all links",
                            "nvml_links": "NVML field values NVLINK_THROUGHPUT_DATA_TX/RX (KiB), summed over links 0..17",
                            "smi": "nvidia-smi nvlink -gt d (KiB per link, summed)"}[mode]
                return
            except Exception as e:  # noqa: BLE001
                self.errors.append(f"{mode}: {type(e).__name__}: {e}")
        self.mode = None
        self.how = "unavailable on this box: " + " | ".join(self.errors)

    @staticmethod
    def _val(v):
        t = int(v.valueType)
        return int({0: v.value.dVal, 1: v.value.uiVal, 2: v.value.ulVal, 3: v.value.ullVal, 4: v.value.sllVal}.get(t, v.value.ullVal))

    def read(self):
        if self.mode is None:
            return None
        if self.mode in ("nvml_sum", "nvml_links"):
            if self.nv is None:
                raise RuntimeError("no NVML")
            nv = self.nv
            ids = [nv.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_TX, nv.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_RX]
            scopes = [0xFFFFFFFF] if self.mode == "nvml_sum" else list(range(18))
            out = []
            for i in ids:
                vals = nv.nvmlDeviceGetFieldValues(self.h, [(i, sc) for sc in scopes])
                tot = 0
                for v in vals:
                    if v.nvmlReturn != 0:
                        raise RuntimeError(f"field {v.fieldId} scope {v.scopeId}: nvmlReturn {v.nvmlReturn}")
                    tot += self._val(v)
                out.append(tot * 1024)
            return tuple(out)
        txt = subprocess.run(["nvidia-smi", "nvlink", "-gt", "d", "-i", str(self.index)], capture_output=True, text=True).stdout
        tx = rx = 0
        for line in txt.splitlines():
            line = line.strip()
            if "Data Tx" in line:
                tx += int(line.split(":")[-1].split()[0])
            elif "Data Rx" in line:
                rx += int(line.split(":")[-1].split()[0])
        if tx == 0 and rx == 0:
            raise RuntimeError("no counters in nvidia-smi output: " + txt[:120].replace("\n", " / "))
        return tx * 1024, rx * 1024


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="gpurun_out/r2_nvlink_bytes.json")
    ap.add_argument("--launches", type=int, default=200)
    ap.add_argument("--bytes", type=int, default=31500000)
    args = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    comm = Comm.from_process_group(arena_bytes=96 << 20)
    comm.set_param(_lib.PARAM_TIMEOUT_MS, 5000)
    nvls = comm.enable_nvls(512 << 20)
    ctr = LinkCounters(local) if rank == 0 else None
    S, N = args.bytes, args.launches
    rows = []

    def measure(name, fn, algorithmic_tx, algorithmic_rx, note, size=S):
        for _ in range(5):
            fn()
        torch.cuda.synchronize()
        dist.barrier()
        c0 = ctr.read() if ctr else None
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(N):
            fn()
        b.record()
        torch.cuda.synchronize()
        dist.barrier()
        c1 = ctr.read() if ctr else None
        us = a.elapsed_time(b) * 1e3 / N
        if rank == 0:
            have = c0 is not None and c1 is not None
            tx, rx = ((c1[0] - c0[0]) / N, (c1[1] - c0[1]) / N) if have else (float("nan"), float("nan"))
            row = {"kernel": name, "bytes": size, "world": world, "us_per_launch_back_to_back": us,
                   "nvlink_tx_bytes_per_launch": tx, "nvlink_rx_bytes_per_launch": rx,
                   "algorithmic_tx_bytes": algorithmic_tx, "algorithmic_rx_bytes": algorithmic_rx,
                   "tx_over_algorithmic": tx / algorithmic_tx if algorithmic_tx else None,
                   "rx_over_algorithmic": rx / algorithmic_rx if algorithmic_rx else None,
                   "tx_GBps": tx / us / 1e3, "rx_GBps": rx / us / 1e3, "note": note}
            rows.append(row)
            print(json.dumps(row), flush=True)

    W = world
    x = torch.randn(S // 4, device=dev)
    assert comm.register(x)
    p2p = 2.0 * (W - 1) / W * S
    measure(f"k_allreduce_twoshot_zc<f32,W={W}>", lambda: comm.allreduce(x, algo=ALGO_TWOSHOT), (W - 1) / W * S * 1.0 + 0.0, (W - 1) / W * S,
            "zero-copy two-shot: a GPU pulls (W-1)/W*S (its shard from W-1 peers: RX) and pushes its reduced shard to W-1 peers ((W-1)/W*S: TX); "
            "peers pulling this GPU's data add (W-1)/W*S TX and their pushes (W-1)/W*S RX => 2(W-1)/W*S per direction in total")
    rows and rows[-1].update({"algorithmic_tx_bytes": p2p, "algorithmic_rx_bytes": p2p,
                              "tx_over_algorithmic": rows[-1]["nvlink_tx_bytes_per_launch"] / p2p,
                              "rx_over_algorithmic": rows[-1]["nvlink_rx_bytes_per_launch"] / p2p})
    if nvls:
        y = comm.nvls_empty(S // 4)
        y.normal_()
        wire = (1.0 + 1.0 / W) * S
        measure("k_nvls_inplace<f32>", lambda: comm.allreduce(y, algo=ALGO_NVLS), wire, wire,
                "in-switch: TX = S (serving every rank's multimem.ld_reduce of this GPU's copy) + S/W (multimem.st of the own shard); "
                "RX = S/W (reduced shard) + S (every rank's multimem.st lands here)")
        yb = comm.nvls_empty(S // 2, torch.bfloat16)
        yb.normal_()
        measure("k_nvls_inplace<bf16>", lambda: comm.allreduce(yb, algo=ALGO_NVLS), wire, wire, "as above, bf16 elements, fp32 accumulate in the switch")
        comm.set_param(_lib.PARAM_NVLS_LEAN, 0)
        measure("k_allreduce_nvls<f32> (round 1 kernel)", lambda: comm.allreduce(y, algo=ALGO_NVLS), wire, wire, "same data movement, unicast flag barrier")
        comm.set_param(_lib.PARAM_NVLS_LEAN, 1)
        z = comm.nvls_empty(S // 16 * 16, torch.uint8)
        measure("k_broadcast_mc (in place)", lambda: comm.broadcast(z, root=0), float(S // 16 * 16), float(S // 16 * 16) if False else 0.0,
                "root = this GPU: one multimem.st stream of S bytes; the switch replicates it (the root's own copy comes back as RX)")
        u = torch.empty(S, dtype=torch.uint8, device=dev)
        measure("k_broadcast_mc (staged)", lambda: comm.broadcast(u, root=0), float(S), 0.0, "root = this GPU, staging slot in the arena, receivers copy locally")
    comm.set_param(_lib.PARAM_BCAST_MC, 0)
    u2 = torch.empty(S, dtype=torch.uint8, device=dev)
    measure("k_broadcast (pull from root)", lambda: comm.broadcast(u2, root=0), float((W - 1) * S), 0.0, "root = this GPU: every peer pulls S bytes from it")
    comm.set_param(_lib.PARAM_BCAST_MC, 1)
    small = torch.randn(27882, device=dev)
    s_b = 27882 * 4
    measure(f"k_allreduce_oneshot<f32,W={W}>", lambda: comm.allreduce(small, algo=ALGO_ONESHOT), float((W - 1) * s_b), float((W - 1) * s_b),
            "staged one-shot on the 111,528-byte MLP bucket: every rank pulls all W-1 peer copies", size=s_b)
    yn = torch.randn(S // 4, device=dev)
    measure("ncclDevKernel_AllReduce (NCCL 2.28, comparator)", lambda: dist.all_reduce(yn), p2p, p2p, "whatever algorithm NCCL picks at this size")
    comm.check_alive()
    if rank == 0:
        with open(args.out, "w") as f:
            json.dump({"world": world, "how": ctr.how, "launches_per_measurement": N, "rows": rows}, f, indent=1)
        print("NVLINK_BYTES_OK", flush=True)
    dist.barrier()
    comm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
