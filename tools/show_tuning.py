"""Print the measured crossover table of this world size (torchrun): Comm.autotune() as the accelerators run it at set-up."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ray_lightning_accelerators_b200 import _lib  # noqa: E402
from ray_lightning_accelerators_b200.comm import Comm  # noqa: E402
from ray_lightning_accelerators_b200.ddp_hook import _want_nvls  # noqa: E402

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
local = int(os.environ.get("LOCAL_RANK", rank))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
comm = Comm.from_process_group(arena_bytes=64 << 20)
if _want_nvls(comm) or os.environ.get("B2C_NVLS") == "1":
    comm.enable_nvls(128 << 20)
t = comm.autotune()
if rank == 0:
    t = dict(t, nvls=comm.nvls_enabled, params={"NVLS_MIN_BYTES": comm.get_param(_lib.PARAM_NVLS_MIN_BYTES),
                                                "ZEROCOPY_MIN_BYTES": comm.get_param(_lib.PARAM_ZEROCOPY_MIN_BYTES),
                                                "CROSSOVER_BYTES": comm.get_param(_lib.PARAM_CROSSOVER_BYTES)})
    print(json.dumps(t))
    out = sys.argv[1] if len(sys.argv) > 1 else None
    if out:
        with open(out, "w") as f:
            json.dump(t, f, indent=1)
comm.check_alive()
dist.barrier()
comm.close()
dist.destroy_process_group()
