"""FusedOptimizer — the optimizer step folded into the gradient reduction.

The reference reaches ``optimizer.step()`` in the PTL train loop right after the gradient allreduce
(entered at ray_lightning/ray_ddp.py:218-219 / ray_horovod.py:159): every rank reads all averaged gradients,
all parameters and all optimizer state, and writes parameters and state back.  Here each gradient bucket's
launch does both: the rank that reduces a shard applies the optimizer to it (its state exists on that rank only,
1/W of the total) and the updated weights are written into every rank's parameter bucket (b2c_fused_step).
With the multicast arena up the module-buffer broadcast of the next forward (C4) rides in the last bucket's launch.

Wraps a ``torch.optim.SGD`` / ``Adam`` / ``AdamW`` and keeps its interface: hyper-parameters are read from
``param_groups`` at every launch, so LR schedulers keep working.  ``step()`` itself launches nothing — by the
time it is called the update has happened inside backward.  What is NOT supported: gradient clipping between
backward and step (the averaged gradient never exists as a whole on any rank), amsgrad / maximize, and
optimizer ``state_dict`` round trips (the state is sharded).
"""
import torch

from . import _lib


def _kind_of(opt):
    if type(opt) is torch.optim.SGD:
        return _lib.OPT_SGD
    if type(opt) is torch.optim.AdamW:
        return _lib.OPT_ADAMW
    if type(opt) is torch.optim.Adam:
        return _lib.OPT_ADAM
    return None


def supported(opt):
    if _kind_of(opt) is None:
        return False
    return not any(g.get("amsgrad") or g.get("maximize") for g in opt.param_groups)


class FusedOptimizer:
    def __init__(self, optimizer, ddp):
        """ddp: a ddp_hook.B2cDataParallel (owns the reducer, the comm and the buffer sync)."""
        if not supported(optimizer):
            raise TypeError(f"fused step supports torch.optim.SGD / Adam / AdamW without amsgrad / maximize, got {type(optimizer)}")
        self._opt = optimizer
        self._ddp = ddp
        self._red = ddp.reducer
        self._comm = ddp.comm
        self._kind = _kind_of(optimizer)
        self._steps = []
        group_of = {p: g for g in optimizer.param_groups for p in g["params"]}
        comm = self._comm
        for b in self._red.buckets:
            if b.flat.dtype != torch.float32:
                raise TypeError("fused step needs fp32 gradient buckets")
            groups = {id(group_of.get(p)) for p in b.params}
            if None in [group_of.get(p) for p in b.params] or len(groups) != 1:
                raise ValueError("fused step: all parameters of a gradient bucket must belong to one param_group")
            b.opt_state = {"group": group_of[b.params[0]], "t": 0}
            n = b.flat.numel()
            # parameters become views of a flat bucket with the gradient bucket's layout
            pflat = None
            if b.in_arena:
                pflat = comm.nvls_empty(n, torch.float32).zero_()
            else:
                pflat = torch.zeros(n, dtype=torch.float32, device=b.flat.device)
                if comm.world > 1:
                    comm.register(pflat)
            with torch.no_grad():
                for p, off in zip(b.params, b.offsets):
                    view = pflat[off:off + p.numel()].view_as(p)
                    view.copy_(p.data)
                    p.data = view
            b.param_flat = pflat
            _, shard = comm.fused_shard(n)
            need1 = self._kind != _lib.OPT_SGD or any(g.get("momentum", 0) for g in optimizer.param_groups)
            need2 = self._kind != _lib.OPT_SGD
            b.opt_state["s1"] = torch.zeros(max(shard, 4), dtype=torch.float32, device=b.flat.device) if need1 else None
            b.opt_state["s2"] = torch.zeros(max(shard, 4), dtype=torch.float32, device=b.flat.device) if need2 else None
        # C4 fold: the flat module-buffer tensor lives in the arena -> multicast it with the last bucket
        bs = ddp.bufsync
        self._fold = (bs.flat is not None and comm.world > 1 and comm.in_nvls_arena(bs.flat) and
                      bs.flat.numel() % 16 == 0 and all(b.in_arena for b in self._red.buckets))
        ddp._c4_folded = self._fold
        self._red.launch_fn = self._launch

    # ------------------------------------------------------------------ per bucket, on the comm's side stream
    def _launch(self, b, stream):
        g = b.opt_state["group"]
        b.opt_state["t"] += 1
        t = b.opt_state["t"]
        o = _lib.Opt()
        o.kind = self._kind
        o.lr = float(g["lr"])
        o.weight_decay = float(g.get("weight_decay", 0.0))
        if self._kind == _lib.OPT_SGD:
            o.momentum = float(g.get("momentum", 0.0))
            o.dampening = float(g.get("dampening", 0.0))
            o.nesterov = int(bool(g.get("nesterov", False)))
            o.first_step = int(t == 1)
        else:
            b1, b2 = g["betas"]
            o.beta1, o.beta2, o.eps = float(b1), float(b2), float(g["eps"])
            o.bias_correction1 = 1.0 - b1 ** t
            o.bias_correction2 = 1.0 - b2 ** t
        last = b.index == len(self._red.buckets) - 1
        bc = self._ddp.bufsync.flat if (self._fold and last) else None
        self._comm.fused_step(b.flat, b.param_flat, o, b.opt_state["s1"], b.opt_state["s2"], scale=self._red.scale,
                              bcast=bc, bcast_root=self._ddp.bufsync.root, stream=stream)
        if bc is not None:
            self._ddp._buffers_synced = True

    # ------------------------------------------------------------------ optimizer surface
    def step(self, closure=None):
        """The update already happened inside backward; only the stream dependency is left to state."""
        self._red.wait()
        return None

    def zero_grad(self, set_to_none=False):
        self._red.zero_grad()

    @property
    def param_groups(self):
        return self._opt.param_groups

    @property
    def state(self):
        return self._opt.state

    def state_dict(self):
        return self._opt.state_dict()

    def load_state_dict(self, sd):
        return self._opt.load_state_dict(sd)

    def __getattr__(self, name):
        return getattr(self._opt, name)
