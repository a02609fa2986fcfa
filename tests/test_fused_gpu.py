"""GPU: the fused optimizer step (b2c_fused_step / b2c_group_fused_step) through the C-ABI.

Oracle: the CPU restatement of the gradient average (oracle/b2c_oracle.c, pinned to gloo) followed by the real
``torch.optim`` update on the host — what the reference path computes (allreduce, then optimizer.step() in the PTL
loop behind ray_ddp.py:218-219).  Bar: every updated weight within 1e-6 relative (plus 1e-6 of one learning-rate
step for weights near zero) after EVERY step of a multi-step run, so the sharded optimizer state is checked too;
replicas bit-identical.
"""
import numpy as np
import pytest
import torch

from oracle import oracle as O

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    from ray_lightning_accelerators_b200 import _lib
    from ray_lightning_accelerators_b200.comm import Comm, LocalGroup, fused_shard

DEV = "cuda:0"

CASES = [
    ("sgd", dict(lr=0.1, momentum=0.9)),
    ("sgd", dict(lr=0.05, momentum=0.0, weight_decay=1e-3)),
    ("sgd", dict(lr=0.1, momentum=0.8, dampening=0.1, weight_decay=1e-4)),
    ("sgd", dict(lr=0.1, momentum=0.9, nesterov=True)),
    ("adam", dict(lr=0.1, betas=(0.9, 0.999), eps=1e-8)),
    ("adam", dict(lr=1e-3, betas=(0.8, 0.99), eps=1e-6, weight_decay=1e-2)),
    ("adamw", dict(lr=1e-4, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2)),
]


def _torch_opt(kind, p, cfg):
    cls = {"sgd": torch.optim.SGD, "adam": torch.optim.Adam, "adamw": torch.optim.AdamW}[kind]
    return cls([p], **cfg)


def _opt_struct(kind, cfg, t):
    o = _lib.Opt()
    o.kind = {"sgd": _lib.OPT_SGD, "adam": _lib.OPT_ADAM, "adamw": _lib.OPT_ADAMW}[kind]
    o.lr = cfg["lr"]
    o.weight_decay = cfg.get("weight_decay", 0.0)
    if kind == "sgd":
        o.momentum, o.dampening = cfg.get("momentum", 0.0), cfg.get("dampening", 0.0)
        o.nesterov, o.first_step = int(cfg.get("nesterov", False)), int(t == 1)
    else:
        b1, b2 = cfg["betas"]
        o.beta1, o.beta2, o.eps = b1, b2, cfg["eps"]
        o.bias_correction1, o.bias_correction2 = 1.0 - b1 ** t, 1.0 - b2 ** t
    return o


def _close(got, ref, lr):
    got, ref = got.astype(np.float64), ref.astype(np.float64)
    return bool((np.abs(got - ref) <= 1e-6 * np.abs(ref) + 1e-6 * lr).all())


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("kind,cfg", CASES)
def test_group_fused_step_matches_allreduce_then_torch_optim(world, kind, cfg):
    torch.cuda.set_device(0)
    n = 27884  # the MNIST-MLP bucket padded to whole vectors; shards are ragged for W = 8
    g = LocalGroup(world, device=0, arena_bytes=4 << 20)
    g.set_param(_lib.PARAM_TIMEOUT_MS, 3000)
    w0 = O.payload("randn", n, 77)
    ref_p = torch.nn.Parameter(torch.from_numpy(w0.copy()))
    ref_opt = _torch_opt(kind, ref_p, cfg)
    rw, rs1, rs2 = w0.copy(), np.zeros(n, np.float32), np.zeros(n, np.float32)   # the oracle's restatement
    params = [torch.from_numpy(w0.copy()).to(DEV) for _ in range(world)]
    shards = [fused_shard(g._lib, n, world, r) for r in range(world)]
    s1 = [torch.zeros(max(sh[1], 4), device=DEV) for sh in shards]
    s2 = [torch.zeros(max(sh[1], 4), device=DEV) for sh in shards]
    for t in range(1, 5):
        grads = [O.payload("randn", n, r, seed=1000 * t) for r in range(world)]
        dev_grads = [torch.from_numpy(x).to(DEV) for x in grads]
        g.fused_step(dev_grads, params, _opt_struct(kind, cfg, t), s1, s2)
        torch.cuda.synchronize()
        g.check_alive()
        avg = O.b2c_allreduce_f32(grads)                            # the contract's average (== gloo at these W)
        ref_p.grad = torch.from_numpy(avg.copy())
        ref_opt.step()
        want = ref_p.detach().numpy()
        O.fused_step_f32(kind, cfg, t, avg, rw, rs1, rs2)
        if kind == "sgd":   # same operations in the same order: the kernel equals the restatement (and torch) bit for bit
            assert np.array_equal(params[0].cpu().numpy().view(np.uint32), rw.view(np.uint32)), (cfg, world, t)
        for r in range(world):
            got = params[r].cpu().numpy()
            assert _close(got, want, cfg["lr"]), (kind, cfg, world, t, r, float(np.abs(got - want).max()))
            assert torch.equal(params[r].view(torch.int32), params[0].view(torch.int32)), "replicas differ"
        for r in range(world):   # the gradient buckets are inputs only
            assert np.array_equal(dev_grads[r].cpu().numpy().view(np.uint32), grads[r].view(np.uint32))
    g.close()


@pytest.mark.parametrize("kind,cfg", CASES)
@pytest.mark.parametrize("n", [5, 27882, 1 << 20])
def test_single_rank_fused_step_is_torch_optim(kind, cfg, n):
    """world = 1: no peer, the whole step is the local streaming kernel (any length, ragged tail included)."""
    torch.cuda.set_device(0)
    comm = Comm(0, 1, device=0, arena_bytes=1 << 20)
    w0 = O.payload("randn", n, 5)
    ref_p = torch.nn.Parameter(torch.from_numpy(w0.copy()))
    ref_opt = _torch_opt(kind, ref_p, cfg)
    p = torch.from_numpy(w0.copy()).to(DEV)
    s1, s2 = torch.zeros(n + 4, device=DEV)[:n], torch.zeros(n + 4, device=DEV)[:n]
    for t in range(1, 4):
        grad = O.payload("wide" if t == 2 else "randn", n, 9, seed=31 * t)
        comm.fused_step(torch.from_numpy(grad).to(DEV), p, _opt_struct(kind, cfg, t), s1, s2)
        torch.cuda.synchronize()
        ref_p.grad = torch.from_numpy(grad.copy())
        ref_opt.step()
        assert _close(p.cpu().numpy(), ref_p.detach().numpy(), cfg["lr"]), (kind, cfg, n, t)
    assert comm.stats()["fused_steps"] == 3
    comm.close()


def test_fused_step_argument_errors():
    torch.cuda.set_device(0)
    g = LocalGroup(2, device=0, arena_bytes=1 << 20)
    x = [torch.zeros(8, device=DEV) for _ in range(2)]
    s = [torch.zeros(8, device=DEV) for _ in range(2)]
    o = _lib.Opt()
    o.kind = 9
    with pytest.raises(_lib.B2cError) as e:
        g.fused_step(x, x, o, s, s)
    assert e.value.code == _lib.E_UNSUPPORTED
    o.kind = _lib.OPT_ADAM   # bias corrections missing
    with pytest.raises(_lib.B2cError) as e:
        g.fused_step(x, x, o, s, s)
    assert e.value.code == _lib.E_INVAL
    y = [torch.zeros(6, device=DEV) for _ in range(2)]   # not a whole number of vectors
    o = _opt_struct("sgd", dict(lr=0.1, momentum=0.9), 1)
    with pytest.raises(_lib.B2cError):
        g.fused_step(y, y, o, s, s)
    g.close()


def test_data_parallel_wrapper_with_fused_optimizer_follows_plain_training():
    """One process, one GPU, through the plugin objects a worker uses: B2cDataParallel + fuse_optimizer against the
    same model trained with the bare torch optimizer on the GPU (torch's foreach Adam) — 30 steps, loss within 1e-4
    relative at every step (the per-step weight bar is 1e-6, checked above; Adam at lr = 0.1 amplifies last-bit
    differences between two correct implementations over a run)."""
    from examples import workloads
    from ray_lightning_accelerators_b200.ddp_hook import B2cDataParallel
    torch.cuda.set_device(0)
    torch.backends.cudnn.deterministic = True
    comm = Comm(0, 1, device=0, arena_bytes=1 << 20)

    def run(fused):
        model = workloads.build_model("mnist_mlp", seed=0).to(DEV)
        opt = model.configure_optimizers()          # Adam(lr=0.1), the reference example's optimizer
        net = model
        if fused:
            net = B2cDataParallel(model, comm)
            opt = net.fuse_optimizer(opt)
            assert type(opt).__name__ == "FusedOptimizer"
        losses = []
        for step in range(30):
            x, y = workloads.synthetic_batch("mnist_mlp", 32, 0, step)
            x, y = x.to(DEV), y.to(DEV)
            opt.zero_grad(set_to_none=True) if not fused else opt.zero_grad()
            loss = model.loss_fn(net(x), y)
            loss.backward()
            opt.step()
            losses.append(float(loss.item()))
        return np.array(losses)

    ref, got = run(False), run(True)
    rel = np.abs(got - ref) / np.maximum(np.abs(ref), 1e-12)
    assert rel.max() <= 1e-4, rel.max()
    assert comm.stats()["fused_steps"] > 0
    comm.close()
