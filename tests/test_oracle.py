"""CPU: pin the oracle (oracle/b2c_oracle.c) against the golden vectors produced by the real
upstream path (torch gloo, tests/golden/make_golden.py).  Tolerances are north_star's:
bit-exact broadcast; allreduce <= 1e-6 relative in fp32, one bf16 ulp in bf16."""
import os
import re

import numpy as np
import pytest

from oracle import oracle as O

WORLDS = (2, 3, 4, 8)


def _load(golden_dir, world):
    return np.load(os.path.join(golden_dir, f"gloo_collectives_w{world}.npz"))


def _inputs(kind, n, world):
    return [O.payload(kind, n, r) for r in range(world)]


@pytest.mark.parametrize("world", WORLDS)
def test_ddp_allreduce_f32_matches_gloo(golden_dir, world):
    g = _load(golden_dir, world)
    names = [k for k in g.files if k.startswith("ar_f32_") and "special" not in k]
    assert names
    for name in names:
        _, _, kind, n = name.split("_")
        grads = _inputs(kind, int(n), world)
        ref = g[name]
        got = O.ddp_allreduce_f32(grads)
        if kind == "int" and world in (2, 4, 8):
            # every term k/W and every partial sum is exact -> order cannot matter
            assert np.array_equal(got.view(np.uint32), ref.view(np.uint32)), name
        elif world == 2:
            # a+b is commutative: the restatement must be bit-exact
            assert np.array_equal(got.view(np.uint32), ref.view(np.uint32)), name
        else:
            # gloo's ring order is not rank order: 1e-6 relative to sum_r |g_r|/W (fp32 tolerance)
            denom = sum(np.abs(x.astype(np.float64)) for x in grads) / world
            assert O.rel_err(got, ref, denom) <= 1e-6, name
        # the kernels' contract (sum in rank order, then * 1/W) vs the reference semantics
        mine = O.b2c_allreduce_f32(grads)
        denom = sum(np.abs(x.astype(np.float64)) for x in grads) / world
        assert O.rel_err(mine, ref, denom) <= 1e-6, name
        if world in (2, 4, 8) and kind != "wide":
            # 1/W exact, no denormal results in these payloads: pre- vs post-scaling agree bit for bit at W=2
            if world == 2:
                assert np.array_equal(mine.view(np.uint32), ref.view(np.uint32)), name
        # and both against the exact fp64 answer
        exact = O.exact_mean_f64(grads)
        assert O.rel_err(mine, exact, denom) <= 1e-6
        assert O.rel_err(ref, exact, denom) <= 1e-6


@pytest.mark.parametrize("world", WORLDS)
def test_ddp_allreduce_bf16_matches_gloo(golden_dir, world):
    g = _load(golden_dir, world)
    names = [k for k in g.files if k.startswith("ar_bf16_")]
    assert names
    for name in names:
        _, _, kind, n = name.split("_")
        bits = [O.to_bf16_bits(x) for x in _inputs(kind, int(n), world)]
        ref = g[name]
        got = O.ddp_allreduce_bf16(bits)
        if kind == "int" and world in (2, 4, 8):
            assert np.array_equal(got, ref), name
            assert np.array_equal(O.b2c_allreduce_bf16(bits), ref), name
        elif world == 2:
            assert np.array_equal(got, ref), name
        # fp32-accumulate / single rounding (the kernels) is the more exact of the two: it must be
        # within half a bf16 ulp of the fp64 answer, and gloo's multi-rounding result within
        # W roundings of it
        mine = O.b2c_allreduce_bf16(bits)
        exact = O.exact_mean_bf16_f64(bits)
        nearest = O.to_bf16_bits(exact.astype(np.float32))
        if world in (2, 4, 8):
            assert O.bf16_ulp_diff(mine, nearest).max() <= 1, name
        else:
            # 1/W is inexact: cancellation leaves residue; bound relative to sum|g|/W (bf16 eps 2^-8)
            denom = sum(np.abs(O.from_bf16_bits(b).astype(np.float64)) for b in bits) / world
            assert O.rel_err(O.from_bf16_bits(mine), exact, denom) <= 2.0 ** -8, name
        if world == 2:
            assert O.bf16_ulp_diff(mine, ref).max() <= 1, name  # north_star: one bf16 ulp


@pytest.mark.parametrize("world", WORLDS)
def test_specials_propagate_like_gloo(golden_dir, world):
    g = _load(golden_dir, world)
    grads = _inputs("special", 24, world)
    ref = g["ar_f32_special_24"]
    mine = O.b2c_allreduce_f32(grads)
    assert np.array_equal(np.isnan(mine), np.isnan(ref))
    assert np.array_equal(np.isinf(mine), np.isinf(ref))
    m = np.isfinite(ref)
    if world == 2:
        # denormal inputs: (a/2 + b/2) and (a+b)/2 differ by at most one denormal ulp
        assert np.all(np.abs(mine[m].astype(np.float64) - ref[m].astype(np.float64)) <= 1.5e-45 + 1e-6 * np.abs(ref[m]))
        assert np.array_equal(np.signbit(mine[m & (ref != 0)]), np.signbit(ref[m & (ref != 0)]))


@pytest.mark.parametrize("world", WORLDS)
def test_broadcast_matches_gloo(golden_dir, world):
    g = _load(golden_dir, world)
    names = [k for k in g.files if k.startswith("bc_")]
    assert len(names) == 5
    for name in names:
        n, root = map(int, re.match(r"bc_(\d+)_root(\d+)", name).groups())
        bufs = [np.random.default_rng(77 + r).integers(0, 256, size=n, dtype=np.uint8) for r in range(world)]
        outs = O.broadcast(bufs, root)
        for o in outs:
            assert np.array_equal(o, g[name]), name


def test_hvd_average_equals_ddp_for_pow2():
    for world in (2, 4, 8):
        grads = _inputs("randn", 513, world)
        a = O.hvd_average_f32(grads)
        b = O.b2c_allreduce_f32(grads)
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32))


def test_bf16_conversion_roundtrip_matches_torch():
    import torch
    x = O.payload("wide", 4096, 0)
    bits = O.to_bf16_bits(x)
    ref = torch.from_numpy(x).to(torch.bfloat16).view(torch.uint16).numpy()
    assert np.array_equal(bits, ref)
    assert np.array_equal(O.from_bf16_bits(bits), torch.from_numpy(ref.copy()).view(torch.bfloat16).float().numpy())


def test_wire_bf16_matches_torch_compress_semantics():
    import torch
    world = 4
    grads = _inputs("randn", 1000, world)
    mine = O.b2c_allreduce_f32_bf16wire(grads)
    # torch bf16_compress_hook (default_hooks.py:57-93): cast, div, allreduce in bf16, cast back
    acc = None
    for x in grads:
        t = torch.from_numpy(x).to(torch.bfloat16).div_(world)
        acc = t if acc is None else acc + t
    ref = acc.float().numpy()
    q = [O.from_bf16_bits(O.to_bf16_bits(x)) for x in grads]
    denom = sum(np.abs(x.astype(np.float64)) for x in q) / world
    # torch rounds W times in bf16 (eps 2^-8 each), the kernel once; under cancellation only a
    # bound relative to sum|g|/W is meaningful
    assert O.rel_err(mine, ref, denom) <= world * 2.0 ** -8
    exact = O.exact_mean_f64(q)
    assert O.rel_err(mine, exact, denom) <= 2.0 ** -8
    assert O.rel_err(mine, exact, denom) <= O.rel_err(ref, exact, denom) + 1e-12


# ------------------------------------------------------------------ fused optimizer step: restatement pinned to torch.optim
OPT_CASES = [
    ("sgd", dict(lr=0.1, momentum=0.9)),
    ("sgd", dict(lr=0.05, momentum=0.0, weight_decay=1e-3)),
    ("sgd", dict(lr=0.1, momentum=0.8, dampening=0.1, weight_decay=1e-4)),
    ("sgd", dict(lr=0.1, momentum=0.9, nesterov=True)),
    ("adam", dict(lr=0.1, betas=(0.9, 0.999), eps=1e-8)),
    ("adam", dict(lr=1e-3, betas=(0.8, 0.99), eps=1e-6, weight_decay=1e-2)),
    ("adamw", dict(lr=1e-4, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2)),
]


@pytest.mark.parametrize("kind,cfg", OPT_CASES)
def test_fused_step_restatement_is_pinned_to_torch_optim(kind, cfg):
    """oracle.fused_step_f32 restates the optimizer epilogue of the fused kernels; the real upstream optimizers
    (what the reference's PTL loop calls after the allreduce) pin it: SGD bit for bit, Adam / AdamW within 1e-6
    relative (+ 1e-6 of a learning-rate step) over five steps."""
    import torch
    n = 27884
    w0 = O.payload("randn", n, 77)
    p = torch.nn.Parameter(torch.from_numpy(w0.copy()))
    opt = {"sgd": torch.optim.SGD, "adam": torch.optim.Adam, "adamw": torch.optim.AdamW}[kind]([p], **cfg)
    w, s1, s2 = w0.copy(), np.zeros(n, np.float32), np.zeros(n, np.float32)
    for t in range(1, 6):
        g = O.payload("wide" if t == 3 else "randn", n, 3, seed=100 * t)
        p.grad = torch.from_numpy(g.copy())
        opt.step()
        O.fused_step_f32(kind, cfg, t, g, w, s1, s2)
        ref = p.detach().numpy()
        if kind == "sgd":
            assert np.array_equal(w.view(np.uint32), ref.view(np.uint32)), (kind, cfg, t)
        else:
            err = np.abs(w.astype(np.float64) - ref.astype(np.float64))
            assert (err <= 1e-6 * np.abs(ref) + 1e-6 * cfg["lr"]).all(), (kind, cfg, t, float(err.max()))
