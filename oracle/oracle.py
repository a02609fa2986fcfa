"""oracle/oracle.py — TEST INFRASTRUCTURE, NOT PRODUCT.

numpy front-end over oracle/b2c_oracle.c (see that file's header for what is restated,
from where, and how it is pinned).  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline leg may import this module; the product package never does.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liboracle.so")
_lib = None


def build():
    subprocess.run(["make", "-C", _HERE], check=True, capture_output=True)
    return _SO


def _load():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        _lib = ctypes.CDLL(_SO)
    return _lib


def _ptr_array(arrs):
    return (ctypes.c_void_p * len(arrs))(*[a.ctypes.data for a in arrs])


def _prep(grads, dtype):
    arrs = [np.ascontiguousarray(g, dtype=dtype) for g in grads]
    n = arrs[0].size
    assert all(a.size == n for a in arrs), "all ranks must contribute the same count"
    return arrs, n


def _order(order, W):
    if order is None:
        return None
    assert sorted(order) == list(range(W))
    return (ctypes.c_int * W)(*order)


# ------------------------------------------------------------------ payloads
def payload(kind, n, rank, seed=1234):
    """Deterministic per-rank gradient payloads (SURVEY.md §8c known-answer families)."""
    rng = np.random.default_rng(seed + rank)
    if kind == "int":  # every partial sum exact in fp32 AND bf16 -> bit-exact in any order
        return rng.integers(-8, 9, size=n).astype(np.float32)
    if kind == "randn":
        return rng.standard_normal(n).astype(np.float32)
    if kind == "wide":  # wide dynamic range: stresses summation order
        return (rng.standard_normal(n) * np.exp2(rng.integers(-20, 20, size=n))).astype(np.float32)
    if kind == "special":
        base = np.array([0.0, -0.0, np.inf, -np.inf, np.nan, 1e-42, -1e-42, 3.3895314e38,
                         1.0, -1.0, 65504.0, 1.1754944e-38], dtype=np.float32)
        out = np.resize(base, n).astype(np.float32)
        if rank % 2:
            out = np.roll(out, 1)
        return out
    raise ValueError(kind)


def to_bf16_bits(x):
    x = np.ascontiguousarray(x, dtype=np.float32)
    out = np.empty(x.size, dtype=np.uint16)
    _load().oracle_f32_to_bf16(ctypes.c_void_p(x.ctypes.data), ctypes.c_size_t(x.size), ctypes.c_void_p(out.ctypes.data))
    return out.reshape(x.shape)


def from_bf16_bits(h):
    h = np.ascontiguousarray(h, dtype=np.uint16)
    out = np.empty(h.size, dtype=np.float32)
    _load().oracle_bf16_to_f32(ctypes.c_void_p(h.ctypes.data), ctypes.c_size_t(h.size), ctypes.c_void_p(out.ctypes.data))
    return out.reshape(h.shape)


# ------------------------------------------------------------------ reference semantics
def ddp_allreduce_f32(grads, order=None):
    """div_(W) then SUM (torch default_hooks.py:18-32), summation in `order`."""
    arrs, n = _prep(grads, np.float32)
    out = np.empty(n, np.float32)
    _load().oracle_ddp_allreduce_f32(_ptr_array(arrs), len(arrs), ctypes.c_size_t(n), _order(order, len(arrs)),
                                     ctypes.c_void_p(out.ctypes.data))
    return out


def ddp_allreduce_bf16(grads_bits, order=None):
    arrs, n = _prep(grads_bits, np.uint16)
    out = np.empty(n, np.uint16)
    _load().oracle_ddp_allreduce_bf16(_ptr_array(arrs), len(arrs), ctypes.c_size_t(n), _order(order, len(arrs)),
                                      ctypes.c_void_p(out.ctypes.data))
    return out


def hvd_average_f32(grads):
    arrs, n = _prep(grads, np.float32)
    out = np.empty(n, np.float32)
    _load().oracle_hvd_average_f32(_ptr_array(arrs), len(arrs), ctypes.c_size_t(n), ctypes.c_void_p(out.ctypes.data))
    return out


# ------------------------------------------------------------------ the kernels' contract
def b2c_allreduce_f32(grads, scale=None):
    arrs, n = _prep(grads, np.float32)
    scale = 1.0 / len(arrs) if scale is None else scale
    out = np.empty(n, np.float32)
    _load().oracle_b2c_allreduce_f32(_ptr_array(arrs), len(arrs), ctypes.c_size_t(n), ctypes.c_float(scale),
                                     ctypes.c_void_p(out.ctypes.data))
    return out


def b2c_allreduce_bf16(grads_bits, scale=None):
    arrs, n = _prep(grads_bits, np.uint16)
    scale = 1.0 / len(arrs) if scale is None else scale
    out = np.empty(n, np.uint16)
    _load().oracle_b2c_allreduce_bf16(_ptr_array(arrs), len(arrs), ctypes.c_size_t(n), ctypes.c_float(scale),
                                      ctypes.c_void_p(out.ctypes.data))
    return out


def b2c_allreduce_f32_bf16wire(grads, scale=None):
    arrs, n = _prep(grads, np.float32)
    scale = 1.0 / len(arrs) if scale is None else scale
    out = np.empty(n, np.float32)
    _load().oracle_b2c_allreduce_f32_bf16wire(_ptr_array(arrs), len(arrs), ctypes.c_size_t(n), ctypes.c_float(scale),
                                              ctypes.c_void_p(out.ctypes.data))
    return out


def exact_mean_f64(grads):
    arrs, n = _prep(grads, np.float32)
    out = np.empty(n, np.float64)
    _load().oracle_exact_mean_f64(_ptr_array(arrs), len(arrs), ctypes.c_size_t(n), ctypes.c_void_p(out.ctypes.data))
    return out


def exact_mean_bf16_f64(grads_bits):
    arrs, n = _prep(grads_bits, np.uint16)
    out = np.empty(n, np.float64)
    _load().oracle_exact_mean_bf16_f64(_ptr_array(arrs), len(arrs), ctypes.c_size_t(n), ctypes.c_void_p(out.ctypes.data))
    return out


def broadcast(bufs, root=0):
    """Every rank ends with root's bytes (bit-exact)."""
    src = np.ascontiguousarray(bufs[root]).view(np.uint8).copy()
    return [src.copy() for _ in bufs]


# ------------------------------------------------------------------ tolerances (north_star)
def rel_err(got, ref, denom=None):
    """max |got-ref| / max(denom, tiny) over finite positions; `denom` defaults to |ref|."""
    got = np.asarray(got, np.float64)
    ref = np.asarray(ref, np.float64)
    m = np.isfinite(ref) & np.isfinite(got)
    if not m.any():
        return 0.0
    d = np.abs(ref) if denom is None else np.asarray(denom, np.float64)
    d = np.maximum(d, np.finfo(np.float32).tiny)
    return float(np.max(np.abs(got - ref)[m] / d[m]))


def bf16_ulp_diff(a_bits, b_bits):
    """ulp distance between two bf16 bit arrays (sign-magnitude -> ordered ints)."""
    def key(x):
        x = np.asarray(x, np.uint16).astype(np.int32)
        return np.where(x & 0x8000, 0x8000 - x, x)  # negative values mirror below zero
    return np.abs(key(a_bits) - key(b_bits))


# ------------------------------------------------------------------ fused optimizer step (restatement)
def _f32(x):
    return np.asarray(x, dtype=np.float32)


def _fma(a, b, c):
    """fp32 fused multiply-add: the product of two fp32 values is exact in fp64."""
    return (a.astype(np.float64) * np.float64(b) + c.astype(np.float64)).astype(np.float32)


def fused_step_f32(kind, cfg, t, g, w, s1, s2):
    """CPU restatement of libb2c.so's optimizer epilogue (csrc/b2c_kernels.cuh: opt_apply), which follows
    torch.optim.SGD / Adam / AdamW (single-tensor form) operation by operation in fp32.  Pinned against the real
    torch optimizers by tests/test_oracle.py.  g: averaged gradient; w, s1, s2: fp32 arrays, updated in place;
    t: 1-based step count.  Scalars are formed in double and rounded to fp32 once, as the host side does."""
    g, lr, wd = _f32(g).copy(), cfg["lr"], cfg.get("weight_decay", 0.0)
    if kind == "sgd":
        mom, damp = cfg.get("momentum", 0.0), cfg.get("dampening", 0.0)
        if wd != 0.0:
            g = _fma(w, np.float32(wd), g)
        if mom != 0.0:
            if t == 1:
                s1[:] = g
            else:
                s1[:] = _fma(g, np.float32(1.0 - damp), (s1 * np.float32(mom)).astype(np.float32))
            g = _fma(s1, np.float32(mom), g) if cfg.get("nesterov", False) else s1
        w[:] = _fma(g, np.float32(-lr), w)
        return w
    b1, b2 = cfg["betas"]
    if kind == "adamw":
        w[:] = w * np.float32(1.0 - lr * wd)
    elif wd != 0.0:
        g = _fma(w, np.float32(wd), g)
    s1[:] = _fma((g - s1).astype(np.float32), np.float32(1.0 - b1), s1)
    s2[:] = ((g * np.float32(1.0 - b2)).astype(np.float32).astype(np.float64) * g.astype(np.float64) +
             (s2 * np.float32(b2)).astype(np.float32).astype(np.float64)).astype(np.float32)
    step_size = np.float32(lr / (1.0 - b1 ** t))
    bc2_sqrt = np.float32(np.sqrt(1.0 - b2 ** t))
    denom = (np.sqrt(s2).astype(np.float32) / bc2_sqrt).astype(np.float32) + np.float32(cfg["eps"])
    w[:] = _fma((s1 / denom).astype(np.float32), -step_size, w)
    return w
