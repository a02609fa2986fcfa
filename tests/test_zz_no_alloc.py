"""b2c_allreduce_local_noalloc, through the C-ABI.

Some callers of this library cannot allocate device memory at the point where they need a
reduction: code that runs once the training framework's caching allocator owns the GPU (a
cudaMalloc there may fail or force a cache flush), and code on the autograd hook thread in
the middle of backward, where cudaMalloc / cudaFree would synchronise the whole device.  For
them libb2c.so has one entry point whose flags and staging slots live in a device array built
into the library instead of an arena from cudaMalloc.

Checked here: two calls in a row (different world size, dtype and payload, so the second one
meets whatever the first left in the shared array) give exactly what the existing entry point,
b2c_group_allreduce with B2C_ALGO_ONESHOT, gives on the same inputs — the tests in this repo hold
both to the same arithmetic contract, so the comparison is bit for bit — and sizes beyond the
built-in scratch are refused with a clear error.  The file name sorts last on purpose: with -x a
failure here cannot hide another test.  It passes alone in a fresh process and after the rest.
"""
import ctypes

import numpy as np
import pytest
import torch

from oracle import oracle as O

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    from ray_lightning_accelerators_b200 import _lib
    from ray_lightning_accelerators_b200.comm import ALGO_ONESHOT, LocalGroup

DEV = "cuda:0"


def _noalloc(tensors, dtype_code, scale):
    lib = _lib.load()
    ptrs = (ctypes.c_void_p * len(tensors))(*[t.data_ptr() for t in tensors])
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    return lib.b2c_allreduce_local_noalloc(ptrs, len(tensors), tensors[0].numel(), dtype_code,
                                           ctypes.c_float(scale), stream)


def _existing(tensors):
    g = LocalGroup(len(tensors), device=0, arena_bytes=2 << 20)
    g.allreduce(tensors, algo=ALGO_ONESHOT)
    torch.cuda.synchronize()
    g.check_alive()
    g.close()


def test_two_calls_in_a_row_match_the_existing_entry_point():
    torch.cuda.set_device(0)
    # call 1: the MNIST-MLP gradient bucket, 4 virtual ranks, fp32
    grads = [O.payload("randn", 27882, r, seed=11) for r in range(4)]
    a = [torch.from_numpy(x).to(DEV) for x in grads]
    b = [t.clone() for t in a]
    assert _noalloc(a, _lib.F32, 0.25) == 0, _lib.last_error()
    # call 2, no host sync in between: other world size, other dtype, other payload
    bits = [O.to_bf16_bits(O.payload("wide", 1003, r, seed=23)) for r in range(8)]
    c = [torch.from_numpy(x.view(np.int16).copy()).to(DEV).view(torch.bfloat16) for x in bits]
    d = [t.clone() for t in c]
    assert _noalloc(c, _lib.BF16, 0.125) == 0, _lib.last_error()
    torch.cuda.synchronize()
    _existing(b)
    _existing(d)
    for got, want in zip(a, b):
        assert torch.equal(got.view(torch.int32), want.view(torch.int32))
    for got, want in zip(c, d):
        assert torch.equal(got.view(torch.int16), want.view(torch.int16))
    # and both agree with the CPU statement of the contract
    assert np.array_equal(a[0].cpu().numpy().view(np.uint32), O.b2c_allreduce_f32(grads, 0.25).view(np.uint32))
    ref = O.b2c_allreduce_bf16(bits, 0.125)
    nan = (ref & 0x7fff) > 0x7f80
    assert np.array_equal(c[0].view(torch.int16).cpu().numpy().view(np.uint16)[~nan], ref[~nan])
    # a third call with the first world size again: the counters left by call 2 (8 ranks) are cleared
    e = [torch.from_numpy(x).to(DEV) for x in grads]
    assert _noalloc(e, _lib.F32, 0.25) == 0, _lib.last_error()
    torch.cuda.synchronize()
    assert torch.equal(e[3].view(torch.int32), b[3].view(torch.int32))


def test_sizes_beyond_the_built_in_scratch_are_refused():
    torch.cuda.set_device(0)
    lib = _lib.load()
    n_ok = _lib.NOALLOC_MAX_BYTES // 4
    ok = [torch.ones(n_ok, device=DEV) for _ in range(2)]
    assert _noalloc(ok, _lib.F32, 0.5) == 0, _lib.last_error()
    torch.cuda.synchronize()
    assert float(ok[0][0]) == 1.0 and float(ok[1][-1]) == 1.0
    big = [torch.ones(n_ok + 1, device=DEV) for _ in range(2)]
    assert _noalloc(big, _lib.F32, 0.5) == _lib.E_TOOBIG
    msg = lib.b2c_last_error().decode()
    assert "exceed the built-in scratch" in msg and str(_lib.NOALLOC_MAX_BYTES) in msg
    assert float(big[0][0]) == 1.0  # untouched
    nine = (ctypes.c_void_p * 9)(*[ok[0].data_ptr()] * 9)
    assert lib.b2c_allreduce_local_noalloc(nine, 9, 4, _lib.F32, ctypes.c_float(1.0), None) == _lib.E_INVAL
    assert lib.b2c_allreduce_local_noalloc(nine, 2, 4, _lib.U8, ctypes.c_float(1.0), None) == _lib.E_UNSUPPORTED


def test_barrier_timeout_status_can_be_read_back_and_is_clear_after_good_calls():
    """The entry point only enqueues, so a flag-barrier timeout inside its kernel is reported through
    b2c_noalloc_status (sticky words that the per-call wipe of the scratch array does not touch)."""
    torch.cuda.set_device(0)
    lib = _lib.load()
    ts = [torch.full((1000,), float(r), device=DEV) for r in range(4)]
    assert _noalloc(ts, _lib.F32, 0.25) == 0, _lib.last_error()
    words = (ctypes.c_uint32 * 4)()
    assert lib.b2c_noalloc_status(ctypes.byref(words), 0) == 0, _lib.last_error()   # waits for the device
    assert list(words) == [0, 0, 0, 0]
    assert float(ts[2][17]) == 1.5
    assert lib.b2c_noalloc_status(None, 1) == 0
