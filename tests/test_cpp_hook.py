"""CPU: plumbing of the optional C++ DDP hook (csrc/b2c_torch_hook.cpp): it registers on a real c10d
Reducer, is called once per bucket with the bucket's buffer, and DDP uses what it leaves there.  The
callee is a ctypes stub here (there is no GPU); on a GPU box the callee is b2c_allreduce."""
import ctypes

import pytest
import torch
import torch.distributed as dist

from ray_lightning_accelerators_b200 import _lib


def test_cpp_hook_runs_on_a_real_reducer():
    mod = _lib.load_torch_hook()
    if mod is None:
        pytest.skip("optional torch hook not built (python -c 'import __graft_entry__ as g; g.build()')")
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:29779", rank=0, world_size=1)
    try:
        calls = []
        fn_t = ctypes.CFUNCTYPE(ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int,
                                ctypes.c_int, ctypes.c_float, ctypes.c_int, ctypes.c_void_p)

        def stub(comm, buf, count, dtype, wire, scale, algo, stream):
            arr = (ctypes.c_float * count).from_address(buf)
            for i in range(count):
                arr[i] *= 3.0
            calls.append((comm, count, dtype, wire, scale, algo, stream))
            return 0 if len(calls) < 3 else -4

        cb = fn_t(stub)
        torch.manual_seed(0)
        net = torch.nn.Linear(4, 2)
        ref = torch.nn.Linear(4, 2)
        ref.load_state_dict(net.state_dict())
        ddp = torch.nn.parallel.DistributedDataParallel(net)
        # error-string getter: the real library passes the address of b2c_last_error; a ctypes callback that
        # returns char* is not safe to call from C++, so the stub run passes none
        hid = mod.register_hook(ddp.reducer, ctypes.cast(cb, ctypes.c_void_p).value, 0, 1234, 0.5, 2, 1 << 20)
        x = torch.ones(3, 4)
        ddp(x).sum().backward()
        ref(x).sum().backward()
        assert calls == [(1234, 10, 0, 0, 0.5, 2, None)] and mod.buckets_seen(hid) == 1
        assert torch.allclose(net.weight.grad, 3 * ref.weight.grad) and torch.allclose(net.bias.grad, 3 * ref.bias.grad)
        ddp.zero_grad()
        ddp(x).sum().backward()
        assert len(calls) == 2
        with pytest.raises(RuntimeError, match=r"b2c_allreduce failed \(-4\)"):   # a failing b2c call surfaces
            ddp.zero_grad()
            ddp(x).sum().backward()
    finally:
        dist.destroy_process_group()
