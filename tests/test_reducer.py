"""CPU: host logic of the shared bucket reducer (ray_lightning_accelerators_b200/reducer.py) — bucket planning,
gradients as views of the buckets, launch order, unused parameters, re-arming after zero_grad(set_to_none=True).
The collective itself is replaced by a recorder here; the real kernels are covered by the GPU tests."""
import torch

from ray_lightning_accelerators_b200.reducer import B2cReducer, plan_buckets


def _model():
    torch.manual_seed(0)
    return torch.nn.Sequential(torch.nn.Linear(8, 16), torch.nn.ReLU(), torch.nn.Linear(16, 16), torch.nn.ReLU(),
                               torch.nn.Linear(16, 4))


def test_plan_is_reverse_order_with_a_small_first_bucket():
    m = _model()
    params = list(m.parameters())
    runs = plan_buckets(params, cap_bytes=16 * 16 * 4 + 64, first_bytes=100)
    flat = [p for run in runs for p in run]
    assert flat == list(reversed(params))
    assert sum(p.numel() * 4 for p in runs[0]) <= 100 or len(runs[0]) == 1
    for run in runs[1:]:
        assert len(run) == 1 or sum(p.numel() * 4 for p in run) <= 16 * 16 * 4 + 64


def test_gradients_are_views_and_buckets_launch_in_index_order():
    m = _model()
    calls = []
    red = B2cReducer(m.parameters(), None, bucket_cap_bytes=600, first_bucket_bytes=100, world=2,
                     fallback_allreduce=lambda flat, scale: calls.append((flat.data_ptr(), scale)))
    assert len(red.buckets) >= 3
    for b in red.buckets:
        for p, off in zip(b.params, b.offsets):
            assert p.grad.data_ptr() == b.flat.data_ptr() + off * 4 and off * 4 % 16 == 0
    m(torch.randn(3, 8)).sum().backward()
    assert [c[0] for c in calls] == [b.flat.data_ptr() for b in red.buckets]   # strict index order
    assert all(c[1] == 0.5 for c in calls)
    for b in red.buckets:   # autograd accumulated straight into the bucket
        for p in b.params:
            assert p.grad.data_ptr() >= b.flat.data_ptr() and torch.count_nonzero(p.grad) > 0
    assert red.passes == 1 and not any(b.launched for b in red.buckets)


def test_unused_parameters_are_reduced_as_zeros_at_the_end_of_backward():
    m = _model()
    extra = torch.nn.Linear(4, 4)                       # never used in forward
    calls = []
    red = B2cReducer(list(m.parameters()) + list(extra.parameters()), None, bucket_cap_bytes=600, first_bucket_bytes=100,
                     world=2, fallback_allreduce=lambda flat, scale: calls.append(flat.data_ptr()))
    m(torch.randn(3, 8)).sum().backward()
    assert calls == [b.flat.data_ptr() for b in red.buckets]   # every bucket, still in order
    assert float(extra.weight.grad.abs().sum()) == 0.0


def test_rearm_points_none_gradients_back_at_zeroed_slices():
    m = _model()
    red = B2cReducer(m.parameters(), None, bucket_cap_bytes=600, first_bucket_bytes=100, world=1)
    m(torch.randn(3, 8)).sum().backward()
    for p in m.parameters():
        p.grad = None                                   # optimizer.zero_grad(set_to_none=True)
    red.rearm()
    for p in m.parameters():
        assert p.grad is p._b2c_grad_view and float(p.grad.abs().sum()) == 0.0
    g1 = [p.grad.data_ptr() for p in m.parameters()]
    m(torch.randn(3, 8)).sum().backward()
    assert [p.grad.data_ptr() for p in m.parameters()] == g1    # no fresh tensors, no copies


def test_no_sync_accumulates_without_launching():
    m = _model()
    calls = []
    red = B2cReducer(m.parameters(), None, bucket_cap_bytes=1 << 20, world=2,
                     fallback_allreduce=lambda flat, scale: calls.append(1))
    red._enabled = False
    m(torch.randn(3, 8)).sum().backward()
    assert calls == []
    g = m[0].weight.grad.clone()
    red._enabled = True
    m(torch.randn(3, 8)).sum().backward()
    assert len(calls) == len(red.buckets)
    assert not torch.equal(m[0].weight.grad, g)         # second pass accumulated on top of the first


def test_fuse_optimizer_declines_when_a_bucket_spans_param_groups():
    """The fused step takes one set of hyper-parameters per bucket: an optimizer whose param_groups cut through a
    bucket keeps its own step() (with a warning) instead of failing."""
    import warnings
    from ray_lightning_accelerators_b200.ddp_hook import B2cDataParallel
    m = _model()
    w = B2cDataParallel.__new__(B2cDataParallel)       # host logic only: no communicator on the CPU box
    torch.nn.Module.__init__(w)
    w.module, w._c4_folded = m, False
    w.reducer = B2cReducer(m.parameters(), None, world=1)
    opt = torch.optim.SGD([{"params": m[0].parameters()}, {"params": list(m[2].parameters()) + list(m[4].parameters()), "lr": 0.01}], lr=0.1)
    with warnings.catch_warnings(record=True) as rec:
        warnings.simplefilter("always")
        assert w.fuse_optimizer(opt) is opt
    assert any("not applicable" in str(r.message) for r in rec)
    assert w.fuse_optimizer(torch.optim.RMSprop(m.parameters())) is not None   # unsupported kind: returned unchanged
