"""CPU: size-independent properties of the oracle (the same ones the GPU suite uses at full bucket sizes)."""
import numpy as np
from hypothesis import given, settings
from hypothesis import strategies as st

from oracle import oracle as O


@settings(max_examples=40, deadline=None)
@given(world=st.sampled_from([1, 2, 3, 4, 8]), n=st.integers(1, 300), seed=st.integers(0, 10_000))
def test_integer_payloads_do_not_depend_on_the_order(world, n, seed):
    grads = [O.payload("int", n, r, seed=seed) for r in range(world)]
    order = list(np.random.default_rng(seed).permutation(world))
    a = O.ddp_allreduce_f32(grads)
    b = O.ddp_allreduce_f32(grads, order=[int(i) for i in order])
    if world in (1, 2, 4, 8):  # k/W and every partial sum exact
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
        assert np.array_equal(a.view(np.uint32), O.b2c_allreduce_f32(grads).view(np.uint32))
        assert np.array_equal(a.view(np.uint32), O.hvd_average_f32(grads).view(np.uint32))
    else:
        assert np.allclose(a, b, rtol=1e-6, atol=1e-6)


@settings(max_examples=40, deadline=None)
@given(world=st.sampled_from([2, 4, 8]), n=st.integers(1, 300), seed=st.integers(0, 10_000))
def test_contract_is_within_tolerance_of_the_exact_mean(world, n, seed):
    grads = [O.payload("wide", n, r, seed=seed) for r in range(world)]
    got = O.b2c_allreduce_f32(grads)
    exact = O.exact_mean_f64(grads)
    denom = sum(np.abs(g.astype(np.float64)) for g in grads) / world
    assert O.rel_err(got, exact, denom) <= 1e-6
    bits = [O.to_bf16_bits(g) for g in grads]
    gotb = O.b2c_allreduce_bf16(bits)
    exactb = O.exact_mean_bf16_f64(bits)
    nearest = O.to_bf16_bits(exactb.astype(np.float32))
    assert O.bf16_ulp_diff(gotb, nearest).max() <= 1


@settings(max_examples=25, deadline=None)
@given(n=st.integers(1, 200), seed=st.integers(0, 10_000))
def test_two_equal_ranks_are_a_fixed_point_and_scaling_is_linear(n, seed):
    x = O.payload("randn", n, 0, seed=seed)
    assert np.array_equal(O.b2c_allreduce_f32([x, x]).view(np.uint32), x.view(np.uint32))
    y = O.payload("randn", n, 1, seed=seed)
    s = O.b2c_allreduce_f32([x, y], scale=1.0)
    h = O.b2c_allreduce_f32([x, y], scale=0.5)
    assert np.array_equal((s * np.float32(0.5)).view(np.uint32), h.view(np.uint32))  # power-of-two scale commutes


@settings(max_examples=20, deadline=None)
@given(world=st.integers(1, 8), n=st.integers(0, 100), root=st.integers(0, 7), seed=st.integers(0, 1000))
def test_broadcast_is_idempotent_and_exact(world, n, root, seed):
    root %= world
    bufs = [np.random.default_rng(seed + r).integers(0, 256, n, dtype=np.uint8) for r in range(world)]
    once = O.broadcast(bufs, root)
    twice = O.broadcast(once, root)
    assert all(np.array_equal(a, bufs[root]) for a in once) and all(np.array_equal(a, b) for a, b in zip(once, twice))
