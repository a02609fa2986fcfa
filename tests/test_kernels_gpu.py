"""GPU parity tests, through the C-ABI (libb2c.so via ctypes).

One B200 is enough: a LocalGroup runs W virtual ranks of the SAME kernels and flag protocol
in one cooperative launch (grid.y = W).  The cross-process CUDA-IPC path over NVLink is
covered by tests/test_multigpu.py (needs >= 2 GPUs).

Bars (north_star): broadcast bit-exact; allreduce bit-exact against the kernels' contract
(oracle b2c_*), bit-exact against real gloo output where the reduction order cannot matter
(W = 2, or integer payloads at power-of-two W), <= 1e-6 relative (fp32) / one bf16 ulp (bf16)
otherwise.
"""
import os

import numpy as np
import pytest
import torch

from oracle import oracle as O

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    from ray_lightning_accelerators_b200 import _lib
    from ray_lightning_accelerators_b200.comm import (ALGO_AUTO, ALGO_ONESHOT, ALGO_TWOSHOT, LocalGroup)
else:  # collected on CPU boxes, skipped by the marker
    ALGO_AUTO, ALGO_ONESHOT, ALGO_TWOSHOT = 0, 1, 2

DEV = "cuda:0"
# (algo, variant): variant 0 / 1 = staged two-shot (push / pull), 2 = zero-copy two-shot in place
ALGOS = [(ALGO_ONESHOT, 0), (ALGO_TWOSHOT, 0), (ALGO_TWOSHOT, 1), (ALGO_TWOSHOT, 2)]
ALGO_IDS = ["oneshot", "twoshot-push", "twoshot-pull", "twoshot-zerocopy"]


def _select(group, variant):
    group.set_param(_lib.PARAM_ZEROCOPY, 1 if variant == 2 else 0)
    group.set_param(_lib.PARAM_TWOSHOT_VARIANT, variant if variant < 2 else 0)


@pytest.fixture(scope="module")
def groups():
    made = {}

    def get(world, arena_bytes=64 << 20):
        key = (world, arena_bytes)
        if key not in made:
            made[key] = LocalGroup(world, device=0, arena_bytes=arena_bytes)
            made[key].set_param(_lib.PARAM_TIMEOUT_MS, 3000)
        return made[key]

    yield get
    for g in made.values():
        torch.cuda.synchronize()
        g.check_alive()
        g.close()


def _run(group, arrs, algo, variant, wire=None, scale=None, dtype=torch.float32, offset=0):
    """arrs: per-rank numpy fp32 (or uint16 bf16 bits).  Returns per-rank numpy outputs."""
    _select(group, variant)
    ts = []
    for a in arrs:
        if dtype == torch.bfloat16:
            full = torch.zeros(a.size + offset + 8, dtype=torch.int16, device=DEV)
            full[offset:offset + a.size] = torch.from_numpy(a.view(np.int16).copy()).to(DEV)
            t = full.view(torch.bfloat16)[offset:offset + a.size]
        else:
            full = torch.full((a.size + offset + 8,), 7.0, dtype=torch.float32, device=DEV)
            full[offset:offset + a.size] = torch.from_numpy(a).to(DEV)
            t = full[offset:offset + a.size]
        ts.append((full, t))
    group.allreduce([t for _, t in ts], scale=scale, wire_dtype=wire, algo=algo)
    torch.cuda.synchronize()
    group.check_alive()
    outs = []
    for full, t in ts:
        if dtype == torch.bfloat16:
            outs.append(t.view(torch.int16).cpu().numpy().view(np.uint16))
            guard = full[offset + t.numel():].cpu().numpy()
            assert (guard == 0).all(), "wrote past the end of the buffer"
        else:
            outs.append(t.cpu().numpy())
            guard = full[offset + t.numel():].cpu().numpy()
            assert (guard == 7.0).all(), "wrote past the end of the buffer"
            assert (full[:offset].cpu().numpy() == 7.0).all(), "wrote before the buffer"
    return outs


def _assert_f32_bits(got, ref, msg=None):
    """bit-exact, except that a NaN only has to be a NaN (x86 keeps payloads, the GPU
    returns the canonical one)."""
    nan = np.isnan(ref)
    assert np.array_equal(np.isnan(got), nan), msg
    assert np.array_equal(got[~nan].view(np.uint32), ref[~nan].view(np.uint32)), msg


def _assert_replicas(outs):
    for o in outs[1:]:
        assert np.array_equal(o.view(np.uint8), outs[0].view(np.uint8)), "replicas differ bitwise"


# ------------------------------------------------------------------ fp32
@pytest.mark.parametrize("algo,variant", ALGOS, ids=ALGO_IDS)
@pytest.mark.parametrize("world", [1, 2, 3, 4, 5, 8])
def test_allreduce_f32_matches_contract_bitexact(groups, world, algo, variant):
    g = groups(world)
    for kind in ("int", "randn", "wide", "special"):
        for n in (1, 3, 4, 5, 17, 1000, 27882, 262147):
            grads = [O.payload(kind, n, r) for r in range(world)]
            outs = _run(g, grads, algo, variant)
            _assert_replicas(outs)
            _assert_f32_bits(outs[0], O.b2c_allreduce_f32(grads), (kind, n))


@pytest.mark.parametrize("algo,variant", ALGOS, ids=ALGO_IDS)
@pytest.mark.parametrize("world", [2, 3, 4, 8])
def test_allreduce_f32_vs_gloo_golden(groups, golden_dir, world, algo, variant):
    gold = np.load(os.path.join(golden_dir, f"gloo_collectives_w{world}.npz"))
    g = groups(world)
    for name in [k for k in gold.files if k.startswith("ar_f32_")]:
        _, _, kind, n = name.split("_")
        grads = [O.payload(kind, int(n), r) for r in range(world)]
        outs = _run(g, grads, algo, variant)
        ref = gold[name]
        if kind == "special":
            assert np.array_equal(np.isnan(outs[0]), np.isnan(ref))
            assert np.array_equal(np.isinf(outs[0]), np.isinf(ref))
            if world == 2:
                m = ~np.isnan(ref)
                assert np.array_equal(outs[0][m].view(np.uint32), ref[m].view(np.uint32))
            continue
        if world == 2 or (kind == "int" and world in (4, 8)):
            assert np.array_equal(outs[0].view(np.uint32), ref.view(np.uint32)), name  # bit-exact
        else:
            denom = sum(np.abs(x.astype(np.float64)) for x in grads) / world
            assert O.rel_err(outs[0], ref, denom) <= 1e-6, name  # fp32 tolerance of north_star


@pytest.mark.parametrize("offset", [1, 2, 3])
def test_allreduce_f32_misaligned_base(groups, offset):
    g = groups(4)
    for algo, variant in ALGOS:
        grads = [O.payload("randn", 1031, r) for r in range(4)]
        outs = _run(g, grads, algo, variant, offset=offset)
        assert np.array_equal(outs[0].view(np.uint32), O.b2c_allreduce_f32(grads).view(np.uint32))
        _assert_replicas(outs)


def test_allreduce_empty_is_noop(groups):
    g = groups(2)
    ts = [torch.empty(0, device=DEV), torch.empty(0, device=DEV)]
    g.allreduce(ts)
    torch.cuda.synchronize()


def test_custom_scale_sum(groups):
    g = groups(4)
    grads = [O.payload("randn", 4099, r) for r in range(4)]
    outs = _run(g, grads, ALGO_TWOSHOT, 0, scale=1.0)
    assert np.array_equal(outs[0].view(np.uint32), O.b2c_allreduce_f32(grads, 1.0).view(np.uint32))


# ------------------------------------------------------------------ bf16 and bf16 wire
@pytest.mark.parametrize("algo,variant", ALGOS, ids=ALGO_IDS)
@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_allreduce_bf16_matches_contract_bitexact(groups, world, algo, variant):
    g = groups(world)
    for kind in ("int", "randn", "wide", "special"):
        for n in (1, 7, 8, 9, 1000, 27882, 131075):
            bits = [O.to_bf16_bits(O.payload(kind, n, r)) for r in range(world)]
            outs = _run(g, bits, algo, variant, dtype=torch.bfloat16)
            _assert_replicas(outs)
            ref = O.b2c_allreduce_bf16(bits)
            nan = (ref & 0x7fff) > 0x7f80
            assert np.array_equal(outs[0][~nan], ref[~nan]), (kind, n)
            assert np.array_equal((outs[0] & 0x7fff) > 0x7f80, nan), (kind, n)


@pytest.mark.parametrize("world", [2, 4, 8])
def test_allreduce_bf16_vs_gloo_golden(groups, golden_dir, world):
    gold = np.load(os.path.join(golden_dir, f"gloo_collectives_w{world}.npz"))
    g = groups(world)
    for name in [k for k in gold.files if k.startswith("ar_bf16_")]:
        _, _, kind, n = name.split("_")
        bits = [O.to_bf16_bits(O.payload(kind, int(n), r)) for r in range(world)]
        for algo, variant in ALGOS:
            outs = _run(g, bits, algo, variant, dtype=torch.bfloat16)
            if kind == "int" or world == 2:
                assert O.bf16_ulp_diff(outs[0], gold[name]).max() <= (0 if kind == "int" else 1), name
            else:
                denom = sum(np.abs(O.from_bf16_bits(b).astype(np.float64)) for b in bits) / world
                assert O.rel_err(O.from_bf16_bits(outs[0]), O.from_bf16_bits(gold[name]), denom) <= world * 2.0 ** -8


@pytest.mark.parametrize("algo,variant", ALGOS, ids=ALGO_IDS)
@pytest.mark.parametrize("world", [2, 8])
def test_allreduce_f32_on_bf16_wire(groups, world, algo, variant):
    g = groups(world)
    for n in (1, 8, 13, 4099, 27882):
        grads = [O.payload("randn", n, r) for r in range(world)]
        outs = _run(g, grads, algo, variant, wire=torch.bfloat16)
        _assert_replicas(outs)
        ref = O.b2c_allreduce_f32_bf16wire(grads)
        assert np.array_equal(outs[0].view(np.uint32), ref.view(np.uint32)), n


# ------------------------------------------------------------------ broadcast
@pytest.mark.parametrize("world", [2, 3, 8])
def test_broadcast_bit_exact(groups, golden_dir, world):
    g = groups(world)
    rng = np.random.default_rng(5)
    for n in (1, 15, 16, 17, 4099, 111528, 1 << 20):
        for root in {0, world - 1, n % world}:
            for offset in (0, 1):
                bufs = [rng.integers(0, 256, size=n + offset + 16, dtype=np.uint8) for _ in range(world)]
                full = [torch.from_numpy(b.copy()).to(DEV) for b in bufs]
                g.broadcast([f[offset:offset + n] for f in full], root=root)
                torch.cuda.synchronize()
                g.check_alive()
                want = bufs[root][offset:offset + n]
                for r in range(world):
                    got = full[r].cpu().numpy()
                    assert np.array_equal(got[offset:offset + n], want), (n, root, r)
                    assert np.array_equal(got[:offset], bufs[r][:offset])
                    assert np.array_equal(got[offset + n:], bufs[r][offset + n:]), "wrote past the end"
    # the gloo golden vectors (same seeds as tests/golden/make_golden.py)
    gold = np.load(os.path.join(golden_dir, f"gloo_collectives_w{world}.npz"))
    for name in [k for k in gold.files if k.startswith("bc_")]:
        n = int(name.split("_")[1])
        root = int(name.split("root")[1])
        ts = [torch.from_numpy(np.random.default_rng(77 + r).integers(0, 256, size=n, dtype=np.uint8)).to(DEV)
              for r in range(world)]
        g.broadcast(ts, root=root)
        torch.cuda.synchronize()
        for t in ts:
            assert np.array_equal(t.cpu().numpy(), gold[name]), name


def test_broadcast_nan_payload_bits_survive(groups):
    g = groups(4)
    x = torch.tensor([0x7fc00001, 0xffc12345, 0x7f800001, 0x00000001], dtype=torch.int64).to(torch.int32)
    ts = [x.clone().to(DEV) if r == 2 else torch.zeros(4, dtype=torch.int32, device=DEV) for r in range(4)]
    g.broadcast([t.view(torch.float32) for t in ts], root=2)
    torch.cuda.synchronize()
    for t in ts:
        assert torch.equal(t.cpu(), x)


# ------------------------------------------------------------------ protocol: epochs, double buffering, chunking
def test_back_to_back_calls_change_payload_every_time(groups):
    """Stale-read / ABA detector: 60 calls, no host sync in between, payload, size and
    algorithm change every call; every result is checked afterwards."""
    world = 4
    g = groups(world)
    rng = np.random.default_rng(0)
    sizes = [int(s) for s in rng.integers(1, 300000, size=60)]
    todo = []
    for i, n in enumerate(sizes):
        algo, variant = ALGOS[i % 4]
        _select(g, variant)
        grads = [O.payload("randn", n, r, seed=100 * i) for r in range(world)]
        ts = [torch.from_numpy(x).to(DEV, non_blocking=False) for x in grads]
        g.allreduce(ts, algo=algo)
        if i % 7 == 3:
            g.barrier()
        if i % 11 == 5:
            bs = [torch.full((1000,), float(r), device=DEV) for r in range(world)]
            g.broadcast(bs, root=i % world)
            todo.append(("b", bs, i % world))
        todo.append(("a", ts, grads))
    torch.cuda.synchronize()
    g.check_alive()
    for kind, ts, ref in todo:
        if kind == "a":
            want = O.b2c_allreduce_f32(ref)
            for t in ts:
                assert np.array_equal(t.cpu().numpy().view(np.uint32), want.view(np.uint32))
        else:
            for t in ts:
                assert (t.cpu().numpy() == float(ref)).all()


def test_message_larger_than_arena_slot_is_chunked(groups):
    world = 2
    g = groups(world, arena_bytes=1 << 20)  # 512 KiB slots
    n = (3 << 20) // 4 + 5  # 3 MiB + tail
    grads = [O.payload("randn", n, r) for r in range(world)]
    for algo, variant in ALGOS:
        outs = _run(g, grads, algo, variant)
        assert np.array_equal(outs[0].view(np.uint32), O.b2c_allreduce_f32(grads).view(np.uint32))
    bufs = [torch.from_numpy(np.random.default_rng(r).integers(0, 256, size=(3 << 20) + 3, dtype=np.uint8)).to(DEV)
            for r in range(world)]
    want = bufs[1].clone()
    g.broadcast(bufs, root=1)
    torch.cuda.synchronize()
    assert torch.equal(bufs[0], want) and torch.equal(bufs[1], want)


def test_auto_algo_follows_crossover(groups):
    g = groups(4)
    g.set_param(_lib.PARAM_CROSSOVER_BYTES, 64 * 1024)
    small = [torch.ones(1024, device=DEV) for _ in range(4)]
    big = [torch.ones(1 << 20, device=DEV) for _ in range(4)]
    g.allreduce(small, algo=ALGO_AUTO)
    assert g.stats()["last_algo"] == ALGO_ONESHOT
    g.allreduce(big, algo=ALGO_AUTO)
    assert g.stats()["last_algo"] == ALGO_TWOSHOT
    torch.cuda.synchronize()
    assert torch.equal(big[0], torch.ones(1 << 20, device=DEV))
    g.set_param(_lib.PARAM_CROSSOVER_BYTES, 512 * 1024)


# ------------------------------------------------------------------ full-size properties (ResNet-50 buckets)
RESNET50_BUCKET_ELEMS = [3_103_744, 7_875_587, 7_416_832, 6_754_305, 406_568]  # ~ResNet-50's 11.84/30.04/28.29/25.77/1.55 MiB buckets (two made ragged)


@pytest.mark.parametrize("world", [2, 8])
def test_full_size_buckets_properties(groups, world):
    """At BASELINE sizes the CPU oracle is too slow to run per case; use size-independent
    properties instead: (1) the torch fp32 restatement of the contract on the GPU
    (((g0*s)+(g1*s))+...) must match bit for bit, (2) identical inputs are a fixed point at
    W = 2, (3) replicas are bitwise identical."""
    g = groups(world, arena_bytes=160 << 20)
    for n in RESNET50_BUCKET_ELEMS:
        gen = torch.Generator(device=DEV).manual_seed(n)
        grads = [torch.randn(n, device=DEV, generator=gen) for _ in range(world)]
        s = 1.0 / world
        want = grads[0] * s
        for x in grads[1:]:
            want = want + x * s
        for algo, variant in ALGOS:
            _select(g, variant)
            ts = [x.clone() for x in grads]
            g.allreduce(ts, algo=algo)
            torch.cuda.synchronize()
            for t in ts:
                assert torch.equal(t.view(torch.int32), want.view(torch.int32)), (n, algo, variant)
        if world == 2:  # x/2 + x/2 == x exactly; for W > 2 partial sums like 3x/W need extra bits
            same = [grads[0].clone() for _ in range(world)]
            g.allreduce(same, algo=ALGO_TWOSHOT)
            torch.cuda.synchronize()
            assert torch.equal(same[0], grads[0]) and torch.equal(same[-1], grads[0])
    g.check_alive()


def test_bf16_full_size_bucket(groups):
    world = 8
    g = groups(world, arena_bytes=160 << 20)
    n = 23_440_896  # BERT-base bf16 embedding bucket (~44.7 MiB)
    gen = torch.Generator(device=DEV).manual_seed(1)
    grads = [torch.randn(n, device=DEV, generator=gen).to(torch.bfloat16) for _ in range(world)]
    acc = grads[0].float() * 0.125
    for x in grads[1:]:
        acc = acc + x.float() * 0.125
    want = acc.to(torch.bfloat16)
    ts = [x.clone() for x in grads]
    g.allreduce(ts, algo=ALGO_TWOSHOT)
    torch.cuda.synchronize()
    for t in ts:
        assert torch.equal(t.view(torch.int16), want.view(torch.int16))


# ------------------------------------------------------------------ liveness
def test_missing_peer_times_out_instead_of_hanging():
    """Two ranks in this process; only rank 0 ever launches.  The bounded spin must give up,
    mark the comm dead (sticky B2C_ETIMEOUT) and leave the GPU usable."""
    import ctypes
    lib = _lib.load()
    hs = []
    for r in range(2):
        h = ctypes.c_void_p()
        _lib.check(lib.b2c_comm_create(r, 2, 0, 1 << 20, ctypes.byref(h)))
        hs.append(h)
    blobs = b""
    for h in hs:
        buf = ctypes.create_string_buffer(_lib.EXPORT_BLOB_BYTES)
        n = ctypes.c_size_t(_lib.EXPORT_BLOB_BYTES)
        _lib.check(lib.b2c_comm_export(h, buf, ctypes.byref(n)))
        blobs += buf.raw
    for h in hs:
        _lib.check(lib.b2c_comm_import(h, blobs, _lib.EXPORT_BLOB_BYTES))
    _lib.check(lib.b2c_set_param(hs[0], _lib.PARAM_TIMEOUT_MS, 50))
    t = torch.ones(1000, device=DEV)
    s = torch.cuda.current_stream().cuda_stream
    _lib.check(lib.b2c_allreduce(hs[0], ctypes.c_void_p(t.data_ptr()), 1000, _lib.F32, _lib.F32,
                                 ctypes.c_float(0.5), ALGO_ONESHOT, ctypes.c_void_p(s)))
    torch.cuda.synchronize()
    st = _lib.Stats()
    _lib.check(lib.b2c_get_stats(hs[0], ctypes.byref(st)))
    assert st.device_status == -_lib.E_TIMEOUT
    assert st.device_status_detail[2] == 1  # was waiting for rank 1
    rc = lib.b2c_allreduce(hs[0], ctypes.c_void_p(t.data_ptr()), 1000, _lib.F32, _lib.F32,
                           ctypes.c_float(0.5), ALGO_ONESHOT, ctypes.c_void_p(s))
    assert rc == _lib.E_TIMEOUT and b"timed out" in lib.b2c_last_error()
    for h in hs:
        _lib.check(lib.b2c_comm_destroy(h))
    assert float((torch.ones(4, device=DEV) * 2).sum().item()) == 8.0  # GPU still healthy


def test_error_paths_do_not_throw():
    import ctypes
    lib = _lib.load()
    h = ctypes.c_void_p()
    assert lib.b2c_comm_create(0, 9, 0, 1 << 20, ctypes.byref(h)) == _lib.E_INVAL
    assert lib.b2c_comm_create(3, 2, 0, 1 << 20, ctypes.byref(h)) == _lib.E_INVAL
    _lib.check(lib.b2c_comm_create(0, 2, 0, 1 << 20, ctypes.byref(h)))
    t = torch.ones(8, device=DEV)
    rc = lib.b2c_allreduce(h, ctypes.c_void_p(t.data_ptr()), 8, _lib.F32, _lib.F32, ctypes.c_float(1.0), 0, None)
    assert rc == _lib.E_STATE and b"import" in lib.b2c_last_error()
    _lib.check(lib.b2c_comm_destroy(h))
    g = LocalGroup(2, device=0, arena_bytes=1 << 20)
    with pytest.raises(TypeError):
        g.allreduce([torch.ones(4, device=DEV, dtype=torch.float16)] * 2)
    with pytest.raises(ValueError):
        g.allreduce([torch.ones(4), torch.ones(4)])  # CPU tensors: there is no CPU path
    with pytest.raises(_lib.B2cError):
        g.set_param(_lib.PARAM_THREADS, 100)
    g.close()


def test_single_rank_average_is_a_no_op_without_a_launch():
    """world = 1, scale = 1: the average over one rank is the input bit for bit (NaN payloads included), so the call
    returns without launching anything; a custom scale still runs the streaming scale pass."""
    from ray_lightning_accelerators_b200.comm import Comm
    comm = Comm(0, 1, device=0, arena_bytes=1 << 20)
    x = torch.from_numpy(O.payload("special", 4099, 0)).to(DEV)
    before = x.clone()
    l0 = comm.stats()["launches"]
    comm.allreduce(x)
    torch.cuda.synchronize()
    st = comm.stats()
    assert st["launches"] == l0 and st["noop_calls"] == 1
    assert torch.equal(x.view(torch.int32), before.view(torch.int32))
    y = torch.full((4099,), 3.0, device=DEV)
    comm.allreduce(y, scale=0.5)
    torch.cuda.synchronize()
    assert comm.stats()["launches"] == l0 + 1 and bool((y == 1.5).all())
    comm.close()


def test_phase_ring_keeps_the_last_kernels_in_launch_order(groups):
    """b2c_get_phase_ring: four %globaltimer stamps per kernel for the last 32 kernels of a comm, oldest first —
    what bench.py uses to split the in-step time of a bucket into 'waiting for peers' and 'moving data'."""
    g = groups(4)
    ts = [torch.ones(50000, device=DEV) for _ in range(4)]
    base = g.phase_ring_us()[1]
    for _ in range(40):
        g.allreduce(ts, algo=ALGO_TWOSHOT)
    rows, total = g.phase_ring_us()
    assert total == base + 40 and len(rows) == 32
    starts = [r[3] for r in rows]
    assert starts == sorted(starts) and len(set(starts)) == 32
    assert all(r[0] >= 0 and r[1] >= 0 and r[0] + r[1] < 5000 for r in rows)   # microseconds, sane
