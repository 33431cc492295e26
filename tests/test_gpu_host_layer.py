"""GPU: the C++ host layer (rela_b200._C, the pybind module carrying rela's
class names) against the golden vectors recorded from the UNMODIFIED reference
and against the oracle. Priorities go in RAW, as the reference's callers pass
them; ids, gathered bytes, ring state, sum_ and stored weights must be
bit-exact, IS weights within 1e-6 relative."""
import numpy as np
import pytest

from harness import (GOLDEN_CASES, CppReplay, CudaReplay, Oracle, apply_trace, check_against_golden,
                     compare_records, ff_schema, gen_trace)

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", sorted(GOLDEN_CASES))
def test_host_layer_matches_reference_golden(name):
    impl = check_against_golden(name, lambda *a: CppReplay(*a), raw_priority=True)
    assert impl.sync() == 0
    st = impl.stats()
    assert st["samples_ordered"] == 0 and st["device_errors"] == 0


@pytest.mark.parametrize("name", sorted(GOLDEN_CASES))
def test_c_abi_matches_reference_golden(name):
    impl = check_against_golden(name, lambda *a: CudaReplay(*a))
    assert impl.sync() == 0


def test_host_layer_python_surface():
    """Names, constructor and methods of rela/pybind.cc:37-59; error behaviour of
    prioritized_replay.h:203-206,236-237 as Python exceptions."""
    import torch
    from rela_b200 import _C

    r = _C.FFPrioritizedReplay(64, 1, 0.6, 0.4, 0)
    assert r.size() == 0 and r.num_add() == 0
    with pytest.raises(RuntimeError, match="empty"):
        r.sample(4, "cuda:0")
    K = 16
    tr = _C.FFTransition({"s": torch.randint(0, 255, (K, 4, 84, 84), dtype=torch.uint8)},
                         {"a": torch.arange(K)}, torch.randn(K), torch.zeros(K, dtype=torch.bool),
                         torch.ones(K), {"s": torch.randint(0, 255, (K, 4, 84, 84), dtype=torch.uint8)})
    r.add(tr, torch.rand(K) + 0.1)
    assert r.size() == K and r.num_add() == K
    batch, w = r.sample(8, "cuda:0")
    assert isinstance(batch, _C.FFTransition)
    assert batch.obs["s"].shape == (8, 4, 84, 84) and batch.obs["s"].is_cuda
    assert batch.action["a"].dtype == torch.int64 and w.shape == (8,)
    ids = batch.action["a"].cpu()
    assert torch.equal(batch.obs["s"].cpu(), tr.obs["s"][ids])
    assert torch.equal(batch.next_obs["s"].cpu(), tr.next_obs["s"][ids])
    with pytest.raises(RuntimeError, match="not been updated"):
        r.sample(8, "cuda:0")
    with pytest.raises(RuntimeError):
        r.update_priority(torch.rand(7))
    with pytest.raises(RuntimeError, match="1-D"):
        r.update_priority(torch.rand(8, 1))
    r.update_priority(torch.rand(8))
    b2, w2 = r.sample(8, "cpu")  # the reference's device="cpu" still works: batch is copied back
    assert not b2.obs["s"].is_cuda and not w2.is_cuda
    r.update_priority(torch.rand(8).cuda())  # device-resident priorities (device pow)
    r.sync()


@pytest.mark.parametrize("issuer", ["1", "0"])
@pytest.mark.parametrize("prefetch", [1, 3])
def test_prefetch_through_the_host_layer_matches_oracle(monkeypatch, prefetch, issuer):
    """prefetch > 0 through the C++ host layer, with its issuer thread (rela_b200/cpp/issuer_core.cc:
    batches drawn ahead and host-priority updates are launched by one background thread in call
    order) and without it (RELA_B200_ISSUER_THREAD=0: everything on the calling thread): the
    reference's prefetch semantics either way -- ids, bytes, weights, ring state, sum_."""
    monkeypatch.setenv("RELA_B200_ISSUER_THREAD", issuer)
    schema = ff_schema(frame=(1, 12, 12), n_actions=4)
    cap, alpha = 256, 0.6
    ops = gen_trace(schema, cap, 19 + prefetch, 80, 20, 32)
    orc = Oracle(schema, cap, 3, alpha, 0.4, prefetch)
    cu = CppReplay(schema, cap, 3, alpha, 0.4, prefetch)
    ra = apply_trace(orc, ops, alpha, False)
    rb = apply_trace(cu, ops, alpha, True)
    assert cu.sync() == 0
    compare_records(ra, rb, "oracle prefetch=%d vs host layer (issuer thread %s)" % (prefetch, issuer))
    assert orc.state() == cu.state()
    st = cu.stats()
    assert st["samples_ordered"] == 0 and st["device_errors"] == 0


@pytest.mark.parametrize("issuer", ["1", "0"])
def test_prefetch_errors_surface_at_the_call_that_caused_them(monkeypatch, issuer):
    """The reference asserts in sample() / updatePriority() (prioritized_replay.h:203-206,236-237);
    with batches drawn ahead -- by the issuer thread or not -- the host layer raises at the same call."""
    import torch
    from rela_b200 import _C

    monkeypatch.setenv("RELA_B200_ISSUER_THREAD", issuer)
    r = _C.FFPrioritizedReplay(64, 1, 0.6, 0.4, 2)
    with pytest.raises(RuntimeError, match="empty"):
        r.sample(4, "cuda:0")
    K = 16
    tr = _C.FFTransition({"s": torch.randint(0, 255, (K, 1, 8, 8), dtype=torch.uint8)}, {"a": torch.arange(K)},
                         torch.randn(K), torch.zeros(K, dtype=torch.bool), torch.ones(K),
                         {"s": torch.randint(0, 255, (K, 1, 8, 8), dtype=torch.uint8)})
    r.add(tr, torch.rand(K) + 0.1)
    with pytest.raises(RuntimeError):
        r.update_priority(torch.rand(8))  # nothing sampled yet
    batch, w = r.sample(8, "cuda:0")
    ids = batch.action["a"].cpu()
    assert torch.equal(batch.obs["s"].cpu(), tr.obs["s"][ids])
    with pytest.raises(RuntimeError, match="not been updated"):
        r.sample(8, "cuda:0")
    with pytest.raises(RuntimeError):
        r.update_priority(torch.rand(7))
    with pytest.raises(RuntimeError, match="1-D"):
        r.update_priority(torch.rand(8, 1))
    r.update_priority(torch.rand(8))
    with pytest.raises(RuntimeError, match="differs"):
        r.sample(4, "cuda:0")  # batches of 8 are already drawn ahead
    for _ in range(5):
        batch, w = r.sample(8, "cuda:0")
        ids = batch.action["a"].cpu()
        assert torch.equal(batch.next_obs["s"].cpu(), tr.next_obs["s"][ids])
        r.update_priority(torch.rand(8).cuda() if _ % 2 else torch.rand(8))  # device / host priorities mixed
    r.sync()
    del r  # destruction with batches still drawn ahead


def test_device_pow_tolerance():
    """Default C ABI mode (device pow, no RR_PRIO_IS_WEIGHT): stored weights are
    within 1e-6 relative of torch.pow (north_star tolerance for priorities)."""
    from harness import torch_pow

    schema = ff_schema(frame=(1, 4, 4), n_actions=2)
    cu = CudaReplay(schema, 4096, 3, 0.6, 0.4, 0, is_weight=False)
    ops = gen_trace(schema, 4096, 21, 0, 256, 64, prefill=4096)
    apply_trace(cu, ops, 0.6, True)
    assert cu.sync() == 0
    w, _ = cu.weights()
    want = np.concatenate([torch_pow(op[2], 0.6) for op in ops if op[0] == "add"])
    np.testing.assert_allclose(w[: len(want)], want, rtol=1e-6, atol=0)
    mism = int((w[: len(want)] != want).sum())
    print("device pow vs torch.pow: %d of %d weights differ by 1 ulp" % (mism, len(want)))


def test_batch_pool_never_recycles_live_tensors():
    """The host layer pools batch tensors; a batch (or a view of it) that the
    caller still holds must never be overwritten by a later sample."""
    import torch
    from rela_b200 import _C

    r = _C.FFPrioritizedReplay(256, 3, 1.0, 1.0, 0)
    K = 64
    tr = _C.FFTransition({"s": torch.randint(0, 255, (K, 4, 20, 20), dtype=torch.uint8)}, {"a": torch.arange(K)},
                         torch.arange(K).float(), torch.zeros(K, dtype=torch.bool), torch.ones(K), {})
    r.add(tr, torch.rand(K) + 0.1)
    held, held_w = r.sample(16, "cuda:0")
    snap = held.obs["s"].clone()
    snap_a = held.action["a"].clone()
    view = None
    ptrs = set()
    for k in range(12):
        r.update_priority(torch.rand(16) + 0.1)
        b, w = r.sample(16, "cuda:0")
        ptrs.add(b.obs["s"].data_ptr())
        assert b.obs["s"].data_ptr() != held.obs["s"].data_ptr()
        assert torch.equal(b.obs["s"].cpu(), tr.obs["s"][b.action["a"].cpu()])
        if k == 3:
            view = b.reward[2:5]  # a view keeps the storage alive after `b` is rebound
            view_snap = view.clone()
            view_ptr = b.reward.data_ptr()
        elif view is not None:
            assert b.reward.data_ptr() != view_ptr
    torch.cuda.synchronize()
    assert torch.equal(held.obs["s"], snap) and torch.equal(held.action["a"], snap_a)
    assert torch.equal(view, view_snap)
    assert len(ptrs) <= 4, "released batches are recycled"
    r.update_priority(torch.rand(16) + 0.1)
    r.sync()
