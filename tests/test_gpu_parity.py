"""GPU parity: the CUDA path through the C ABI vs the C oracle on identical
scripted traces. ids and gathered bytes bit-exact; IS weights <= 1e-6 rel."""
import numpy as np
import pytest

from harness import (CudaReplay, Oracle, apply_trace, compare_records, ff_schema, gen_trace, rnn_schema)

pytestmark = pytest.mark.gpu


def _run_pair(schema, cap, seed, alpha, beta, prefetch, ops, cuda_cls=CudaReplay, **cuda_kw):
    orc = Oracle(schema, cap, seed, alpha, beta, prefetch)
    cu = cuda_cls(schema, cap, seed, alpha, beta, prefetch, **cuda_kw)
    ra = apply_trace(orc, ops, alpha, False)
    rb = apply_trace(cu, ops, alpha, False)
    assert cu.sync() == 0
    compare_records(ra, rb, "oracle vs cuda")
    sa, sb = orc.state(), cu.state()
    assert sa == sb, (sa, sb)
    wa, ea = orc.weights()
    wb, eb = cu.weights()
    live = np.zeros(orc.S, dtype=bool)
    idx = (sa["head"] + np.arange(sa["safe_size"])) % orc.S
    live[idx] = True
    np.testing.assert_array_equal(wa[live], wb[live])
    np.testing.assert_array_equal(ea[live], eb[live])
    assert (wb[~live] == 0).all(), "dead slots must hold weight 0"
    check_tiles(cu, wb)
    return cu


def check_tiles(cu, w):
    """The incrementally maintained hierarchy entries (level 0: groups, level 1: leaves) equal a
    rebuild from the weights, bit for bit, and (mins always, sums when exactly representable) a
    host recomputation."""
    inc = [cu.tiles(level=lv) for lv in (0, 1)]  # both read before the rebuild overwrites them
    for lv in (0, 1):
        s_inc, m_inc, ts = inc[lv]
        s_full, m_full, _ = cu.tiles(rebuild=True, level=lv)
        np.testing.assert_array_equal(s_inc.view(np.uint64), s_full.view(np.uint64))
        np.testing.assert_array_equal(m_inc.view(np.uint32), m_full.view(np.uint32))
        wp = np.zeros(len(s_inc) * ts, dtype=np.float32)
        wp[: len(w)] = w
        wt = wp.reshape(len(s_inc), ts)
        nz = np.where(wt != 0, np.abs(wt), np.float32(np.finfo(np.float32).max))
        np.testing.assert_array_equal(nz.min(axis=1), m_inc)
        np.testing.assert_allclose(wt.astype(np.float64).sum(axis=1), s_inc, rtol=1e-12, atol=0)


@pytest.mark.parametrize("cap,K,B,alpha,beta,prefetch", [
    (256, 20, 32, 0.6, 0.4, 0),
    (1000, 80, 128, 1.0, 1.0, 0),
    (3000, 80, 512, 0.6, 0.4, 0),
    (256, 20, 32, 0.6, 0.4, 1),
    (256, 20, 32, 0.6, 0.4, 2),
    (40, 10, 64, 0.6, 0.4, 0),   # B > size: many repeated ids in one batch
    (40, 10, 64, 0.6, 0.4, 1),
])
def test_ff_trace(cap, K, B, alpha, beta, prefetch):
    schema = ff_schema(frame=(4, 12, 12), n_actions=6)
    ops = gen_trace(schema, cap, 3, 60, K, B)
    cu = _run_pair(schema, cap, 7, alpha, beta, prefetch, ops)
    st = cu.stats()
    assert st["samples_ordered"] == 0 and st["samples_exact"] > 0


def test_ff_full_frames():
    schema = ff_schema()
    ops = gen_trace(schema, 400, 5, 25, 80, 128)
    _run_pair(schema, 400, 11, 0.6, 0.4, 0, ops)


def test_ff_device_rows():
    schema = ff_schema()
    ops = gen_trace(schema, 300, 6, 20, 40, 64)
    _run_pair(schema, 300, 5, 0.6, 0.4, 0, ops, device_rows=True)


def test_ordered_walk_matches(monkeypatch):
    monkeypatch.setenv("RR_FORCE_ORDERED", "1")
    schema = ff_schema(frame=(4, 12, 12), n_actions=6)
    ops = gen_trace(schema, 3000, 9, 30, 80, 256)
    cu = _run_pair(schema, 3000, 7, 0.6, 0.4, 0, ops)
    assert cu.stats()["samples_exact"] == 0


def test_wide_dynamic_range_takes_ordered_walk():
    """Weights spanning > 2^29 relative to the total are not exactly summable:
    the kernel must detect it and still match the reference walk."""
    schema = ff_schema(frame=(1, 4, 4), n_actions=2)

    def prio(rng, n):
        return (10.0 ** rng.uniform(-12, 3, size=n)).astype(np.float32)

    ops = gen_trace(schema, 2000, 13, 30, 50, 64, prio=prio)
    cu = _run_pair(schema, 2000, 3, 1.0, 0.5, 0, ops)
    assert cu.stats()["samples_ordered"] > 0


def test_zero_priorities_and_leading_zero_mass():
    schema = ff_schema(frame=(1, 4, 4), n_actions=2)
    ops = gen_trace(schema, 500, 17, 40, 25, 32, zero_frac=0.3)
    _run_pair(schema, 500, 3, 1.0, 0.5, 0, ops)


def test_rnn_trace():
    schema = rnn_schema(T=7, frame=(2, 6, 6), n_actions=4, hid=16)
    ops = gen_trace(schema, 64, 5, 30, 8, 16)
    _run_pair(schema, 64, 3, 0.9, 0.6, 0, ops)


def test_rnn_full_size_rows():
    schema = rnn_schema()  # T=123, 4x84x84 frames, hid 512
    ops = gen_trace(schema, 24, 5, 4, 4, 8, prefill=12)
    _run_pair(schema, 24, 3, 0.9, 0.6, 0, ops)


@pytest.mark.parametrize("group_log2,B", [(0, 200), (1, 200), (3, 64), (5, 700)])
def test_small_groups_many_groups(monkeypatch, group_log2, B):
    """The group level of the hierarchy at a small capacity: groups of 1 ... 32 leaves give 44 ... 2
    groups for 11,250 slots, so the scan over group sums, the head offset inside a group, the
    per-group maintenance of update_kernel (both candidate modes: fewer groups than batch elements
    and more) and the group refresh of commit / eviction all see many groups."""
    monkeypatch.setenv("RR_GROUP_LOG2", str(group_log2))
    schema = ff_schema(frame=(1, 4, 4), n_actions=2)
    ops = gen_trace(schema, 9000, 31 + group_log2, 40, 80, B, zero_frac=0.05)
    cu = _run_pair(schema, 9000, 5, 0.6, 0.4, 0, ops)
    assert cu.stats()["samples_ordered"] == 0


def test_more_groups_than_batch_elements(monkeypatch):
    """ng > B: update_kernel's candidates are batch elements, not groups."""
    monkeypatch.setenv("RR_GROUP_LOG2", "0")
    schema = ff_schema(frame=(1, 4, 4), n_actions=2)
    ops = gen_trace(schema, 20000, 91, 30, 400, 24, zero_frac=0.02)
    _run_pair(schema, 20000, 5, 0.6, 0.4, 0, ops)


def test_headline_capacity_selection_differential():
    """BASELINE.json configs[1] geometry -- capacity 2^21, 2,621,440 slots = 160 groups of 64
    leaves, batch 512 -- with tiny rows, so that the ORACLE holds it too: every sampled id is
    compared, not only properties (the oracle walks 2 M weights per sample: a few cycles)."""
    schema = [f for f in ff_schema(frame=(1, 1, 1), n_actions=1) if f.name in ("action.a", "reward")]
    cap = 1 << 21
    ops = gen_trace(schema, cap, 77, 6, 100_000, 512, prefill=cap, adds_per_cycle=(1, 1))
    cu = _run_pair(schema, cap, 9, 0.6, 0.4, 0, ops)
    st = cu.stats()
    assert st["samples_ordered"] == 0 and st["pops"] > 0


def test_six_million_slots():
    """capacity 6 M -> 7.5 M slots -> 458 groups; selection only (tiny rows), a few cycles."""
    schema = [f for f in ff_schema(frame=(1, 1, 1), n_actions=1) if f.name in ("action.a", "reward")]
    cap = 6_000_000
    ops = gen_trace(schema, cap, 77, 4, 200_000, 512, prefill=cap, adds_per_cycle=(1, 1))
    _run_pair(schema, cap, 9, 0.6, 0.4, 0, ops)


def test_selection_overrun_is_reported_at_the_next_call():
    """The reference asserts when the walk runs past the last element
    (prioritized_replay.h:297-302). Here the kernel raises a sticky error that the NEXT
    rr_sample / rr_update_priority returns without any synchronisation, until rr_sync clears it.
    Provoked through the public API: all priorities zero -> sum_ == 0 -> no slot reaches any target."""
    from harness import make_rows
    from rela_b200 import capi

    schema = ff_schema(frame=(1, 4, 4), n_actions=2)
    cu = CudaReplay(schema, 64, 1, 0.6, 0.4, 0)
    rng = np.random.default_rng(0)
    cu.add(make_rows(schema, 32, rng), np.zeros(32, np.float32))
    rc, _, _, _ = cu.sample(8)   # launches; the overrun is raised on the device
    assert rc == 0
    import torch

    torch.cuda.synchronize()
    assert cu.update(np.ones(8, np.float32)) == capi.RR_ERR_OVERRUN
    assert b"reference asserts" in cu.lib.rr_last_error()
    rc, _, _, _ = cu.sample(8)
    assert rc == capi.RR_ERR_OVERRUN or rc == capi.RR_ERR_STATE
    assert cu.sync() == capi.RR_ERR_OVERRUN   # reported and cleared
    assert cu.sync() == 0


@pytest.mark.parametrize("prefetch", [1, 2, 3])
def test_pipelined_full_frames(prefetch):
    """Batches drawn ahead gather on the two gather streams while the control plane (selection,
    update, commit) runs beside them: full 4x84x84 rows so that the gather really is the long
    kernel, adds between samples so that evicted rows are re-reserved while earlier gathers may
    still be queued."""
    from harness import CudaReplayHostPrefetch

    schema = ff_schema()
    cap, alpha = 600, 0.6
    ops = gen_trace(schema, cap, 19 + prefetch, 80, 60, 128, adds_per_cycle=(0, 3))
    orc = Oracle(schema, cap, 3, alpha, 0.4, prefetch)
    cu = CudaReplayHostPrefetch(schema, cap, 3, alpha, 0.4, prefetch)
    ra = apply_trace(orc, ops, alpha, False)
    rb = apply_trace(cu, ops, alpha, False)
    assert cu.sync() == 0
    compare_records(ra, rb, "oracle prefetch=%d vs pipelined" % prefetch)
    assert orc.state() == cu.state()


@pytest.mark.parametrize("prefetch", [1, 2])
@pytest.mark.parametrize("own_stream", [False, True])
def test_library_prefetch_queue(prefetch, own_stream):
    """rr_create(prefetch > 0): the library's own queue (device-to-device hand-out), from a
    foreign stream and from the replay's own stream."""
    schema = ff_schema(frame=(4, 12, 12), n_actions=6)
    cap, alpha = 256, 0.6
    ops = gen_trace(schema, cap, 3, 80, 20, 32)
    _run_pair(schema, cap, 7, alpha, 0.4, prefetch, ops, own_stream=own_stream)


def test_batch_not_multiple_of_cta_targets():
    schema = ff_schema(frame=(1, 4, 4), n_actions=2)
    for B in (1, 31, 33, 700):
        ops = gen_trace(schema, 2000, 41 + B, 10, 40, B)
        _run_pair(schema, 2000, 5, 0.6, 0.4, 0, ops)


@pytest.mark.parametrize("cap,K,B", [(1, 1, 3), (2, 1, 4), (5, 1, 7), (3, 3, 1)])
def test_tiny_capacities(cap, K, B):
    """S = int(1.25 * capacity) is 1, 2, 6, 3 slots: every add wraps, every sample evicts or repeats ids."""
    schema = ff_schema(frame=(1, 4, 4), n_actions=2)
    ops = gen_trace(schema, cap, 50 + cap, 30, K, B, prefill=min(cap, K), adds_per_cycle=(0, 2))
    _run_pair(schema, cap, 3, 0.6, 0.4, 0, ops)


def test_empty_and_oversized_adds_and_protocol_errors():
    import ctypes as C

    from rela_b200 import capi

    schema = ff_schema(frame=(1, 4, 4), n_actions=2)
    cu = CudaReplay(schema, 8, 1, 0.6, 0.4, 0)
    L = cu.lib
    rng = np.random.default_rng(0)
    from harness import make_rows

    # sampling an empty buffer: the reference asserts (prioritized_replay.h:297-302)
    rc, _, _, _ = cu.sample(4)
    assert rc == capi.RR_ERR_EMPTY
    # add of zero rows is a no-op
    cu.add(make_rows(schema, 0, rng), np.zeros(0, np.float32))
    assert cu.size() == 0 and cu.num_add() == 0
    # a block larger than the ring can never be appended (the reference would block forever)
    rows = make_rows(schema, 11, rng)
    with pytest.raises(capi.ReplayError):
        cu.add(rows, np.ones(11, np.float32))
    cu.add(make_rows(schema, 6, rng), np.ones(6, np.float32))
    rc, _, _, _ = cu.sample(4)
    assert rc == 0
    # second sample before update_priority (lines 203-206)
    rc, _, _, _ = cu.sample(4)
    assert rc == capi.RR_ERR_STATE and b"not been updated" in L.rr_last_error()
    # wrong number of priorities (line 237)
    assert cu.update(np.ones(3, np.float32)) == capi.RR_ERR_ARG
    assert cu.update(np.ones(4, np.float32)) == 0
    # update without an outstanding batch
    assert cu.update(np.ones(4, np.float32)) == capi.RR_ERR_STATE
    # batch above the documented bound
    outs = (C.c_void_p * len(schema))()
    assert L.rr_sample(cu.h, 1 << 20, outs, C.c_void_p(1), None) == capi.RR_ERR_ARG
    assert cu.sync() == 0


def test_reference_test_schema_empty_dicts():
    """rela/tests/test_prioritized_replay.cc: obs / action / next_obs are EMPTY dicts and
    reward, terminal, bootstrap are int32 -- the schema is whatever the first add brings."""
    from harness import Field

    schema = [Field("reward", "int32", ()), Field("terminal", "int32", ()), Field("bootstrap", "int32", ())]
    ops = gen_trace(schema, 32, 5, 20, 8, 40)
    _run_pair(schema, 32, 42, 1.0, 1.0, 0, ops)


@pytest.mark.parametrize("device_rows", [False, True])
def test_caller_on_replay_stream(device_rows):
    """rr_get_stream: a caller that passes the replay's own stream gets plain stream order
    (no event hand-shakes; the sample / update kernels chain by PDL) and the same results."""
    schema = ff_schema(frame=(1, 12, 12), n_actions=4)
    ops = gen_trace(schema, 700, 31, 120, 40, 96)
    _run_pair(schema, 700, 5, 0.6, 0.4, 0, ops, own_stream=True, device_rows=device_rows)


@pytest.mark.parametrize("B", [1500, 4096, 16384])
def test_large_batches(B):
    """Batches whose update needs more shared memory than the default per-kernel bound, up to kMaxBatch."""
    from harness import Field

    schema = [Field("action.a", "int64", ()), Field("reward", "float32", (3,))]
    cap = 40000
    ops = gen_trace(schema, cap, 13, 6, 5000, B, prefill=cap - 2000, adds_per_cycle=(1, 2))
    _run_pair(schema, cap, 21, 0.6, 0.4, 0, ops)


@pytest.mark.parametrize("B,group_log2", [(700, 6), (1024, 1), (2000, 0)])
def test_large_batches_drawn_ahead_see_evictions(monkeypatch, B, group_log2):
    """Batches above kUpdStaged (512) are not staged in shared memory by update_kernel: its group
    CTAs read ids, priorities and eviction flags from L2 and CTA 0 works in a device scratch area.
    With batches drawn ahead, later evictions hit slots of a pending batch, so the update must skip
    them (rela/prioritized_replay.h:110-112) on that path too -- in both candidate modes (groups of
    64 leaves: a few groups; groups of 1-2 leaves: more groups than a CTA's share)."""
    from harness import CudaReplayHostPrefetch, Field

    monkeypatch.setenv("RR_GROUP_LOG2", str(group_log2))
    schema = [Field("action.a", "int64", ()), Field("reward", "float32", (3,))]
    cap, alpha, prefetch = 20000, 0.6, 2
    ops = gen_trace(schema, cap, 23 + B, 16, 3000, B, prefill=cap - 1000, adds_per_cycle=(1, 2), zero_frac=0.02)
    cu = _run_pair(schema, cap, 3, alpha, 0.4, prefetch, ops, cuda_cls=CudaReplayHostPrefetch)
    assert cu.stats()["samples_ordered"] == 0


def test_large_batch_more_groups_than_elements(monkeypatch):
    """n > kUpdStaged AND ng > n: unstaged update_kernel whose candidates are batch elements."""
    from harness import Field

    monkeypatch.setenv("RR_GROUP_LOG2", "0")
    schema = [Field("action.a", "int64", ()), Field("reward", "float32", (3,))]
    cap = 300000  # 375,000 slots = 1,465 groups of one leaf > 700
    ops = gen_trace(schema, cap, 5, 8, 20000, 700, prefill=cap - 10000, adds_per_cycle=(1, 2))
    _run_pair(schema, cap, 21, 0.6, 0.4, 0, ops)


@pytest.mark.parametrize("prefetch", [1, 2, 3])
def test_host_managed_prefetch_matches_oracle(prefetch):
    """rr_prefetch_into / rr_take_prefetched (zero-copy prefetch, as the C++ host layer uses them)
    give the reference's prefetch semantics: batches are drawn before the pending updatePriority."""
    from harness import CudaReplayHostPrefetch

    schema = ff_schema(frame=(1, 12, 12), n_actions=4)
    cap, alpha = 256, 0.6
    ops = gen_trace(schema, cap, 19, 60, 20, 32)
    orc = Oracle(schema, cap, 3, alpha, 0.4, prefetch)
    cu = CudaReplayHostPrefetch(schema, cap, 3, alpha, 0.4, prefetch)
    ra = apply_trace(orc, ops, alpha, False)
    rb = apply_trace(cu, ops, alpha, False)
    assert cu.sync() == 0
    compare_records(ra, rb, "oracle prefetch=%d vs host-managed prefetch" % prefetch)
    # protocol errors
    from rela_b200 import capi

    rc, _, _, _ = cu.sample(32)
    assert rc == 0
    assert cu.lib.rr_take_prefetched(cu.h, None) == capi.RR_ERR_STATE  # previous batch not updated
    outs = (__import__("ctypes").c_void_p * len(schema))()
    assert cu.lib.rr_sample(cu.h, 32, outs, 1, None) == capi.RR_ERR_STATE
    assert cu.update(np.ones(32, np.float32)) == 0
    assert cu.lib.rr_sample(cu.h, 32, outs, 1, None) == capi.RR_ERR_STATE  # queued batches pending
    # a checkpoint may be taken with batches drawn ahead (they are saved with it): tests/test_gpu_checkpoint.py
