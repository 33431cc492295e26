"""GPU parity: the CUDA path through the C ABI vs the C oracle on identical
scripted traces. ids and gathered bytes bit-exact; IS weights <= 1e-6 rel."""
import numpy as np
import pytest

from harness import (CudaReplay, Oracle, apply_trace, compare_records, ff_schema, gen_trace, rnn_schema)

pytestmark = pytest.mark.gpu


def _run_pair(schema, cap, seed, alpha, beta, prefetch, ops, **cuda_kw):
    orc = Oracle(schema, cap, seed, alpha, beta, prefetch)
    cu = CudaReplay(schema, cap, seed, alpha, beta, prefetch, **cuda_kw)
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
    return cu


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


@pytest.mark.parametrize("tile_log2", [9, 10, 11])
def test_larger_prefix_tiles(monkeypatch, tile_log2):
    """Replays above 4 M slots use tiles of 512 ... 2048 slots (at most 16 K tile
    prefixes fit in shared memory); force them at a small capacity."""
    monkeypatch.setenv("RR_TILE_LOG2", str(tile_log2))
    schema = ff_schema(frame=(1, 4, 4), n_actions=2)
    ops = gen_trace(schema, 9000, 31 + tile_log2, 40, 80, 200, zero_frac=0.05)
    cu = _run_pair(schema, 9000, 5, 0.6, 0.4, 0, ops)
    assert cu.stats()["samples_ordered"] == 0


def test_six_million_slots_natural_tile_size():
    """capacity 6 M -> 7.5 M slots -> 512-slot tiles chosen by rr_create; selection
    only (tiny rows), a few cycles (the oracle walks 6 M weights per sample)."""
    schema = [f for f in ff_schema(frame=(1, 1, 1), n_actions=1) if f.name in ("action.a", "reward")]
    cap = 6_000_000
    ops = gen_trace(schema, cap, 77, 4, 200_000, 512, prefill=cap, adds_per_cycle=(1, 1))
    _run_pair(schema, cap, 9, 0.6, 0.4, 0, ops)


def test_batch_not_multiple_of_cta_targets():
    schema = ff_schema(frame=(1, 4, 4), n_actions=2)
    for B in (1, 31, 33, 700):
        ops = gen_trace(schema, 2000, 41 + B, 10, 40, B)
        _run_pair(schema, 2000, 5, 0.6, 0.4, 0, ops)
