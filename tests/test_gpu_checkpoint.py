"""Checkpoint / restore (rr_save / rr_load): a restored replay continues
bit-identically to the uninterrupted run -- checked against the oracle run over the whole trace."""
import os

import numpy as np
import pytest

from harness import CudaReplay, Oracle, apply_trace, compare_records, ff_schema, gen_trace, rnn_schema
from test_gpu_parity import check_tiles

pytestmark = pytest.mark.gpu


def _split(ops, frac):
    """Cut after an ("update", ...) op: no batch is outstanding there."""
    cuts = [i + 1 for i, op in enumerate(ops) if op[0] == "update"]
    k = cuts[int(len(cuts) * frac)]
    return ops[:k], ops[k:]


@pytest.mark.parametrize("kind,cap,K,B,frac", [
    ("ff", 300, 25, 48, 0.1),    # saved before the ring has wrapped
    ("ff", 300, 25, 48, 0.6),    # saved after wrap-around and evictions (head != 0)
    ("rnn", 60, 6, 8, 0.5),
])
def test_restore_continues_bit_identically(tmp_path, kind, cap, K, B, frac):
    from rela_b200 import capi

    schema = ff_schema(frame=(1, 10, 10), n_actions=3) if kind == "ff" else rnn_schema(T=7, frame=(1, 6, 6), n_actions=3, hid=8)
    ops = gen_trace(schema, cap, 77, 90, K, B)
    first, second = _split(ops, frac)
    orc = Oracle(schema, cap, 11, 0.6, 0.4, 0)
    want = apply_trace(orc, ops, 0.6, False)

    a = CudaReplay(schema, cap, 11, 0.6, 0.4, 0)
    got = apply_trace(a, first, 0.6, False)
    path = str(tmp_path / "replay.ckpt")
    capi.check(a.lib.rr_save(a.h, path.encode()))
    state_a, (w_a, e_a) = a.state(), a.weights()
    a.close()

    b = CudaReplay(schema, cap, 999, 0.6, 0.4, 0)  # the seed comes back with the RNG state
    capi.check(b.lib.rr_load(b.h, path.encode()))
    assert b.state() == state_a and b.num_add() == sum(len(op[2]) for op in first if op[0] == "add")
    w_b, e_b = b.weights()
    np.testing.assert_array_equal(w_a, w_b)
    np.testing.assert_array_equal(e_a, e_b)
    check_tiles(b, w_b)
    got += apply_trace(b, second, 0.6, False)
    assert b.sync() == 0
    compare_records(want, got, "oracle (uninterrupted) vs cuda (saved + restored)")
    assert orc.state() == b.state()


def test_save_and_load_refuse_bad_states(tmp_path):
    from rela_b200 import capi

    schema = ff_schema(frame=(1, 4, 4), n_actions=2)
    ops = gen_trace(schema, 64, 3, 6, 8, 16)
    a = CudaReplay(schema, 64, 1, 0.6, 0.4, 0)
    apply_trace(a, ops, 0.6, False)
    path = str(tmp_path / "r.ckpt").encode()
    rc, _, _, _ = a.sample(16)
    assert rc == 0
    assert a.lib.rr_save(a.h, path) == capi.RR_ERR_STATE  # a batch awaits its priorities
    assert a.update(np.ones(16, np.float32)) == 0
    capi.check(a.lib.rr_save(a.h, path))
    assert a.lib.rr_load(a.h, path) == capi.RR_ERR_STATE  # not freshly created

    other = CudaReplay(schema, 128, 1, 0.6, 0.4, 0)
    assert other.lib.rr_load(other.h, path) == capi.RR_ERR_ARG and b"capacity" in other.lib.rr_last_error()
    narrow = CudaReplay(ff_schema(frame=(1, 4, 5), n_actions=2), 64, 1, 0.6, 0.4, 0)
    assert narrow.lib.rr_load(narrow.h, path) == capi.RR_ERR_ARG and b"schema" in narrow.lib.rr_last_error()

    raw = open(path, "rb").read()
    cut = str(tmp_path / "cut.ckpt").encode()
    open(cut, "wb").write(raw[: len(raw) - 100])
    fresh = CudaReplay(schema, 64, 1, 0.6, 0.4, 0)
    assert fresh.lib.rr_load(fresh.h, cut) == capi.RR_ERR_ARG and b"truncated" in fresh.lib.rr_last_error()
    junk = str(tmp_path / "junk.ckpt").encode()
    open(junk, "wb").write(b"\0" * 8192)
    fresh2 = CudaReplay(schema, 64, 1, 0.6, 0.4, 0)
    assert fresh2.lib.rr_load(fresh2.h, junk) == capi.RR_ERR_ARG
    assert fresh2.lib.rr_load(fresh2.h, str(tmp_path / "missing").encode()) == capi.RR_ERR_ARG


def test_python_layer_save_load(tmp_path):
    """rela_b200.rela.FFPrioritizedReplay.save / .load: the schema sidecar lets a brand-new
    replay (which has never seen an add) take the checkpoint."""
    import torch

    import rela_b200.rela as rela

    def transition(n, g):
        obs = {"s": torch.randint(0, 256, (n, 2, 8, 8), dtype=torch.uint8, generator=g)}
        act = {"a": torch.randint(0, 4, (n,), dtype=torch.int64, generator=g)}
        nxt = {"s": torch.randint(0, 256, (n, 2, 8, 8), dtype=torch.uint8, generator=g)}
        return rela.FFTransition(obs, act, torch.randn(n, generator=g), torch.zeros(n, dtype=torch.bool),
                                 torch.ones(n), nxt)

    g = torch.Generator().manual_seed(5)
    a = rela.FFPrioritizedReplay(100, 3, 0.6, 0.4, 0)
    for _ in range(7):  # 112 rows: above capacity (the sample evicts 12), below the 125 slots (add never blocks)
        a.add(transition(16, g), torch.rand(16, generator=g) + 0.01)
    batch, w = a.sample(32, "cuda:0")
    a.update_priority(torch.rand(32, generator=g) + 0.01)
    path = str(tmp_path / "py.ckpt")
    a.save(path)
    assert os.path.exists(path + ".schema")

    b = rela.FFPrioritizedReplay(100, 12345, 0.6, 0.4, 0)
    b.load(path)
    assert b.size() == a.size() and b.num_add() == a.num_add()
    prio = torch.rand(32, generator=g) + 0.01
    for _ in range(3):
        ba, wa = a.sample(32, "cuda:0")
        bb, wb = b.sample(32, "cuda:0")
        assert torch.equal(wa, wb)
        assert torch.equal(ba.obs["s"], bb.obs["s"]) and torch.equal(ba.action["a"], bb.action["a"])
        assert torch.equal(ba.reward, bb.reward) and torch.equal(ba.next_obs["s"], bb.next_obs["s"])
        a.update_priority(prio)
        b.update_priority(prio)


@pytest.mark.parametrize("prefetch", [1, 2])
def test_save_load_with_batches_drawn_ahead(tmp_path, prefetch):
    """prefetch > 0 (the reference's main.py default is 1): after the first sample() there are
    always batches drawn ahead. They are saved with the checkpoint -- ids, weights, gathered rows --
    and a restored replay hands out the very batches the uninterrupted run does, also when the
    slots they came from have been overwritten since."""
    import torch

    import rela_b200.rela as rela

    def transition(n, g, serial):
        obs = {"s": torch.randint(0, 256, (n, 2, 8, 8), dtype=torch.uint8, generator=g)}
        act = {"a": torch.arange(serial, serial + n, dtype=torch.int64)}
        nxt = {"s": torch.randint(0, 256, (n, 2, 8, 8), dtype=torch.uint8, generator=g)}
        return rela.FFTransition(obs, act, torch.randn(n, generator=g), torch.zeros(n, dtype=torch.bool),
                                 torch.ones(n), nxt)

    g = torch.Generator().manual_seed(5)
    a = rela.FFPrioritizedReplay(100, 3, 0.6, 0.4, prefetch)
    serial = 0
    for _ in range(7):
        a.add(transition(16, g, serial), torch.rand(16, generator=g) + 0.01)
        serial += 16
    for _ in range(3):
        a.sample(32, "cuda:0")
        a.update_priority(torch.rand(32, generator=g) + 0.01)
        a.add(transition(16, g, serial), torch.rand(16, generator=g) + 0.01)  # overwrites evicted slots
        serial += 16
    path = str(tmp_path / "pf.ckpt")
    a.save(path)   # `prefetch` batches are queued at this point
    b = rela.FFPrioritizedReplay(100, 777, 0.6, 0.4, prefetch)
    b.load(path)
    assert b.size() == a.size() and b.num_add() == a.num_add()
    for k in range(6):
        ba, wa = a.sample(32, "cuda:0")
        bb, wb = b.sample(32, "cuda:0")
        assert torch.equal(wa, wb), k
        assert torch.equal(ba.action["a"], bb.action["a"]), k
        assert torch.equal(ba.obs["s"], bb.obs["s"]) and torch.equal(ba.next_obs["s"], bb.next_obs["s"])
        assert torch.equal(ba.reward, bb.reward)
        prio = torch.rand(32, generator=g) + 0.01
        a.update_priority(prio)
        b.update_priority(prio)
        if k % 2 == 0:
            t = transition(16, g, serial)
            p = torch.rand(16, generator=g) + 0.01
            a.add(t, p)
            b.add(t, p)
            serial += 16
    a.sync()
    b.sync()


def test_save_is_atomic(tmp_path):
    """rr_save writes <path>.tmp and renames: an existing checkpoint survives a failed save."""
    from rela_b200 import capi

    schema = ff_schema(frame=(1, 4, 4), n_actions=2)
    a = CudaReplay(schema, 64, 1, 0.6, 0.4, 0)
    apply_trace(a, gen_trace(schema, 64, 3, 6, 8, 16), 0.6, False)
    path = tmp_path / "r.ckpt"
    capi.check(a.lib.rr_save(a.h, str(path).encode()))
    good = path.read_bytes()
    assert not (tmp_path / "r.ckpt.tmp").exists()
    # a save that cannot create its temporary file fails and leaves the checkpoint alone
    (tmp_path / "r.ckpt.tmp").mkdir()
    assert a.lib.rr_save(a.h, str(path).encode()) == capi.RR_ERR_ARG
    assert path.read_bytes() == good
