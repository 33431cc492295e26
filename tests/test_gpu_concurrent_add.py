"""GPU: the multi-producer protocol of ConcurrentQueue::blockAppend
(rela/prioritized_replay.h:43-78), after the reference's own
rela/tests/test_concurrent_queue.cc:39-119: many threads append blocks
concurrently; size and sum_ must come out exact, blocks must stay contiguous,
and adders blocked on a full ring must be released by the sampler's eviction."""
import ctypes as C
import threading

import numpy as np
import pytest

from harness import CudaReplay, Field

pytestmark = pytest.mark.gpu

SCHEMA = [Field("action.a", "int64", ()), Field("reward", "float32", (3,))]


def _rows(tag, n):
    a = np.full(n, tag, dtype=np.int64)
    r = np.tile(np.arange(n, dtype=np.float32)[:, None], (1, 3))
    return [a, r]


def test_concurrent_appends_exact_size_and_sum():
    """test_concurrent_queue.cc:39-70: 5,000 concurrent appends of 10, all weights one."""
    n_threads, per_thread, blk = 16, 320, 10
    total = n_threads * per_thread * blk
    cu = CudaReplay(SCHEMA, total, 1, 1.0, 1.0, 0)
    errs = []

    def work(t):
        try:
            for k in range(per_thread):
                cu.add(_rows(t * per_thread + k, blk), np.ones(blk, np.float32))
        except Exception as e:  # noqa: BLE001
            errs.append(e)

    th = [threading.Thread(target=work, args=(t,)) for t in range(n_threads)]
    [x.start() for x in th]
    [x.join() for x in th]
    assert not errs, errs
    assert cu.sync() == 0
    st = cu.state()
    assert st["size"] == st["safe_size"] == total and cu.size() == total and cu.num_add() == total
    assert st["sum"] == float(total)
    w, _ = cu.weights()
    assert (w[:total] == 1).all()
    # every block landed contiguous and intact: sample everything back by id
    rc, rows, ws, ids = cu.sample(4096)
    assert rc == 0
    a, r = rows
    # slot s belongs to block s // blk, row s % blk of it
    np.testing.assert_array_equal(r[:, 0], (ids % blk).astype(np.float32))
    tags_by_block = {}
    for s, tag in zip(ids, a):
        tags_by_block.setdefault(int(s) // blk, set()).add(int(tag))
    assert all(len(v) == 1 for v in tags_by_block.values()), "a block was interleaved with another"


def test_blocked_adders_are_released_by_eviction():
    """test_concurrent_queue.cc:72-119: producers outrun the capacity and block on
    the full ring (line 47) until the consumer's sample evicts (lines 311-315)."""
    cap, blk, n_threads, per_thread = 400, 10, 8, 60   # 4,800 rows through a ring of 500 slots
    cu = CudaReplay(SCHEMA, cap, 1, 1.0, 1.0, 0)
    done = [0]
    lock = threading.Lock()

    def work(t):
        for k in range(per_thread):
            cu.add(_rows(t, blk), np.ones(blk, np.float32))
        with lock:
            done[0] += 1

    th = [threading.Thread(target=work, args=(t,), daemon=True) for t in range(n_threads)]
    [x.start() for x in th]
    import time

    t0 = time.time()
    samples = 0
    while done[0] < n_threads:
        assert time.time() - t0 < 120, "producers stayed blocked"
        if cu.size() == 0:
            time.sleep(0.001)
            continue
        rc, _, w, ids = cu.sample(32, want_rows=False)
        assert rc == 0
        assert cu.update(np.ones(32, np.float32)) == 0
        samples += 1
        assert cu.size() <= int(1.25 * cap)
    [x.join() for x in th]
    assert cu.sync() == 0
    total = n_threads * per_thread * blk
    assert cu.num_add() == total
    # drain: one more sample evicts down to capacity; all weights are one, so sum_ == size
    rc, _, _, _ = cu.sample(32, want_rows=False)
    assert rc == 0 and cu.update(np.ones(32, np.float32)) == 0
    st = cu.state()
    assert st["safe_size"] == cap and st["sum"] == float(cap), st
    assert cu.stats()["pops"] == total - cap
