"""GPU: parity at BASELINE.json's full sizes.

configs[0] (capacity 2^16, batch 128, full 4x84x84 rows, 820 adds of K=80 so that
eviction is exercised): differential against the oracle, bit-exact.

configs[1] (capacity 2^21 x 56,617 B = 148 GB of slabs, batch 512): the oracle
cannot hold it; checked through size-independent properties instead --
(1) every gathered row equals the row that was inserted under its serial
    (frames are a pure function of the serial, regenerated on the device);
(2) sampled ids are non-decreasing in ring order (the reference's walk never rewinds)
    and stratified: id_i's prefix interval contains target_i's segment;
(3) sum_ is the reference's running total, reproduced from the stored weights:
    float32 sequential sum per add block, accumulated in double
    (prioritized_replay.h:58-73), minus the popped weights (:84-96), plus the
    float32 differences of update (:105-118); the updated slots hold the new weights;
(4) importance weights: max == 1, all in (0, 1], equal to (size*w/sum)^-beta / max within 1e-6;
(5) eviction brings size back to capacity and zeroes exactly the oldest slots.
"""
import ctypes as C

import numpy as np
import pytest
import torch

from harness import CudaReplay, Oracle, apply_trace, compare_records, ff_schema, gen_trace

pytestmark = pytest.mark.gpu


def test_config1_full_size_differential():
    schema = ff_schema()
    cap, K, B = 1 << 16, 80, 128
    ops = gen_trace(schema, cap, 1, 24, K, B, prefill=cap - 4 * K, adds_per_cycle=(1, 3))
    orc = Oracle(schema, cap, 7, 0.6, 0.4, 0)
    cu = CudaReplay(schema, cap, 7, 0.6, 0.4, 0)
    ra = apply_trace(orc, ops, 0.6, False)
    rb = apply_trace(cu, ops, 0.6, False)
    assert cu.sync() == 0
    compare_records(ra, rb, "config 1")
    assert orc.state() == cu.state()
    assert cu.stats()["pops"] > 0 and cu.stats()["samples_ordered"] == 0


def _frames(serial, salt):
    """uint8 [n, 4, 84, 84] as a pure function of the serial (int64 [n], device)."""
    col = torch.arange(28224, device=serial.device, dtype=torch.int64)[None, :]
    v = (serial[:, None] * 2654435761 + col * 40503 + salt * 97) >> 7
    return (v & 255).to(torch.uint8).reshape(-1, 4, 84, 84)


def test_config2_full_size_properties():
    free, _ = torch.cuda.mem_get_info()
    if free < 165e9:
        pytest.skip("needs 165 GB of free HBM")
    import rela_b200.rela as rela
    from rela_b200 import capi

    dev = torch.device("cuda", 0)
    cap, B, alpha, beta = 1 << 21, 512, 0.6, 0.4
    S = int(1.25 * cap)
    r = rela.FFPrioritizedReplay(cap, 7, alpha, beta, 0, 0)
    g = torch.Generator(device=dev)
    g.manual_seed(3)
    blk = 4096
    extra = 3 * blk  # > capacity: the first sample evicts
    n_total = cap + extra
    prio_all = torch.empty(n_total, device=dev)
    for s0 in range(0, n_total, blk):
        serial = torch.arange(s0, s0 + blk, device=dev, dtype=torch.int64)
        p = torch.rand(blk, device=dev, generator=g) + 0.01
        prio_all[s0:s0 + blk] = p
        tr = rela.FFTransition({"s": _frames(serial, 1), "eps": (serial % 1000).float()[:, None] / 1000,
                                "legal_move": torch.ones((blk, 18), device=dev)},
                               {"a": serial}, (serial % 97).float(), (serial % 200) == 0,
                               ((serial % 200) != 0).float(),
                               {"s": _frames(serial, 2), "eps": (serial % 1000).float()[:, None] / 1000,
                                "legal_move": torch.ones((blk, 18), device=dev)})
        r.add(tr, p)
    r.sync()
    assert r.size() == n_total and r.num_add() == n_total
    L = capi.lib()
    h = C.c_void_p(r.handle())

    def state():
        v = [C.c_int() for _ in range(4)]
        s = C.c_double()
        capi.check(L.rr_debug_state(h, *[C.addressof(x) for x in v], C.addressof(s)))
        w = np.zeros(S, np.float32)
        e = np.zeros(S, np.uint8)
        capi.check(L.rr_debug_weights(h, w.ctypes.data, e.ctypes.data))
        return dict(head=v[0].value, size=v[3].value, sum=s.value), w, e

    st, w0, _ = state()
    assert st["head"] == 0 and st["size"] == n_total
    expect_sum = 0.0
    for s0 in range(0, n_total, blk):  # (3) blockAppend: float block sum in index order, then sum_ += (double) it
        expect_sum += float(np.cumsum(w0[s0:s0 + blk], dtype=np.float32)[-1])
    assert st["sum"] == expect_sum
    for it in range(3):
        st_before, w_before, _ = state()
        batch, wt = r.sample(B, "cuda:0")
        ids = np.zeros(B, np.int32)
        capi.check(L.rr_last_sample_ids(h, B, ids.ctypes.data))
        st, w1, ev = state()
        # (5) eviction to capacity on the first sample
        if it == 0:
            expect_sum += -float(np.sum(w_before[:extra].astype(np.float64)))  # blockPop: diff -= w, sum_ += diff
            assert st["sum"] == expect_sum
            assert st["size"] == cap and st["head"] == extra
            assert (w1[:extra] == 0).all() and (ev[:extra] == 1).all() and (w1[extra:n_total] > 0).all()
        else:
            assert st["size"] == cap
        # (1) gathered rows == inserted rows
        serial = batch.action["a"]
        assert torch.equal(batch.obs["s"], _frames(serial, 1))
        assert torch.equal(batch.next_obs["s"], _frames(serial, 2))
        assert torch.equal(batch.reward, (serial % 97).float())
        assert torch.equal(batch.terminal, (serial % 200) == 0)
        assert torch.equal(batch.obs["eps"], (serial % 1000).float()[:, None] / 1000)
        np.testing.assert_array_equal(serial.cpu().numpy() % S, ids)  # slot of serial s is s mod S (no wrap yet)
        # (2) ring order + stratification against the pre-sample weights
        head = st_before["head"]
        logical = (ids - head) % S
        assert (np.diff(logical) >= 0).all()
        live = np.roll(w_before.astype(np.float64), -head)[: st_before["size"]]
        prefix = np.cumsum(live)
        total = np.float32(st_before["sum"])
        seg = np.float32(total / np.float32(B))
        for i in (0, 1, B // 2, B - 1):
            lo_t, hi_t = np.float64(seg) * i, min(np.float64(seg) * (i + 1), float(total))
            k = int(logical[i])
            below = prefix[k - 1] if k > 0 else 0.0
            assert below < hi_t * (1 + 1e-6) and prefix[k] >= lo_t * (1 - 1e-6), (i, below, prefix[k], lo_t, hi_t)
        # (4) importance weights
        wsel = w_before[ids].astype(np.float64)
        size_is = st_before["size"]
        raw = (size_is * (wsel / float(total))) ** (-beta)
        got = wt.cpu().numpy()
        np.testing.assert_allclose(got, raw / raw.max(), rtol=2e-6)
        assert got.max() == 1.0 and (got > 0).all()
        # (3) update round trip
        newp = torch.rand(B, generator=torch.Generator().manual_seed(it)) + 0.01
        r.update_priority(newp)
        st2, w2, _ = state()
        neww = torch.pow(newp, alpha).numpy()
        keep = ev[ids] == 0
        cur = w1.copy()
        diff = 0.0
        for s, x, k in zip(ids, neww, keep):  # ConcurrentQueue::update, in order, skipping evicted ids
            if k:
                diff += float(np.float32(np.float32(x) - cur[s]))
                cur[s] = np.float32(x)
        expect_sum += diff
        assert st2["sum"] == expect_sum
        np.testing.assert_array_equal(w2, cur)
    r.sync()
