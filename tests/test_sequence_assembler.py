"""R2D2 sequence assembly on the device
(rela_b200::SequenceAssembler) against the reference's R2D2TransitionBuffer
(rela/r2d2_actor.h:10-187, compiled where it lies into tests/dropin/_build/
seq_fixture.so). The same stream of K-env steps -- random episode ends, hidden
states that count the steps since the last reset -- goes through both into two
B200 replays with the same seed; what a learner then samples must be
identical: every field bit for bit, seq_len, h0, importance weights.

Per-step priorities are multiples of 1/64 and eta = 0.5, so the eta-mix is
exact in float32 whatever the summation order; burn-in = 0 cases compare the
carried-over priorities too (with burn-in > 0 the reference reads them out of
bounds, r2d2_actor.h:139 -- see rela_b200/sequence_assembler.h)."""
import os
import sys

import numpy as np
import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
BUILD = os.path.join(HERE, "dropin", "_build")
HAVE = os.path.exists(os.path.join(BUILD, "seq_fixture.so")) and os.path.exists(os.path.join(BUILD, "rela.so"))

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not HAVE, reason="tests/dropin/_build/seq_fixture.so missing")]


def _steps(K, n_steps, seed, p_term, frame=(2, 6, 6), hid=8):
    g = torch.Generator().manual_seed(seed)
    since = torch.zeros(K)
    out = []
    for t in range(n_steps):
        obs = {"s": torch.randint(0, 256, (K,) + frame, dtype=torch.uint8, generator=g),
               "eps": torch.rand((K, 1), generator=g)}
        action = {"a": torch.randint(0, 5, (K,), generator=g)}
        reward = torch.randn(K, generator=g)
        terminal = torch.rand(K, generator=g) < p_term
        bootstrap = 1.0 - terminal.float()
        priority = torch.randint(16, 256, (K,), generator=g).float() / 64
        hidden = {"h0": since.view(1, K, 1).expand(1, K, hid).clone(),
                  "c0": (2 * since).view(1, K, 1).expand(1, K, hid).clone()}
        out.append((obs, action, reward, terminal, bootstrap, priority, hidden))
        since = torch.where(terminal, torch.zeros(K), since + 1)
    return out


@pytest.mark.parametrize("seq_len,burn_in,multi_step,p_term,compare_weights", [
    (8, 0, 2, 0.07, True),
    (6, 0, 3, 0.15, True),
    (8, 4, 2, 0.07, False),
    (4, 3, 3, 0.10, False),   # burnin + multiStep > seqLen: the carry-over overlaps itself
    (5, 2, 1, 0.30, False),   # many short episodes: second sequences from carried terminals
    (12, 6, 3, 0.0, False),   # no terminals: only full windows and carry-overs
])
def test_device_assembly_matches_reference_buffer(seq_len, burn_in, multi_step, p_term, compare_weights):
    if BUILD not in sys.path:
        sys.path.insert(0, BUILD)
    import rela
    import seq_fixture

    K, n_steps, eta, B = 6, 160, 0.5, 32
    T = burn_in + seq_len + multi_step
    steps = _steps(K, n_steps, 100 + seq_len + burn_in, p_term)
    ref = rela.RNNPrioritizedReplay(4096, 3, 1.0, 0.6, 0)
    dev = rela.RNNPrioritizedReplay(4096, 3, 1.0, 0.6, 0)
    n_ref = seq_fixture.feed_reference(ref, K, multi_step, seq_len, burn_in, eta, steps)
    n_dev = seq_fixture.feed_assembler(dev, K, multi_step, seq_len, burn_in, eta, steps)
    assert n_ref == n_dev and n_ref > 5
    assert ref.size() == dev.size() == n_ref and ref.num_add() == dev.num_add()
    if compare_weights:
        draws = 6
        for _ in range(draws):
            ba, wa = ref.sample(B, "cuda:0")
            bb, wb = dev.sample(B, "cuda:0")
            _same(ba, bb, T, B)
            assert torch.allclose(wa, wb, rtol=1e-6, atol=0)
            p = torch.rand(B) + 0.1
            ref.update_priority(p)
            dev.update_priority(p)
    else:
        # the sequence priorities may differ (reference defect): compare the stored
        # sequences themselves by drawing every one of them through uniform priorities
        seen_a, seen_b = {}, {}
        for rep, seen in ((ref, seen_a), (dev, seen_b)):
            for _ in range(60):
                b, w = rep.sample(B, "cuda:0")
                for k in range(B):
                    key = (b.obs["s"][:, k].cpu().numpy().tobytes(), float(b.seq_len[k]))
                    seen[key] = (b.h0["h0"][:, k].cpu(), b.h0["c0"][:, k].cpu(), b.action["a"][:, k].cpu(),
                                 b.reward[:, k].cpu(), b.terminal[:, k].cpu(), b.bootstrap[:, k].cpu(),
                                 b.obs["eps"][:, k].cpu())
                rep.update_priority(torch.ones(B))
        assert len(seen_a) > 5
        common = set(seen_a) & set(seen_b)
        assert len(common) >= 0.8 * min(len(seen_a), len(seen_b)), (len(seen_a), len(seen_b), len(common))
        only = (set(seen_a) ^ set(seen_b))
        # after the first update every weight is 1, so 60 x 32 draws see (almost) everything; whatever
        # was seen on both sides must agree in every other field
        for key in common:
            for x, y in zip(seen_a[key], seen_b[key]):
                assert torch.equal(x, y)
        assert len(only) <= 0.2 * max(len(seen_a), len(seen_b))


def _same(ba, bb, T, B):
    assert tuple(ba.obs["s"].shape)[:2] == (T, B)
    for name in ("s", "eps"):
        assert torch.equal(ba.obs[name], bb.obs[name]), name
    for name in ("h0", "c0"):
        assert torch.equal(ba.h0[name], bb.h0[name]), name
    assert torch.equal(ba.action["a"], bb.action["a"])
    assert torch.equal(ba.reward, bb.reward) and torch.equal(ba.terminal, bb.terminal)
    assert torch.equal(ba.bootstrap, bb.bootstrap) and torch.equal(ba.seq_len, bb.seq_len)


@pytest.mark.parametrize("seq_len,burn_in,multi_step,p_term", [(8, 0, 2, 0.07), (6, 0, 3, 0.15)])
def test_device_priorities_mix_on_the_device(seq_len, burn_in, multi_step, p_term):
    """Per-step priorities handed over as CUDA tensors stay in a device window and the eta-mix is
    computed by the K9 kernel (rr_aggregate_priority) inside popInto: same sequences, same
    priorities (exact here: multiples of 1/64, eta 0.5, alpha 1), same samples as the reference."""
    if BUILD not in sys.path:
        sys.path.insert(0, BUILD)
    import rela
    import seq_fixture

    K, n_steps, eta, B = 6, 160, 0.5, 32
    T = burn_in + seq_len + multi_step
    steps = _steps(K, n_steps, 300 + seq_len, p_term)
    steps_dev = [s[:5] + (s[5].cuda(),) + s[6:] for s in steps]
    ref = rela.RNNPrioritizedReplay(4096, 3, 1.0, 0.6, 0)
    dev = rela.RNNPrioritizedReplay(4096, 3, 1.0, 0.6, 0)
    n_ref = seq_fixture.feed_reference(ref, K, multi_step, seq_len, burn_in, eta, steps)
    n_dev = seq_fixture.feed_assembler(dev, K, multi_step, seq_len, burn_in, eta, steps_dev)
    assert n_ref == n_dev and n_ref > 5
    for _ in range(6):
        ba, wa = ref.sample(B, "cuda:0")
        bb, wb = dev.sample(B, "cuda:0")
        _same(ba, bb, T, B)
        assert torch.allclose(wa, wb, rtol=1e-6, atol=0)
        p = torch.rand(B) + 0.1
        ref.update_priority(p)
        dev.update_priority(p)
