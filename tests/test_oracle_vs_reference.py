"""CPU: the C oracle against the UNMODIFIED reference run live (oracle/_ref/libref_replay.so, built
from /root/reference by oracle/Makefile) on randomized traces -- a wider net than the committed
golden vectors: many seeds, alpha / beta pairs, prefetch 0 and 1, zero and wide-range priorities,
batches larger than the buffer, rings that wrap and evict many times, RNN records, n-step returns.
Skipped where the reference build is absent (it does not exist on the GPU box unless it travelled).

Everything is bit-exact except the IS weights (<= 1e-6 relative, north_star's bound)."""
import numpy as np
import pytest

from harness import (Oracle, RefReplay, apply_trace, compare_records, ff_schema, gen_trace, have_ref, oracle_lib,
                     ref_lib, rnn_schema)

pytestmark = [pytest.mark.ref, pytest.mark.skipif(not have_ref(), reason="oracle/_ref not built")]


def _diff(schema, cap, seed, alpha, beta, prefetch, ops, ff=True):
    ref = RefReplay(schema, cap, seed, alpha, beta, prefetch)
    orc = Oracle(schema, cap, seed, alpha, beta, prefetch)
    ra = apply_trace(ref, ops, alpha, True)    # the reference raises priorities itself
    rb = apply_trace(orc, ops, alpha, False)   # the oracle is handed torch.pow(priority, alpha)
    assert len(ra) == len(rb) > 0
    compare_records(ra, rb, "reference vs oracle")
    assert ref.num_add() == orc.num_add()
    if ff:
        sa, sb = ref.state(), orc.state()
        assert sa == sb, (sa, sb)  # head, tail, size, safe_size and sum_ (double) exactly
        wa, ea = ref.weights()
        wb, eb = orc.weights()
        live = (sa["head"] + np.arange(sa["safe_size"])) % ref.S
        np.testing.assert_array_equal(wa[live], wb[live])
        np.testing.assert_array_equal(ea[live], eb[live])


@pytest.mark.parametrize("seed", range(6))
@pytest.mark.parametrize("alpha,beta", [(0.6, 0.4), (1.0, 1.0), (0.9, 0.6), (0.0, 0.4)])
def test_ff_random_traces(seed, alpha, beta):
    rng = np.random.default_rng(1000 + seed)
    cap = int(rng.integers(24, 400))
    K = int(rng.integers(1, 40))
    B = int(rng.integers(1, 96))
    schema = ff_schema(frame=(1, 5, 5), n_actions=3)
    ops = gen_trace(schema, cap, 77 + seed, 70, K, B, adds_per_cycle=(0, 4))
    _diff(schema, cap, 5 + seed, alpha, beta, 0, ops)


@pytest.mark.parametrize("seed", range(4))
def test_ff_prefetch1(seed):
    """prefetch = 1 is the deepest deterministic setting of the reference."""
    schema = ff_schema(frame=(1, 4, 4), n_actions=2)
    cap = 90 + 37 * seed
    ops = gen_trace(schema, cap, 300 + seed, 60, 15, 24)
    _diff(schema, cap, 9, 0.6, 0.4, 1, ops)


@pytest.mark.parametrize("zero_frac", [0.3, 0.9])
def test_ff_zero_priorities(zero_frac):
    """Zero weights: skipped by `accSum > 0`, and a sample may land on one only through the
    `sum - 0.2` clamp (rela/prioritized_replay.h:281-296)."""
    schema = ff_schema(frame=(1, 4, 4), n_actions=2)
    ops = gen_trace(schema, 200, 41, 60, 20, 32, zero_frac=zero_frac)
    _diff(schema, 200, 3, 0.6, 0.4, 0, ops)


def test_ff_wide_range_priorities():
    """Priorities over 15 decades: the double accumulation of float weights really rounds."""
    schema = ff_schema(frame=(1, 4, 4), n_actions=2)
    prio = lambda rng, n: (10.0 ** rng.uniform(-12, 3, n)).astype(np.float32)  # noqa: E731
    ops = gen_trace(schema, 300, 43, 80, 25, 48, prio=prio)
    _diff(schema, 300, 17, 1.0, 0.5, 0, ops)
    _diff(schema, 300, 17, 0.6, 0.4, 0, ops)


def test_ff_batch_larger_than_buffer():
    schema = ff_schema(frame=(1, 4, 4), n_actions=2)
    ops = gen_trace(schema, 16, 45, 50, 4, 128, prefill=6)
    _diff(schema, 16, 1, 0.6, 0.4, 0, ops)


def test_ff_many_wraps():
    """A 40-slot ring written ~2500 times: every slot is reused and evicted dozens of times."""
    schema = ff_schema(frame=(1, 4, 4), n_actions=2)
    ops = gen_trace(schema, 32, 47, 400, 6, 16, adds_per_cycle=(1, 1))
    _diff(schema, 32, 2, 0.6, 0.4, 0, ops)


@pytest.mark.parametrize("seed", range(3))
def test_rnn_random_traces(seed):
    schema = rnn_schema(T=5 + seed, frame=(1, 4, 4), n_actions=3, hid=8)
    cap = 48 + 20 * seed
    ops = gen_trace(schema, cap, 500 + seed, 40, 6, 12)
    _diff(schema, cap, 4 + seed, 0.9, 0.6, 0, ops, ff=False)


@pytest.mark.parametrize("n", [1, 2, 3, 5])
@pytest.mark.parametrize("p_term", [0.0, 0.05, 0.5, 1.0])
def test_nstep_random(n, p_term):
    """MultiStepTransitionBuffer::popTransition (rela/dqn_actor.h:58-106), bit-exact with the
    reference build's FMA contraction."""
    rng = np.random.default_rng(n * 100 + int(p_term * 10))
    K = 257
    reward = rng.standard_normal((n + 1, K)).astype(np.float32)
    terminal = (rng.random((n + 1, K)) < p_term).astype(np.uint8)
    a = [np.zeros(K, np.float32), np.zeros(K, np.float32), np.zeros(K, np.uint8)]
    b = [np.zeros(K, np.float32), np.zeros(K, np.float32), np.zeros(K, np.uint8)]
    ref_lib().ref_nstep_pop(n, K, np.float32(0.997), reward.ctypes.data, terminal.ctypes.data,
                            a[0].ctypes.data, a[1].ctypes.data, a[2].ctypes.data)
    oracle_lib().orc_nstep_return(n, K, 0.997, reward.ctypes.data, terminal.ctypes.data,
                                  b[0].ctypes.data, b[1].ctypes.data, b[2].ctypes.data, 1)
    for x, y in zip(a, b):
        np.testing.assert_array_equal(x.view(np.uint8), y.view(np.uint8))
