"""CPU: the C oracle (oracle/replay_oracle.c) against the golden vectors that
tests/golden/make_golden.py recorded from the UNMODIFIED reference. This is
what pins the oracle (the reference ships no KATs)."""
import numpy as np
import pytest

from harness import GOLDEN_CASES, GOLDEN_DIR, Oracle, check_against_golden, oracle_lib

import os


@pytest.mark.parametrize("name", sorted(GOLDEN_CASES))
def test_oracle_matches_reference_golden(name):
    check_against_golden(name, lambda *a: Oracle(*a))


def test_nstep_golden():
    g = np.load(os.path.join(GOLDEN_DIR, "nstep.npz"))
    reward, terminal = g["reward"], g["terminal"]
    n, K = reward.shape[0] - 1, reward.shape[1]
    o_r, o_b, o_t = np.zeros(K, np.float32), np.zeros(K, np.float32), np.zeros(K, np.uint8)
    oracle_lib().orc_nstep_return(n, K, float(g["gamma"]), reward.ctypes.data, terminal.ctypes.data,
                                  o_r.ctypes.data, o_b.ctypes.data, o_t.ctypes.data, 1)
    np.testing.assert_array_equal(o_r, g["out_reward"])  # bit-exact with the FMA build
    np.testing.assert_array_equal(o_b, g["out_bootstrap"])
    np.testing.assert_array_equal(o_t, g["out_terminal"])


def test_actor_math_golden():
    g = np.load(os.path.join(GOLDEN_DIR, "actor_math.npz"))
    pr, sl = np.ascontiguousarray(g["priority"]), np.ascontiguousarray(g["seq_len"])
    out = np.zeros(pr.shape[0], np.float32)
    oracle_lib().orc_aggregate_priority(pr.shape[0], pr.shape[1], pr.ctypes.data, sl.ctypes.data,
                                        float(g["eta"]), int(g["burn_in"]), out.ctypes.data)
    np.testing.assert_allclose(out, g["agg"], rtol=1e-6, atol=0)
    K = g["reward"].shape[0]
    td = np.zeros(K, np.float32)
    arrs = [np.ascontiguousarray(g[k]) for k in ("reward", "bootstrap", "q_next", "q_sa")]
    oracle_lib().orc_td_priority(K, *[a.ctypes.data for a in arrs], float(np.float32(g["gamma_n"])), td.ctypes.data)
    np.testing.assert_allclose(td, g["td"], rtol=1e-6, atol=1e-7)
