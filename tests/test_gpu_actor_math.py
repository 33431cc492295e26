"""GPU: actor-side transition arithmetic (K8 n-step return, K9 TD priority and
eta-mix) through the C ABI vs golden vectors recorded from the reference
(MultiStepTransitionBuffer::popTransition, rela/dqn_actor.h:58-106; the torch
expressions of pyrela/apex.py:44-45 and pyrela/r2d2.py:113-120) and vs the C
oracle at BASELINE config 4 size (80 threads x 80 envs)."""
import os

import numpy as np
import pytest
import torch

from harness import GOLDEN_DIR, oracle_lib

pytestmark = pytest.mark.gpu


def test_nstep_golden_bit_exact():
    import rela_b200

    g = np.load(os.path.join(GOLDEN_DIR, "nstep.npz"))
    R, b, t = rela_b200.nstep_return(torch.from_numpy(g["reward"]).cuda(),
                                     torch.from_numpy(g["terminal"]).cuda(), float(g["gamma"]))
    np.testing.assert_array_equal(R.cpu().numpy(), g["out_reward"])  # same FMA contraction as the reference build
    np.testing.assert_array_equal(b.cpu().numpy(), g["out_bootstrap"])
    np.testing.assert_array_equal(t.cpu().numpy().astype(np.uint8), g["out_terminal"])


def test_actor_math_golden():
    import rela_b200

    g = np.load(os.path.join(GOLDEN_DIR, "actor_math.npz"))
    agg = rela_b200.aggregate_priority(torch.from_numpy(g["priority"]).cuda(), torch.from_numpy(g["seq_len"]).cuda(),
                                       float(g["eta"]), int(g["burn_in"]))
    np.testing.assert_allclose(agg.cpu().numpy(), g["agg"], rtol=1e-6, atol=0)
    td = rela_b200.td_priority(*[torch.from_numpy(np.ascontiguousarray(g[k])).cuda()
                                 for k in ("reward", "bootstrap", "q_next", "q_sa")], float(np.float32(g["gamma_n"])))
    np.testing.assert_allclose(td.cpu().numpy(), g["td"], rtol=1e-6, atol=1e-7)


@pytest.mark.parametrize("n,K", [(3, 6400), (1, 80), (5, 1000), (3, 1)])
def test_nstep_vs_oracle(n, K):
    import rela_b200

    rng = np.random.default_rng(n * 1000 + K)
    r = rng.standard_normal((n + 1, K)).astype(np.float32)
    t = (rng.random((n + 1, K)) < 0.1).astype(np.uint8)
    o_r, o_b, o_t = np.zeros(K, np.float32), np.zeros(K, np.float32), np.zeros(K, np.uint8)
    for fma in (1, 0):
        oracle_lib().orc_nstep_return(n, K, 0.997, r.ctypes.data, t.ctypes.data, o_r.ctypes.data, o_b.ctypes.data,
                                      o_t.ctypes.data, fma)
        R, b, tt = rela_b200.nstep_return(torch.from_numpy(r).cuda(), torch.from_numpy(t).cuda(), 0.997,
                                          contract_fma=bool(fma))
        np.testing.assert_array_equal(R.cpu().numpy(), o_r)
        np.testing.assert_array_equal(b.cpu().numpy(), o_b)
        np.testing.assert_array_equal(tt.cpu().numpy().astype(np.uint8), o_t)


def test_aggregate_vs_oracle_config4():
    import rela_b200

    rng = np.random.default_rng(5)
    nseq, L = 6400, 80
    pr = rng.random((nseq, L)).astype(np.float32)
    sl = rng.integers(41, 121, size=nseq).astype(np.float32)
    # padding beyond seq_len is already zero in the reference (r2d2_actor.h:82)
    pr[np.arange(L)[None, :] >= sl[:, None]] = 0
    out = np.zeros(nseq, np.float32)
    oracle_lib().orc_aggregate_priority(nseq, L, pr.ctypes.data, sl.ctypes.data, 0.9, 40, out.ctypes.data)
    got = rela_b200.aggregate_priority(torch.from_numpy(pr).cuda(), torch.from_numpy(sl).cuda(), 0.9, 40)
    np.testing.assert_allclose(got.cpu().numpy(), out, rtol=1e-6, atol=0)
