"""Generate tests/golden/*.npz by running the UNMODIFIED reference
(oracle/_ref/libref_replay.so, built from /root/reference by oracle/Makefile)
on scripted traces. Run in the build container only:

    python tests/golden/make_golden.py

Each file stores the trace definition (so the inputs can be regenerated
deterministically from the seed) and everything the reference produced:
sampled slot ids, IS weights, sha256 of every gathered field, size() after each
sample, and the final ring state / weights / evicted flags.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from harness import (GOLDEN_CASES as CASES, RefReplay, apply_trace, digest, gen_trace, ref_lib,  # noqa: E402
                     schema_of)

def main():
    info = ref_lib().ref_build_info().decode()
    for name, (kind, kw, cap, seed, alpha, beta, pf, tseed, cycles, K, B) in CASES.items():
        schema = schema_of(kind, kw)
        ops = gen_trace(schema, cap, tseed, cycles, K, B)
        ref = RefReplay(schema, cap, seed, alpha, beta, pf)
        rec = apply_trace(ref, ops, alpha, True)
        out = dict(
            ids=np.stack([r["ids"] for r in rec]),
            w=np.stack([r["w"] for r in rec]),
            size=np.array([r["size"] for r in rec], dtype=np.int64),
            rows_sha=np.stack([np.stack([digest(x) for x in r["rows"]]) for r in rec]),
            num_add=np.int64(ref.num_add()),
            build=np.bytes_(info.encode()),
        )
        if kind == "ff":
            st = ref.state()
            wts, ev = ref.weights()
            out.update(head=st["head"], tail=st["tail"], rsize=st["size"], safe_size=st["safe_size"],
                       sum=np.float64(st["sum"]), weights=wts, evicted=ev)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
        print(name, "samples", len(rec), "bytes", os.path.getsize(os.path.join(HERE, name + ".npz")))

    # n-step return golden: reference MultiStepTransitionBuffer::popTransition
    rng = np.random.default_rng(21)
    n, K = 3, 640
    reward = rng.standard_normal((n + 1, K)).astype(np.float32)
    terminal = (rng.random((n + 1, K)) < 0.08).astype(np.uint8)
    o_r, o_b, o_t = np.zeros(K, np.float32), np.zeros(K, np.float32), np.zeros(K, np.uint8)
    ref_lib().ref_nstep_pop(n, K, np.float32(0.997), reward.ctypes.data, terminal.ctypes.data,
                            o_r.ctypes.data, o_b.ctypes.data, o_t.ctypes.data)
    np.savez_compressed(os.path.join(HERE, "nstep.npz"), reward=reward, terminal=terminal, gamma=np.float32(0.997),
                        out_reward=o_r, out_bootstrap=o_b, out_terminal=o_t)

    # eta-mix aggregate + TD priority: the reference's Python expressions
    # (pyrela/r2d2.py:113-120, pyrela/apex.py:44-45) evaluated with eager torch
    import torch

    pr = torch.rand(64, 80, generator=torch.Generator().manual_seed(5))
    sl = torch.randint(41, 121, (64,), generator=torch.Generator().manual_seed(6)).float()
    eta, burn_in = 0.9, 40
    mask = torch.arange(0, pr.size(1))
    mask = (mask.unsqueeze(0) < sl.unsqueeze(1)).float()
    p = pr * mask
    p_mean = p.sum(1) / (sl - burn_in)
    p_max = p.max(1)[0]
    agg = eta * p_max + (1.0 - eta) * p_mean
    g = torch.Generator().manual_seed(7)
    reward_t, q_next, q_sa = torch.randn(640, generator=g), torch.randn(640, generator=g), torch.randn(640, generator=g)
    boot = (torch.rand(640, generator=g) > 0.1).float()
    gamma_n = 0.997 ** 3
    target = reward_t + boot * gamma_n * q_next
    td = (target - q_sa).abs()
    np.savez_compressed(os.path.join(HERE, "actor_math.npz"), priority=pr.numpy(), seq_len=sl.numpy(),
                        eta=np.float32(eta), burn_in=np.int64(burn_in), agg=agg.numpy(),
                        reward=reward_t.numpy(), bootstrap=boot.numpy(), q_next=q_next.numpy(), q_sa=q_sa.numpy(),
                        gamma_n=np.float64(gamma_n), td=td.numpy())
    print("actor math golden written")


if __name__ == "__main__":
    main()
