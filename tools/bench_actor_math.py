"""BASELINE.json configs[3] / SURVEY 8(d) config 4: actor-side n-step return + priority arithmetic
for 80 actor threads x K = 80 envs (6,400 envs per call), n = 3, gamma 0.997; eta-mix over
[6400, 80] per-step priorities, eta 0.9, burn-in 40.

  reference leg (CPU, runs anywhere oracle/_ref is built): MultiStepTransitionBuffer::popTransition
      (rela/dqn_actor.h:58-106) for K = 80 envs on ONE core -- what each of the reference's actor
      threads does per step -- and the eta-mix / TD expressions of pyrela (r2d2.py:113-120,
      apex.py:44-45) evaluated with eager torch on the host.
  device leg (needs a B200): rr_nstep_return / rr_td_priority / rr_aggregate_priority on all 6,400
      envs in one launch each, inputs resident in HBM, CUDA events around `iters` back-to-back calls.

Prints one JSON line per leg. `python tools/bench_actor_math.py [--cpu-only] [--iters N]`
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

THREADS, K, N_STEP, GAMMA, L, ETA, BURN = 80, 80, 3, 0.997, 80, 0.9, 40


def cpu_leg(iters):
    import numpy as np
    import torch
    from harness import have_ref, ref_lib

    torch.set_num_threads(1)
    rng = np.random.default_rng(0)
    out = {"leg": "reference-cpu", "cores": 1, "envs_per_call": K}
    if have_ref():
        reward = rng.standard_normal((N_STEP + 1, K)).astype(np.float32)
        terminal = (rng.random((N_STEP + 1, K)) < 0.01).astype(np.uint8)
        ref_lib().ref_nstep_loop(N_STEP, K, np.float32(GAMMA), reward.ctypes.data, terminal.ctypes.data, 50)
        s = ref_lib().ref_nstep_loop(N_STEP, K, np.float32(GAMMA), reward.ctypes.data, terminal.ctypes.data, iters)
        out["nstep_pop_us_per_call"] = s / iters * 1e6
        out["nstep_env_steps_per_s_per_core"] = K * iters / s
        out["nstep_env_steps_per_s_80_threads_ideal"] = THREADS * K * iters / s
    else:
        out["nstep"] = "oracle/_ref not built"
    # eta-mix (r2d2.py:113-120) and TD priority (apex.py:44-45), eager torch, one actor's K envs
    pr, sl = torch.rand(K, L), torch.randint(41, 121, (K,)).float()
    r, qn, qs, bs = torch.randn(K), torch.randn(K), torch.randn(K), (torch.rand(K) > 0.01).float()

    def agg():
        mask = (torch.arange(0, pr.size(1)).unsqueeze(0) < sl.unsqueeze(1)).float()
        p = pr * mask
        return ETA * p.max(1)[0] + (1.0 - ETA) * (p.sum(1) / (sl - BURN))

    def td():
        return (r + bs * (GAMMA ** N_STEP) * qn - qs).abs()

    for name, fn in (("aggregate", agg), ("td_priority", td)):
        for _ in range(20):
            fn()
        t0 = time.perf_counter()
        for _ in range(iters):
            fn()
        s = time.perf_counter() - t0
        out[name + "_us_per_call"] = s / iters * 1e6
        out[name + "_envs_per_s_per_core"] = K * iters / s
    print(json.dumps(out), flush=True)


def device_leg(iters):
    import torch
    import rela_b200

    dev = torch.device("cuda", 0)
    n_env = THREADS * K
    g = torch.Generator(device=dev).manual_seed(0)
    reward = torch.randn((N_STEP + 1, n_env), device=dev, generator=g)
    terminal = torch.rand((N_STEP + 1, n_env), device=dev, generator=g) < 0.01
    pr = torch.rand((n_env, L), device=dev, generator=g)
    sl = torch.randint(41, 121, (n_env,), device=dev, generator=g).float()
    r, qn, qs = (torch.randn(n_env, device=dev, generator=g) for _ in range(3))
    bs = (torch.rand(n_env, device=dev, generator=g) > 0.01).float()
    legs = {
        "nstep": lambda: rela_b200.nstep_return(reward, terminal, GAMMA),
        "aggregate": lambda: rela_b200.aggregate_priority(pr, sl, ETA, BURN),
        "td_priority": lambda: rela_b200.td_priority(r, bs, qn, qs, GAMMA ** N_STEP),
    }
    out = {"leg": "b200", "envs_per_call": n_env}
    for name, fn in legs.items():
        for _ in range(20):
            fn()
        torch.cuda.synchronize(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize(dev)
        ms = e0.elapsed_time(e1) / iters
        out[name + "_us_per_call"] = ms * 1e3
        out[name + "_envs_per_s"] = n_env / (ms * 1e-3)
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--cpu-only", action="store_true")
    ap.add_argument("--iters", type=int, default=2000)
    a = ap.parse_args()
    cpu_leg(a.iters)
    if not a.cpu_only:
        import torch

        if not torch.cuda.is_available():
            raise SystemExit("bench_actor_math: no CUDA device (use --cpu-only for the reference leg)")
        device_leg(a.iters)
