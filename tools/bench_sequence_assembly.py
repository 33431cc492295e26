"""R2D2 sequence assembly, reference vs device.
One actor thread's worth of work: K envs x n steps of 4x84x84 frames through
(a) the reference's R2D2TransitionBuffer -> torch::stack of T=123 steps per
sequence on the host -> add (H2D of the 3.5 MB sequence), and (b)
rela_b200::SequenceAssembler -> one H2D per step, windows in HBM, device-to-
device add. Both feed a B200 replay (tests/dropin/_build/seq_fixture.so)."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests", "dropin", "_build"))
import torch

import rela
import seq_fixture

K, n_steps, seq_len, burn_in, multi_step = 40, 400, 80, 40, 3
g = torch.Generator().manual_seed(1)
since = torch.zeros(K)
steps = []
for t in range(n_steps):
    term = torch.rand(K, generator=g) < 0.004
    steps.append(({"s": torch.randint(0, 256, (K, 4, 84, 84), dtype=torch.uint8, generator=g),
                   "eps": torch.rand((K, 1), generator=g), "legal_move": torch.ones(K, 18)},
                  {"a": torch.randint(0, 18, (K,), generator=g)}, torch.randn(K, generator=g), term,
                  1.0 - term.float(), torch.rand(K, generator=g) + 0.1,
                  {"h0": since.view(1, K, 1).expand(1, K, 512).clone(), "c0": since.view(1, K, 1).expand(1, K, 512).clone()}))
    since = torch.where(term, torch.zeros(K), since + 1)
out = {}
for name, fn in (("reference_buffer", seq_fixture.feed_reference), ("device_assembler", seq_fixture.feed_assembler)):
    rep = rela.RNNPrioritizedReplay(2000, 1, 0.9, 0.6, 0)
    fn(rep, K, multi_step, seq_len, burn_in, 0.9, steps[:20])  # warm-up: schema, allocations
    rep = rela.RNNPrioritizedReplay(2000, 1, 0.9, 0.6, 0)
    torch.cuda.synchronize()
    t0 = time.time()
    n = fn(rep, K, multi_step, seq_len, burn_in, 0.9, steps)
    torch.cuda.synchronize()
    dt = time.time() - t0
    out[name] = {"seconds": dt, "env_steps_per_s": K * n_steps / dt, "sequences": n}
out["speedup"] = out["reference_buffer"]["seconds"] / out["device_assembler"]["seconds"]
out["workload"] = "K=%d envs x %d steps, T=%d, 4x84x84 frames, LSTM 512, one producer thread" % (
    K, n_steps, burn_in + seq_len + multi_step)
print(json.dumps(out))
