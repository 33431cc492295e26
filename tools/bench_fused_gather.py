"""gather + the learner's `s.float() / 255`, separate vs fused in the gather.
Usage: python tools/bench_fused_gather.py [capacity] [batch]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import rela_b200.rela as rela
from bench import fill_fast

dev = torch.device("cuda", 0)
cap = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
B = int(sys.argv[2]) if len(sys.argv) > 2 else 512
K = 2000
prio = torch.rand((16, B), device=dev) + 0.01


def run(mode):
    r = rela.FFPrioritizedReplay(cap, 7, 0.6, 0.4, 0, 0)
    fill_fast(rela, torch, r, dev, cap)
    if mode != "separate":
        dt = torch.float32 if mode == "fused_f32" else torch.bfloat16
        r.set_output_format("obs.s", dt)
        r.set_output_format("next_obs.s", dt)
    keep = None

    def cycle(k):
        nonlocal keep
        batch, w = r.sample(B, "cuda:0")
        if mode == "separate":
            keep = (batch.obs["s"].float() / 255, batch.next_obs["s"].float() / 255)
        else:
            keep = (batch.obs["s"], batch.next_obs["s"])
        r.update_priority(prio[k % 16])

    for k in range(100):
        cycle(k)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for k in range(K):
        cycle(k)
    e1.record()
    torch.cuda.synchronize()
    r.sync()
    return e0.elapsed_time(e1) / K * 1e3


out = {"capacity": cap, "batch": B, "what": "sample + both frame leaves as byte/255 + update_priority, us per cycle"}
for mode in ("separate", "fused_f32", "fused_bf16"):
    out[mode] = run(mode)
print(json.dumps(out))
