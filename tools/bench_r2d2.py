"""BASELINE.json configs[2] (R2D2 sequence replay) on one B200: seq_len 80,
burn-in 40, 3-step => T = 123 steps per sequence, LSTM h/c 512, batch 64,
alpha 0.9 / beta 0.6 (pyrela/scripts/ref_run_r2d2.sh:15-16). One sequence is
3,487,091 B, so 100 k sequences are 349 GB (436 GB with the reference's 25 %
slack slots) and do NOT fit one GPU: this run holds --capacity sequences (default
36,000 = 157 GB of slabs) and says so. Prints sample+update cycles/s,
sequences/s and the gather's algorithmic GB/s (2 x 3,487,091 B per sequence)."""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import rela_b200.rela as rela

ap = argparse.ArgumentParser()
ap.add_argument("--capacity", type=int, default=36000)
ap.add_argument("--batch", type=int, default=64)
ap.add_argument("--steps", type=int, default=300)
args = ap.parse_args()
dev = torch.device("cuda", 0)
T, HID, NA = 123, 512, 18
SEQ_BYTES = 3487091
r = rela.RNNPrioritizedReplay(args.capacity, 7, 0.9, 0.6, 0, 0)
g = torch.Generator(device=dev)
g.manual_seed(1)
t0 = time.time()
n_add = 0
blk = 16
while n_add < args.capacity + 8:
    els = []
    for _ in range(blk):
        els.append(rela.RNNTransition(
            {"s": torch.randint(0, 256, (T, 4, 84, 84), dtype=torch.uint8, device=dev, generator=g),
             "eps": torch.rand((T, 1), device=dev, generator=g), "legal_move": torch.ones((T, NA), device=dev)},
            {"h0": torch.randn((1, HID), device=dev, generator=g), "c0": torch.randn((1, HID), device=dev, generator=g)},
            {"a": torch.randint(0, NA, (T,), device=dev, generator=g)},
            torch.randn(T, device=dev, generator=g), torch.zeros(T, dtype=torch.bool, device=dev),
            torch.ones(T, device=dev), torch.randint(41, 121, (1,), device=dev, generator=g).float()))
    r.add(els, torch.rand(blk, device=dev, generator=g) + 0.01)
    n_add += blk
r.sync()
fill_s = time.time() - t0
B = args.batch
prio = torch.rand((16, B), device=dev, generator=g) + 0.01
for k in range(5):
    b, w = r.sample(B, "cuda:0")
    r.update_priority(prio[k % 16])
assert b.obs["s"].shape == (T, B, 4, 84, 84) and b.h0["h0"].shape == (1, B, HID) and b.seq_len.shape == (B,)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for k in range(args.steps):
    b, w = r.sample(B, "cuda:0")
    r.update_priority(prio[k % 16])
e1.record()
torch.cuda.synchronize()
r.sync()
ms = e0.elapsed_time(e1) / args.steps
print(json.dumps({"workload": "R2D2 sequence replay T=123, batch %d, %d sequences held (%.0f GB of slabs; 100k = 349 GB "
                              "does not fit one GPU)" % (B, args.capacity, int(1.25 * args.capacity) * SEQ_BYTES / 1e9),
                  "ms_per_cycle": ms, "sequences_per_s": B / (ms * 1e-3), "env_steps_per_s": B * 80 / (ms * 1e-3),
                  "cycle_algorithmic_gbs": 2 * SEQ_BYTES * B / (ms * 1e-3) / 1e9, "fill_s": fill_s,
                  "size": r.size()}))
