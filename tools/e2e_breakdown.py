"""Host-side breakdown of the e2e cycle (sample -> weights.cpu() -> update_priority)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import rela_b200.rela as rela
from bench import fill_fast
dev = torch.device("cuda", 0)
cap = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 18
pf = int(sys.argv[2]) if len(sys.argv) > 2 else 0
r = rela.FFPrioritizedReplay(cap, 7, 0.6, 0.4, pf, 0)
fill_fast(rela, torch, r, dev, cap)
B = 512
prio = (torch.rand((16, B)) + 0.01).pin_memory()
acc = [0.0] * 4
N = 3000
for k in range(N + 100):
    t0 = time.perf_counter()
    batch, w = r.sample(B, "cuda:0")
    t1 = time.perf_counter()
    wh = w.cpu()
    t2 = time.perf_counter()
    r.update_priority(prio[k % 16])
    t3 = time.perf_counter()
    if k >= 100:
        acc[0] += t1 - t0; acc[1] += t2 - t1; acc[2] += t3 - t2
tot = sum(acc[:3])
print("sample() host %.1f us | weights.cpu() %.1f us | update_priority() host %.1f us | total %.1f us" %
      tuple(x / N * 1e6 for x in (acc[0], acc[1], acc[2], tot)))
