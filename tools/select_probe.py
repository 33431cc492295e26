"""Development probe: phase timing of select_kernel at the bench configuration."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import rela_b200
from rela_b200 import capi
from bench import fill_device

cap = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 21
B = 512
dev = torch.device("cuda", 0)
# small rows: the selection does not depend on the row size
r = rela_b200.FFPrioritizedReplay(cap, 7, 0.6, 0.4, 0, device=dev)
n = 0
g = torch.Generator(device=dev); g.manual_seed(1)
while n < cap + 64:
    k = min(65536, cap + 64 - n)
    tr = rela_b200.FFTransition({"s": torch.zeros((k, 16), dtype=torch.uint8, device=dev)}, {"a": torch.arange(n, n + k, device=dev)},
                                torch.zeros(k, device=dev), torch.zeros(k, dtype=torch.bool, device=dev), torch.ones(k, device=dev), {})
    r.add(tr, torch.rand(k, device=dev, generator=g) + 0.01)
    n += k
L = capi.lib()
acc = np.zeros(4)
tail = 0.0
N = 50
r2 = rela_b200.replay  # noqa: F841
for it in range(N + 5):
    b, w = r.sample(B, dev)
    r.update_priority(torch.rand(B) + 0.01)
    pr = np.zeros(16, dtype=np.uint64)
    capi.check(L.rr_debug_select_probe(r._h, pr.ctypes.data))
    if it >= 5:
        acc += np.diff(pr[:5].astype(np.int64))
        tail += float(int(pr[7]) - int(pr[6]))
names = ["loads + min + targets (mt19937)", "tile-prefix scan", "offset + tile search", "in-tile search + ids"]
for nm, v in zip(names, acc / N):
    print(f"{nm:34s} {v:9.0f} cyc  {v/1.9e3:6.2f} us")
print(f"{'tail of the last CTA (evict, IS)':34s} {tail / N:9.0f} cyc  {tail / N / 1.9e3:6.2f} us")
print("CTA 0 span", acc.sum() / N / 1.9e3, "us + tail")
