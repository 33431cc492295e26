"""Development probe (GPU box): GPU-side timeline of the selection / update / gather kernels in the
pipelined cycle (rr_debug_timeline stamps), to see what waits for what.

Usage: python tools/timeline_probe.py [capacity=2^21] [depth=2] [cycles=12]
"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import rela_b200
from rela_b200 import capi
from cycle_probe import fill

NAMES = {1: "S+", 2: "S-", 3: "U+", 4: "U-", 5: "G+", 6: "G-", 7: "Uall-", 8: "Gall-", 9: "Ux", 10: "S0", 11: "U0",
         12: "G0", 13: "Gsig", 14: "Gdone"}


def main():
    a = [int(x) for x in sys.argv[1:]]
    capacity = a[0] if len(a) > 0 else 1 << 21
    depth = a[1] if len(a) > 1 else 2
    cycles = a[2] if len(a) > 2 else 12
    B = 512
    dev = torch.device("cuda", 0)
    L = capi.lib()
    replay = rela_b200.FFPrioritizedReplay(capacity, 7, 0.6, 0.4, 0, 0)
    fill(replay, dev, capacity)
    h = C.c_void_p(replay.handle())
    schema = replay.schema()
    R = 8
    outs = [[torch.empty((B,) + tuple(t[2]), dtype=t[1], device=dev) for t in schema] for _ in range(R)]
    ws = [torch.empty(B, dtype=torch.float32, device=dev) for _ in range(R)]
    prio = torch.rand((16, B), device=dev) + 0.01
    nf = len(schema)
    flat = (C.c_void_p * (R * nf))(*[t.data_ptr() for o in outs for t in o])
    wsp = (C.c_void_p * R)(*[t.data_ptr() for t in ws])
    psp = (C.c_void_p * 16)(*[prio[i].data_ptr() for i in range(16)])
    ms, nl = C.c_double(), C.c_uint64()
    for d in (0, depth):
        capi.check(L.rr_cycle_bench(h, B, d, 300, 50, R, flat, wsp, 16, psp, capi.RR_PRIO_DEVICE, C.byref(ms), C.byref(nl)))
        print("depth %d: %.2f us per cycle" % (d, ms.value * 1e3))
        buf = np.zeros(1024, dtype=np.uint64)
        L.rr_debug_timeline(h, 1, buf.ctypes.data, 0)   # switch on
        capi.check(L.rr_cycle_bench(h, B, d, cycles, 0, R, flat, wsp, 16, psp, capi.RR_PRIO_DEVICE, C.byref(ms), C.byref(nl)))
        n = L.rr_debug_timeline(h, 0, buf.ctypes.data, 1024)
        ev = sorted((int(v) >> 8, int(v) & 15, (int(v) >> 4) & 15) for v in buf[:n])
        t0 = ev[0][0]
        print("  " + "  ".join("%s:%x@%.1f" % (NAMES.get(c, "?"), g, (t - t0) / 1e3) for t, c, g in ev))


if __name__ == "__main__":
    main()
