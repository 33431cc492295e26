"""Development probe (GPU box): the sample / update cycle at the headline configuration, by mode.

  serial     rr_sample + rr_update_priority(device priorities) on the replay's own stream
  pipelined  rr_take_prefetched + rr_update_priority + rr_prefetch_into, depth D: the gathers of
             batches drawn ahead run on the gather streams beside the control plane
  gather     rr_gather_bench on 1 stream (back to back) and on the 2 gather streams
  select     clock64 phase stamps of select_kernel (CTA 0)

Usage: python tools/cycle_probe.py [capacity=2^21] [batch=512] [steps=3000]
"""
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import rela_b200
from rela_b200 import capi

FRAME = (4, 84, 84)


def fill(replay, dev, capacity, block=8192):
    g = torch.Generator(device=dev)
    g.manual_seed(1)
    done = 0
    while done < capacity + 64:
        n = min(block, capacity + 64 - done)
        obs = {"s": torch.randint(0, 256, (n,) + FRAME, dtype=torch.uint8, device=dev, generator=g),
               "eps": torch.rand((n, 1), device=dev, generator=g), "legal_move": torch.ones((n, 18), device=dev)}
        nxt = {"s": torch.randint(0, 256, (n,) + FRAME, dtype=torch.uint8, device=dev, generator=g),
               "eps": torch.rand((n, 1), device=dev, generator=g), "legal_move": torch.ones((n, 18), device=dev)}
        term = torch.rand(n, device=dev, generator=g) < 0.005
        tr = rela_b200.FFTransition(obs, {"a": torch.arange(done, done + n, dtype=torch.int64, device=dev)},
                                    torch.randn(n, device=dev, generator=g), term, 1.0 - term.float(), nxt)
        replay.add(tr, torch.rand(n, device=dev, generator=g) + 0.01)
        done += n
    replay.sync()


def main():
    a = [int(x) for x in sys.argv[1:]]
    capacity = a[0] if len(a) > 0 else 1 << 21
    B = a[1] if len(a) > 1 else 512
    steps = a[2] if len(a) > 2 else 3000
    dev = torch.device("cuda", 0)
    L = capi.lib()
    replay = rela_b200.FFPrioritizedReplay(capacity, 7, 0.6, 0.4, 0, 0)
    fill(replay, dev, capacity)
    h = C.c_void_p(replay.handle())
    schema = replay.schema()
    own = C.c_void_p()
    capi.check(L.rr_get_stream(h, C.byref(own)))
    own_stream = torch.cuda.ExternalStream(own.value, device=dev)
    R = 8
    outs = [[torch.empty((B,) + tuple(t[2]), dtype=t[1], device=dev) for t in schema] for _ in range(R)]
    ptrs = [(C.c_void_p * len(o))(*[t.data_ptr() for t in o]) for o in outs]
    ws = [torch.empty(B, dtype=torch.float32, device=dev) for _ in range(R)]
    prio = torch.rand((16, B), device=dev) + 0.01
    res = {"capacity": capacity, "batch": B, "steps": steps}

    def timed(fn, n):
        torch.cuda.synchronize(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(own_stream)
        for k in range(n):
            fn(k)
        capi.check(L.rr_stream_fence(h, own))
        e1.record(own_stream)
        torch.cuda.synchronize(dev)
        return e0.elapsed_time(e1) / n * 1e3

    def serial(k):
        capi.check(L.rr_sample(h, B, ptrs[k % R], C.c_void_p(ws[k % R].data_ptr()), own))
        capi.check(L.rr_update_priority(h, B, C.c_void_p(prio[k % 16].data_ptr()), capi.RR_PRIO_DEVICE, own))

    for k in range(50):
        serial(k)
    res["serial_us"] = timed(serial, steps)

    # select phases (last call)
    pr = np.zeros(16, dtype=np.uint64)
    acc = np.zeros(4)
    for k in range(20):
        serial(k)
        capi.check(L.rr_debug_select_probe(h, pr.ctypes.data))
        acc += np.diff(pr[:5].astype(np.int64))
    res["select_phases_us"] = dict(zip(["loads+min+targets", "group scan", "offset+group search", "leaf+slot search+ids"],
                                       [round(float(v) / 20 / 1.9e3, 2) for v in acc]))

    for depth in (1, 2, 3):
        issued = [0]

        def issue():
            k = issued[0]
            capi.check(L.rr_prefetch_into(h, B, ptrs[k % R], C.c_void_p(ws[k % R].data_ptr()), own))
            issued[0] += 1

        def cycle(k):
            capi.check(L.rr_take_prefetched(h, own))
            capi.check(L.rr_update_priority(h, B, C.c_void_p(prio[k % 16].data_ptr()), capi.RR_PRIO_DEVICE, own))
            issue()

        for _ in range(depth):
            issue()
        for k in range(50):
            cycle(k)
        res["pipelined_depth%d_us" % depth] = timed(cycle, steps)
        for k in range(depth):
            capi.check(L.rr_take_prefetched(h, own))
            capi.check(L.rr_update_priority(h, B, C.c_void_p(prio[0].data_ptr()), capi.RR_PRIO_DEVICE, own))

    # the same loops from inside the library (no ctypes per call)
    nfl = len(schema)
    flat_all = (C.c_void_p * (R * nfl))(*[t.data_ptr() for o in outs for t in o])
    wsp = (C.c_void_p * R)(*[t.data_ptr() for t in ws])
    psp = (C.c_void_p * 16)(*[prio[i].data_ptr() for i in range(16)])
    for depth in (0, 1, 2, 3):
        ms, nl = C.c_double(), C.c_uint64()
        capi.check(L.rr_cycle_bench(h, B, depth, steps, 50, R, flat_all, wsp, 16, psp, capi.RR_PRIO_DEVICE,
                                    C.byref(ms), C.byref(nl)))
        res["c_loop_depth%d_us" % depth] = ms.value * 1e3
        if depth == 2:  # phases of the last selection that ran beside a gather
            capi.check(L.rr_debug_select_probe(h, pr.ctypes.data))
            d = np.diff(pr[:5].astype(np.int64))
            res["select_phases_pipelined_us"] = [round(float(v) / 1.9e3, 2) for v in d]

    # gather kernel alone
    capi.check(L.rr_set_profiling(h, 1, B))
    for k in range(16):
        serial(k)
    capi.check(L.rr_set_profiling(h, 0, B))
    nf = len(schema)
    flat = (C.c_void_p * (R * nf))(*[t.data_ptr() for o in outs for t in o])
    for streams in (1, 2):
        ms, n, b = C.c_double(), C.c_uint64(), C.c_uint64()
        capi.check(L.rr_gather_bench(h, R, flat, max(1, steps // 16), streams, C.byref(ms), C.byref(n), C.byref(b)))
        res["gather_%dstream_us" % streams] = ms.value * 1e3
        res["gather_%dstream_gbs" % streams] = b.value / (ms.value * 1e-3) / 1e9
    capi.check(L.rr_sync(h))
    st = capi.RRStats()
    capi.check(L.rr_get_stats(h, C.byref(st)))
    res["samples_ordered"] = st.samples_ordered
    res["device_errors"] = st.device_errors
    print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
