"""Pipelined sample / update cycle with and without $RR_GATHER_STREAM.

The loop is the prefetching learner of the C++ host layer, driven through the C ABI from a
FOREIGN stream with the new priorities already in HBM:

    rr_take_prefetched(k) -> rr_update_priority(k, device priorities) -> rr_prefetch_into(k + depth)

With one internal stream the kernels run select -> gather -> update in a chain (~31 us at the
headline configuration). With RR_GATHER_STREAM=1 the gathers queue on their own stream and a
cycle costs max(select + update, gather). The switch is read by rr_create, so each setting gets
its own replay (filled again). Usage (GPU box):

    python tools/bench_gather_stream.py [capacity=2^20] [batch=512] [depth=2] [steps=5000]

Prints one JSON line per setting. UNMEASURED at the time of writing (written after the round's
GPU budget was spent); parity of the switch: RELA_B200_EXPERIMENTAL=1 pytest tests/test_gpu_gather_stream.py
"""
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))


def run(two_stream, capacity, B, depth, steps):
    import torch
    import rela_b200.rela
    from bench import fill_device
    from rela_b200 import capi

    os.environ["RR_GATHER_STREAM"] = "1" if two_stream else "0"
    dev = torch.device("cuda", 0)
    L = capi.lib()
    replay = rela_b200.rela.FFPrioritizedReplay(capacity, 7, 0.6, 0.4, 0, 0)
    fill_device(replay, torch, dev, capacity, 64)
    h = C.c_void_p(replay.handle())
    schema = replay.schema()  # (name, dtype, shape, chunks, row_bytes, time_major)
    R = depth + 2
    outs = [[torch.empty((B,) + tuple(t[2]), dtype=t[1], device=dev) for t in schema] for _ in range(R)]
    ptrs = [(C.c_void_p * len(o))(*[t.data_ptr() for t in o]) for o in outs]
    ws = [torch.empty(B, dtype=torch.float32, device=dev) for _ in range(R)]
    prio = torch.rand((16, B), device=dev) + 0.01
    stream = torch.cuda.current_stream(dev)
    sp = C.c_void_p(stream.cuda_stream)
    issued = 0

    def issue():
        nonlocal issued
        capi.check(L.rr_prefetch_into(h, B, ptrs[issued % R], C.c_void_p(ws[issued % R].data_ptr()), sp))
        issued += 1

    def cycle(k):
        capi.check(L.rr_take_prefetched(h, sp))
        capi.check(L.rr_update_priority(h, B, C.c_void_p(prio[k % 16].data_ptr()), capi.RR_PRIO_DEVICE, sp))
        issue()

    for _ in range(depth):
        issue()
    for k in range(200):
        cycle(k)
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for k in range(steps):
        cycle(k)
    capi.check(L.rr_stream_fence(h, sp))
    e1.record(stream)
    torch.cuda.synchronize(dev)
    ms = e0.elapsed_time(e1) / steps
    # drain the queue so that the replay can be destroyed cleanly
    for k in range(depth):
        capi.check(L.rr_take_prefetched(h, sp))
        capi.check(L.rr_update_priority(h, B, C.c_void_p(prio[0].data_ptr()), capi.RR_PRIO_DEVICE, sp))
    replay.sync()
    print(json.dumps({"gather_stream": bool(two_stream), "capacity": capacity, "batch": B, "depth": depth,
                      "steps": steps, "us_per_cycle": ms * 1e3, "transitions_per_s": B / (ms * 1e-3)}), flush=True)
    del replay
    torch.cuda.empty_cache()


if __name__ == "__main__":
    a = [int(x) for x in sys.argv[1:]]
    capacity = a[0] if len(a) > 0 else 1 << 20
    B = a[1] if len(a) > 1 else 512
    depth = a[2] if len(a) > 2 else 2
    steps = a[3] if len(a) > 3 else 5000
    for ts in (0, 1):
        run(ts, capacity, B, depth, steps)
