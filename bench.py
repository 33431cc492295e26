#!/usr/bin/env python
"""bench.py -- replay sample+gather throughput (BASELINE.json metric).

A "step" is one pass of the hot path over one batch: sample(B) (prefix
hierarchy + stratified search + IS weights + gather of B transitions into the
learner batch) followed by update_priority(B). Workload (N=1): BASELINE.json
configs[1], Ape-X 3-step replay, capacity 2^21 transitions of 56,617 B, batch
512, one B200.

  value     whole-job transitions/s with priorities already in HBM, C ABI calls
  e2e       same cycle through the public Python API (rela_b200.FFPrioritizedReplay,
            the mirror of rela.FFPrioritizedReplay) with the new priorities coming
            from pinned HOST memory every step and the IS weights read back D2H
  roofline  gather kernel: 2 * 56,617 B * 512 per launch / CUDA-event time of the
            kernel (events on the replay's own stream), vs MEASURED_PEAKS.json
  cpu_baseline  the reference's own C++ replay (oracle/_ref) timed on this box's host

`--impl reference` times only the reference's CPU implementation.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

FRAME = (4, 84, 84)
N_ACT = 18
ROW_BYTES = 56617  # bytes per Ape-X transition


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20000)
    ap.add_argument("--warmup", type=int, default=100)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--capacity", type=int, default=1 << 21)
    ap.add_argument("--batch", type=int, default=512)
    ap.add_argument("--alpha", type=float, default=0.6)
    ap.add_argument("--beta", type=float, default=0.4)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-sharded", action="store_true")
    ap.add_argument("--cpu-cycles", type=int, default=300)
    return ap.parse_args()


# ------------------------------------------------------------------ clocks
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.lines = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            parts = [p.strip() for p in ln.split(",")]
            if len(parts) < 8:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
            except ValueError:
                continue
            for nm, val in zip(names, parts[4:8]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------- reference CPU arm
def reference_fill(capacity, B, alpha, beta, prefetch):
    """The unmodified reference (oracle/_ref) on this host, filled to capacity
    with K=80 adds aliasing a pool of 64 pre-generated batches (the reference
    stores tensor views of the actor's batch, rela/types.cc:48-67, so this is
    what it holds in training too)."""
    import numpy as np
    from harness import RefReplay, ff_schema, make_rows

    schema = ff_schema(FRAME, N_ACT)
    rng = np.random.default_rng(1)
    K = 80
    pool = [make_rows(schema, K, rng, 0) for _ in range(64)]
    ref = RefReplay(schema, capacity, 7, alpha, beta, prefetch)
    t0 = time.time()
    for i in range(capacity // K + 1):
        ref.add(pool[i % 64], rng.random(K).astype(np.float32) + 0.01, alias=True)
    ref.fill_s = time.time() - t0
    ref.prio = rng.random((8, B)).astype(np.float32) + 0.01
    ref.prefetch = prefetch
    return ref


def reference_best(args, cycles, warm, prefetches=(0, 2), probe=15):
    """Probe prefetch in {0, 2} x OMP threads in {1 (README.md:125-130), all}
    for `probe` cycles each, then time `cycles` sample(B,"cpu") +
    updatePriority cycles on the fastest combination."""
    import torch

    all_threads = torch.get_num_threads()
    refs = [reference_fill(args.capacity, args.batch, args.alpha, args.beta, pf) for pf in prefetches]
    runs = []
    for ref in refs:
        for nthreads in sorted({1, all_threads}):
            torch.set_num_threads(nthreads)
            ref.cycle_loop(args.batch, 3, ref.prio)
            dt = ref.cycle_loop(args.batch, probe, ref.prio)
            runs.append({"prefetch": ref.prefetch, "omp_threads": nthreads,
                         "transitions_per_s": args.batch * probe / dt, "ms_per_step": dt / probe * 1e3,
                         "threads_used": (ref.prefetch + 1) if nthreads == 1 else max(nthreads, ref.prefetch + 1),
                         "fill_s": ref.fill_s, "ref": ref})
    best = max(runs, key=lambda r: r["transitions_per_s"])
    torch.set_num_threads(best["omp_threads"])
    best["ref"].cycle_loop(args.batch, warm, best["ref"].prio)
    dt = best["ref"].cycle_loop(args.batch, cycles, best["ref"].prio)
    torch.set_num_threads(all_threads)
    best = dict(best, transitions_per_s=args.batch * cycles / dt, ms_per_step=dt / cycles * 1e3, seconds=dt)
    for r in runs:
        r.pop("ref")
    best.pop("ref")
    for ref in refs:
        ref.close()
    return best, runs


def cpu_baseline_obj(best, runs, cycles):
    from harness import ref_lib

    return {"value": best["transitions_per_s"], "unit": "transitions/s", "cores": best["threads_used"],
            "kind": "reference",
            "sample": "%d sample+update_priority cycles at full capacity (filled by K=80 adds aliasing 64 "
                      "batches) with prefetch=%d, %d OMP thread(s): the fastest of the probed "
                      "prefetch x threads combinations %s; %s"
                      % (cycles, best["prefetch"], best["omp_threads"],
                         [(r["prefetch"], r["omp_threads"]) for r in runs],
                         ref_lib().ref_build_info().decode()),
            "host_threads_available": os.cpu_count(),
            "runs": [{k: r[k] for k in ("prefetch", "omp_threads", "transitions_per_s", "ms_per_step")}
                     for r in runs]}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from harness import have_ref

    # a reference cycle costs 7-12 ms: bound the run to a few minutes whatever --steps says
    cycles = max(min(args.steps, 3000), 1)
    warm = max(min(args.warmup, 20), 1)
    if not have_ref():
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libref_replay.so not built"}))
        return
    best, runs = reference_best(args, cycles, warm)
    tps = best["transitions_per_s"]
    line = {
        "impl": "reference", "metric": "replay sample+gather transitions/s", "value": tps,
        "unit": "transitions/s", "n_gpus": args.gpus, "steps": cycles, "warmup": warm,
        "ms_per_step": best["ms_per_step"], "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": workload_config(args, "host-cpu"),
        "cpu_baseline": cpu_baseline_obj(best, runs, cycles),
        "e2e": {"value": tps, "unit": "transitions/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


def workload_config(args, where):
    return {"workload": "Ape-X 3-step replay, capacity %d transitions x %d B (4x84x84 u8 obs+next_obs), "
                        "batch %d, alpha %.1f beta %.1f, sample+update_priority cycle"
                        % (args.capacity, ROW_BYTES, args.batch, args.alpha, args.beta),
            "capacity": args.capacity, "batch": args.batch, "row_bytes": ROW_BYTES,
            "l2_policy": "inputs larger than L2 (random rows of a %.0f GB slab; 8 rotating output batches)"
                         % (args.capacity * 1.25 * ROW_BYTES / 1e9),
            "parallelism": where}


# ------------------------------------------------------------- our arm
def fill_device(replay, torch, dev, capacity, extra, block=8192, seed=1):
    """Write real bytes to every slot: fresh random frames generated on the
    device, added through add() in blocks (device-resident rows -> K1 insert)."""
    from rela_b200.rela import FFTransition  # the C++ host layer's record type

    g = torch.Generator(device=dev)
    g.manual_seed(seed)
    total = capacity + extra
    done = 0
    pairs = []
    while done < total:
        n = min(block, total - done)
        obs = {"s": torch.randint(0, 256, (n,) + FRAME, dtype=torch.uint8, device=dev, generator=g),
               "eps": torch.rand((n, 1), device=dev, generator=g),
               "legal_move": torch.ones((n, N_ACT), device=dev)}
        nxt = {"s": torch.randint(0, 256, (n,) + FRAME, dtype=torch.uint8, device=dev, generator=g),
               "eps": torch.rand((n, 1), device=dev, generator=g),
               "legal_move": torch.ones((n, N_ACT), device=dev)}
        term = torch.rand(n, device=dev, generator=g) < 0.005
        tr = FFTransition(obs, {"a": torch.arange(done, done + n, dtype=torch.int64, device=dev)},
                          torch.randn(n, device=dev, generator=g), term, 1.0 - term.float(), nxt)
        prio = torch.rand(n, device=dev, generator=g) + 0.01
        timed = n == block and len(pairs) < 64 and done >= 4 * block
        if timed:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        replay.add(tr, prio)
        if timed:
            e1.record()
            pairs.append((e0, e1))
        done += n
    replay.sync()
    # K1 insert: device-resident rows -> slab rows, 2 x row bytes per transition
    # (read the producer's row, write the slot), CUDA events around add()
    if pairs:
        ms = sum(a.elapsed_time(b) for a, b in pairs) / len(pairs)
        return {"kernel": "row_copy_kernel (insert) + commit_kernel", "rows_per_add": block,
                "us_per_add": ms * 1e3, "adds_timed": len(pairs),
                "achieved_gbs": 2 * ROW_BYTES * block / (ms * 1e-3) / 1e9,
                "transitions_per_s": block / (ms * 1e-3)}
    return None


def run_sharded(args, torch, dist, dev, rank, world, local_rank):
    """BASELINE.json configs[4]: ONE replay of world * capacity transitions whose
    payload is partitioned over the GPUs (rela_b200/sharded.py); every batch is
    assembled on the learner GPU (rank 0) by bulk stores over NVLink. Returns the
    object reported under "sharded" (informational; `value` stays the
    independent-shards number)."""
    from rela_b200.sharded import FF_LEAVES, ShardedPrioritizedReplay

    B, K, W = args.batch, min(args.steps, 5000), max(args.warmup, 3)
    total_cap = args.capacity * world
    rep = ShardedPrioritizedReplay(total_cap, 7, args.alpha, args.beta, FF_LEAVES, device=dev, batch_sets=2)
    g = torch.Generator(device=dev)
    g.manual_seed(100 + rank)
    t0 = time.time()
    block = 2048  # rows per rank per collective add
    rounds = (total_cap + 64 + block * world - 1) // (block * world)
    for it in range(rounds):
        n = block
        term = torch.rand(n, device=dev, generator=g) < 0.005
        leaves = [torch.randint(0, 256, (n,) + FRAME, dtype=torch.uint8, device=dev, generator=g),
                  torch.rand((n, 1), device=dev, generator=g), torch.ones((n, N_ACT), device=dev),
                  torch.arange(n, dtype=torch.int64, device=dev) + (it * world + rank) * n,
                  torch.randn(n, device=dev, generator=g), term, 1.0 - term.float(),
                  torch.randint(0, 256, (n,) + FRAME, dtype=torch.uint8, device=dev, generator=g),
                  torch.rand((n, 1), device=dev, generator=g), torch.ones((n, N_ACT), device=dev)]
        if rep._size + n * world > rep.S:
            break
        rep.add(leaves, torch.rand(n, device=dev, generator=g) + 0.01)
    torch.cuda.synchronize(dev)
    fill_s = time.time() - t0
    prio = torch.rand((16, B), device=dev, generator=g) + 0.01

    def cycle(k):
        tens, w = rep.sample(B, 0)
        rep.update_priority(prio[k % 16] if rank == 0 else None)

    for k in range(W):
        cycle(k)
    dist.barrier()
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for k in range(K):
        cycle(k)
    e1.record()
    dist.barrier()
    torch.cuda.synchronize(dev)
    t = torch.tensor([e0.elapsed_time(e1)], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    rep.engine.sync()
    size = rep.size()
    completion = rep.engine.fused_completion
    rep.engine.close()
    tps = B * K / (ms * 1e-3)
    return {"workload": "one replay of %d transitions striped over %d GPUs (%.1f GB of slabs per GPU), batch %d "
                        "assembled on GPU 0 over NVLink, sample+update_priority" %
                        (total_cap, world, int(1.25 * total_cap) / world * ROW_BYTES / 1e9, B),
            "value": tps, "unit": "transitions/s", "ms_per_step": ms / K, "steps": K, "size": size,
            "learner_nvlink_ingress_gbs": tps * ROW_BYTES * (world - 1) / world / 1e9,
            "nvlink_peak_gbs_per_direction": 900.0, "fill_s": fill_s,
            "completion": "flag stores from the gather kernels into learner memory + one-warp acquire kernel"
                          if completion else "4-byte NCCL all-reduce",
            "collectives_per_step": ("none: rows, completion flags and the %d-byte priority push are one-sided stores "
                                     "over NVLink peer memory" % (4 * B)) if completion else
                                    ("1 x 4-byte all-reduce + 1 x %d-byte broadcast (NCCL)" % (4 * B))}



def run_ours(args):
    import numpy as np
    import torch
    import rela_b200
    import rela_b200.rela
    from rela_b200 import capi

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; rela_b200 has no CPU path")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        # one slice of the host cores per rank: the cycle is a chain of short kernels, so a
        # descheduled launch thread shows up directly in the max-over-ranks time
        try:
            cpus = sorted(os.sched_getaffinity(0))
            per = max(1, len(cpus) // world)
            os.sched_setaffinity(0, set(cpus[local_rank * per:(local_rank + 1) * per]) or set(cpus))
        except (AttributeError, OSError):
            pass
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=dev)

    B, K, W = args.batch, args.steps, max(args.warmup, 3)
    L = capi.lib()
    sharded = None
    if world > 1 and not args.no_sharded:
        sharded = run_sharded(args, torch, dist, dev, rank, world, local_rank)
        torch.cuda.synchronize(dev)
        torch.cuda.empty_cache()
    # One replay, built and filled through the C++ host layer (rela_b200/_C.so,
    # rela's class names); the device-resident leg below drives the same object
    # through the C ABI by its handle.
    replay = rela_b200.rela.FFPrioritizedReplay(args.capacity, 7 + rank, args.alpha, args.beta, 0, local_rank)
    t0 = time.time()
    insert = fill_device(replay, torch, dev, args.capacity, 64, seed=1 + rank)
    fill_s = time.time() - t0
    assert replay.size() == args.capacity + 64

    class Leaf:
        def __init__(self, name, dtype, shape, chunks, row_bytes, time_major=False):
            self.name, self.dtype, self.shape, self.chunks = name, dtype, tuple(shape), chunks

        def batch_shape(self, n):
            return (n,) + self.shape if self.chunks == 1 else (self.shape[0], n) + self.shape[1:]

    schema = [Leaf(*t) for t in replay.schema()]
    assert sum(t[4] for t in replay.schema()) == ROW_BYTES
    stream = torch.cuda.current_stream(dev)
    sp = C.c_void_p(stream.cuda_stream)
    h = C.c_void_p(replay.handle())

    def get_stats():
        st = capi.RRStats()
        capi.check(L.rr_get_stats(h, C.byref(st)))
        return {n: getattr(st, n) for n, _ in st._fields_}

    # 8 rotating output batches (232 MB > L2) and device-resident priorities
    R = 8
    outs = [[torch.empty(lf.batch_shape(B), dtype=lf.dtype, device=dev) for lf in schema] for _ in range(R)]
    out_ptrs = [(C.c_void_p * len(o))(*[t.data_ptr() for t in o]) for o in outs]
    out_w = [torch.empty(B, dtype=torch.float32, device=dev) for _ in range(R)]
    g = torch.Generator(device=dev)
    g.manual_seed(99 + rank)
    prio_dev = torch.rand((16, B), device=dev, generator=g) + 0.01
    prio_host = (torch.rand((16, B)) + 0.01).pin_memory()

    # The device-resident loop is a learner whose own kernels run on the replay's
    # stream (rr_get_stream): plain stream order, no event hand-shake, and the
    # select -> gather -> update kernels chain by programmatic dependent launch.
    # The same loop from a foreign stream (two event hand-shakes per call) is
    # reported beside it as value_foreign_stream.
    own = C.c_void_p()
    capi.check(L.rr_get_stream(h, C.byref(own)))
    own_stream = torch.cuda.ExternalStream(own.value, device=dev)

    def make_cycle(stream_ptr):
        def cycle(k):
            capi.check(L.rr_sample(h, B, out_ptrs[k % R], C.c_void_p(out_w[k % R].data_ptr()), stream_ptr))
            capi.check(L.rr_update_priority(h, B, C.c_void_p(prio_dev[k % 16].data_ptr()), capi.RR_PRIO_DEVICE,
                                            stream_ptr))
        return cycle

    cycle_dev = make_cycle(own)
    cycle_foreign = make_cycle(sp)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def timed(fn, steps, on=None):
        on = stream if on is None else on
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(on)
        for k in range(steps):
            fn(k)
        capi.check(L.rr_stream_fence(h, C.c_void_p(on.cuda_stream)))
        e1.record(on)
        barrier()
        ms = e0.elapsed_time(e1)
        if dist is not None:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    # ---- value: device-resident inputs, C ABI
    for k in range(W):
        cycle_dev(k)
    st0 = get_stats()
    clocks = ClockSampler(local_rank)
    if rank == 0:  # the line is rank 0's; one nvidia-smi poller per box is enough
        clocks.start()
    ms = timed(cycle_dev, K, own_stream)
    st1 = get_stats()
    launches = st1["kernel_launches"] - st0["kernel_launches"]
    ms_foreign = timed(cycle_foreign, max(1, K // 4))

    # ---- roofline: record the id sets of 16 real samples, then re-launch the
    # gather kernel for them back to back (K launches, 8 rotating output
    # batches) inside one CUDA-event pair on the replay's own stream
    capi.check(L.rr_set_profiling(h, 1, B))
    for k in range(16):
        cycle_dev(k)
    capi.check(L.rr_set_profiling(h, 0, B))
    nf = len(schema)
    flat = (C.c_void_p * (R * nf))(*[t.data_ptr() for o in outs for t in o])
    ms_pl, n_l, bpl = C.c_double(), C.c_uint64(), C.c_uint64()
    barrier()
    capi.check(L.rr_gather_bench(h, R, flat, max(1, K // 16), C.byref(ms_pl), C.byref(n_l), C.byref(bpl)))
    gather_us = ms_pl.value * 1e3

    # the same kernel on a batch 8x larger: how much of the gap to the HBM peak at
    # batch 512 is the fixed per-launch latency (id load, first-byte latency, store drain)
    big = None
    if world == 1 and B <= 1024:
        B2 = 8 * B
        outs2 = [[torch.empty(lf.batch_shape(B2), dtype=lf.dtype, device=dev) for lf in schema] for _ in range(2)]
        ptrs2 = [(C.c_void_p * len(o))(*[t.data_ptr() for t in o]) for o in outs2]
        w2 = torch.empty(B2, dtype=torch.float32, device=dev)
        prio2 = torch.rand(B2, device=dev, generator=g) + 0.01
        torch.cuda.synchronize(dev)
        capi.check(L.rr_set_profiling(h, 1, B2))
        for k in range(4):
            capi.check(L.rr_sample(h, B2, ptrs2[k % 2], C.c_void_p(w2.data_ptr()), own))
            capi.check(L.rr_update_priority(h, B2, C.c_void_p(prio2.data_ptr()), capi.RR_PRIO_DEVICE, own))
        capi.check(L.rr_set_profiling(h, 0, B2))
        flat2 = (C.c_void_p * (2 * nf))(*[t.data_ptr() for o in outs2 for t in o])
        ms2, n2, b2 = C.c_double(), C.c_uint64(), C.c_uint64()
        barrier()
        capi.check(L.rr_gather_bench(h, 2, flat2, max(1, K // 128), C.byref(ms2), C.byref(n2), C.byref(b2)))
        big = {"batch": B2, "kernel_us": ms2.value * 1e3, "launches_timed": int(n2.value),
               "bytes_per_launch": 2 * ROW_BYTES * B2}
        del outs2, ptrs2, flat2

    # ---- e2e: public API, priorities from pinned host memory, weights read back
    dev_str = "cuda:%d" % local_rank

    w_pinned = [torch.empty(B, dtype=torch.float32).pin_memory() for _ in range(2)]

    def cycle_e2e(k):
        batch, w = replay.sample(B, dev_str)
        # D2H read of the step's result into pinned memory (also the learner's sync point)
        w_host = w_pinned[k & 1]
        w_host.copy_(w, non_blocking=True)
        stream.synchronize()
        replay.update_priority(prio_host[k % 16])
        return w_host

    for k in range(W):
        cycle_e2e(k)
    ms_e2e = timed(cycle_e2e, K)
    replay.sync()
    st2 = get_stats()
    assert st2["samples_ordered"] == 0, "benchmark weights must take the exact parallel selection"
    assert st2["device_errors"] == 0

    # same loop with prefetch = 2, the configuration the reference arm is fastest in
    # (rela/prioritized_replay.h:215-230: batches drawn ahead of the pending updatePriority);
    # a second replay, since prefetch is a constructor argument
    e2e_runs = [{"prefetch": 0, "ms_per_step": ms_e2e / K}]
    import gc

    replay = None
    gc.collect()
    torch.cuda.empty_cache()
    replay = rela_b200.rela.FFPrioritizedReplay(args.capacity, 7 + rank, args.alpha, args.beta, 2, local_rank)
    fill_device(replay, torch, dev, args.capacity, 64, seed=1 + rank)
    h = C.c_void_p(replay.handle())
    for k in range(W):
        cycle_e2e(k)
    ms_e2e_pf = timed(cycle_e2e, K)
    replay.sync()
    st3 = get_stats()
    assert st3["samples_ordered"] == 0 and st3["device_errors"] == 0
    e2e_runs.append({"prefetch": 2, "ms_per_step": ms_e2e_pf / K})
    e2e_prefetch = 0
    if ms_e2e_pf < ms_e2e:
        ms_e2e, e2e_prefetch = ms_e2e_pf, 2
    clk = clocks.stop()

    peaks = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except OSError:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
    gather_bytes = 2 * ROW_BYTES * B
    traffic = None  # dram__bytes_read.sum + dram__bytes_write.sum per launch, from the committed ncu --set full capture
    try:
        with open(os.path.join(ROOT, "profiles", "r1c_gather_traffic.json")) as f:
            tj = json.load(f)
        if B == 512:
            traffic = tj["dram_bytes_read_per_launch"] + tj["dram_bytes_write_per_launch"]
    except (OSError, KeyError, ValueError):
        pass
    achieved = gather_bytes / (gather_us * 1e-6) / 1e9 if gather_us > 0 else 0.0

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return
    value = world * B * K / (ms * 1e-3)
    e2e = world * B * K / (ms_e2e * 1e-3)
    line = {
        "metric": "replay sample+gather transitions/s", "value": value, "unit": "transitions/s",
        "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms / K, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": workload_config(args, "1 GPU" if world == 1 else "%d independent capacity shards" % world),
        "e2e": {"value": e2e, "unit": "transitions/s", "h2d_bytes_per_step": 4 * B, "d2h_bytes_per_step": 4 * B,
                "ms_per_step": ms_e2e / K, "prefetch": e2e_prefetch,
                "runs": [dict(r, transitions_per_s=world * B / (r["ms_per_step"] * 1e-3)) for r in e2e_runs],
                "api": "rela_b200.rela.FFPrioritizedReplay (C++ host layer, rela/pybind.cc names): sample(B,'cuda:0') + weights -> pinned host copy + stream sync + update_priority(pinned cpu tensor)"},
        "gpu_launches": int(launches),
        "value_foreign_stream": {"value": world * B * max(1, K // 4) / (ms_foreign * 1e-3), "unit": "transitions/s",
                                 "ms_per_step": ms_foreign / max(1, K // 4),
                                 "note": "same loop, caller on its own stream: event hand-shake in and out of "
                                         "every rr_sample / rr_update_priority"},
        "clocks": clk,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic,
                     "traffic_note": "ncu capture profiles/r1c_gather_ncu_full_details.txt: DRAM reads 29.57 MB = the "
                                     "algorithmic read half; the 29 MB of stores land in L2 and drain after the kernel",
                     "kernel": "row_copy_kernel (gather)",
                     "kernel_us": gather_us, "launches_timed": int(n_l.value),
                     "method": "K gather launches over 16 recorded id sets, back to back, one CUDA-event pair",
                     "bytes_per_launch": gather_bytes, "peak_source": peak_src,
                     "step_share": gather_us * 1e-3 / (ms / K)},
        "fill_s": fill_s, "slab_gb": st2["slab_bytes"] / 1e9,
    }
    if big is not None:
        big["achieved"] = big["bytes_per_launch"] / (big["kernel_us"] * 1e-6) / 1e9
        big["frac"] = big["achieved"] / peak
        line["roofline"]["large_batch"] = big
    if insert is not None:
        insert["frac_of_hbm_peak"] = insert["achieved_gbs"] / peak
        line["insert"] = insert
    if sharded is not None:
        line["sharded"] = sharded
    if world == 1 and not args.no_cpu_baseline:
        from harness import have_ref

        if have_ref():
            best, runs = reference_best(args, args.cpu_cycles, 5)
            line["cpu_baseline"] = cpu_baseline_obj(best, runs, args.cpu_cycles)
        else:
            line["cpu_baseline"] = {"value": None, "unit": "transitions/s", "cores": 0, "kind": "reference",
                                    "sample": "oracle/_ref/libref_replay.so missing"}
    print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
