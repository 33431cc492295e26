#!/usr/bin/env python
"""bench.py -- replay sample+gather throughput (BASELINE.json metric).

A "step" is one pass of the hot path over one batch: sample(B) (stratified
selection + importance weights + gather of B transitions into the learner
batch) followed by update_priority(B).

  --gpus 1   BASELINE.json configs[1]: Ape-X 3-step replay, capacity 2^21
             transitions of 56,617 B (148 GB of slabs), batch 512, one B200.
  --gpus N   BASELINE.json configs[4]: ONE replay of N * 2^21 transitions striped
             over the N GPUs (one process per GPU, torchrun), every GPU drawing
             its 512 rows of one stratified batch of N * 512 per step (rows
             travel GPU to GPU over NVLink; rela_b200/cpp/sharded_core.cc).

Before anything is timed the run checks itself ("parity"): the replay and the
reference (oracle/_ref, the unmodified rela sources; the C oracle when that is
absent or the buffer is striped) are fed the same adds and priorities and run
the same sample / update cycles with the same seed; sampled ids, gathered bytes,
importance weights and the final weights / sum_ are compared. A failed check
aborts the run without a number.

  value     whole-job transitions/s with the new priorities already in HBM
            (device-resident inputs), fastest of prefetch 0 / prefetch 2
  e2e       the same cycle through the public Python API (the reference's class
            names) with the new priorities coming from pinned HOST memory every
            step and the importance weights read back D2H every step
  roofline  gather kernel: 2 * 56,617 B * rows per launch / CUDA-event time of
            the kernel on the stream(s) it is launched on, vs MEASURED_PEAKS.json
  cpu_baseline  the reference's own C++ replay (oracle/_ref) timed on this host
  config1 / r2d2 / actor_math   BASELINE.json configs[0], [2], [3], both arms

`--impl reference` times only the reference's CPU implementation.
"""
import argparse
import ctypes as C
import gc
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
K_ADD = 80         # rows per add of the reference's actors


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=50)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--capacity", type=int, default=1 << 21, help="transitions per GPU")
    ap.add_argument("--batch", type=int, default=512, help="rows per learner GPU per step")
    ap.add_argument("--alpha", type=float, default=0.6)
    ap.add_argument("--beta", type=float, default=0.4)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the config1 / r2d2 / actor_math legs")
    ap.add_argument("--no-parity", action="store_true", help="development only: the line then carries parity: null")
    ap.add_argument("--parity-cycles", type=int, default=32)
    ap.add_argument("--cpu-cycles", type=int, default=200)
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


class PhaseWatchdog:
    """A striped run is N processes whose kernels wait for one another across GPUs; if one phase of it
    ever hangs, the run must still end within minutes and report what it has. Every rank runs one of
    these: `enter(name, budget_s)` arms a deadline for the phase that starts, `done()` disarms it; a
    daemon thread calls `on_timeout(name)` when a phase overruns (the callback prints what has been
    measured so far on rank 0 and ends the process; it does not return). Every call the main thread
    can hang in releases the GIL (ctypes, pybind call guards, torch synchronisation), so the thread
    gets to run."""

    def __init__(self, on_timeout, poll_s=0.25, log=None):
        self.on_timeout = on_timeout
        self.poll_s = poll_s
        self.log = log  # called with one line per phase entered (a trail on stderr, should all else fail)
        self.lock = threading.Lock()
        self.phase, self.deadline = None, None
        self.history = []
        self.thread = threading.Thread(target=self._run, daemon=True)
        self.thread.start()

    def enter(self, name, budget_s):
        with self.lock:
            now = time.monotonic()
            if self.phase is not None:
                self.history.append((self.phase, round(now - self.t0, 2)))
            self.phase, self.deadline, self.t0 = name, now + budget_s, now
        if self.log is not None:
            self.log("bench.py phase: %s (budget %d s)" % (name, budget_s))

    def done(self):
        with self.lock:
            if self.phase is not None:
                self.history.append((self.phase, round(time.monotonic() - self.t0, 2)))
            self.phase, self.deadline = None, None

    def _run(self):
        while True:
            time.sleep(self.poll_s)
            with self.lock:
                late = self.deadline is not None and time.monotonic() > self.deadline
                name = self.phase
                if late:
                    self.deadline = None
            if late:
                self.on_timeout(name)


def host_info():
    model = "?"
    try:
        for ln in open("/proc/cpuinfo"):
            if ln.startswith("model name"):
                model = ln.split(":", 1)[1].strip()
                break
    except OSError:
        pass
    return {"nproc": os.cpu_count(), "cpu_model": model}


def workload_config(total_capacity, global_batch, alpha, beta):
    """The same dict in both arms (the driver compares them): what is stored and what a step draws."""
    return {"workload": "Ape-X 3-step replay, %d transitions x %d B (4x84x84 u8 obs + next_obs), batch %d per step, "
                        "alpha %.1f beta %.1f, sample + update_priority cycle"
                        % (total_capacity, ROW_BYTES, global_batch, alpha, beta),
            "capacity": total_capacity, "batch": global_batch, "row_bytes": ROW_BYTES,
            "l2_policy": "inputs larger than L2 (random rows of %.0f GB of slabs; 8 rotating output batches)"
                         % (int(total_capacity * 1.25) * ROW_BYTES / 1e9)}


# ------------------------------------------- the add stream both arms are fed
def priority_stream(n_adds, seed=1):
    """Priorities of the fill, one row of K_ADD per add: rand + 0.01."""
    import numpy as np

    return (np.random.default_rng(seed).random((n_adds, K_ADD), dtype=np.float32) + np.float32(0.01))


def n_fill_adds(capacity):
    return capacity // K_ADD + 1  # 26,215 adds at 2^21: 48 rows above capacity, the first sample evicts


# ------------------------------------------------------- reference CPU arm
def reference_fill(capacity, B, alpha, beta, prefetch, prios):
    """The unmodified reference (oracle/_ref) on this host, filled with K=80 adds
    aliasing a pool of 64 pre-generated batches (the reference stores tensor views
    of the actor's batch, rela/types.cc:48-67, so this is what it holds in training
    too), with the priorities of `prios`."""
    import numpy as np
    from harness import RefReplay, ff_schema, make_rows

    schema = ff_schema(FRAME, N_ACT)
    rng = np.random.default_rng(1)
    pool = [make_rows(schema, K_ADD, rng, 0) for _ in range(64)]
    ref = RefReplay(schema, capacity, 7, alpha, beta, prefetch)
    t0 = time.time()
    for i in range(prios.shape[0]):
        ref.add(pool[i % 64], prios[i], alias=True)
    ref.fill_s = time.time() - t0
    ref.prio = rng.random((8, B)).astype(np.float32) + 0.01
    ref.prefetch = prefetch
    return ref


def reference_runs(refs, B, cycles, warm, probe=10, devices=("cpu",)):
    """Probe prefetch x OMP threads {1 (README.md:125-130), all} x sample() device for `probe`
    cycles each, then time `cycles` sample(B, device) + updatePriority cycles on the fastest
    "cpu" combination (the line's value) and, if asked, on the fastest "cuda:0" one."""
    import torch

    all_threads = torch.get_num_threads()
    runs = []
    for ref in refs:
        for dev in devices:
            for nthreads in sorted({1, all_threads}):
                torch.set_num_threads(nthreads)
                ref.cycle_loop(B, 2, ref.prio, dev)
                dt = ref.cycle_loop(B, probe, ref.prio, dev)
                runs.append({"prefetch": ref.prefetch, "omp_threads": nthreads, "device": dev,
                             "transitions_per_s": B * probe / dt, "ms_per_step": dt / probe * 1e3,
                             "threads_used": (ref.prefetch + 1) if nthreads == 1 else max(nthreads, ref.prefetch + 1),
                             "ref": ref})
    best = {}
    for dev in devices:
        cand = max((r for r in runs if r["device"] == dev), key=lambda r: r["transitions_per_s"])
        torch.set_num_threads(cand["omp_threads"])
        cand["ref"].cycle_loop(B, warm, cand["ref"].prio, dev)
        dt = cand["ref"].cycle_loop(B, cycles, cand["ref"].prio, dev)
        best[dev] = dict(cand, transitions_per_s=B * cycles / dt, ms_per_step=dt / cycles * 1e3, seconds=dt)
        best[dev].pop("ref")
    torch.set_num_threads(all_threads)
    for r in runs:
        r.pop("ref")
    return best, runs


def cpu_baseline_obj(best, runs, cycles, fill_note):
    from harness import ref_lib

    b = best["cpu"]
    out = {"value": b["transitions_per_s"], "unit": "transitions/s", "cores": b["threads_used"], "kind": "reference",
           "sample": "%d sample + update_priority cycles at full capacity (%s) with prefetch=%d, %d OMP thread(s), "
                     "sample(B, 'cpu'): the fastest of the probed prefetch x threads combinations; %s"
                     % (cycles, fill_note, b["prefetch"], b["omp_threads"], ref_lib().ref_build_info().decode()),
           "host": host_info(),
           "runs": [{k: r[k] for k in ("prefetch", "omp_threads", "device", "transitions_per_s", "ms_per_step")}
                    for r in runs]}
    if "cuda:0" in best:
        c = best["cuda:0"]
        out["device_cuda0"] = {"value": c["transitions_per_s"], "ms_per_step": c["ms_per_step"],
                               "prefetch": c["prefetch"], "omp_threads": c["omp_threads"],
                               "note": "sample(B, 'cuda:0') + weights.cpu(): the reference assembles the batch on the "
                                       "host and copies it to the learner GPU from pageable memory (types.cc:34-43)"}
    return out


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    from harness import have_ref

    world = args.gpus
    total_cap, Bg = args.capacity * world, args.batch * world
    # a reference cycle costs 6-12 ms at 2^21: bound the run to a few minutes whatever --steps says
    cycles = max(min(args.steps, 200 if world > 1 else 1000), 1)
    warm = max(min(args.warmup, 10), 1)
    if not have_ref():
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libref_replay.so not built"}))
        return
    prios = priority_stream(n_fill_adds(total_cap))
    refs = [reference_fill(total_cap, Bg, args.alpha, args.beta, pf, prios) for pf in ((0, 2) if world <= 2 else (2,))]
    devices = ("cpu", "cuda:0") if (torch.cuda.is_available() and world == 1) else ("cpu",)
    best, runs = reference_runs(refs, Bg, cycles, warm, devices=devices)
    for r in refs:
        r.close()
    tps = best["cpu"]["transitions_per_s"]
    line = {
        "impl": "reference", "metric": "replay sample+gather transitions/s", "value": tps,
        "unit": "transitions/s", "n_gpus": args.gpus, "steps": cycles, "warmup": warm,
        "ms_per_step": best["cpu"]["ms_per_step"], "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": workload_config(total_cap, Bg, args.alpha, args.beta),
        "parallelism": "host cpu (one reference replay holds all %d transitions)" % total_cap,
        "cpu_baseline": cpu_baseline_obj(best, runs, cycles, "filled by K=80 adds aliasing 64 batches"),
        "e2e": {"value": tps, "unit": "transitions/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


# ------------------------------------------------------------- our arm
def frames_of(torch, serial, salt):
    """uint8 [n, 4, 84, 84] as a pure function of the serial (int64 [n], device)."""
    col = torch.arange(28224, device=serial.device, dtype=torch.int64)[None, :]
    v = (serial[:, None] * 2654435761 + col * 40503 + salt * 97) >> 7
    return (v & 255).to(torch.uint8).reshape(-1, 4, 84, 84)


def block_of(torch, dev, s0, n):
    """Rows s0 .. s0 + n - 1 of the synthetic stream: every leaf is a function of the serial."""
    serial = torch.arange(s0, s0 + n, device=dev, dtype=torch.int64)
    eps = ((serial % 1000).float() / 1000)[:, None]
    lm = torch.ones((n, N_ACT), device=dev)
    term = (serial % 200) == 0
    return {"obs": {"s": frames_of(torch, serial, 1), "eps": eps, "legal_move": lm},
            "action": {"a": serial}, "reward": (serial % 97).float(), "terminal": term,
            "bootstrap": 1.0 - term.float(),
            "next_obs": {"s": frames_of(torch, serial, 2), "eps": eps.clone(), "legal_move": lm.clone()}}


def transition_of(rela, blk, lo=None, hi=None):
    sl = slice(lo, hi)
    return rela.FFTransition({k: v[sl] for k, v in blk["obs"].items()}, {"a": blk["action"]["a"][sl]},
                             blk["reward"][sl], blk["terminal"][sl], blk["bootstrap"][sl],
                             {k: v[sl] for k, v in blk["next_obs"].items()})


def check_rows(torch, batch, serial):
    """Gathered rows vs the rows that were inserted under these serials, byte for byte."""
    dev = batch.action["a"].device
    want = block_of(torch, dev, 0, 1)  # shapes only
    del want
    s = serial.to(dev)
    eps = ((s % 1000).float() / 1000)[:, None]
    term = (s % 200) == 0
    ok = bool(torch.equal(batch.action["a"], s))
    ok = ok and bool(torch.equal(batch.obs["s"], frames_of(torch, s, 1)))
    ok = ok and bool(torch.equal(batch.next_obs["s"], frames_of(torch, s, 2)))
    ok = ok and bool(torch.equal(batch.obs["eps"], eps)) and bool(torch.equal(batch.next_obs["eps"], eps))
    ok = ok and bool(torch.equal(batch.reward, (s % 97).float())) and bool(torch.equal(batch.terminal, term))
    ok = ok and bool(torch.equal(batch.bootstrap, 1.0 - term.float()))
    ok = ok and bool((batch.obs["legal_move"] == 1).all()) and bool((batch.next_obs["legal_move"] == 1).all())
    return ok


def fill_k80(rela, torch, replay, dev, prios, block=8000):
    """The reference actors' add pattern: K=80 rows per add (one float block sum per add goes
    into sum_, so the partition matters for parity), host priorities (raised with torch::pow on
    the CPU like the reference does). Rows are generated on the device in blocks of 100 adds
    and handed over as views. Returns serial_of_slot (numpy int64 [S])."""
    import numpy as np

    n_adds = prios.shape[0]
    S = int(1.25 * replay_capacity(replay))
    serial_of_slot = np.full(S, -1, dtype=np.int64)
    per = block // K_ADD
    tail = 0
    for a0 in range(0, n_adds, per):
        a1 = min(n_adds, a0 + per)
        blk = block_of(torch, dev, a0 * K_ADD, (a1 - a0) * K_ADD)
        for a in range(a0, a1):
            lo = (a - a0) * K_ADD
            replay.add(transition_of(rela, blk, lo, lo + K_ADD), torch.from_numpy(prios[a]))
        n = (a1 - a0) * K_ADD
        serial_of_slot[(tail + np.arange(n)) % S] = np.arange(a0 * K_ADD, a0 * K_ADD + n)
        tail = (tail + n) % S
    replay.sync()
    return serial_of_slot


_CAP = {}


def replay_capacity(replay):
    return _CAP[id(replay)]


def fill_fast(rela, torch, replay, dev, capacity, extra=64, block=8192, time_inserts=False):
    """Real bytes in every slot, added in blocks of 8192 device-resident rows (K1 insert)."""
    done, total = 0, capacity + extra
    g = torch.Generator(device=dev)
    g.manual_seed(3)
    pairs = []
    while done < total:
        n = min(block, total - done)
        blk = block_of(torch, dev, done, n)
        prio = torch.rand(n, device=dev, generator=g) + 0.01
        timed = time_inserts and n == block and len(pairs) < 64 and done >= 4 * block
        if timed:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        replay.add(transition_of(rela, blk), prio)
        if timed:
            e1.record()
            pairs.append((e0, e1))
        done += n
    replay.sync()
    if pairs:
        ms = sum(a.elapsed_time(b) for a, b in pairs) / len(pairs)
        return {"kernel": "row_copy_kernel (insert) + commit_kernel", "rows_per_add": block,
                "us_per_add": ms * 1e3, "adds_timed": len(pairs),
                "achieved_gbs": 2 * ROW_BYTES * block / (ms * 1e-3) / 1e9,
                "transitions_per_s": block / (ms * 1e-3)}
    return None


def last_ids(capi, L, h, n):
    import numpy as np

    ids = np.zeros(n, dtype=np.int32)
    capi.check(L.rr_last_sample_ids(h, n, ids.ctypes.data))
    return ids


def get_stats(capi, L, h):
    st = capi.RRStats()
    capi.check(L.rr_get_stats(h, C.byref(st)))
    return {n: getattr(st, n) for n, _ in st._fields_}


def pow_ulp_check(rela, torch, capi, L, dev, alpha, n=16384):
    """How far the kernel's own priority ^ alpha (fp64 pow rounded once: the path of DEVICE-resident
    priorities) is from torch.pow on the CPU (the function the reference calls; host priorities
    are raised with exactly that and are bit-identical by construction)."""
    import numpy as np

    r = rela.FFPrioritizedReplay(n, 3, alpha, 0.4, 0, dev.index)
    g = torch.Generator().manual_seed(11)
    p = torch.rand(n, generator=g) + 0.01
    tr = rela.FFTransition({}, {"a": torch.arange(n, device=dev)}, torch.zeros(n, device=dev),
                           torch.zeros(n, dtype=torch.bool, device=dev), torch.ones(n, device=dev), {})
    r.add(tr, p.to(dev))  # device priorities: raised by the kernel
    r.sync()
    w = np.zeros(int(1.25 * n), dtype=np.float32)
    capi.check(L.rr_debug_weights(C.c_void_p(r.handle()), w.ctypes.data, None))
    want = torch.pow(p, float(np.float32(alpha))).numpy()
    got = w[:n]
    diff = got != want
    rel = np.abs(got.astype(np.float64) - want) / want
    return {"n": n, "mismatches": int(diff.sum()), "max_rel": float(rel.max()),
            "note": "device-resident priorities only (the `value` loop's updates); host priorities use torch.pow"}


def parity_single(args, rela, torch, capi, L, replay, dev, prios, serial_of_slot, ref):
    """>= 32 sample / update cycles on the GPU replay and on the reference with one seed: ids,
    importance weights, gathered bytes; then the whole weight array, the eviction flags and sum_."""
    import numpy as np

    B = args.batch
    h = C.c_void_p(replay.handle())
    rng = np.random.default_rng(5)
    dev_str = "cuda:%d" % dev.index
    out = {"cycles": args.parity_cycles, "against": ref.kind_name, "ids_equal": True, "rows_equal": True,
           "weights_max_rel": 0.0, "batch": B}
    head = 0
    for c in range(args.parity_cycles):
        batch, w = replay.sample(B, dev_str)
        ids = last_ids(capi, L, h, B)
        rc, _, w_ref, ids_ref = ref.sample(B, False)
        assert rc == 0
        if not np.array_equal(ids, ids_ref):
            out["ids_equal"] = False
            out["first_mismatch"] = {"cycle": c, "index": int(np.argmax(ids != ids_ref))}
            break
        wg = w.cpu().numpy()
        out["weights_max_rel"] = max(out["weights_max_rel"], float(np.max(np.abs(wg - w_ref) / w_ref)))
        if not check_rows(torch, batch, torch.from_numpy(serial_of_slot[ids])):
            out["rows_equal"] = False
            break
        del batch
        p = rng.random(B).astype(np.float32) + np.float32(0.01)
        replay.update_priority(torch.from_numpy(p))
        ref.update(p)
    replay.sync()
    st = get_stats(capi, L, h)
    out["samples_ordered"] = int(st["samples_ordered"])
    out["device_errors"] = int(st["device_errors"])
    if out["ids_equal"] and out["rows_equal"]:
        S = int(1.25 * args.capacity)
        wg = np.zeros(S, dtype=np.float32)
        eg = np.zeros(S, dtype=np.uint8)
        capi.check(L.rr_debug_weights(h, wg.ctypes.data, eg.ctypes.data))
        wr, er = ref.weights()
        v = [C.c_int() for _ in range(4)]
        s = C.c_double()
        capi.check(L.rr_debug_state(h, *[C.addressof(x) for x in v], C.addressof(s)))
        rs = ref.state()
        live = np.zeros(S, dtype=bool)
        live[(rs["head"] + np.arange(rs["safe_size"])) % S] = True
        out["stored_weights_bitwise_mismatches"] = int((wg[live] != wr[live]).sum())
        out["evicted_flags_equal"] = bool(np.array_equal(eg[live], er[live]))
        out["sum_equal"] = bool(s.value == rs["sum"])
        out["ring_equal"] = (v[0].value, v[1].value, v[2].value, v[3].value) == (
            rs["head"], rs["tail"], rs["size"], rs["safe_size"])
        del head
    out["pow_ulp"] = pow_ulp_check(rela, torch, capi, L, dev, args.alpha)
    out["ok"] = bool(out["ids_equal"] and out["rows_equal"] and out["weights_max_rel"] <= 1e-6 and
                     out["samples_ordered"] == 0 and out["device_errors"] == 0 and
                     out.get("stored_weights_bitwise_mismatches", 1) == 0 and out.get("evicted_flags_equal") and
                     out.get("sum_equal") and out.get("ring_equal") and out["pow_ulp"]["max_rel"] <= 1e-6)
    return out


class OracleAsRef:
    """The C oracle behind the reference's test interface (used when oracle/_ref is absent)."""
    kind_name = "oracle/replay_oracle.c (C restatement; oracle/_ref absent on this box)"

    def __init__(self, capacity, seed, alpha, beta, prefetch, prios):
        from harness import Field, Oracle, torch_pow

        self.alpha, self.tp = alpha, torch_pow
        self.o = Oracle([Field("action.a", "int64", ())], capacity, seed, alpha, beta, prefetch)
        import numpy as np

        for i in range(prios.shape[0]):
            self.o.add([np.arange(i * K_ADD, (i + 1) * K_ADD, dtype=np.int64)], torch_pow(prios[i], alpha))

    def sample(self, B, want_rows):
        return self.o.sample(B, False)

    def update(self, p):
        return self.o.update(self.tp(p, self.alpha))

    def weights(self):
        return self.o.weights()

    def state(self):
        return self.o.state()

    def close(self):
        self.o.close()


def peaks():
    p = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
    except OSError:
        pass
    peak = float(p.get("hbm_gbs", 6650.0))
    return peak, ("measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in p else
                  "fallback 6650 GB/s (B200_PROFILING.md)")


def traffic_from_profiles():
    for name in ("r2_gather_traffic.json", "r1c_gather_traffic.json"):
        try:
            with open(os.path.join(ROOT, "profiles", name)) as f:
                tj = json.load(f)
            return tj["dram_bytes_read_per_launch"] + tj["dram_bytes_write_per_launch"], name
        except (OSError, KeyError, ValueError):
            continue
    return None, None


class Loops:
    """The timed loops over one replay, C ABI (device-resident priorities) and public API."""

    def __init__(self, torch, capi, L, replay, dev, B, dist=None):
        self.torch, self.capi, self.L, self.replay, self.dev, self.B, self.dist = torch, capi, L, replay, dev, B, dist
        self.h = C.c_void_p(replay.handle())
        self.schema = replay.schema()  # (name, dtype, shape, chunks, row_bytes, time_major)
        own = C.c_void_p()
        capi.check(L.rr_get_stream(self.h, C.byref(own)))
        self.own = own
        self.own_stream = torch.cuda.ExternalStream(own.value, device=dev)
        self.stream = torch.cuda.current_stream(dev)
        self.R = 8  # rotating output batches: 232 MB > L2
        self.outs = [[torch.empty(self._shape(t, B), dtype=t[1], device=dev) for t in self.schema]
                     for _ in range(self.R)]
        self.ptrs = [(C.c_void_p * len(o))(*[t.data_ptr() for t in o]) for o in self.outs]
        self.w = [torch.empty(B, dtype=torch.float32, device=dev) for _ in range(self.R)]
        g = torch.Generator(device=dev)
        g.manual_seed(99)
        self.prio_dev = torch.rand((16, B), device=dev, generator=g) + 0.01
        self.prio_host = (torch.rand((16, B)) + 0.01).pin_memory()
        self.w_pinned = [torch.empty(B, dtype=torch.float32).pin_memory() for _ in range(2)]

    @staticmethod
    def _shape(t, B):
        shape = tuple(t[2])
        return (B,) + shape if not t[5] else (shape[0], B) + shape[1:]

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()
        self.torch.cuda.synchronize(self.dev)

    def timed(self, fn, steps, on=None):
        on = self.stream if on is None else on
        self.barrier()
        e0, e1 = self.torch.cuda.Event(enable_timing=True), self.torch.cuda.Event(enable_timing=True)
        e0.record(on)
        for k in range(steps):
            fn(k)
        self.capi.check(self.L.rr_stream_fence(self.h, C.c_void_p(on.cuda_stream)))
        e1.record(on)
        self.barrier()
        ms = e0.elapsed_time(e1)
        if self.dist is not None:
            t = self.torch.tensor([ms], device=self.dev)
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    # -- C ABI, priorities already in HBM, caller on the replay's own stream
    def serial(self, k):
        c, L, h, R, B = self.capi, self.L, self.h, self.R, self.B
        c.check(L.rr_sample(h, B, self.ptrs[k % R], C.c_void_p(self.w[k % R].data_ptr()), self.own))
        c.check(L.rr_update_priority(h, B, C.c_void_p(self.prio_dev[k % 16].data_ptr()), c.RR_PRIO_DEVICE, self.own))

    def serial_foreign(self, k):
        c, L, h, R, B = self.capi, self.L, self.h, self.R, self.B
        sp = C.c_void_p(self.stream.cuda_stream)
        c.check(L.rr_sample(h, B, self.ptrs[k % R], C.c_void_p(self.w[k % R].data_ptr()), sp))
        c.check(L.rr_update_priority(h, B, C.c_void_p(self.prio_dev[k % 16].data_ptr()), c.RR_PRIO_DEVICE, sp))

    def pipelined(self, depth):
        """prefetch = depth through the C ABI: the batch handed out was drawn `depth` samples ago
        (the reference's futures, prioritized_replay.h:215-230); its successors are drawn before its
        update_priority, their gathers run on the gather streams beside the control plane."""
        c, L, h, R, B = self.capi, self.L, self.h, self.R, self.B
        state = {"issued": 0}

        def issue():
            k = state["issued"]
            c.check(L.rr_prefetch_into(h, B, self.ptrs[k % R], C.c_void_p(self.w[k % R].data_ptr()), self.own))
            state["issued"] += 1

        def cycle(k):
            c.check(L.rr_take_prefetched(h, self.own))
            c.check(L.rr_update_priority(h, B, C.c_void_p(self.prio_dev[k % 16].data_ptr()), c.RR_PRIO_DEVICE,
                                         self.own))
            issue()

        def start():
            for _ in range(depth):
                issue()

        def drain():
            for _ in range(depth):
                c.check(L.rr_take_prefetched(h, self.own))
                c.check(L.rr_update_priority(h, B, C.c_void_p(self.prio_dev[0].data_ptr()), c.RR_PRIO_DEVICE,
                                             self.own))

        return start, cycle, drain

    # -- public API: host priorities in, weights read back
    def e2e(self, k):
        batch, w = self.replay.sample(self.B, "cuda:%d" % self.dev.index)
        w_host = self.w_pinned[k & 1]
        w_host.copy_(w, non_blocking=True)   # D2H read of the step's result (also the learner's sync point)
        self.stream.synchronize()
        self.replay.update_priority(self.prio_host[k % 16])
        return w_host

    def c_loop(self, depth, steps, warmup):
        """The same calls in a loop inside the library (rr_cycle_bench): no ctypes overhead per call,
        like the reference arm's loop (oracle/ref_driver.cc). Returns (ms per step, kernel launches)."""
        c, L, h, R, B = self.capi, self.L, self.h, self.R, self.B
        nf = len(self.schema)
        flat = (C.c_void_p * (R * nf))(*[t.data_ptr() for o in self.outs for t in o])
        ws = (C.c_void_p * R)(*[t.data_ptr() for t in self.w])
        ps = (C.c_void_p * 16)(*[self.prio_dev[i].data_ptr() for i in range(16)])
        ms, n = C.c_double(), C.c_uint64()
        self.barrier()
        c.check(L.rr_cycle_bench(h, B, depth, steps, warmup, R, flat, ws, 16, ps, c.RR_PRIO_DEVICE, C.byref(ms),
                                 C.byref(n)))
        ms_v = ms.value
        if self.dist is not None:
            t = self.torch.tensor([ms_v], device=self.dev)
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
            ms_v = float(t.item())
        return ms_v, int(n.value)

    def gather_bench(self, reps, streams, outs=None, ptr_sets=None):
        c, L, h = self.capi, self.L, self.h
        outs = self.outs if outs is None else outs
        nf = len(self.schema)
        flat = (C.c_void_p * (len(outs) * nf))(*[t.data_ptr() for o in outs for t in o])
        ms, n, b = C.c_double(), C.c_uint64(), C.c_uint64()
        self.barrier()
        c.check(L.rr_gather_bench(h, len(outs), flat, max(1, reps), streams, C.byref(ms), C.byref(n), C.byref(b)))
        return ms.value * 1e3, int(n.value), int(b.value)


def time_value_modes(lp, K, W, stats):
    """`value`: prefetch 0 (strictly serial: a batch is drawn after the previous update) and
    prefetch 2 (drawn ahead, like the reference's futures), device-resident priorities, the C ABI
    calls issued from a loop inside the library; the same loops driven from Python through ctypes
    are listed beside them (`ms_per_step_ctypes`)."""
    runs = []
    for depth in (0, 2):
        ms, launches = lp.c_loop(depth, K, W)
        run = {"prefetch": depth, "ms_per_step": ms, "gpu_launches": launches}
        if depth == 0:
            for k in range(W):
                lp.serial(k)
            run["ms_per_step_ctypes"] = lp.timed(lp.serial, K, lp.own_stream) / K
        else:
            start, cycle, drain = lp.pipelined(depth)
            start()
            for k in range(W):
                cycle(k)
            run["ms_per_step_ctypes"] = lp.timed(cycle, K, lp.own_stream) / K
            drain()
        runs.append(run)
    return runs


def run_single(args):
    import numpy as np
    import torch
    import rela_b200
    import rela_b200.rela as rela
    from rela_b200 import capi
    from harness import have_ref

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; rela_b200 has no CPU path")
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    B, K, W = args.batch, args.steps, max(args.warmup, 3)
    L = capi.lib()
    peak, peak_src = peaks()
    t_start = time.time()

    # ---- the replay, filled the way the reference's actors fill theirs (K=80 adds)
    prios = priority_stream(n_fill_adds(args.capacity))
    replay = rela.FFPrioritizedReplay(args.capacity, 7, args.alpha, args.beta, 0, 0)
    _CAP[id(replay)] = args.capacity
    t0 = time.time()
    if args.no_parity:  # development / profiling runs: 257 adds of 8192 rows instead of 26,215 of 80
        serial_of_slot = None
        fill_fast(rela, torch, replay, dev, args.capacity, extra=prios.size - args.capacity)
    else:
        serial_of_slot = fill_k80(rela, torch, replay, dev, prios)
    fill_s = time.time() - t0
    h = C.c_void_p(replay.handle())
    assert replay.size() == prios.size and sum(t[4] for t in replay.schema()) == ROW_BYTES

    # ---- parity: checked before anything is timed
    parity = None
    refs = {}
    if not args.no_parity:
        if have_ref():
            ref = reference_fill(args.capacity, B, args.alpha, args.beta, 0, prios)
            ref.kind_name = "reference (oracle/_ref: unmodified rela/prioritized_replay.h + types.cc)"
            refs[0] = ref
        else:
            ref = OracleAsRef(args.capacity, 7, args.alpha, args.beta, 0, prios)
        parity = parity_single(args, rela, torch, capi, L, replay, dev, prios, serial_of_slot, ref)
        if not have_ref():
            ref.close()
        if not parity["ok"]:
            print(json.dumps({"error": "parity check failed; no number reported", "parity": parity}))
            raise SystemExit(2)
    del serial_of_slot

    stats = lambda: get_stats(capi, L, h)  # noqa: E731
    lp = Loops(torch, capi, L, replay, dev, B)
    clocks = ClockSampler(0)
    clocks.start()

    # ---- value
    value_runs = time_value_modes(lp, K, W, stats)
    ms_foreign = lp.timed(lp.serial_foreign, max(1, K // 4))

    # ---- roofline: the gather kernel over 16 recorded id sets
    capi.check(L.rr_set_profiling(h, 1, B))
    for k in range(16):
        lp.serial(k)
    capi.check(L.rr_set_profiling(h, 0, B))
    g1_us, g1_n, gbytes = lp.gather_bench(max(16, K // 16), 1)   # >= 256 launches whatever --steps says
    g2_us, g2_n, _ = lp.gather_bench(max(16, K // 16), 2)
    big = None
    if B <= 1024:
        B2 = 8 * B
        outs2 = [[torch.empty(Loops._shape(t, B2), dtype=t[1], device=dev) for t in lp.schema] for _ in range(2)]
        ptrs2 = [(C.c_void_p * len(o))(*[t.data_ptr() for t in o]) for o in outs2]
        w2 = torch.empty(B2, dtype=torch.float32, device=dev)
        prio2 = torch.rand(B2, device=dev) + 0.01
        torch.cuda.synchronize(dev)
        capi.check(L.rr_set_profiling(h, 1, B2))
        for k in range(4):
            capi.check(L.rr_sample(h, B2, ptrs2[k % 2], C.c_void_p(w2.data_ptr()), lp.own))
            capi.check(L.rr_update_priority(h, B2, C.c_void_p(prio2.data_ptr()), capi.RR_PRIO_DEVICE, lp.own))
        capi.check(L.rr_set_profiling(h, 0, B2))
        us2, n2, b2 = lp.gather_bench(max(2, K // 128), 1, outs2)
        big = {"batch": B2, "kernel_us": us2, "launches_timed": n2, "bytes_per_launch": b2,
               "achieved": b2 / (us2 * 1e-6) / 1e9, "frac": b2 / (us2 * 1e-6) / 1e9 / peak}
        del outs2, ptrs2

    # ---- e2e, prefetch 0 (this replay)
    for k in range(W):
        lp.e2e(k)
    ms_e2e0 = lp.timed(lp.e2e, K)
    replay.sync()
    st = stats()
    assert st["samples_ordered"] == 0 and st["device_errors"] == 0
    slab_gb = st["slab_bytes"] / 1e9
    e2e_runs = [{"prefetch": 0, "ms_per_step": ms_e2e0 / K}]

    # ---- e2e, prefetch 2: a second replay (prefetch is a constructor argument), filled in big blocks
    del lp
    replay = None
    gc.collect()
    torch.cuda.empty_cache()
    replay = rela.FFPrioritizedReplay(args.capacity, 7, args.alpha, args.beta, 2, 0)
    insert = fill_fast(rela, torch, replay, dev, args.capacity, time_inserts=True)
    h = C.c_void_p(replay.handle())
    lp = Loops(torch, capi, L, replay, dev, B)
    for k in range(W):
        lp.e2e(k)
    ms_e2e2 = lp.timed(lp.e2e, K)
    replay.sync()
    st = stats()
    assert st["samples_ordered"] == 0 and st["device_errors"] == 0
    e2e_runs.append({"prefetch": 2, "ms_per_step": ms_e2e2 / K})
    clk = clocks.stop()
    del lp
    replay = None
    gc.collect()
    torch.cuda.empty_cache()

    best_v = min(value_runs, key=lambda r: r["ms_per_step"])
    best_e = min(e2e_runs, key=lambda r: r["ms_per_step"])
    # the gather as the chosen `value` run launches it: back to back on the internal stream
    # (prefetch 0) or alternating over the two gather streams (batches drawn ahead)
    g_us, g_n, method = (g2_us, g2_n, "two gather streams") if best_v["prefetch"] > 0 else (g1_us, g1_n, "one stream")
    achieved = gbytes / (g_us * 1e-6) / 1e9
    traffic, traffic_src = traffic_from_profiles()
    line = {
        "metric": "replay sample+gather transitions/s", "value": B / (best_v["ms_per_step"] * 1e-3),
        "unit": "transitions/s", "n_gpus": 1, "steps": K, "warmup": W, "ms_per_step": best_v["ms_per_step"],
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": workload_config(args.capacity, B, args.alpha, args.beta),
        "parallelism": "1 GPU",
        "value_runs": [dict(r, transitions_per_s=B / (r["ms_per_step"] * 1e-3)) for r in value_runs],
        "value_note": "C ABI calls on the replay's own stream issued from a loop inside the library (rr_cycle_bench; "
                      "ms_per_step_ctypes: the same loop driven from Python), new priorities already in HBM; prefetch 0 = "
                      "every batch drawn after the previous update_priority; prefetch 2 = two batches drawn ahead "
                      "(rr_prefetch_into / rr_take_prefetched, the reference's futures): the fastest is `value`",
        "e2e": {"value": B / (best_e["ms_per_step"] * 1e-3), "unit": "transitions/s", "h2d_bytes_per_step": 4 * B,
                "d2h_bytes_per_step": 4 * B, "ms_per_step": best_e["ms_per_step"], "prefetch": best_e["prefetch"],
                "runs": [dict(r, transitions_per_s=B / (r["ms_per_step"] * 1e-3)) for r in e2e_runs],
                "api": "rela_b200.rela.FFPrioritizedReplay (C++ host layer, rela/pybind.cc names): sample(B,'cuda:0') + "
                       "weights -> pinned host copy + stream sync + update_priority(pinned cpu tensor)"},
        "gpu_launches": best_v["gpu_launches"],
        "value_foreign_stream": {"value": B * max(1, K // 4) / (ms_foreign * 1e-3), "unit": "transitions/s",
                                 "ms_per_step": ms_foreign / max(1, K // 4),
                                 "note": "prefetch 0 from torch's stream: event hand-shake in and out of every call"},
        "parity": parity, "clocks": clk,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic,
                     "traffic_note": "dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed ncu "
                                     "--set full capture profiles/%s (not from this run); the stores of a launch land "
                                     "in L2 and drain after it" % traffic_src if traffic_src else None,
                     "kernel": "row_copy_kernel (gather)", "kernel_us": g_us, "launches_timed": g_n,
                     "method": "K gather launches over 16 recorded id sets inside one CUDA-event pair, issued the way "
                               "the chosen `value` run issues them: " + method,
                     "methods": {"one_stream_us": g1_us, "one_stream_frac": gbytes / (g1_us * 1e-6) / 1e9 / peak,
                                 "two_gather_streams_us": g2_us,
                                 "two_gather_streams_frac": gbytes / (g2_us * 1e-6) / 1e9 / peak},
                     "bytes_per_launch": gbytes, "peak_source": peak_src,
                     "frac_of_nominal_8tbs": achieved / 8000.0,
                     "step_share": g_us * 1e-3 / best_v["ms_per_step"],
                     "cycle_frac_of_peak": gbytes / (best_v["ms_per_step"] * 1e-3) / 1e9 / peak},
        "fill_s": fill_s, "slab_gb": slab_gb,
    }
    if big is not None:
        line["roofline"]["large_batch"] = big
    if insert is not None:
        insert["frac_of_hbm_peak"] = insert["achieved_gbs"] / peak
        line["insert"] = insert

    # ---- the other BASELINE.json configurations, both arms (measured once each, for the record)
    if not args.no_extras:
        for name, fn in (("config1", extra_config1), ("r2d2", extra_r2d2), ("actor_math", extra_actor_math)):
            try:
                line[name] = fn(args, torch, rela, capi, L, dev, peak)
            except Exception as e:  # noqa: BLE001  (the headline must not depend on a side leg)
                line[name] = {"error": "%s: %s" % (type(e).__name__, e)}
            gc.collect()
            torch.cuda.empty_cache()

    # ---- the reference on this host's cores
    if not args.no_cpu_baseline:
        if have_ref():
            if 0 not in refs:
                refs[0] = reference_fill(args.capacity, B, args.alpha, args.beta, 0, prios)
            refs[2] = reference_fill(args.capacity, B, args.alpha, args.beta, 2, prios)
            best, runs = reference_runs([refs[0], refs[2]], B, args.cpu_cycles, 5, devices=("cpu", "cuda:0"))
            line["cpu_baseline"] = cpu_baseline_obj(best, runs, args.cpu_cycles, "filled by K=80 adds aliasing 64 batches")
        else:
            line["cpu_baseline"] = {"value": None, "unit": "transitions/s", "cores": 0, "kind": "reference",
                                    "sample": "oracle/_ref/libref_replay.so missing"}
    for r in refs.values():
        r.close()
    line["wall_s"] = time.time() - t_start
    print(json.dumps(line))


# ---------------------------------------------------- the other configurations
def extra_config1(args, torch, rela, capi, L, dev, peak):
    """BASELINE.json configs[0] (SURVEY.md 8d "Config 1"): capacity 2^16, batch 128, alpha 0.6,
    beta 0.4, prefetch 0, 820 adds of K=80; 200 cycles after 20 warm-up, both arms."""
    import numpy as np
    from harness import have_ref

    cap, B, n_adds = 1 << 16, 128, 820
    prios = priority_stream(n_adds, seed=2)
    replay = rela.FFPrioritizedReplay(cap, 7, 0.6, 0.4, 0, dev.index)
    _CAP[id(replay)] = cap
    serial_of_slot = fill_k80(rela, torch, replay, dev, prios)
    lp = Loops(torch, capi, L, replay, dev, B)
    out = {"workload": "capacity 65536, batch 128, alpha 0.6 beta 0.4, prefetch 0, 820 adds of K=80, full 4x84x84 rows",
           "cycles": 200, "warmup": 20}
    ref = None
    if have_ref():
        ref = reference_fill(cap, B, 0.6, 0.4, 0, prios)
        # differential check first: 8 cycles, ids + weights + rows
        h = lp.h
        rng = np.random.default_rng(9)
        ok, wrel = True, 0.0
        for c in range(8):
            batch, w = replay.sample(B, "cuda:%d" % dev.index)
            ids = last_ids(capi, L, h, B)
            _, _, w_ref, ids_ref = ref.sample(B, False)
            ok = ok and bool(np.array_equal(ids, ids_ref)) and check_rows(torch, batch, torch.from_numpy(serial_of_slot[ids]))
            wrel = max(wrel, float(np.max(np.abs(w.cpu().numpy() - w_ref) / w_ref)))
            p = rng.random(B).astype(np.float32) + np.float32(0.01)
            replay.update_priority(torch.from_numpy(p))
            ref.update(p)
        out["parity"] = {"cycles": 8, "ids_and_rows_equal": ok, "weights_max_rel": wrel}
        if not ok or wrel > 1e-6:
            raise RuntimeError("config 1 parity failed: %r" % out["parity"])
    for k in range(20):
        lp.e2e(k)
    ms = lp.timed(lp.e2e, 200)
    out["e2e"] = {"ms_per_step": ms / 200, "transitions_per_s": B * 200 / (ms * 1e-3),
                  "api": "public API, host priorities in, weights read back, prefetch 0"}
    for k in range(20):
        lp.serial(k)
    ms = lp.timed(lp.serial, 200, lp.own_stream)
    out["value"] = {"ms_per_step": ms / 200, "transitions_per_s": B * 200 / (ms * 1e-3),
                    "api": "C ABI, device-resident priorities, prefetch 0"}
    replay.sync()
    del lp, replay
    if ref is not None:
        import torch as _t

        _t.set_num_threads(1)  # README.md:125-130
        ref.cycle_loop(B, 20, ref.prio)
        dt = ref.cycle_loop(B, 200, ref.prio)
        _t.set_num_threads(os.cpu_count() or 1)
        out["reference_cpu"] = {"ms_per_step": dt / 200 * 1e3, "transitions_per_s": B * 200 / dt, "cores": 1,
                                "note": "unmodified reference, sample(B,'cpu') + updatePriority, OMP_NUM_THREADS=1"}
        out["e2e_speedup"] = out["e2e"]["transitions_per_s"] / out["reference_cpu"]["transitions_per_s"]
        ref.close()
    return out


def extra_r2d2(args, torch, rela, capi, L, dev, peak):
    """BASELINE.json configs[2]: R2D2 sequence replay, T = 40 + 80 + 3 = 123, LSTM h/c 512, batch 64,
    alpha 0.9 / beta 0.6, priorities from the eta-mix (eta 0.9) of per-step values. 100 k sequences
    are 349 GB: one GPU holds 36,000 (157 GB of slabs incl. the 25 % slack slots); the sequence
    count is stated in the result."""
    import numpy as np
    from harness import RefReplay, have_ref, make_rows, rnn_schema

    T, Bq, hid = 123, 64, 512
    seq_bytes = 3487091
    free, _ = torch.cuda.mem_get_info(dev)
    cap = int(min(36000, (free - 6e9) / (1.25 * seq_bytes)))
    if cap < 2000:
        raise RuntimeError("not enough free HBM for the R2D2 leg")
    replay = rela.RNNPrioritizedReplay(cap, 7, 0.9, 0.6, 0, dev.index)
    g = torch.Generator(device=dev)
    g.manual_seed(4)
    blk = 64
    t0 = time.time()
    done = 0
    while done < cap:
        n = min(blk, cap - done)
        seq_len = torch.randint(41, 121, (n,), device=dev, generator=g).float()
        per_step = torch.rand((n, 80), device=dev, generator=g)
        prio = rela_b200_aggregate(per_step, seq_len, 0.9, 40)
        term = torch.zeros((T, n), dtype=torch.bool, device=dev)
        elems = []
        obs_s = torch.randint(0, 256, (n, T) + FRAME, dtype=torch.uint8, device=dev, generator=g)
        eps = torch.rand((n, T, 1), device=dev, generator=g)
        lm = torch.ones((n, T, N_ACT), device=dev)
        h0 = torch.randn((n, 1, hid), device=dev, generator=g)
        c0 = torch.randn((n, 1, hid), device=dev, generator=g)
        act = torch.randint(0, N_ACT, (n, T), device=dev, generator=g)
        rew = torch.randn((n, T), device=dev, generator=g)
        boot = torch.ones((n, T), device=dev)
        for i in range(n):
            elems.append(rela.RNNTransition({"s": obs_s[i], "eps": eps[i], "legal_move": lm[i]},
                                            {"h0": h0[i], "c0": c0[i]}, {"a": act[i]}, rew[i], term[:, i],
                                            boot[i], seq_len[i:i + 1]))
        replay.add(elems, prio)
        done += n
    replay.sync()
    fill_s = time.time() - t0
    per_step = torch.rand((16, Bq, 80), device=dev, generator=g)
    lens = torch.randint(41, 121, (16, Bq), device=dev, generator=g).float()
    prio_dev = torch.stack([rela_b200_aggregate(per_step[i], lens[i], 0.9, 40) for i in range(16)])
    prio_host = prio_dev.cpu().pin_memory()
    dev_str = "cuda:%d" % dev.index
    stream = torch.cuda.current_stream(dev)
    w_pinned = [torch.empty(Bq, dtype=torch.float32).pin_memory() for _ in range(2)]

    def e2e(k):
        batch, w = replay.sample(Bq, dev_str)
        w_pinned[k & 1].copy_(w, non_blocking=True)
        stream.synchronize()
        replay.update_priority(prio_host[k % 16])

    def dev_cycle(k):
        batch, w = replay.sample(Bq, dev_str)
        replay.update_priority(prio_dev[k % 16])

    def timed(fn, n):
        torch.cuda.synchronize(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for k in range(n):
            fn(k)
        e1.record()
        torch.cuda.synchronize(dev)
        return e0.elapsed_time(e1) / n

    n_cyc = 300
    for k in range(10):
        e2e(k)
    ms_e2e = timed(e2e, n_cyc)
    for k in range(10):
        dev_cycle(k)
    ms_dev = timed(dev_cycle, n_cyc)
    replay.sync()
    st = get_stats(capi, L, C.c_void_p(replay.handle()))
    out = {"workload": "R2D2 sequences: T 123 (burn-in 40 + 80 + 3), 4x84x84 u8 frames, LSTM h0/c0 512, batch 64, "
                       "alpha 0.9 beta 0.6, eta 0.9; batch gathered time-major [T, B, ...]",
           "sequences_held": cap, "sequences_in_config": 100000, "bytes_per_sequence": seq_bytes,
           "slab_gb": st["slab_bytes"] / 1e9, "fill_s": fill_s, "cycles": n_cyc,
           "note": "100,000 sequences = 349 GB do not fit one GPU; one B200 holds %d (striping over 4 GPUs holds all)" % cap,
           "value": {"ms_per_step": ms_dev, "sequences_per_s": Bq / (ms_dev * 1e-3),
                     "env_steps_per_s": Bq * T / (ms_dev * 1e-3),
                     "cycle_gbs": 2 * seq_bytes * Bq / (ms_dev * 1e-3) / 1e9,
                     "cycle_frac_of_hbm_peak": 2 * seq_bytes * Bq / (ms_dev * 1e-3) / 1e9 / peak,
                     "api": "public API, device-resident eta-mix priorities (rr_aggregate_priority), prefetch 0"},
           "e2e": {"ms_per_step": ms_e2e, "sequences_per_s": Bq / (ms_e2e * 1e-3),
                   "api": "public API, host priorities in, weights read back, prefetch 0"},
           "samples_ordered": int(st["samples_ordered"])}
    del replay
    gc.collect()
    torch.cuda.empty_cache()
    if have_ref():
        schema = rnn_schema()
        rng = np.random.default_rng(3)
        pool = [make_rows(schema, 4, rng, 0) for _ in range(4)]
        ref_cap = cap
        ref = RefReplay(schema, ref_cap, 7, 0.9, 0.6, 0)
        for i in range(ref_cap // 4):
            ref.add(pool[i % 4], rng.random(4).astype(np.float32) + 0.01, alias=True)
        pr = rng.random((4, Bq)).astype(np.float32) + 0.01
        ref.cycle_loop(Bq, 2, pr)
        n_ref = 12
        dt = ref.cycle_loop(Bq, n_ref, pr)
        out["reference_cpu"] = {"ms_per_step": dt / n_ref * 1e3, "sequences_per_s": Bq * n_ref / dt, "cycles": n_ref,
                                "cores": os.cpu_count(),
                                "note": "unmodified reference RNNPrioritizedReplay holding the same %d sequences (views "
                                        "of a pool of 16), sample(64,'cpu') + updatePriority" % ref_cap}
        out["e2e_speedup"] = out["e2e"]["sequences_per_s"] / out["reference_cpu"]["sequences_per_s"]
        ref.close()
    return out


def rela_b200_aggregate(per_step, seq_len, eta, burn_in):
    import rela_b200

    return rela_b200.aggregate_priority(per_step.contiguous(), seq_len.contiguous(), eta, burn_in)


def extra_actor_math(args, torch, rela, capi, L, dev, peak):
    """BASELINE.json configs[3]: actor-side batched n-step return + priorities for 80 threads x
    K = 80 envs (6,400 envs), n = 3, gamma 0.997: one launch each on the device vs the reference's
    MultiStepTransitionBuffer::popTransition (dqn_actor.h:58-106) per actor thread of 80 envs."""
    import numpy as np
    import rela_b200
    from harness import have_ref, ref_lib

    K, n, gamma, eta, burn = 6400, 3, 0.997, 0.9, 40
    g = torch.Generator(device=dev)
    g.manual_seed(6)
    reward = torch.randn((n + 1, K), device=dev, generator=g)
    terminal = torch.rand((n + 1, K), device=dev, generator=g) < 0.01
    pr = torch.rand((K, 80), device=dev, generator=g)
    sl = torch.randint(41, 121, (K,), device=dev, generator=g).float()
    r, bs, qn, qs = (torch.randn(K, device=dev, generator=g) for _ in range(4))
    fns = {"nstep_return": lambda: rela_b200.nstep_return(reward, terminal, gamma),
           "aggregate_priority": lambda: rela_b200.aggregate_priority(pr, sl, eta, burn),
           "td_priority": lambda: rela_b200.td_priority(r, bs, qn, qs, gamma ** n)}
    out = {"workload": "6,400 envs (80 actor threads x K = 80), n = 3, gamma 0.997, eta 0.9, burn-in 40", "device": {}}
    iters = 2000
    for name, fn in fns.items():
        for _ in range(20):
            fn()
        torch.cuda.synchronize(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize(dev)
        us = e0.elapsed_time(e1) / iters * 1e3
        out["device"][name] = {"us_per_call": us, "envs_per_s": K / (us * 1e-6), "calls_per_s": 1e6 / us}
    if have_ref():
        lib = ref_lib()
        Kt = 80
        rr_ = np.random.default_rng(0).standard_normal((n + 1, Kt)).astype(np.float32)
        tt = (np.random.default_rng(1).random((n + 1, Kt)) < 0.01).astype(np.uint8)
        loops = 400
        dt = lib.ref_nstep_loop(n, Kt, gamma, rr_.ctypes.data, tt.ctypes.data, loops)
        us = dt / loops * 1e6
        out["reference_cpu"] = {"nstep_pop_us_per_call_of_80_envs": us, "env_steps_per_s_per_core": Kt / (us * 1e-6),
                                "env_steps_per_s_80_threads_ideal": 80 * Kt / (us * 1e-6), "cores": 1,
                                "note": "MultiStepTransitionBuffer::popTransition for one actor thread's 80 envs, timed "
                                        "on one core; 80 threads ideal = x80 (the box has %d threads)" % (os.cpu_count() or 0)}
        out["nstep_speedup_vs_80_ideal_threads"] = (out["device"]["nstep_return"]["envs_per_s"] /
                                                    out["reference_cpu"]["env_steps_per_s_80_threads_ideal"])
    return out


# ------------------------------------------------------------- N GPUs: one striped replay
def run_striped(args):
    """BASELINE.json configs[4]: one replay of world * capacity transitions striped over the GPUs,
    every GPU a learner drawing `batch` rows of one stratified batch of world * batch per step."""
    import numpy as np
    import torch
    import torch.distributed as dist
    import rela_b200
    import rela_b200.rela as rela
    from rela_b200 import capi
    from harness import Field, Oracle, torch_pow

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; rela_b200 has no CPU path")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    try:  # one slice of the host cores per rank: the cycle is a chain of short kernels
        cpus = sorted(os.sched_getaffinity(0))
        per = max(1, len(cpus) // world)
        os.sched_setaffinity(0, set(cpus[local_rank * per:(local_rank + 1) * per]) or set(cpus))
    except (AttributeError, OSError):
        pass
    B, K, W = args.batch, args.steps, max(args.warmup, 3)
    Bg = B * world
    total_cap = args.capacity * world
    peak, peak_src = peaks()
    t_start = time.time()
    state = {"args": args, "world": world, "B": B, "Bg": Bg, "K": K, "W": W, "total_cap": total_cap, "parity": None,
             "runs": [], "peak": (peak, peak_src), "t_start": t_start}

    def end_run(why):
        striped_end_run(state, rank, world, why)

    wd = PhaseWatchdog(end_run, log=(lambda m: print(m, file=sys.stderr, flush=True)) if rank == 0 else None)
    state["wd"] = wd
    _STRIPED["end_run"], _STRIPED["rank"] = end_run, rank
    wd.enter("NCCL process group of %d ranks (torch.distributed)" % world, 240)
    dist.init_process_group("nccl", device_id=dev)  # barriers and the max-over-ranks of the timings only
    L = capi.lib()
    group = "bench%s" % os.environ.get("MASTER_PORT", "0")
    n_local = 2048                       # rows per rank per collective add
    n_block = n_local * world
    n_adds = (total_cap + 64 + n_block - 1) // n_block

    def make(prefetch, learner):
        return rela.FFPrioritizedReplay(total_cap, 7, args.alpha, args.beta, prefetch, local_rank, rank, world, learner,
                                        group)

    def fill(replay, oracle=None):
        """Collective adds of n_local rows per rank: block b holds serials b * n_block + [0, n_block),
        rank r contributing [r * n_local, (r + 1) * n_local). Priorities come from one seeded stream
        every rank regenerates (host priorities: raised with torch::pow on the CPU)."""
        rng = np.random.default_rng(21)
        for b in range(n_adds):
            p = rng.random(n_block, dtype=np.float32) + np.float32(0.01)
            s0 = b * n_block + rank * n_local
            blk = block_of(torch, dev, s0, n_local)
            replay.add(transition_of(rela, blk), torch.from_numpy(p[rank * n_local:(rank + 1) * n_local]))
            if oracle is not None:
                oracle.add([np.arange(b * n_block, (b + 1) * n_block, dtype=np.int64)], torch_pow(p, args.alpha))
        replay.sync()

    def serial_of(ids_oracle_rows):
        return torch.from_numpy(np.ascontiguousarray(ids_oracle_rows))

    # ---- replay 1: every rank a learner, prefetch 0: parity, value (serial), e2e (prefetch 0)
    wd.enter("set-up and fill of the first replay (with the oracle beside it)", 240)
    replay = make(0, -1)
    parity = None
    orc = None
    if not args.no_parity:
        orc = Oracle([Field("action.a", "int64", ())], total_cap, 7, args.alpha, args.beta, 0)
    t0 = time.time()
    fill(replay, orc)
    fill_s = time.time() - t0
    state["fill_s"] = fill_s
    wd.enter("parity cycles against the oracle", 120)
    h = C.c_void_p(replay.handle())
    dev_str = "cuda:%d" % local_rank
    if orc is not None:
        rng = np.random.default_rng(5)
        cycles = max(4, min(args.parity_cycles, 8))
        parity = {"cycles": cycles, "against": "oracle/replay_oracle.c: ONE single-buffer oracle of %d transitions drawing "
                                               "batches of %d" % (total_cap, Bg),
                  "ids_equal": True, "rows_equal": True, "weights_max_rel": 0.0, "batch": Bg}
        per_shard = -(-int(1.25 * total_cap) // world)  # slots per shard (rr_set_shard)
        local_rows = 0
        for c in range(cycles):
            batch, w = replay.sample(B, dev_str)
            # this rank's rows are chunk last_chunk() of the stratified batch of world * B: the
            # chunks are assigned by locality (rr_route::my_chunk), a rotation of the ranks
            chunk = replay.last_chunk()
            sl = slice(chunk * B, (chunk + 1) * B)
            ids = last_ids(capi, L, h, Bg)
            local_rows += int(np.sum(ids[sl] // per_shard == rank))
            rc, rows_o, w_o, ids_o = orc.sample(Bg, True)
            assert rc == 0
            if not np.array_equal(ids, ids_o):
                parity["ids_equal"] = False
                break
            torch.cuda.synchronize(dev)
            parity["weights_max_rel"] = max(parity["weights_max_rel"],
                                            float(np.max(np.abs(w.cpu().numpy() - w_o[sl]) / w_o[sl])))
            if not check_rows(torch, batch, serial_of(rows_o[0][sl])):
                parity["rows_equal"] = False
                break
            p = rng.random(Bg).astype(np.float32) + np.float32(0.01)
            replay.update_priority(torch.from_numpy(p[sl].copy()))
            orc.update(torch_pow(p, args.alpha))
        replay.sync()
        st = get_stats(capi, L, h)
        v = [C.c_int() for _ in range(4)]
        s = C.c_double()
        capi.check(L.rr_debug_state(h, *[C.addressof(x) for x in v], C.addressof(s)))
        os_ = orc.state()
        parity["sum_equal"] = bool(s.value == os_["sum"])
        parity["samples_ordered"] = int(st["samples_ordered"])
        parity["device_errors"] = int(st["device_errors"])
        ok = (parity["ids_equal"] and parity["rows_equal"] and parity["weights_max_rel"] <= 1e-6 and parity["sum_equal"]
              and parity["samples_ordered"] == 0 and parity["device_errors"] == 0)
        t = torch.tensor([1.0 if ok else 0.0], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        parity["ok_on_every_rank"] = bool(t.item() == 1.0)
        t = torch.tensor([float(local_rows)], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        parity["rows_local_frac"] = float(t.item()) / (cycles * Bg)  # rows that did not cross NVLink
        state["parity"] = parity
        orc.close()
        if not parity["ok_on_every_rank"]:
            if rank == 0:
                print(json.dumps({"error": "parity check failed; no number reported", "parity": parity}))
            dist.destroy_process_group()
            raise SystemExit(2)

    own = C.c_void_p()
    capi.check(L.rr_get_stream(h, C.byref(own)))
    stream = torch.cuda.current_stream(dev)
    g = torch.Generator(device=dev)
    g.manual_seed(99 + rank)
    prio_dev = torch.rand((16, B), device=dev, generator=g) + 0.01
    prio_host = (torch.rand((16, B)) + 0.01).pin_memory()
    w_pinned = [torch.empty(B, dtype=torch.float32).pin_memory() for _ in range(2)]

    def barrier():
        dist.barrier()
        torch.cuda.synchronize(dev)

    def timed(rep, fn, steps, on):
        hh = C.c_void_p(rep.handle())
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(on)
        for k in range(steps):
            fn(k)
        capi.check(L.rr_stream_fence(hh, C.c_void_p(on.cuda_stream)))
        e1.record(on)
        barrier()
        t = torch.tensor([e0.elapsed_time(e1)], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def loops(rep):
        own_h = C.c_void_p()
        capi.check(L.rr_get_stream(C.c_void_p(rep.handle()), C.byref(own_h)))
        own_stream = torch.cuda.ExternalStream(own_h.value, device=dev)

        def value_cycle(k):  # device-resident priorities, the caller working on the replay's own stream
            with torch.cuda.stream(own_stream):
                batch, w = rep.sample(B, dev_str)
                rep.update_priority(prio_dev[k % 16])

        def e2e_cycle(k):
            batch, w = rep.sample(B, dev_str)
            if w is not None:  # a learner: read the step's result back (also its sync point)
                w_pinned[k & 1].copy_(w, non_blocking=True)
            stream.synchronize()
            rep.update_priority(prio_host[k % 16])

        return value_cycle, e2e_cycle, own_stream

    def timed_cpp(rep, steps, on):
        """The value loop issued by a C++ caller (pybind `cycle_loop`: the typed sample() +
        updatePriority() of the host layer in a C++ for loop), device-resident priorities."""
        hh = C.c_void_p(rep.handle())
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(on):
            e0.record(on)
            rep.cycle_loop(B, dev_str, prio_dev, steps)
            capi.check(L.rr_stream_fence(hh, C.c_void_p(on.cuda_stream)))
            e1.record(on)
        barrier()
        t = torch.tensor([e0.elapsed_time(e1)], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def measure(rep, learners, prefetch, publish):
        """One replay's numbers, published as each one is known (a later leg that hangs must not take
        the earlier ones with it): value from the interpreter's loop, then from the C++ loop, then e2e."""
        value_cycle, e2e_cycle, own_stream = loops(rep)
        hh = C.c_void_p(rep.handle())
        tag = "prefetch %d, %d learner(s)" % (prefetch, learners)
        res = {"prefetch": prefetch, "learners": learners}
        wd.enter("warm-up + value loop from the interpreter (%s)" % tag, 60)
        for k in range(W):
            value_cycle(k)
        def healthy(what):
            # a leg during which the replay raised a device error (a wait that timed out) is not
            # a measurement: kernels behind the failure no longer wait for anything
            st = get_stats(capi, L, hh)
            if st["device_errors"] != 0 or st["samples_ordered"] != 0:
                raise RuntimeError("%s (%s): device_errors=%d samples_ordered=%d" % (
                    what, tag, st["device_errors"], st["samples_ordered"]))
            return st

        ms_py = timed(rep, value_cycle, max(1, K // 4), own_stream) / max(1, K // 4)
        st = healthy("value loop from the interpreter")
        res.update(value_ms=ms_py, value_ms_python_loop=ms_py, value_loop="interpreter", e2e_ms=None,
                   gpu_launches=None, slab_gb=st["slab_bytes"] / 1e9)
        publish(dict(res))
        wd.enter("value loop from C++ (%s)" % tag, 60)
        s0 = get_stats(capi, L, hh)
        ms_v = timed_cpp(rep, K, own_stream)
        s1 = healthy("value loop from C++")
        res.update(value_ms=ms_v / K, value_loop="C++", gpu_launches=int(s1["kernel_launches"] - s0["kernel_launches"]))
        publish(dict(res))
        wd.enter("e2e loop (%s)" % tag, 60)
        for k in range(W):
            e2e_cycle(k)
        ms_e = timed(rep, e2e_cycle, K, stream)
        rep.sync()
        healthy("e2e loop")
        res.update(e2e_ms=ms_e / K)
        publish(dict(res))
        # (the caller enters the next phase: the replay's teardown is a collective step too)
        wd.enter("teardown of the replay (%s)" % tag, 120)
        return res

    def kernel_timeline(rep, cycles=16):
        """GPU-side durations of this rank's kernels in the pipelined value loop (rr_debug_timeline:
        globaltimer stamps of CTA 0 after its dependency wait and at its end, and of the last CTA of
        the update / gather kernels): what a step is made of."""
        value_cycle, _, own_stream = loops(rep)
        hh = C.c_void_p(rep.handle())
        buf = np.zeros(1024, dtype=np.uint64)
        barrier()
        L.rr_debug_timeline(hh, 1, buf.ctypes.data, 0)
        with torch.cuda.stream(own_stream):
            rep.cycle_loop(B, dev_str, prio_dev, cycles)
        barrier()
        n = L.rr_debug_timeline(hh, 0, buf.ctypes.data, 1024)
        raw = sorted((int(v) >> 8, int(v) & 15, (int(v) >> 4) & 15) for v in buf[:max(n, 0)])
        ev = [(t, c) for t, c, _ in raw]
        def spans(a, b):
            out, t0 = [], None
            for t, c in ev:
                if c == a:
                    t0 = t
                elif c == b and t0 is not None:
                    out.append((t - t0) / 1e3)
                    t0 = None
            return out
        def med(x):
            x = sorted(x)
            return round(x[len(x) // 2], 2) if x else None
        starts = [t for t, c in ev if c == 1]
        return {"select_us": med(spans(1, 2)), "update_cta0_us": med(spans(3, 4)),
                "update_wait_for_all_shards_us": med(spans(3, 9)), "update_all_ctas_us": med(spans(3, 7)),
                "gather_cta0_us": med(spans(5, 6)), "gather_all_ctas_us": med(spans(5, 8)),
                "select_to_select_us": med([(b - a) / 1e3 for a, b in zip(starts, starts[1:])]),
                "events": " ".join("%s%s@%.1f" % ({1: "S+", 2: "S-", 3: "U+", 4: "U-", 5: "G+", 6: "G-", 7: "Uall-",
                                                    8: "Gall-", 9: "Ux", 10: "S0", 11: "U0", 12: "G0", 13: "Gsig",
                                                    14: "Gdone"}.get(c, str(c)), "" if c in (7, 8) else ":%x" % g,
                                                   (t - raw[0][0]) / 1e3)
                                   for t, c, g in raw[len(raw) // 2:len(raw) // 2 + 96]) if raw else "",
                "events_legend": "name:tag@us -- tag = low 4 bits of the batch's epoch (S, G) or of the exchange "
                                 "sequence (U); S0/U0/G0 = CTA 0 resident (before its dependency wait), S+/U+/G+ after "
                                 "the wait, Ux = all shards' priorities in, Gsig = this shard's rows stored and "
                                 "signalled, Gdone = all shards' rows of the batch in",
                "note": "rank 0, pipelined value loop, medians over %d steps; select/update from the kernel's "
                        "dependency wait to CTA 0's end, *_all_ctas to the last CTA's end" % cycles}

    # ---- the measured replays, the most exercised configuration first. `state` always holds what is
    # known so far: if a later phase overruns its budget the watchdog prints the line built from it
    # (marked "incomplete") instead of leaving the run without a number.
    state["runs"] = []

    def publish_run(res):
        runs = state["runs"]
        for i, r in enumerate(runs):
            if r["prefetch"] == res["prefetch"] and r["learners"] == res["learners"]:
                runs[i] = res
                return
        runs.append(res)

    def publish_single(res):
        state["single"] = res

    def fresh(prefetch, learner):
        wd.enter("fill of the replay with prefetch %d, learner %d" % (prefetch, learner), 120)
        gc.collect()
        torch.cuda.empty_cache()
        barrier()
        rep = make(prefetch, learner)
        fill(rep)
        return rep

    clocks = ClockSampler(local_rank)
    state["clocks"] = clocks
    if rank == 0:
        clocks.start()
    # ---- replay 1 (already filled, parity-checked): every rank a learner, prefetch 0
    m0 = measure(replay, world, 0, publish_run)
    state["slab_gb"] = m0["slab_gb"]
    replay = None
    # ---- replay 2: every rank a learner, prefetch 2
    replay = fresh(2, -1)
    measure(replay, world, 2, publish_run)
    wd.enter("kernel timeline (prefetch 2)", 60)
    state["timeline"] = kernel_timeline(replay)
    replay = None
    # ---- replay 2b: prefetch 4. A batch's path -- selection, gather over NVLink, completion flags of
    # all shards, the all-gather of its priorities -- crosses GPUs twice; more batches in flight
    # hide more of that latency (the reference's prefetch is a constructor argument, any depth)
    replay = fresh(4, -1)
    measure(replay, world, 4, publish_run)
    replay = None
    # ---- replay 3: ONE learner (GPU 0) receiving whole batches of `batch` rows over NVLink, prefetch 4
    replay = fresh(4, 0)
    measure(replay, 1, 4, publish_single)
    replay = None
    wd.enter("teardown", 60)
    gc.collect()
    torch.cuda.empty_cache()
    if rank == 0:
        print(json.dumps(striped_line(state, None)), flush=True)
    state["printed"] = True
    wd.enter("process group teardown", 30)  # (a hang here must not turn a finished run into a timeout)
    dist.destroy_process_group()
    wd.done()


_STRIPED = {}  # run_striped's exit hook, for main()


def striped_end_run(state, rank, world, why):
    """A phase of a striped run overran its budget (or raised): rank 0 prints the line of what has
    been measured, or an error object if there is nothing yet; every rank ends its process (no
    clean-up: kernels of this or another rank may be waiting for one another). Does not return."""
    rc = 0
    if rank == 0 and not state.get("printed"):
        try:
            if state["runs"]:
                print(json.dumps(striped_line(state, why)), flush=True)
            else:
                wd = state.get("wd")
                print(json.dumps({"error": "striped run ended without a number: %s" % why, "n_gpus": world,
                                  "phases_s": wd.history if wd is not None else None, "parity": state["parity"]}),
                      flush=True)
                rc = 3
        except Exception as e:  # noqa: BLE001  (never hang in the exit path)
            print("bench.py: could not build the partial line: %r" % (e,), file=sys.stderr, flush=True)
            rc = 3
    sys.stderr.write("bench.py rank %d: ending the run: %s\n" % (rank, why))
    sys.stderr.flush()
    os._exit(rc)


def striped_line(state, incomplete):
    """The JSON line of a striped run from what `state` holds (all of it, or -- when a phase hung and
    the watchdog ends the run -- what had been measured until then; `incomplete` then names the phase)."""
    args, world, B, Bg, K, W = (state[k] for k in ("args", "world", "B", "Bg", "K", "W"))
    total_cap, parity, runs, single = state["total_cap"], state["parity"], state["runs"], state.get("single")
    peak, peak_src = state["peak"]
    clk = state["clocks"].stop() if state.get("clocks") is not None else None
    slab_gb = state.get("slab_gb") or (runs[0]["slab_gb"] if runs else None)

    def tps(ms, rows):
        return rows / (ms * 1e-3)

    best_v = min(runs, key=lambda r: r["value_ms"])
    with_e2e = [r for r in runs if r.get("e2e_ms")]
    best_e = min(with_e2e, key=lambda r: r["e2e_ms"]) if with_e2e else None
    value = tps(best_v["value_ms"], Bg)
    # rows of a learner's chunk that live on another GPU: measured over the parity cycles (chunks are
    # assigned by locality); (N-1)/N without that measurement
    remote = 1.0 - parity["rows_local_frac"] if parity and "rows_local_frac" in parity else (world - 1) / world
    ingress = lambda rate: rate / world * ROW_BYTES * remote / 1e9  # noqa: E731  per-GPU GB/s
    nvlink_cap_rate = world * 900e9 / (ROW_BYTES * max(remote, 1e-9))
    line = {
        "metric": "replay sample+gather transitions/s", "value": value, "unit": "transitions/s", "n_gpus": world,
        "steps": K, "warmup": W, "ms_per_step": best_v["value_ms"], "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": workload_config(total_cap, Bg, args.alpha, args.beta),
        "parallelism": "ONE replay of %d transitions striped over %d GPUs (%.1f GB of slabs each); every GPU draws %d "
                       "rows of one stratified batch of %d per step (the batch's chunks go to the GPUs by locality); "
                       "rows travel GPU to GPU over NVLink" % (total_cap, world, slab_gb or 0.0, B, Bg),
        "value_runs": [{"prefetch": r["prefetch"], "ms_per_step": r["value_ms"], "transitions_per_s": tps(r["value_ms"], Bg),
                        "loop": r["value_loop"], "ms_per_step_python_loop": r["value_ms_python_loop"],
                        "gpu_launches": r["gpu_launches"]} for r in runs],
        "value_note": "sample(B) + update_priority(device priorities) of the C++ host layer on every rank, on the replay's "
                      "own stream, issued by a C++ loop (pybind cycle_loop); ms_per_step_python_loop: the same calls "
                      "issued from the interpreter",
        "e2e": None if best_e is None else {
            "value": tps(best_e["e2e_ms"], Bg), "unit": "transitions/s", "h2d_bytes_per_step": 4 * B * world,
            "d2h_bytes_per_step": 4 * B * world, "ms_per_step": best_e["e2e_ms"], "prefetch": best_e["prefetch"],
            "runs": [{"prefetch": r["prefetch"], "ms_per_step": r["e2e_ms"], "transitions_per_s": tps(r["e2e_ms"], Bg)}
                     for r in with_e2e],
            "api": "rela_b200.rela.FFPrioritizedReplay(..., shard_rank, shard_world) on every rank: sample(B) + weights "
                   "-> pinned host copy + stream sync + update_priority(pinned cpu tensor)"},
        "gpu_launches": best_v["gpu_launches"], "parity": parity, "clocks": clk,
        "roofline": {"bound": "nvlink", "unit": "GB/s", "peak": 900.0,
                     "achieved": ingress(value), "frac": ingress(value) / 900.0, "traffic": None,
                     "peak_source": "NVLink 5, 900 GB/s per direction per GPU (nominal)",
                     "kernel": "row_copy_kernel (routed gather: %.0f %% of every batch crosses NVLink, chunks assigned "
                               "to the GPUs by locality; %.0f %% with a fixed assignment)"
                               % (100 * remote, 100 * (world - 1) / world),
                     "rows_crossing_nvlink_frac": remote,
                     "note": "per-GPU NVLink ingress = remote fraction x 56,617 B x rows per GPU per second; the HBM side "
                             "moves all rows at a fraction of its peak",
                     "linear_cap": {"n_times_nvlink_ingress_cap": nvlink_cap_rate,
                                    "value_frac_of_cap": value / nvlink_cap_rate},
                     "bounded_by": ("NVLink ingress (>= 2/3 of the 900 GB/s port)" if ingress(value) / 900.0 >= 0.66 else
                                    "the control plane's latency chain: selection + update of %d targets and the two "
                                    "flag exchanges per step run back to back on every shard (%.0f us per step), "
                                    "the NVLink port is at %.0f %%" % (Bg, best_v["value_ms"] * 1e3,
                                                                       100 * ingress(value) / 900.0)),
                     "hbm_peak": peak, "hbm_peak_source": peak_src},
        "single_learner": None if not single else {
            "value": tps(single["value_ms"], B), "e2e": tps(single["e2e_ms"], B) if single.get("e2e_ms") else None,
            "unit": "transitions/s", "ms_per_step": single["value_ms"], "prefetch": 4, "batch": B,
            "learner_nvlink_ingress_gbs": tps(single["value_ms"], B) * ROW_BYTES * (world - 1) / world / 1e9,
            "frac_of_900": tps(single["value_ms"], B) * ROW_BYTES * (world - 1) / world / 1e9 / 900.0,
            "note": "the same striped replay with GPU 0 as the only learner: whole batches of %d rows "
                    "assembled on GPU 0" % B},
        "kernel_timeline": state.get("timeline"),
        "phases_s": state["wd"].history if state.get("wd") is not None else None,
        "fill_s": state.get("fill_s"), "slab_gb": slab_gb, "wall_s": time.time() - state["t_start"],
    }
    if incomplete is not None:
        line["incomplete"] = ("the phase '%s' did not finish within its time budget; the run was ended there and this "
                              "line holds what had been measured until then" % incomplete)
    return line


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    elif int(os.environ.get("WORLD_SIZE", "1")) > 1:
        try:
            run_striped(args)
        except SystemExit:
            raise
        except BaseException as e:  # noqa: BLE001
            # One rank failing must not cost the run what it has measured: torchrun tears every rank
            # down as soon as one exits non-zero. The failing rank reports on stderr; rank 0 prints
            # the line of what is known (its watchdog does, if it is waiting for the failed rank).
            import traceback

            traceback.print_exc()
            if "end_run" not in _STRIPED:
                raise
            if _STRIPED["rank"] == 0:
                _STRIPED["end_run"]("rank 0 raised %r" % (e,))
            time.sleep(150)
            os._exit(0)
    else:
        run_single(args)


if __name__ == "__main__":
    main()
