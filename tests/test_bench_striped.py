"""bench.py's striped run (--gpus N) must end within minutes and report what it has even if one of
its phases hangs: the watchdog and the line builder, on the CPU."""
import json
import os
import socket
import subprocess
import sys
import threading
import time

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))

import bench  # noqa: E402


def test_watchdog_fires_only_for_an_overrun_phase():
    fired = []
    ev = threading.Event()

    def on_timeout(name):
        fired.append(name)
        ev.set()

    wd = bench.PhaseWatchdog(on_timeout, poll_s=0.02)
    wd.enter("quick", 5.0)
    time.sleep(0.1)
    wd.enter("slow", 0.15)   # re-arming closes the previous phase
    assert not fired
    assert ev.wait(2.0), "the overrun phase was not reported"
    assert fired == ["slow"]
    time.sleep(0.2)
    assert fired == ["slow"], "one report per overrun"
    wd.enter("ok", 5.0)
    wd.done()
    time.sleep(0.1)
    assert fired == ["slow"]
    assert [n for n, _ in wd.history] == ["quick", "ok"] or [n for n, _ in wd.history] == ["quick", "slow", "ok"]


class _Args:
    alpha, beta = 0.6, 0.4


def _state(runs, single=None, parity=True):
    wd = bench.PhaseWatchdog(lambda name: None, poll_s=10.0)
    return {"args": _Args(), "world": 8, "B": 512, "Bg": 4096, "K": 300, "W": 20, "total_cap": 8 << 21,
            "parity": {"ok_on_every_rank": True, "rows_local_frac": 0.72} if parity else None, "runs": runs,
            "single": single, "peak": (6538.3, "measured"), "t_start": time.time(), "wd": wd, "fill_s": 4.0}


def _run(prefetch, value_ms, e2e_ms, loop="C++"):
    return {"prefetch": prefetch, "learners": 8, "value_ms": value_ms, "value_ms_python_loop": value_ms * 1.1,
            "value_loop": loop, "e2e_ms": e2e_ms, "gpu_launches": 900 if loop == "C++" else None, "slab_gb": 148.4}


def test_striped_line_complete_and_partial():
    full = bench.striped_line(_state([_run(0, 0.13, 0.14), _run(2, 0.08, 0.09)],
                                     single=dict(_run(4, 0.05, 0.06), learners=1)), None)
    json.dumps(full)
    assert full["n_gpus"] == 8 and full["ms_per_step"] == 0.08 and "incomplete" not in full
    assert abs(full["value"] - 4096 / 0.08e-3) < 1e-6 * full["value"]
    assert full["e2e"]["prefetch"] == 2 and full["single_learner"]["batch"] == 512
    # remote fraction comes from the parity cycles' measurement
    assert abs(full["roofline"]["rows_crossing_nvlink_frac"] - 0.28) < 1e-9
    assert full["config"]["capacity"] == 8 << 21 and full["config"]["batch"] == 4096

    # a run that hung after the interpreter's value loop of its first replay: no e2e, no C++ loop yet
    part = bench.striped_line(_state([_run(0, 0.13, None, loop="interpreter")], parity=True), "value loop from C++")
    json.dumps(part)
    assert part["e2e"] is None and part["single_learner"] is None and part["kernel_timeline"] is None
    assert "value loop from C++" in part["incomplete"] and part["value_runs"][0]["loop"] == "interpreter"
    assert part["value"] > 0 and part["gpu_launches"] is None


def test_a_hung_rank_ends_the_run_with_a_partial_line():
    """World 2 over gloo: one rank hangs in a leg, the other waits for it in a collective. Both
    processes must end by themselves with exit code 0 (torchrun tears a job down when a rank exits
    non-zero) and rank 0 must have printed the line of the leg that had been measured."""
    with socket.socket() as sk:
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
    procs = []
    for r in range(2):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable, os.path.join(HERE, "striped_hang_worker.py")], env=env,
                                      stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True))
    outs = []
    for p in procs:
        try:
            outs.append(p.communicate(timeout=120))
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
    assert [p.returncode for p in procs] == [0, 0], outs
    assert "NOT REACHED" not in outs[0][0] + outs[1][0]
    lines = [ln for ln in outs[0][0].splitlines() if ln.startswith("{")]
    assert len(lines) == 1 and outs[1][0].strip() == ""
    line = json.loads(lines[0])
    assert "a leg that hangs" in line["incomplete"] and line["n_gpus"] == 2
    assert line["value"] == 1024 / 0.05e-3 and line["e2e"] is None
    assert [n for n, _ in line["phases_s"]] == ["a leg that works"]
    assert "ending the run" in outs[1][1]


def test_run_striped_control_flow_with_fakes(monkeypatch, capsys):
    """bench.run_striped from its first line to its printed JSON line with everything that needs a
    GPU replaced by a stand-in (the replay class, the C library handle, CUDA events and streams,
    NCCL -> gloo): the orchestration -- phases, the legs published one by one, the four replays,
    the line -- first runs for real on multi-GPU boxes this container does not have."""
    import argparse
    import contextlib
    import types

    import torch
    import torch.distributed as dist

    import rela_b200
    from rela_b200 import capi

    with socket.socket() as sk:
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
    for k, v in (("RANK", "0"), ("LOCAL_RANK", "0"), ("WORLD_SIZE", "1"), ("MASTER_ADDR", "127.0.0.1"),
                 ("MASTER_PORT", str(port))):
        monkeypatch.setenv(k, v)

    real_device = torch.device

    class FakeDeviceMeta(type):
        def __instancecheck__(cls, obj):
            return isinstance(obj, real_device)

    class FakeDevice(metaclass=FakeDeviceMeta):
        def __new__(cls, *a, **k):
            if a and a[0] == "cuda":
                return real_device("cpu")
            return real_device(*a, **k)

    class FakeEvent:
        def __init__(self, enable_timing=False):
            pass

        def record(self, stream=None):
            self.t = time.perf_counter()

        def elapsed_time(self, other):
            return max((other.t - self.t) * 1e3, 1e-3)

    class FakeStream:
        cuda_stream = 0

        def synchronize(self):
            pass

    made = []

    class FakeReplay:
        def __init__(self, cap, seed, alpha, beta, prefetch, local_rank, rank, world, learner, group):
            self.prefetch, self.learner, self.n = prefetch, learner, 0
            made.append(self)

        def add(self, t, p):
            self.n += 1

        def sync(self):
            pass

        def handle(self):
            return 1

        def sample(self, B, dev):
            return object(), torch.zeros(B)

        def update_priority(self, p):
            pass

        def last_chunk(self):
            return 0

        def cycle_loop(self, B, dev, pool, steps):
            for _ in range(steps):
                self.sample(B, dev)

    class FakeL:
        def __getattr__(self, name):
            return lambda *a: 0

    fake_rela = types.SimpleNamespace(FFPrioritizedReplay=FakeReplay, FFTransition=lambda *a: a)
    monkeypatch.setitem(sys.modules, "rela_b200.rela", fake_rela)
    monkeypatch.setattr(rela_b200, "rela", fake_rela, raising=False)
    monkeypatch.setattr(capi, "lib", lambda: FakeL())
    monkeypatch.setattr(torch, "device", FakeDevice)
    monkeypatch.setattr(torch.cuda, "is_available", lambda: True)
    monkeypatch.setattr(torch.cuda, "set_device", lambda d: None)
    monkeypatch.setattr(torch.cuda, "synchronize", lambda d=None: None)
    monkeypatch.setattr(torch.cuda, "empty_cache", lambda: None)
    monkeypatch.setattr(torch.cuda, "current_stream", lambda d=None: FakeStream())
    monkeypatch.setattr(torch.cuda, "ExternalStream", lambda ptr, device=None: FakeStream())
    monkeypatch.setattr(torch.cuda, "stream", lambda s: contextlib.nullcontext())
    monkeypatch.setattr(torch.cuda, "Event", FakeEvent)
    monkeypatch.setattr(torch.Tensor, "pin_memory", lambda self, *a, **k: self)
    monkeypatch.setattr(os, "sched_setaffinity", lambda pid, cpus: None)
    real_init = dist.init_process_group
    monkeypatch.setattr(dist, "init_process_group", lambda backend, device_id=None: real_init("gloo", rank=0, world_size=1))

    args = argparse.Namespace(batch=8, steps=4, warmup=3, capacity=2048, alpha=0.6, beta=0.4, no_parity=True,
                              parity_cycles=8, gpus=1)
    try:
        bench.run_striped(args)
    finally:
        if dist.is_initialized():
            dist.destroy_process_group()
    out = [ln for ln in capsys.readouterr().out.splitlines() if ln.startswith("{")]
    assert len(out) == 1
    line = json.loads(out[0])
    assert "incomplete" not in line and "error" not in line
    assert [(r["prefetch"], r["loop"]) for r in line["value_runs"]] == [(0, "C++"), (2, "C++"), (4, "C++")]
    assert line["e2e"] is not None and len(line["e2e"]["runs"]) == 3
    assert line["single_learner"] is not None and line["kernel_timeline"] is not None
    assert [(m.prefetch, m.learner) for m in made] == [(0, -1), (2, -1), (4, -1), (4, 0)]
    assert all(m.n == made[0].n and m.n > 0 for m in made)          # every replay was filled
    names = [n for n, _ in line["phases_s"]]
    assert names[0].startswith("NCCL process group") and any(n.startswith("kernel timeline") for n in names)
    assert line["steps"] == 4 and line["warmup"] == 3 and line["n_gpus"] == 1
