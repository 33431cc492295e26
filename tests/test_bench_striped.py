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
