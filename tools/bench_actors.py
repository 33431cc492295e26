"""System-level drop-in benchmark, after pyrela/benchmark.py:19-126 (which needs
ALE ROMs): the reference's OWN actor runtime -- Context, BasicThreadLoop,
VectorEnv, ModelLocker, DQNActor with its MultiStepTransitionBuffer -- produces
Ape-X transitions from T threads x K synthetic envs (tests/dropin/counting_env.cc,
4x84x84 uint8 frames) into a replay while a learner loop samples batches on
cuda:0 and updates priorities.

  --impl ours       tests/dropin/_build/rela.so: reference pybind.cc + actors compiled
                    against rela_b200/cpp/dropin/rela/prioritized_replay.h (B200 replay)
  --impl reference  tests/dropin/_build/reference/rela.so: the same sources with the
                    reference's own rela/prioritized_replay.h (host-memory replay)

Prints one JSON line: learner cycles/s, sampled transitions/s, actor adds/s."""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BUILD = os.path.join(ROOT, "tests", "dropin", "_build")
ap = argparse.ArgumentParser()
ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
ap.add_argument("--threads", type=int, default=12)
ap.add_argument("--envs", type=int, default=40)
ap.add_argument("--capacity", type=int, default=1 << 16)
ap.add_argument("--batch", type=int, default=512)
ap.add_argument("--seconds", type=float, default=10.0)
ap.add_argument("--prefetch", type=int, default=0)
args = ap.parse_args()
sys.path.insert(0, os.path.join(BUILD, "reference") if args.impl == "reference" else BUILD)
sys.path.insert(1, BUILD)
sys.path.insert(2, os.path.join(ROOT, "tests"))
import torch  # noqa: E402

import rela  # noqa: E402
import counting_env  # noqa: E402
from test_dropin import CountingAgent  # noqa: E402

assert torch.cuda.is_available()
torch.zeros(1, device="cuda:0")
replay = rela.FFPrioritizedReplay(args.capacity, 1, 0.6, 0.4, args.prefetch)
locker = rela.ModelLocker([CountingAgent() for _ in range(3)], "cpu")
ctx = rela.Context()
actors = []
for t in range(args.threads):
    venv = rela.VectorEnv()
    for e in range(args.envs):
        venv.append(counting_env.CountingEnv(t * args.envs + e, 200 + (t * args.envs + e) % 50, 4, 84, 84))
    actor = rela.DQNActor(locker, 3, args.envs, 0.997, replay)
    actors.append(actor)
    ctx.push_env_thread(rela.BasicThreadLoop(actor, venv, False))
ctx.start()
burn = min(args.capacity, 8 * args.batch)
while replay.size() < burn:
    time.sleep(0.01)
B = args.batch
prio = torch.rand(B) + 0.01
for _ in range(10):
    batch, w = replay.sample(B, "cuda:0")
    replay.update_priority(prio)
torch.cuda.synchronize()
a0, t0 = replay.num_add(), time.time()
act0 = sum(a.num_act() for a in actors)
cycles = 0
while time.time() - t0 < args.seconds:
    batch, w = replay.sample(B, "cuda:0")
    _ = float(w[0])  # the learner reads something from the batch: sync point
    replay.update_priority(prio)
    cycles += 1
torch.cuda.synchronize()
dt = time.time() - t0
adds = replay.num_add() - a0
acts = sum(a.num_act() for a in actors) - act0
print(json.dumps({"impl": args.impl, "threads": args.threads, "envs_per_thread": args.envs, "capacity": args.capacity,
                  "batch": B, "prefetch": args.prefetch, "seconds": dt, "learner_cycles_per_s": cycles / dt,
                  "sampled_transitions_per_s": cycles * B / dt, "actor_adds_per_s": adds / dt,
                  "actor_acts_per_s": acts / dt, "replay_size": replay.size(), "host_cpus": os.cpu_count()}))
sys.stdout.flush()
ctx.terminate()
t1 = time.time()
while not ctx.terminated() and time.time() - t1 < 20:
    replay.sample(B, "cuda:0")
    replay.update_priority(prio)
os._exit(0)
