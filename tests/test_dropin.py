"""Drop-in proof (SURVEY.md section 8b, C++ producer boundary).

tests/dropin/_build/rela.so is the REFERENCE's own rela/pybind.cc + rela/types.cc
compiled against rela_b200/cpp/dropin/rela/prioritized_replay.h: the
reference's DQNActor, MultiStepTransitionBuffer, ModelLocker, VectorEnv,
BasicThreadLoop and Context run unmodified, their threads call
replayBuffer_->add(...) concurrently (rela/dqn_actor.h:189), and Python drives
sample / update_priority exactly as pyrela/main.py:221,245 does. The
environment is tests/dropin/counting_env.cc, whose observations encode
(env id, step), so every sampled transition can be checked end to end:
frames, 3-step return, bootstrap flag and next_obs.
"""
import os
import subprocess
import sys
import time

import numpy as np
import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
BUILD = os.path.join(HERE, "dropin", "_build")
HAVE = os.path.exists(os.path.join(BUILD, "rela.so")) and os.path.exists(os.path.join(BUILD, "counting_env.so"))

N_ACT = 3
FRAME = (4, 84, 84)


class CountingAgent(torch.jit.ScriptModule):
    """The two TorchScript entry points DQNActor calls (rela/dqn_actor.h:162,199),
    as pyrela/apex.py:55-78 defines them; no network."""

    def __init__(self):
        super().__init__()
        self.dummy = torch.nn.Parameter(torch.zeros(1))

    @torch.jit.script_method
    def act(self, obs: dict[str, torch.Tensor]) -> dict[str, torch.Tensor]:
        tag = obs["tag"]
        return {"a": (tag[:, 0] + tag[:, 1]) % 3}

    @torch.jit.script_method
    def compute_priority(self, obs: dict[str, torch.Tensor], action: dict[str, torch.Tensor],
                         reward: torch.Tensor, terminal: torch.Tensor, bootstrap: torch.Tensor,
                         next_obs: dict[str, torch.Tensor]) -> torch.Tensor:
        return reward.abs() + 0.1


HID = 8


class CountingRnnAgent(torch.jit.ScriptModule):
    """The four TorchScript entry points R2D2Actor calls (rela/r2d2_actor.h:221-249,
    304-341), shaped like pyrela/r2d2.py:43-120. The "LSTM state" counts the acts
    since the last reset, so the h0 stored with a sequence is checkable."""

    def __init__(self, eta: float, burn_in: int):
        super().__init__()
        self.dummy = torch.nn.Parameter(torch.zeros(1))
        self.eta = eta
        self.burn_in = burn_in

    @torch.jit.script_method
    def get_h0(self, batchsize: int) -> dict[str, torch.Tensor]:
        return {"h0": torch.zeros(1, batchsize, 8), "c0": torch.zeros(1, batchsize, 8)}

    @torch.jit.script_method
    def act(self, obs: dict[str, torch.Tensor],
            hid: dict[str, torch.Tensor]) -> tuple[dict[str, torch.Tensor], dict[str, torch.Tensor]]:
        tag = obs["tag"]
        return {"a": (tag[:, 0] + tag[:, 1]) % 3}, {"h0": hid["h0"] + 1.0, "c0": hid["c0"] + 2.0}

    @torch.jit.script_method
    def compute_priority(self, obs: dict[str, torch.Tensor], action: dict[str, torch.Tensor],
                         reward: torch.Tensor, terminal: torch.Tensor, bootstrap: torch.Tensor,
                         next_obs: dict[str, torch.Tensor], hid: dict[str, torch.Tensor],
                         next_hid: dict[str, torch.Tensor]) -> torch.Tensor:
        return reward.abs() + 0.1

    @torch.jit.script_method
    def aggregate_priority(self, priority: torch.Tensor, seq_len: torch.Tensor) -> torch.Tensor:
        # pyrela/r2d2.py:103-120
        mask = torch.arange(0, priority.size(1)).unsqueeze(0) < seq_len.unsqueeze(1)
        priority = priority * mask.float()
        p_mean = priority.sum(1) / (seq_len - self.burn_in)
        p_max = priority.max(1)[0]
        return (self.eta * p_max + (1.0 - self.eta) * p_mean).detach().cpu()


def _import_dropin():
    if BUILD not in sys.path:
        sys.path.insert(0, BUILD)
    import rela  # noqa: E402  (tests/dropin/_build/rela.so)
    import counting_env  # noqa: E402

    assert os.path.samefile(os.path.dirname(rela.__file__), BUILD)
    return rela, counting_env


def _period(env_id):
    return PERIOD_BASE + env_id % 4


PERIOD_BASE = 5


def _reward(env_id, step):
    return np.float32(((env_id + step) % 5) * 0.25 - 0.5)


def _expected(env_id, s, n, gamma):
    """(R, bootstrap, terminal0, next_step) of the transition starting at step s,
    following MultiStepTransitionBuffer::popTransition (rela/dqn_actor.h:58-106)."""
    P = _period(env_id)
    steps = [(s + k) % P for k in range(n + 1)]
    term = [st == P - 1 for st in steps]
    bootstrap, last = 1.0, n - 1
    for k in range(n):
        if term[k]:
            bootstrap, last = 0.0, k
            break
    R = np.float32(0.0)
    g = np.float32(gamma)
    for k in range(last, -1, -1):
        R = np.float32(np.float64(_reward(env_id, steps[k])) + np.float64(g) * np.float64(R))  # FMA: one rounding
    return R, bootstrap, term[0], steps[n]


@pytest.mark.skipif(not HAVE, reason="tests/dropin/_build missing (python tests/dropin/build_dropin.py)")
def test_dropin_module_surface():
    """CPU: the module built from the reference's pybind.cc exposes the
    reference's 13 classes (rela/pybind.cc:19-108) and fails loudly without a GPU."""
    rela, counting_env = _import_dropin()
    names = {"Actor", "BasicThreadLoop", "Context", "DQNActor", "Env", "FFPrioritizedReplay", "FFTransition",
             "ModelLocker", "R2D2Actor", "RNNPrioritizedReplay", "RNNTransition", "ThreadLoop", "VectorEnv"}
    assert names <= set(dir(rela))
    rela.ModelLocker([CountingAgent(), CountingAgent()], "cpu")
    if not torch.cuda.is_available():
        with pytest.raises(RuntimeError, match="no CPU path"):
            rela.FFPrioritizedReplay(128, 1, 0.6, 0.4, 0)


@pytest.mark.gpu
@pytest.mark.skipif(not HAVE, reason="tests/dropin/_build missing")
@pytest.mark.parametrize("prefetch", [0, 1])
def test_reference_actors_feed_b200_replay(prefetch):
    """Runs in a child process with a hard timeout: a failure while actor threads
    are blocked in add() on a full ring (rela/prioritized_replay.h:47) would
    otherwise hang in Context::~Context (rela/context.h:26-33), as it does with
    the reference's own replay."""
    out = subprocess.run([sys.executable, os.path.abspath(__file__), "ff", str(prefetch)], capture_output=True,
                         text=True, timeout=300)
    assert out.returncode == 0, (out.stdout[-3000:], out.stderr[-3000:])
    assert "DROPIN_OK" in out.stdout, out.stdout[-2000:]


@pytest.mark.gpu
@pytest.mark.skipif(not HAVE, reason="tests/dropin/_build missing")
def test_reference_r2d2_actors_feed_b200_replay():
    """Same with the reference's R2D2Actor + R2D2TransitionBuffer (sequence assembly,
    burn-in padding, overlap carry-over; rela/r2d2_actor.h) feeding
    rela.RNNPrioritizedReplay through add(std::vector<RNNTransition>, priority).
    Note: with the reference's OWN replay this path throws on torch >= 1.5
    (RNNTransition::makeBatch squeezes a 0-dim seqLen, rela/types.cc:167); the
    B200 replay assembles the batch on the device and returns seq_len as [B]."""
    out = subprocess.run([sys.executable, os.path.abspath(__file__), "rnn", "0"], capture_output=True, text=True,
                         timeout=300)
    assert out.returncode == 0, (out.stdout[-3000:], out.stderr[-3000:])
    assert "DROPIN_OK" in out.stdout, out.stdout[-2000:]


def run_actors(prefetch):
    rela, counting_env = _import_dropin()
    n_threads, n_envs, n_step, gamma = 4, 6, 3, 0.99
    capacity, B = 1500, 64
    replay = rela.FFPrioritizedReplay(capacity, 11, 0.6, 0.4, prefetch)
    locker = rela.ModelLocker([CountingAgent(), CountingAgent()], "cpu")
    ctx = rela.Context()
    actors = []
    for t in range(n_threads):
        venv = rela.VectorEnv()
        for e in range(n_envs):
            env_id = t * n_envs + e
            venv.append(counting_env.CountingEnv(env_id, _period(env_id), *FRAME))
        actor = rela.DQNActor(locker, n_step, n_envs, gamma, replay)
        actors.append(actor)
        ctx.push_env_thread(rela.BasicThreadLoop(actor, venv, False))
    ctx.start()
    t0 = time.time()
    while replay.size() < 4 * B:
        assert time.time() - t0 < 60, "actors did not fill the replay"
        time.sleep(0.01)

    checked = 0
    cycles = 0
    t_start = time.time()
    while (cycles < 120 or replay.num_add() < 4 * capacity) and time.time() - t_start < 120:
        batch, weight = replay.sample(B, "cuda:0")
        assert weight.is_cuda and weight.shape == (B,) and weight.dtype == torch.float32
        w = weight.cpu().numpy()
        assert np.all(w > 0) and np.isclose(w.max(), 1.0)
        s = batch.obs["s"]
        assert s.is_cuda and s.dtype == torch.uint8 and tuple(s.shape) == (B,) + FRAME
        assert tuple(batch.next_obs["s"].shape) == (B,) + FRAME
        tag = batch.obs["tag"].cpu().numpy()
        ntag = batch.next_obs["tag"].cpu().numpy()
        sm = s.reshape(B, FRAME[0], -1)
        lo, hi = sm.amin(dim=2).cpu().numpy(), sm.amax(dim=2).cpu().numpy()
        nm = batch.next_obs["s"].reshape(B, FRAME[0], -1)
        nlo, nhi = nm.amin(dim=2).cpu().numpy(), nm.amax(dim=2).cpu().numpy()
        rew = batch.reward.cpu().numpy()
        boot = batch.bootstrap.cpu().numpy()
        term = batch.terminal.cpu().numpy()
        act = batch.action["a"].cpu().numpy()
        for i in range(B):
            env_id, step = int(tag[i, 0]), int(tag[i, 1])
            R, b, t0_, nstep = _expected(env_id, step, n_step, gamma)
            for c in range(FRAME[0]):
                v = (env_id * 7 + step * 3 + c) & 255
                assert lo[i, c] == v and hi[i, c] == v
                nv = (env_id * 7 + nstep * 3 + c) & 255
                assert nlo[i, c] == nv and nhi[i, c] == nv
            assert ntag[i, 0] == env_id and ntag[i, 1] == nstep
            assert act[i] == (env_id + step) % 3
            assert boot[i] == b and bool(term[i]) == t0_
            assert abs(rew[i] - R) <= 1e-6 * max(1.0, abs(R)), (rew[i], R, env_id, step)
            checked += 1
        replay.update_priority(torch.rand(B) + 0.01)
        assert replay.size() <= int(1.25 * capacity)
        cycles += 1
    assert checked >= 120 * B and replay.num_add() >= 4 * capacity
    assert sum(a.num_act() for a in actors) > 0
    # stop the producers: keep evicting so that adders blocked on a full ring
    # (rela/prioritized_replay.h:47) can leave
    ctx.terminate()
    t0 = time.time()
    while not ctx.terminated():
        assert time.time() - t0 < 60
        replay.sample(B, "cuda:0")
        replay.update_priority(torch.rand(B) + 0.01)
        time.sleep(0.001)


@pytest.mark.gpu
@pytest.mark.skipif(not os.path.exists(os.path.join(BUILD, "test_preplay")), reason="test_preplay not built")
def test_reference_test_preplay_runs_on_b200_replay():
    """rela/tests/test_prioritized_replay.cc compiled AS IS against the drop-in
    header: 10,000 fresh buffers, batch 40 of 32 elements with priorities
    0..31, alpha = beta = 1. The reference prints the empirical frequency vs
    p_i / sum(p) and eyeballs `max diff` (observed 3.9e-4 with its own replay,
    SURVEY.md section 4); the same seeds must give the same figure here."""
    out = subprocess.run([os.path.join(BUILD, "test_preplay")], capture_output=True, text=True, timeout=900)
    assert out.returncode == 0, out.stderr[-2000:]
    last = [ln for ln in out.stdout.splitlines() if ln.startswith("max diff")][-1]
    diff = float(last.split(":")[1])
    assert diff < 6e-4, last


def run_r2d2_actors():
    global PERIOD_BASE
    PERIOD_BASE = 13
    rela, counting_env = _import_dropin()
    n_threads, n_envs, n_step, gamma = 3, 4, 2, 0.99
    seq_len, burn_in, eta = 8, 4, 0.9
    T = burn_in + seq_len + n_step
    frame = (2, 20, 20)
    capacity, B = 300, 16
    replay = rela.RNNPrioritizedReplay(capacity, 5, 0.9, 0.6, 0)
    locker = rela.ModelLocker([CountingRnnAgent(eta, burn_in), CountingRnnAgent(eta, burn_in)], "cpu")
    ctx = rela.Context()
    actors = []
    for t in range(n_threads):
        venv = rela.VectorEnv()
        for e in range(n_envs):
            env_id = t * n_envs + e
            venv.append(counting_env.CountingEnv(env_id, _period(env_id), *frame))
        actor = rela.R2D2Actor(locker, n_step, n_envs, gamma, seq_len, burn_in, replay)
        actors.append(actor)
        ctx.push_env_thread(rela.BasicThreadLoop(actor, venv, False))
    ctx.start()
    t0 = time.time()
    while replay.size() < 2 * B:
        assert time.time() - t0 < 60, "actors did not fill the replay"
        time.sleep(0.01)
    checked = fresh = carried = 0
    cycles = 0
    t_start = time.time()
    while (cycles < 60 or replay.num_add() < 3 * capacity) and time.time() - t_start < 120:
        batch, weight = replay.sample(B, "cuda:0")
        s = batch.obs["s"]
        assert tuple(s.shape) == (T, B) + frame and s.dtype == torch.uint8 and s.is_cuda
        assert tuple(batch.h0["h0"].shape) == (1, B, HID) and tuple(batch.seq_len.shape) == (B,)
        assert tuple(batch.reward.shape) == (T, B) and batch.terminal.dtype == torch.bool
        tag = batch.obs["tag"].cpu().numpy()          # [T, B, 2]
        sm = s.reshape(T, B, frame[0], -1)
        lo, hi = sm.amin(dim=3).cpu().numpy(), sm.amax(dim=3).cpu().numpy()
        rew, boot = batch.reward.cpu().numpy(), batch.bootstrap.cpu().numpy()
        term, act = batch.terminal.cpu().numpy(), batch.action["a"].cpu().numpy()
        h0, c0 = batch.h0["h0"].cpu().numpy(), batch.h0["c0"].cpu().numpy()
        slen = batch.seq_len.cpu().numpy()
        for b in range(B):
            L = int(slen[b])
            assert burn_in < L <= burn_in + seq_len and slen[b] == L
            is_fresh = h0[0, b, 0] == 0
            first = burn_in if is_fresh else 0
            fresh += is_fresh
            carried += not is_fresh
            env_id = int(tag[first, b, 0])
            s0 = int(tag[first, b, 1])
            assert np.all(h0[0, b] == s0) and np.all(c0[0, b] == 2 * s0)
            if is_fresh:
                assert s0 == 0 and not lo[:first, b].any() and not hi[:first, b].any() and not tag[:first, b].any()
            for t in range(first, L):
                step = s0 + (t - first)
                assert tag[t, b, 0] == env_id and tag[t, b, 1] == step
                R, bs, t0_, _ = _expected(env_id, step, n_step, gamma)
                for c in range(frame[0]):
                    v = (env_id * 7 + step * 3 + c) & 255
                    assert lo[t, b, c] == v and hi[t, b, c] == v
                assert act[t, b] == (env_id + step) % 3
                assert boot[t, b] == bs and bool(term[t, b]) == t0_
                assert abs(rew[t, b] - R) <= 1e-6 * max(1.0, abs(R))
                if t0_:
                    assert t == L - 1 or L == burn_in + seq_len
            # beyond the sequence: either the n_step overlap rows or terminal padding
            checked += 1
        replay.update_priority(torch.rand(B) + 0.01)
        cycles += 1
    assert checked >= 60 * B and fresh > 0 and carried > 0, (checked, fresh, carried)
    ctx.terminate()
    t0 = time.time()
    while not ctx.terminated():
        assert time.time() - t0 < 60
        replay.sample(B, "cuda:0")
        replay.update_priority(torch.rand(B) + 0.01)
        time.sleep(0.001)


if __name__ == "__main__":
    code = 1
    try:
        if sys.argv[1] == "rnn":
            run_r2d2_actors()
        else:
            run_actors(int(sys.argv[2]))
        print("DROPIN_OK")
        code = 0
    except BaseException:  # noqa: BLE001
        import traceback

        traceback.print_exc()
    sys.stdout.flush()
    sys.stderr.flush()
    os._exit(code)  # never run Context::~Context with producers blocked on a full ring
