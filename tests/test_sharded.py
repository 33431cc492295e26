"""Capacity-sharded replay: N ranks must reproduce ONE
single-buffer oracle bit for bit.

CPU (`gloo`, world_size 2): the SPMD protocol of rela_b200/sharded.py with the
device engine replaced by an oracle-backed stand-in (tests/sharded_worker.py).
GPU (`nccl`): the real engine; world 1 on any box, world 2 when two GPUs are
visible (gpurun --gpus 2): rows travel over NVLink into the learner's batch."""
import os
import subprocess
import sys

import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
WORKER = os.path.join(HERE, "sharded_worker.py")


def _run(world, backend, port, extra=()):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
           "--master-addr", "127.0.0.1", "--master-port", str(port), WORKER, "--backend", backend, *extra]
    env = dict(os.environ, OMP_NUM_THREADS="1")
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
    assert out.returncode == 0, (out.stdout[-2000:], out.stderr[-4000:])
    assert "SHARDED_OK world=%d" % world in out.stdout, out.stdout[-2000:]


@pytest.mark.parametrize("world,rotate", [(2, False), (3, False), (2, True)])
def test_protocol_gloo(world, rotate):
    _run(world, "gloo", 29640 + world + 10 * rotate, ("--K", str(12 * world)) + (("--rotate-learner",) if rotate else ()))


def test_shard_ranges_cover_the_ring():
    from rela_b200.sharded import shard_range

    for S in (1, 7, 81920, 2621440, 20971520):
        for world in (1, 2, 3, 4, 8):
            cover = []
            for r in range(world):
                lo, hi = shard_range(S, r, world)
                assert 0 <= lo <= hi <= S
                cover.append((lo, hi))
            assert cover[0][0] == 0 and cover[-1][1] == S
            assert all(a[1] == b[0] for a, b in zip(cover, cover[1:]))


@pytest.mark.gpu
def test_sharded_world1_gpu():
    _run(1, "nccl", 29651)


@pytest.mark.gpu
@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs (gpurun --gpus 2)")
@pytest.mark.parametrize("extra", [(), ("--full-frames", "--capacity", "400", "--cycles", "12"), ("--rotate-learner",)])
def test_sharded_world2_gpu(extra):
    _run(2, "nccl", 29652, extra)


@pytest.mark.gpu
@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_replay_on_a_device_that_is_not_current():
    """The C ABI makes its device current per call and restores the caller's:
    a replay on cuda:1 driven from a thread whose current device is cuda:0."""
    import sys

    sys.path.insert(0, HERE)
    from harness import CudaReplay, Oracle, apply_trace, compare_records, ff_schema, gen_trace

    torch.cuda.set_device(0)
    schema = ff_schema(frame=(4, 84, 84))
    ops = gen_trace(schema, 300, 4, 15, 40, 64)
    orc = Oracle(schema, 300, 5, 0.6, 0.4, 0)
    cu = CudaReplay(schema, 300, 5, 0.6, 0.4, 0, device=1)
    ra = apply_trace(orc, ops, 0.6, False)
    rb = apply_trace(cu, ops, 0.6, False)
    assert cu.sync() == 0
    compare_records(ra, rb, "replay on cuda:1")
    assert torch.cuda.current_device() == 0
