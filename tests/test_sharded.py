"""One replay striped over N shard processes: N ranks must reproduce ONE
single-buffer oracle bit for bit.

CPU: the rendezvous the shards set themselves up with (rr_group_*: shared memory, several
processes, stale segments), and the position-aware priority ^ alpha of the host layer.
GPU: the real thing through the C++ host layer (rela_b200/cpp/sharded_core.cc), one process per
rank, ONE GPU PER RANK: the shards' kernels wait for each other's flags, and kernels that wait
on one another must not share a GPU (nothing guarantees that they run at the same time; ranks
sharing one B200 as processes have raised Xid 109 context-switch timeouts, B200_PROFILING.md).
So these tests need as many GPUs as ranks (`gpurun --gpus 2/4`) and are skipped on a smaller
box; on the driver's boxes the same comparison -- ids, bytes and weights of every rank against
ONE oracle -- runs inside `bench.py --gpus N` before anything is timed (`parity` in its line)."""
import ctypes as C
import multiprocessing as mp
import os
import subprocess
import sys

import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
WORKER = os.path.join(HERE, "sharded_worker.py")


def _group_worker(rank, world, name, rounds, q):
    sys.path.insert(0, os.path.dirname(HERE))
    from rela_b200 import capi

    L = capi.lib()
    g = C.c_void_p()
    rc = L.rr_group_create(C.byref(g), name.encode(), rank, world)
    assert rc == 0, (rc, L.rr_last_error())
    assert L.rr_group_rank(g) == rank and L.rr_group_world(g) == world
    for it in range(rounds):
        mine = (C.c_uint8 * 64)(*[(rank * 7 + it + k) % 256 for k in range(64)])
        allb = (C.c_uint8 * (64 * world))()
        assert L.rr_group_allgather(g, mine, 64, allb) == 0
        for r in range(world):
            assert list(allb[r * 64:(r + 1) * 64]) == [(r * 7 + it + k) % 256 for k in range(64)], (rank, it, r)
    assert L.rr_group_barrier(g) == 0
    L.rr_group_destroy(g)
    q.put(rank)


@pytest.mark.parametrize("world,rank0_last", [(2, False), (3, False), (3, True)])
def test_group_rendezvous(world, rank0_last):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    name = "t%d_%d%d" % (os.getpid(), world, rank0_last)
    ps = [ctx.Process(target=_group_worker, args=(r, world, name, 40, q)) for r in range(world)]
    for p in (ps[1:] + ps[:1] if rank0_last else ps):  # the joiners must wait for rank 0's segment
        p.start()
    for p in ps:
        p.join(120)
    assert [p.exitcode for p in ps] == [0] * world
    assert sorted(q.get() for _ in range(world)) == list(range(world))


def test_group_skips_a_stale_segment():
    """A segment left behind by a dead job (magic set, nobody answering) must not be joined."""
    name = "stale%d" % os.getpid()
    path = "/dev/shm/rela_b200." + name
    with open(path, "wb") as f:  # looks initialised, but its rank 0 is gone
        f.write((0x72656c6162323030).to_bytes(8, "little") + b"\0" * 8 + (2).to_bytes(4, "little") + b"\0" * (1 << 21))
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    p1 = ctx.Process(target=_group_worker, args=(1, 2, name, 5, q))
    p1.start()   # opens the stale segment first, gets no answer, retries by name
    import time

    time.sleep(1.0)
    p0 = ctx.Process(target=_group_worker, args=(0, 2, name, 5, q))
    p0.start()
    p0.join(120)
    p1.join(120)
    assert (p0.exitcode, p1.exitcode) == (0, 0)
    assert not os.path.exists(path)  # rank 0 unlinks the name once everybody is in


def test_group_argument_errors():
    sys.path.insert(0, os.path.dirname(HERE))
    from rela_b200 import capi

    L = capi.lib()
    g = C.c_void_p()
    assert L.rr_group_create(C.byref(g), b"x", 3, 2) == capi.RR_ERR_ARG
    assert L.rr_group_create(C.byref(g), b"x", 0, 99) == capi.RR_ERR_ARG
    assert L.rr_group_create(C.byref(g), b"solo%d" % os.getpid(), 0, 1) == 0  # a world of one needs nobody
    buf = (C.c_uint8 * 8)(*range(8))
    out = (C.c_uint8 * 8)()
    assert L.rr_group_allgather(g, buf, 8, out) == 0 and list(out) == list(range(8))
    assert L.rr_group_allgather(g, buf, 1 << 20, out) == capi.RR_ERR_ARG
    L.rr_group_destroy(g)


def _need_gpus(world):
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs: one per rank (gpurun --gpus %d)" % (world, world))


def _run(world, extra=(), timeout=600):
    _need_gpus(world)
    group = "test%d_%d" % (os.getpid(), abs(hash(extra)) % 100000)
    procs = []
    for r in range(world):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE=str(world), OMP_NUM_THREADS="1")
        procs.append(subprocess.Popen([sys.executable, WORKER, "--group", group, *extra], env=env,
                                      stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    outs = []
    for p in procs:
        try:
            out, _ = p.communicate(timeout=timeout)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
        outs.append(out)
    chunks = []
    for r, (p, out) in enumerate(zip(procs, outs)):
        assert p.returncode == 0, "rank %d:\n%s" % (r, out[-4000:])
        assert "SHARDED_OK rank=%d world=%d" % (r, world) in out, out[-2000:]
        line = [ln for ln in out.splitlines() if ln.startswith("SHARDED_CHUNKS rank=%d " % r)][0]
        seq = line.split(" ", 2)[2].strip() if len(line.split(" ", 2)) > 2 else ""
        chunks.append([int(x) for x in seq.split(",") if x != ""])
    return chunks


def _check_chunks(chunks, world):
    """Every rank a learner: at every sample the ranks hold the `world` chunks of ONE stratified batch,
    each exactly once."""
    assert len({len(c) for c in chunks}) == 1 and len(chunks[0]) > 0
    for per_sample in zip(*chunks):
        assert sorted(per_sample) == list(range(world)), per_sample


def test_priorities_are_raised_at_their_position_in_the_block():
    """ATen's CPU pow sends the vectorised body of a tensor through Sleef and its tail through
    libm: the two differ in the last bit for ~1 element in 4. ONE reference replay raises a whole
    collective block at once, so a shard must raise its slice at its position inside the block."""
    from rela_b200 import _C

    import numpy as np

    g = torch.Generator().manual_seed(0)
    differs = 0
    alpha = float(np.float32(0.6))  # the reference's alpha_ is a float (prioritized_replay.h:330)
    for total, world in ((36, 3), (24, 2), (4096, 8), (100, 4), (48, 4)):
        n = total // world
        x = torch.rand(total, generator=g) + 0.01
        full = torch.pow(x, alpha)
        for r in range(world):
            mine = x[r * n:(r + 1) * n].contiguous()
            got = _C.raise_at_position(mine, r * n, total, 0.6)
            assert torch.equal(got, full[r * n:(r + 1) * n]), (total, world, r)
            differs += int((torch.pow(mine, alpha) != full[r * n:(r + 1) * n]).sum())
        assert torch.equal(_C.raise_at_position(x, 0, total, 1.0), x)  # alpha == 1: the identity, like ATen
    assert differs > 0, "the naive slice-wise pow was expected to differ somewhere (else this test is vacuous)"


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 3])
@pytest.mark.parametrize("learner,prefetch", [(-1, 0), (0, 0), (-1, 2), (1, 1), (0, 3)])
def test_striped_replay_matches_one_oracle(world, learner, prefetch):
    """learner -1: every rank draws its share of one stratified batch of world * B (a
    data-parallel learner); learner r: one learner rank. With and without batches drawn ahead
    (one learner with three: its updates rotate through prefetch + 2 slot arrays, because nothing
    but the gathers holds it back and it may be that far ahead of a shard that only receives)."""
    chunks = _run(world, ("--learner", str(learner), "--prefetch", str(prefetch), "--K", str(12 * world)))
    if learner < 0:
        _check_chunks(chunks, world)
        # the trace wraps the ring several times: the head moves, so does the best rotation
        assert len({c for c in chunks[0]}) > 1, "the assignment never changed: the routed path was not exercised"


@pytest.mark.gpu
def test_striped_replay_full_frames_device_rows():
    """4x84x84 rows (the bulk-copy path of the routed gather and of the owner-routed insert),
    rows handed over in device memory."""
    _check_chunks(_run(2, ("--full-frames", "--device-rows", "--capacity", "400", "--cycles", "12", "--K", "40",
                           "--prefetch", "1")), 2)


@pytest.mark.gpu
def test_striped_replay_world4():
    _check_chunks(_run(4, ("--K", "48", "--B", "16", "--cycles", "25")), 4)


def test_route_assignment_keeps_most_rows_local():
    """rr_route_assignment (pure host arithmetic): for every ring range the chosen chunk -> learner
    map is a bijection and the best one (brute force over all permutations: most slots of the live
    range belong to the shard that receives their chunk), and at the benchmark's geometry (size =
    0.8 S) it keeps 50-99 % of the slots local (72-79 % on average over the head positions) where the fixed
    assignment keeps 1 / world."""
    sys.path.insert(0, os.path.dirname(HERE))
    import itertools

    import numpy as np

    from rela_b200 import capi

    L = capi.lib()

    def assignment(head, size, S, world):
        to = (C.c_int * world)()
        capi.check(L.rr_route_assignment(head, size, S, world, to))
        assert sorted(to) == list(range(world)), list(to)
        return list(to)

    def value_table(head, size, S, world):
        per = -(-S // world)
        pos = (head + np.arange(size)) % S
        chunk = np.minimum(np.searchsorted([(c + 1) * size // world for c in range(world)], np.arange(size),
                                           side="right"), world - 1)
        v = np.zeros((world, world), dtype=np.int64)
        np.add.at(v, (chunk, pos // per), 1)
        return v

    rng = np.random.default_rng(3)
    for world in (2, 3, 4, 5):
        for S in (40, 750, 1000, 4093):
            for _ in range(10):
                head = int(rng.integers(0, S))
                size = int(rng.integers(1, S + 1))
                to = assignment(head, size, S, world)
                v = value_table(head, size, S, world)
                got = sum(v[c, to[c]] for c in range(world))
                best = max(sum(v[c, p[c]] for c in range(world)) for p in itertools.permutations(range(world)))
                assert got == best, (world, S, head, size, to, got, best)
    assert assignment(5, 10, 100, 1) == [0] and assignment(0, 0, 100, 4) == [0, 1, 2, 3]
    assert L.rr_route_assignment(0, 10, 100, 9, (C.c_int * 9)()) == capi.RR_ERR_ARG
    for world, floor in ((2, 0.5), (4, 0.58), (8, 0.62)):  # capacity 2^21 per GPU, full buffer, any head
        S = int(1.25 * (world << 21))
        size = world << 21
        per = -(-S // world)
        for head in (0, 64, S // 3, S - 5, per - 1, 3 * per // 2, 5 * per // 7):
            to = assignment(head % S, size, S, world)
            local = 0  # through interval arithmetic (the ranges are too long to enumerate)
            for c in range(world):
                a, b = head + c * size // world, head + (c + 1) * size // world
                lo, hi = to[c] * per, min(S, (to[c] + 1) * per)
                for off in (0, S):
                    local += max(0, min(b, hi + off) - max(a, lo + off))
            assert local / size > floor, (world, head, to, local / size)


@pytest.mark.gpu
def test_shard_ranges_cover_the_ring():
    sys.path.insert(0, os.path.dirname(HERE))
    from rela_b200 import capi

    L = capi.lib()
    for cap in (1, 6, 65536, 2097152):
        S = int(1.25 * cap)
        for world in (1, 2, 3, 4, 8):
            cover = []
            for r in range(world):
                h = C.c_void_p()
                capi.check(L.rr_create(C.byref(h), cap, 1, 0.6, 0.4, 0, 0))
                capi.check(L.rr_set_shard(h, r, world))
                lo, hi = C.c_int(), C.c_int()
                capi.check(L.rr_shard_range(h, C.byref(lo), C.byref(hi)))
                L.rr_destroy(h)
                assert 0 <= lo.value <= hi.value <= S
                cover.append((lo.value, hi.value))
            assert cover[0][0] == 0 and cover[-1][1] == S
            assert all(a[1] == b[0] for a, b in zip(cover, cover[1:]))


@pytest.mark.gpu
@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_replay_on_a_device_that_is_not_current():
    """The C ABI makes its device current per call and restores the caller's:
    a replay on cuda:1 driven from a thread whose current device is cuda:0."""
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
