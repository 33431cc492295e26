"""Worker of the sharded-replay tests (one process per rank).

  torchrun --nproc-per-node N tests/sharded_worker.py --backend nccl   GPU: CapiEngine
  torchrun --nproc-per-node N tests/sharded_worker.py --backend gloo   CPU: protocol only

Every rank replays the same scripted trace. Each add block of K rows is split
rank-major over the ranks (rank r contributes rows [r*K/N, (r+1)*K/N)), which
is exactly how ShardedPrioritizedReplay concatenates them, so ONE single-buffer
oracle fed the unsplit blocks must produce the same ids, bytes and weights.
With gloo the device engine is replaced by OracleShardEngine below (the
oracle as the replicated control plane, payload kept only for owned slots):
that run checks the SPMD protocol -- block order, ownership ranges, broadcast
roots -- not the kernels.
"""
import argparse
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)

from harness import Oracle, ff_schema, gen_trace, torch_pow  # noqa: E402
from rela_b200.sharded import Leaf, ShardedPrioritizedReplay, shard_range  # noqa: E402


class OracleShardEngine:
    """CPU stand-in for CapiEngine used ONLY by the gloo protocol test."""

    def __init__(self, schema, capacity, seed, alpha, beta, rank, world):
        self.fields = schema
        self.orc = Oracle(schema, capacity, seed, alpha, beta, 0)
        self.S = int(1.25 * capacity)
        self.lo, self.hi = shard_range(self.S, rank, world)
        self.rank, self.world = rank, world
        self.device = torch.device("cpu")
        self.schema = None
        self.tail = 0
        self._ids = None

    def set_schema(self, leaves):
        self.schema = leaves

    def add_block(self, n, tensors, weights):
        slots = (self.tail + np.arange(n)) % self.S
        mine = (slots >= self.lo) & (slots < self.hi)
        rows = []
        for t in tensors:
            a = t.numpy().copy()
            a[~mine] = 0  # payload of other shards is not stored here
            rows.append(a)
        self.orc.add(rows, weights.numpy())
        self.tail = (self.tail + n) % self.S

    def sample(self, B, learner):
        rc, rows, w, ids = self.orc.sample(B)
        assert rc == 0
        self._ids = ids
        mine = (ids >= self.lo) & (ids < self.hi)
        out = []
        for a in rows:
            a = a.copy()
            a[~mine] = 0
            t = torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).astype(np.int32))
            dist.reduce(t, dst=learner, op=dist.ReduceOp.SUM)  # exactly one shard contributes each row
            out.append(torch.from_numpy(t.numpy().astype(np.uint8).view(a.dtype).reshape(a.shape)))
        return (out, torch.from_numpy(w)) if self.rank == learner else (None, None)

    def update(self, weights):
        assert self.orc.update(weights.numpy()) == 0

    def size(self):
        return self.orc.size()

    def num_add(self):
        return self.orc.num_add()

    def last_ids(self, B):
        return self._ids


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--backend", default="nccl")
    ap.add_argument("--capacity", type=int, default=600)
    ap.add_argument("--cycles", type=int, default=40)
    ap.add_argument("--K", type=int, default=40)
    ap.add_argument("--B", type=int, default=64)
    ap.add_argument("--full-frames", action="store_true")
    ap.add_argument("--rotate-learner", action="store_true", help="batch k is assembled on rank k % world")
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.backend == "nccl":
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    else:
        dist.init_process_group("gloo")
    alpha, beta, seed = 0.6, 0.4, 7
    schema = ff_schema() if args.full_frames else ff_schema(frame=(4, 12, 12), n_actions=6)
    assert args.K % world == 0
    ops = gen_trace(schema, args.capacity, 3, args.cycles, args.K, args.B)
    names = [f.name for f in schema]
    engine = None
    if args.backend == "gloo":
        engine = OracleShardEngine(schema, args.capacity, seed, alpha, beta, rank, world)
    rep = ShardedPrioritizedReplay(args.capacity, seed, alpha, beta, names, engine=engine,
                                   device=None if engine else "cuda:%d" % local)
    ref = Oracle(schema, args.capacity, seed, alpha, beta, 0)  # every rank checks the batches it learns from
    n_checked = 0
    learner = 0
    for op in ops:
        if op[0] == "add":
            rows, prio = op[1], op[2]
            n = len(prio)
            per = n // world
            if n % world:  # the trace's short prefill block: pad by giving the remainder to nobody
                rows = [r[: per * world] for r in rows]
                prio = prio[: per * world]
                n = per * world
                if n == 0:
                    continue
            sl = slice(rank * per, (rank + 1) * per)
            rep.add([torch.from_numpy(np.ascontiguousarray(r[sl])) for r in rows],
                    torch.from_numpy(np.ascontiguousarray(prio[sl])))
            if ref is not None:
                ref.add(rows, torch_pow(prio, alpha))
        elif op[0] == "sample":
            B = op[1]
            if args.rotate_learner:
                learner = (learner + 1) % world
            tens, w = rep.sample(B, learner)
            ids = rep.engine.last_ids(B)
            if ref is not None:
                rc, rrows, rw, rids = ref.sample(B)
                assert rc == 0
            np.testing.assert_array_equal(ids, rids, err_msg="ids differ from the single-buffer oracle")
            if rank == learner:
                if args.backend == "nccl":
                    torch.cuda.synchronize()
                got = [t.cpu().numpy() for t in tens]
                gw = w.cpu().numpy()
                np.testing.assert_allclose(gw, rw, rtol=1e-6, atol=0)
                for f, (a, b) in enumerate(zip(got, rrows)):
                    assert a.shape == b.shape and a.tobytes() == b.tobytes(), "field %d bytes differ" % f
                n_checked += 1
            else:
                assert tens is None and w is None
            assert rep.size() == (ref.size() if ref is not None else rep.size())
        else:
            prio = op[1]
            rep.update_priority(torch.from_numpy(prio) if rank == learner else None)
            if ref is not None:
                assert ref.update(torch_pow(prio, alpha)) == 0
    if args.backend == "nccl":
        rep.engine.sync()
        st = capi_stats(rep.engine)
        assert st["samples_ordered"] == 0 and st["device_errors"] == 0
    total = torch.tensor([n_checked], device=rep.device if args.backend == "nccl" else "cpu")
    dist.all_reduce(total)
    n_samples = sum(1 for op in ops if op[0] == "sample")
    assert int(total.item()) == n_samples, (int(total.item()), n_samples)
    if rank == 0:
        print("SHARDED_OK world=%d backend=%s samples_checked=%d" % (world, args.backend, n_samples))
    dist.destroy_process_group()


def capi_stats(engine):
    import ctypes as C

    from rela_b200 import capi

    st = capi.RRStats()
    capi.check(engine.L.rr_get_stats(engine.h, C.byref(st)))
    return {n: getattr(st, n) for n, _ in st._fields_}


if __name__ == "__main__":
    main()
