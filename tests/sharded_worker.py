"""Worker of the striped-replay tests: one process per rank, one GPU per rank.

  RANK=r WORLD_SIZE=N python tests/sharded_worker.py --group NAME [--learner -1|r] [--prefetch P] ...

Every rank replays the same scripted trace through the C++ host layer
(rela_b200._C.FFPrioritizedReplay with shard_rank / shard_world, i.e.
rela_b200/cpp/sharded_core.cc). Each add block of K rows is split rank-major over
the ranks (rank r contributes rows [r*K/N, (r+1)*K/N)), which is how the striped
replay orders a collective block, so ONE single-buffer oracle fed the unsplit
blocks must produce the same ids, bytes and weights:

  learner = -1   every rank samples B rows: chunk c = rows [c*B, (c+1)*B) of the oracle's
                 batch of N*B, c = last_chunk() (the chunks are assigned by locality:
                 rr_route::my_chunk); every rank passes the
                 priorities of its rows. The chunk of every sample is printed, the parent
                 test checks that the ranks' chunks are a permutation every time
  learner = r    rank r receives the oracle's whole batch of B
"""
import argparse
import ctypes as C
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)

from harness import Oracle, ff_schema, gen_trace, torch_pow  # noqa: E402


def record(_C, schema, tensors):
    groups = {"obs": {}, "action": {}, "next_obs": {}}
    scalars = {}
    for f, t in zip(schema, tensors):
        grp, _, key = f.name.partition(".")
        if key:
            groups[grp][key] = t
        else:
            scalars[f.name] = t
    return _C.FFTransition(groups["obs"], groups["action"], scalars["reward"], scalars["terminal"],
                           scalars["bootstrap"], groups["next_obs"])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--group", required=True)
    ap.add_argument("--learner", type=int, default=-1)
    ap.add_argument("--prefetch", type=int, default=0)
    ap.add_argument("--capacity", type=int, default=600)
    ap.add_argument("--cycles", type=int, default=40)
    ap.add_argument("--K", type=int, default=24)
    ap.add_argument("--B", type=int, default=32)
    ap.add_argument("--full-frames", action="store_true")
    ap.add_argument("--device-rows", action="store_true")
    a = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    from rela_b200 import _C, capi

    ndev = torch.cuda.device_count()
    assert ndev >= world, "one GPU per rank: the shards' kernels wait on each other's flags"
    dev_index = rank
    torch.cuda.set_device(dev_index)
    dev = torch.device("cuda", dev_index)
    schema = ff_schema() if a.full_frames else ff_schema(frame=(2, 12, 12), n_actions=5)
    alpha, beta, seed = 0.6, 0.4, 11
    n_learn = world if a.learner < 0 else 1
    Btot = a.B * n_learn
    assert a.K % world == 0
    ops = gen_trace(schema, a.capacity, 5, a.cycles, a.K, Btot)
    orc = Oracle(schema, a.capacity, seed, alpha, beta, a.prefetch)
    rep = _C.FFPrioritizedReplay(a.capacity, seed, alpha, beta, a.prefetch, dev_index, rank, world, a.learner, a.group)
    L = capi.lib()
    h = C.c_void_p(rep.handle())
    my_learner = rank if a.learner < 0 else (0 if rank == a.learner else -1)
    n_samples = 0
    chunks = []
    chunk = my_learner
    drawn_from = []  # ring state (head, size) each batch still queued was selected from, oldest first
    for op in ops:
        if op[0] == "add":
            rows, prio = op[1], op[2]
            orc.add(rows, torch_pow(prio, alpha))
            n = len(prio) // world
            lo = rank * n
            tens = [torch.from_numpy(np.ascontiguousarray(r[lo:lo + n])) for r in rows]
            if a.device_rows:
                tens = [t.to(dev) for t in tens]
            rep.add(record(_C, schema, tens), torch.from_numpy(np.ascontiguousarray(prio[lo:lo + n])))
            assert rep.size() == orc.size() and rep.num_add() == orc.num_add()
        elif op[0] == "sample":
            rc, want_rows, want_w, want_ids = orc.sample(Btot)
            assert rc == 0
            # the first sample() selects 1 + prefetch batches, every later one tops the queue up by one;
            # no add runs in between, and an eviction belongs to the first of them: every new
            # selection drew from the ring range the oracle shows now
            st = orc.state()
            while len(drawn_from) < a.prefetch + 1:
                drawn_from.append((st["head"], st["safe_size"]))
            head_sel, size_sel = drawn_from.pop(0)
            batch, w = rep.sample(a.B, "cuda:%d" % dev_index)
            ids = np.zeros(Btot, dtype=np.int32)
            capi.check(L.rr_last_sample_ids(h, Btot, ids.ctypes.data))
            np.testing.assert_array_equal(ids, want_ids, err_msg="rank %d sample %d: ids" % (rank, n_samples))
            assert rep.size() == orc.size()
            if my_learner >= 0:
                chunk = rep.last_chunk()
                assert 0 <= chunk < n_learn
                if a.learner < 0:
                    # the assignment is a pure function of the ring range the batch was selected from
                    # (after that selection's eviction), which every shard holds identically
                    to = (C.c_int * world)()
                    capi.check(L.rr_route_assignment(head_sel, size_sel, int(1.25 * a.capacity), world, to))
                    assert to[chunk] == rank, (rank, chunk, list(to))
                chunks.append(chunk)
                sl = slice(chunk * a.B, (chunk + 1) * a.B)
                torch.cuda.synchronize(dev)
                np.testing.assert_allclose(w.cpu().numpy(), want_w[sl], rtol=1e-6, atol=0,
                                           err_msg="rank %d sample %d: IS weights" % (rank, n_samples))
                attr = {"obs": batch.obs, "action": batch.action, "next_obs": batch.next_obs}
                for f, wr in zip(schema, want_rows):
                    grp, _, key = f.name.partition(".")
                    x = attr[grp][key] if key else getattr(batch, f.name)
                    got = x.cpu().numpy()
                    assert got.shape == wr[sl].shape, (f.name, got.shape, wr[sl].shape)
                    assert got.tobytes() == wr[sl].tobytes(), "rank %d sample %d: field %s bytes differ" % (
                        rank, n_samples, f.name)
            else:
                assert w is None or w.numel() == 0
            n_samples += 1
        else:
            prio = op[1]
            orc.update(torch_pow(prio, alpha))
            if my_learner >= 0:
                mine = prio[chunk * a.B:(chunk + 1) * a.B]
            else:
                mine = np.zeros(0, np.float32)
            rep.update_priority(torch.from_numpy(np.ascontiguousarray(mine)))
    rep.sync()
    # the replicated control state equals the single buffer's
    v = [C.c_int() for _ in range(4)]
    s = C.c_double()
    capi.check(L.rr_debug_state(h, *[C.addressof(x) for x in v], C.addressof(s)))
    st = orc.state()
    assert (v[0].value, v[1].value, v[2].value, v[3].value) == (st["head"], st["tail"], st["size"], st["safe_size"])
    assert s.value == st["sum"], (s.value, st["sum"])
    del rep
    print("SHARDED_CHUNKS rank=%d %s" % (rank, ",".join(map(str, chunks))), flush=True)
    print("SHARDED_OK rank=%d world=%d samples=%d" % (rank, world, n_samples), flush=True)


if __name__ == "__main__":
    main()
