"""Capacity-sharded prioritized replay over the GPUs of one NVSwitch box
(BASELINE.json configs[4]). No reference counterpart:
rela has one host-memory buffer and no communication backend. Parity is
defined against that single buffer: the sharded replay must return the same
ids, bytes and weights as ONE reference replay fed the same elements.

One process per GPU (torchrun), SPMD: every rank calls add / sample /
update_priority collectively, in the same order.

  control plane   replicated. Every shard holds all S weights, evicted flags,
                  the prefix hierarchy, sum_ and the mt19937 (4 B per slot, 10 MB
                  at 2^21) and applies every operation, so all shards select the
                  same ids with NO collective on the sampling path.
  payload         partitioned: shard r stores slots [r * ceil(S/N), ...) of every
                  field slab (56,617 B per slot): 1/N of the bytes per GPU.
  add             the ranks' blocks are concatenated rank-major by an NCCL
                  all-gather (rows + priorities); every shard commits all the
                  priorities and stores the rows it owns.
  sample          every shard runs the selection, then gathers the rows it owns
                  STRAIGHT INTO THE LEARNER GPU'S BATCH TENSOR with bulk stores
                  over NVLink (peer-mapped pointers, rr_ipc_open); one tiny
                  all-reduce marks completion.
  update_priority the learner's priorities are broadcast; every shard applies them.

The engine that touches the device is pluggable so that the SPMD protocol
(ordering, ownership, broadcast roots) can be exercised with `gloo` on CPU in
tests/; the product engine is `CapiEngine` (librela_b200.so) and there is no
CPU engine in this package.
"""
import ctypes as C

import torch
import torch.distributed as dist

from . import capi


class _RawCuda:
    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


class Leaf:
    __slots__ = ("name", "dtype", "shape", "chunks", "row_bytes")

    def __init__(self, name, dtype, shape, chunks=1):
        self.name, self.dtype, self.shape, self.chunks = name, dtype, tuple(int(s) for s in shape), int(chunks)
        n = 1
        for s in self.shape:
            n *= s
        self.row_bytes = n * torch.empty((), dtype=dtype).element_size()

    def batch_shape(self, B):
        return (B,) + self.shape if self.chunks == 1 else (self.shape[0], B) + self.shape[1:]


def shard_range(S, rank, world):
    """Slots [lo, hi) stored by shard `rank` (mirrors rr_set_shard)."""
    per = (S + world - 1) // world
    lo = min(S, rank * per)
    return lo, min(S, lo + per)


class CapiEngine:
    """One shard of the replay in librela_b200.so + the learner batch buffers in peer memory."""

    def __init__(self, capacity, seed, alpha, beta, rank, world, group, device, batch_sets):
        self.L = capi.lib()
        self.rank, self.world, self.group = rank, world, group
        self.device = torch.device(device)
        self.index = self.device.index or 0
        self.batch_sets = batch_sets
        h = C.c_void_p()
        capi.check(self.L.rr_create(C.byref(h), int(capacity), int(seed), float(alpha), float(beta), 0, self.index))
        self.h = h
        capi.check(self.L.rr_set_shard(h, rank, world))
        self.schema = None
        self._sets = {}      # (learner, B) -> list of sets; a set = (field ptr array, weight ptr, tensors or None)
        self._owned = []     # (device, ptr) allocated here
        self._opened = []    # peer mappings
        self._scratch_w = None
        self._token = torch.zeros(1, dtype=torch.int32, device=self.device)
        self._count = 0
        self._flags = {}     # learner -> base pointer of its `world` completion flags (local or peer-mapped)
        # completion of a batch: flag stores from the gather kernels (default) or a 4-byte all-reduce
        import os

        self.fused_completion = os.environ.get("RELA_B200_SHARD_ALLREDUCE", "0") != "1"
        # The steady-state step (select, gather, flag wait, priority push, update) runs
        # on the replay's own stream: plain stream order, no event hand-shake per call.
        # Only the learner touches torch's stream: once after the batch is complete
        # (rr_stream_fence) and once before its new priorities are pushed (rr_stream_wait).
        self._own = C.c_void_p()
        capi.check(self.L.rr_get_stream(h, C.byref(self._own)))
        self.own_stream = self.fused_completion and os.environ.get("RELA_B200_SHARD_OWN_STREAM", "1") != "0"

    def close(self):
        if getattr(self, "h", None):
            torch.cuda.synchronize(self.device)
            for p in self._opened:
                self.L.rr_ipc_close(self.index, p)
            self.L.rr_destroy(self.h)
            for p in self._owned:
                self.L.rr_device_free(self.index, p)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _step_stream(self):
        return self._own if (self.own_stream and self.world > 1) else self._stream()

    def set_schema(self, leaves):
        self.schema = leaves
        arr = (capi.RRField * max(1, len(leaves)))()
        self._names = [lf.name.encode() for lf in leaves]
        for i, lf in enumerate(leaves):
            arr[i].name, arr[i].row_bytes, arr[i].chunks = self._names[i], lf.row_bytes, lf.chunks
        capi.check(self.L.rr_set_schema(self.h, len(leaves), arr))

    def add_block(self, n, tensors, weights):
        ptrs = (C.c_void_p * max(1, len(tensors)))(*[t.data_ptr() for t in tensors])
        flags = capi.RR_ROWS_DEVICE | capi.RR_PRIO_DEVICE | capi.RR_PRIO_IS_WEIGHT
        capi.check(self.L.rr_add(self.h, n, ptrs, None, C.c_void_p(weights.data_ptr()), flags, self._stream()))

    def _make_sets(self, learner, B):
        """Allocate the learner's batch buffers (cudaMalloc on the learner GPU) and
        map them into every other shard's address space (CUDA IPC, NVLink peer access)."""
        nf = len(self.schema)
        sizes = [max(16, lf.row_bytes * B) for lf in self.schema] + [4 * B]
        sets = []
        handles = []
        need_flags = learner not in self._flags
        if self.rank == learner:
            if need_flags:
                p = C.c_void_p()
                capi.check(self.L.rr_device_alloc(self.index, 256, C.byref(p)))
                self._owned.append(p)
                hb = C.create_string_buffer(64)
                capi.check(self.L.rr_ipc_export(self.index, p, hb))
                handles.append(hb.raw)
                self._flags[learner] = p.value
            for _ in range(self.batch_sets):
                ptrs = []
                for nb in sizes:
                    p = C.c_void_p()
                    capi.check(self.L.rr_device_alloc(self.index, nb, C.byref(p)))
                    self._owned.append(p)
                    hb = C.create_string_buffer(64)
                    capi.check(self.L.rr_ipc_export(self.index, p, hb))
                    handles.append(hb.raw)
                    ptrs.append(p.value)
                tens = [torch.as_tensor(_RawCuda(ptrs[i], self.schema[i].row_bytes * B), device=self.device)
                        .view(self.schema[i].dtype).view(self.schema[i].batch_shape(B)) if self.schema[i].row_bytes
                        else torch.empty(self.schema[i].batch_shape(B), dtype=self.schema[i].dtype, device=self.device)
                        for i in range(nf)]
                w = torch.as_tensor(_RawCuda(ptrs[nf], 4 * B), device=self.device).view(torch.float32)
                sets.append(((C.c_void_p * max(1, nf))(*ptrs[:nf]), C.c_void_p(ptrs[nf]), tens, w))
        box = [handles]
        if self.world > 1:
            dist.broadcast_object_list(box, src=dist.get_global_rank(self.group, learner) if self.group else learner,
                                       group=self.group)
        if self.rank != learner:
            hs = box[0]
            k = 0
            if need_flags:
                p = C.c_void_p()
                capi.check(self.L.rr_ipc_open(self.index, hs[0], C.byref(p)))
                self._opened.append(p)
                self._flags[learner] = p.value
                k = 1
            for _ in range(self.batch_sets):
                ptrs = []
                for _nb in sizes:
                    p = C.c_void_p()
                    capi.check(self.L.rr_ipc_open(self.index, hs[k], C.byref(p)))
                    self._opened.append(p)
                    ptrs.append(p.value)
                    k += 1
                sets.append(((C.c_void_p * max(1, nf))(*ptrs[:nf]), None, None, None))
            if self._scratch_w is None or self._scratch_w.numel() < B:
                self._scratch_w = torch.empty(B, dtype=torch.float32, device=self.device)
        self._sets[(learner, B)] = sets

    def sample(self, B, learner):
        key = (learner, B)
        if key not in self._sets:
            self._make_sets(learner, B)
        sets = self._sets[key]
        ptrs, wptr, tens, w = sets[self._count % len(sets)]
        self._count += 1
        if wptr is None:
            wptr = C.c_void_p(self._scratch_w.data_ptr())
        fused = self.fused_completion and self.world > 1
        if fused:
            # completion without a collective: every shard's gather kernel ends by
            # publishing the sample epoch in the learner's memory; the learner's
            # rr_sample appends a one-warp kernel that acquires all `world` flags
            base = self._flags[learner]
            capi.check(self.L.rr_set_completion(self.h, C.c_void_p(base + 8 * self.rank),
                                                C.c_void_p(base) if self.rank == learner else None, self.world))
        capi.check(self.L.rr_sample(self.h, B, ptrs, wptr, self._step_stream()))
        if self.own_stream and self.world > 1 and self.rank == learner:
            capi.check(self.L.rr_stream_fence(self.h, self._stream()))  # the batch, for torch's consumers
        if self.world > 1 and not fused:
            # completion: when this returns on the learner's stream, every shard's
            # gather kernel (ordered before its contribution) has finished its stores
            dist.all_reduce(self._token, group=self.group)
        return (tens, w) if self.rank == learner else (None, None)

    def update(self, weights):
        flags = capi.RR_PRIO_DEVICE | capi.RR_PRIO_IS_WEIGHT
        capi.check(self.L.rr_update_priority(self.h, weights.numel(), C.c_void_p(weights.data_ptr()), flags,
                                             self._stream()))

    INBOX_BYTES = 1 << 20

    def _make_inboxes(self):
        """Every shard owns one inbox in its HBM (8-byte flag, then data at +256),
        mapped by every other shard: the learner pushes the new priorities into
        them with rr_peer_put instead of an NCCL broadcast."""
        p = C.c_void_p()
        capi.check(self.L.rr_device_alloc(self.index, 256 + self.INBOX_BYTES, C.byref(p)))
        self._owned.append(p)
        hb = C.create_string_buffer(64)
        capi.check(self.L.rr_ipc_export(self.index, p, hb))
        handles = [None] * self.world
        dist.all_gather_object(handles, hb.raw, group=self.group)
        self._inbox = [None] * self.world
        self._inbox[self.rank] = p.value
        for r in range(self.world):
            if r != self.rank:
                q = C.c_void_p()
                capi.check(self.L.rr_ipc_open(self.index, handles[r], C.byref(q)))
                self._opened.append(q)
                self._inbox[r] = q.value
        self._err = torch.zeros(1, dtype=torch.int32, device=self.device)
        self._put_epoch = 0

    def exchange(self, weights, learner, n):
        """Learner: push `weights` [n] to every other shard. Others: wait for them.
        Returns the device pointer of the n weights on this shard."""
        if getattr(self, "_inbox", None) is None:
            self._make_inboxes()
        if 4 * n > self.INBOX_BYTES:
            raise capi.ReplayError(capi.RR_ERR_ARG, "update_priority: batch too large for the peer inbox")
        self._put_epoch += 1
        if self.rank == learner:
            others = [r for r in range(self.world) if r != learner]
            dst = (C.c_void_p * len(others))(*[self._inbox[r] + 256 for r in others])
            flg = (C.c_void_p * len(others))(*[self._inbox[r] for r in others])
            if self.own_stream:
                capi.check(self.L.rr_stream_wait(self.h, self._stream()))  # `weights` was produced on torch's stream
            capi.check(self.L.rr_peer_put(self.index, C.c_void_p(weights.data_ptr()), 4 * n, dst, flg, len(others),
                                          self._put_epoch, self._step_stream()))
            return weights.data_ptr()
        mine = self._inbox[self.rank]
        capi.check(self.L.rr_wait_flags(self.index, C.c_void_p(mine), 1, self._put_epoch,
                                        C.c_void_p(self._err.data_ptr()), self._step_stream()))
        return mine + 256

    def update_ptr(self, ptr, n):
        flags = capi.RR_PRIO_DEVICE | capi.RR_PRIO_IS_WEIGHT
        capi.check(self.L.rr_update_priority(self.h, n, C.c_void_p(ptr), flags, self._step_stream()))

    def size(self):
        return self.L.rr_size(self.h)

    def num_add(self):
        return self.L.rr_num_add(self.h)

    def sync(self):
        capi.check(self.L.rr_sync(self.h))
        if getattr(self, "_err", None) is not None and int(self._err.item()) != 0:
            raise capi.ReplayError(capi.RR_ERR_CUDA, "a shard never published its priorities (peer flag timeout)")

    def last_ids(self, B):
        import numpy as np

        ids = np.zeros(B, dtype=np.int32)
        capi.check(self.L.rr_last_sample_ids(self.h, B, ids.ctypes.data))
        return ids


class ShardedPrioritizedReplay:
    """SPMD front end. Collective methods: add, sample, update_priority.

    add(leaves, priority)      leaves: this rank's rows, one [n, ...] tensor per leaf
                               (same n on every rank); priority [n] float32.
                               Global order of the block: rank-major.
    sample(B, learner=0)       -> (list of batch tensors, weights) on `learner`,
                               (None, None) elsewhere. The tensors alias a ring
                               of `batch_sets` buffers: consume them before the
                               batch_sets-th next sample.
    update_priority(priority)  priority given on the learner of the last sample
                               (None elsewhere).
    """

    def __init__(self, capacity, seed, alpha, beta, names, group=None, device=None, batch_sets=2, chunks=None,
                 engine=None):
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.capacity, self.S = int(capacity), int(1.25 * capacity)
        self.alpha = float(alpha)
        self.names = list(names)
        self.chunks = dict(chunks or {})  # leaf name -> T for time-major leaves ([n, T, ...] rows)
        if engine is None:
            if not torch.cuda.is_available():
                raise capi.ReplayError(capi.RR_ERR_CUDA, "no CUDA device: rela_b200 has no CPU path")
            device = device if device is not None else "cuda:%d" % torch.cuda.current_device()
            engine = CapiEngine(capacity, seed, alpha, beta, self.rank, self.world, group, device, batch_sets)
        self.engine = engine
        self.device = engine.device
        self._size = 0
        self._learner = None
        self._B = 0

    def _src(self, rank):
        return dist.get_global_rank(self.group, rank) if self.group is not None else rank

    def _raise(self, priority):
        """priority ** alpha as the reference computes it: torch.pow on the CPU
        for CPU tensors (rela/prioritized_replay.h:188,239)."""
        p = priority.detach()
        if p.dim() != 1 or p.dtype != torch.float32:
            raise capi.ReplayError(capi.RR_ERR_ARG, "priority must be a 1-D float32 tensor")
        w = torch.pow(p, self.alpha) if self.alpha != 1.0 else p
        return w.to(self.device).contiguous()

    def add(self, leaves, priority):
        n = int(priority.numel())
        if self.engine.schema is None:
            specs = [Leaf(name, t.dtype, t.shape[1:], self.chunks.get(name, 1)) for name, t in zip(self.names, leaves)]
            self.engine.set_schema(specs)
        total = n * self.world
        if self._size + total > self.S:
            # the reference blocks here until a sample evicts (prioritized_replay.h:47);
            # in an SPMD program the same thread would have to sample: report it
            raise capi.ReplayError(capi.RR_ERR_STATE, "add: ring is full (%d + %d > %d slots); sample first"
                                   % (self._size, total, self.S))
        w_local = self._raise(priority)
        rows = [t.to(self.device).contiguous() for t in leaves]
        if self.world > 1:
            w_all = torch.empty(total, dtype=torch.float32, device=self.device)
            dist.all_gather_into_tensor(w_all, w_local, group=self.group)
            full = []
            for t in rows:
                out = torch.empty((total,) + tuple(t.shape[1:]), dtype=t.dtype, device=self.device)
                if t.dtype == torch.bool:
                    dist.all_gather_into_tensor(out.view(torch.uint8), t.view(torch.uint8), group=self.group)
                else:
                    dist.all_gather_into_tensor(out, t, group=self.group)
                full.append(out)
        else:
            w_all, full = w_local, rows
        self.engine.add_block(total, full, w_all)
        self._size += total

    def sample(self, batchsize, learner=0):
        if self._learner is not None:
            raise capi.ReplayError(capi.RR_ERR_STATE, "sample: previous samples' priority has not been updated")
        out = self.engine.sample(int(batchsize), learner)
        if self._size > self.capacity:  # eviction inside sample_ (prioritized_replay.h:311-315)
            self._size = self.capacity
        self._learner, self._B = learner, int(batchsize)
        return out

    def update_priority(self, priority=None):
        if self._learner is None:
            raise capi.ReplayError(capi.RR_ERR_STATE, "update_priority: no sampled batch is outstanding")
        if self.rank == self._learner:
            w = self._raise(priority)
            if w.numel() != self._B:
                raise capi.ReplayError(capi.RR_ERR_ARG, "update_priority: %d priorities for a batch of %d"
                                       % (w.numel(), self._B))
        else:
            w = None
        if self.world > 1 and getattr(self.engine, "fused_completion", False):
            # one-sided: the learner stores the weights into every shard's inbox
            ptr = self.engine.exchange(w, self._learner, self._B)
            self._keep = w  # alive until the put kernel has run (stream-ordered free is fine for torch's allocator)
            self.engine.update_ptr(ptr, self._B)
        else:
            if w is None:
                w = torch.empty(self._B, dtype=torch.float32, device=self.device)
            if self.world > 1:
                dist.broadcast(w, src=self._src(self._learner), group=self.group)
            self.engine.update(w)
        self._learner = None

    def size(self):
        return self.engine.size()

    def num_add(self):
        return self.engine.num_add()


FF_LEAVES = ["obs.s", "obs.eps", "obs.legal_move", "action.a", "reward", "terminal", "bootstrap",
             "next_obs.s", "next_obs.eps", "next_obs.legal_move"]
