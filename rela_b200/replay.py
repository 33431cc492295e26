"""Host-side mirror of rela's replay classes over the C ABI.

Names, constructor arguments, method names and error behaviour follow the
reference binding (rela/pybind.cc:19-59) so that pyrela/main.py, apex.py and
r2d2.py can use these classes unchanged:

    replay = FFPrioritizedReplay(capacity, seed, alpha, beta, prefetch)
    batch, weight = replay.sample(batchsize, "cuda:0")
    replay.update_priority(priority_cpu)
    replay.size(); replay.num_add()

`add` is not bound in the reference (actors call it from C++); it is exposed
here with the C++ signatures add(FFTransition batch, priority) and
add(list[RNNTransition], priority) (rela/prioritized_replay.h:186-200).
PyTorch only allocates the tensors; all work is done by librela_b200.so.
"""
import ctypes as C

import torch

from . import capi


class FFTransition:
    """rela.FFTransition (rela/types.h:18-51, pybind.cc:20-26)."""

    __slots__ = ("obs", "action", "reward", "terminal", "bootstrap", "next_obs")

    def __init__(self, obs=None, action=None, reward=None, terminal=None, bootstrap=None, next_obs=None):
        self.obs = obs if obs is not None else {}
        self.action = action if action is not None else {}
        self.reward = reward
        self.terminal = terminal
        self.bootstrap = bootstrap
        self.next_obs = next_obs if next_obs is not None else {}

    def _leaves(self):
        out = [("obs." + k, self.obs[k]) for k in sorted(self.obs)]
        out += [("action." + k, self.action[k]) for k in sorted(self.action)]
        out += [("reward", self.reward), ("terminal", self.terminal), ("bootstrap", self.bootstrap)]
        out += [("next_obs." + k, self.next_obs[k]) for k in sorted(self.next_obs)]
        return out

    @classmethod
    def _from_leaves(cls, leaves):
        t = cls()
        for name, ten in leaves:
            grp, _, key = name.partition(".")
            if key:
                getattr(t, grp)[key] = ten
            else:
                setattr(t, name, ten)
        return t


class RNNTransition:
    """rela.RNNTransition (rela/types.h:53-73, pybind.cc:28-35)."""

    __slots__ = ("obs", "h0", "action", "reward", "terminal", "bootstrap", "seq_len")

    def __init__(self, obs=None, h0=None, action=None, reward=None, terminal=None, bootstrap=None,
                 seq_len=None):
        self.obs = obs if obs is not None else {}
        self.h0 = h0 if h0 is not None else {}
        self.action = action if action is not None else {}
        self.reward = reward
        self.terminal = terminal
        self.bootstrap = bootstrap
        self.seq_len = seq_len

    def _leaves(self):
        out = [("obs." + k, self.obs[k]) for k in sorted(self.obs)]
        out += [("h0." + k, self.h0[k]) for k in sorted(self.h0)]
        out += [("action." + k, self.action[k]) for k in sorted(self.action)]
        out += [("reward", self.reward), ("terminal", self.terminal), ("bootstrap", self.bootstrap),
                ("seq_len", self.seq_len)]
        return out

    _from_leaves = classmethod(FFTransition._from_leaves.__func__)


class _Leaf:
    __slots__ = ("name", "dtype", "shape", "chunks", "row_bytes", "time_major")

    def __init__(self, name, dtype, shape, chunks, time_major=False):
        self.name, self.dtype, self.shape, self.chunks = name, dtype, tuple(shape), chunks
        self.time_major = time_major
        n = 1
        for s in self.shape:
            n *= s
        self.row_bytes = n * torch.empty((), dtype=dtype).element_size()

    def batch_shape(self, B):
        if not self.time_major:
            return (B,) + self.shape
        return (self.shape[0], B) + self.shape[1:]


def _stream_ptr(device):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


class _PrioritizedReplay:
    _transition = None
    _rnn = False

    def __init__(self, capacity, seed, alpha, beta, prefetch, device=None):
        self._lib = capi.lib()
        if not torch.cuda.is_available():
            raise capi.ReplayError(capi.RR_ERR_CUDA, "no CUDA device: rela_b200 has no CPU path")
        self._device = torch.device(device if device is not None else "cuda:%d" % torch.cuda.current_device())
        h = C.c_void_p()
        capi.check(self._lib.rr_create(C.byref(h), int(capacity), int(seed), float(alpha), float(beta),
                                       int(prefetch), self._device.index or 0))
        self._h = h
        self._schema = None

    def __del__(self):
        h = getattr(self, "_h", None)
        if h:
            self._lib.rr_destroy(h)
            self._h = None

    # ---------------------------------------------------------------- schema
    def _set_schema(self, leaves):
        self._schema = leaves
        arr = (capi.RRField * max(1, len(leaves)))()
        self._names = [lf.name.encode() for lf in leaves]
        for i, lf in enumerate(leaves):
            arr[i].name = self._names[i]
            arr[i].row_bytes = lf.row_bytes
            arr[i].chunks = lf.chunks
        capi.check(self._lib.rr_set_schema(self._h, len(leaves), arr))

    # ------------------------------------------------------------------- add
    def _add_flat(self, n, tensors, priority, row_ptrs=None):
        """tensors: one [n, ...] contiguous tensor per leaf (all host or all device)."""
        prio = priority.detach()
        assert prio.dim() == 1 and prio.dtype == torch.float32 and prio.numel() == n
        prio = prio.contiguous()
        flags = 0
        dev_rows = tensors[0].is_cuda if tensors else False
        if dev_rows:
            flags |= capi.RR_ROWS_DEVICE
        if prio.is_cuda:
            flags |= capi.RR_PRIO_DEVICE
        stream = _stream_ptr(self._device) if (dev_rows or prio.is_cuda) else None
        if row_ptrs is None:
            ptrs = (C.c_void_p * max(1, len(tensors)))(*[t.data_ptr() for t in tensors])
            capi.check(self._lib.rr_add(self._h, n, ptrs, None, C.c_void_p(prio.data_ptr()), flags, stream))
        else:
            ptrs = (C.c_void_p * max(1, len(row_ptrs)))(*row_ptrs)
            capi.check(self._lib.rr_add_rows(self._h, n, ptrs, C.c_void_p(prio.data_ptr()), flags, stream))

    # ---------------------------------------------------------------- sample
    def sample(self, batchsize, device):
        """(batch, weights) on `device` (rela/prioritized_replay.h:202-233)."""
        dev = torch.device(device)
        if dev.type != "cuda":
            raise capi.ReplayError(capi.RR_ERR_ARG, "rela_b200 assembles batches on the GPU; device must be cuda")
        if self._schema is None:
            raise capi.ReplayError(capi.RR_ERR_EMPTY, "sample from an empty replay")
        B = int(batchsize)
        outs = [torch.empty(lf.batch_shape(B), dtype=lf.dtype, device=self._device) for lf in self._schema]
        weights = torch.empty(B, dtype=torch.float32, device=self._device)
        ptrs = (C.c_void_p * max(1, len(outs)))(*[t.data_ptr() for t in outs])
        capi.check(self._lib.rr_sample(self._h, B, ptrs, C.c_void_p(weights.data_ptr()),
                                       _stream_ptr(self._device)))
        if self._rnn:
            outs = [t.reshape(B) if lf.name == "seq_len" else t for lf, t in zip(self._schema, outs)]
        batch = self._transition._from_leaves([(lf.name, t) for lf, t in zip(self._schema, outs)])
        if dev != self._device:
            batch = self._transition._from_leaves([(n, t.to(dev)) for n, t in batch._leaves()])
            weights = weights.to(dev)
        return batch, weights

    def update_priority(self, priority):
        """priority: 1-D float32 tensor, CPU (as the reference requires) or CUDA."""
        p = priority.detach()
        if p.dim() != 1 or p.dtype != torch.float32:
            raise capi.ReplayError(capi.RR_ERR_ARG, "priority must be a 1-D float32 tensor")
        p = p.contiguous()
        flags = capi.RR_PRIO_DEVICE if p.is_cuda else 0
        capi.check(self._lib.rr_update_priority(self._h, p.numel(), C.c_void_p(p.data_ptr()), flags,
                                                _stream_ptr(self._device)))

    def size(self):
        return self._lib.rr_size(self._h)

    def num_add(self):
        return self._lib.rr_num_add(self._h)

    def sync(self):
        capi.check(self._lib.rr_sync(self._h))

    def stats(self):
        st = capi.RRStats()
        capi.check(self._lib.rr_get_stats(self._h, C.byref(st)))
        return {n: getattr(st, n) for n, _ in st._fields_}


class FFPrioritizedReplay(_PrioritizedReplay):
    """rela.FFPrioritizedReplay (pybind.cc:37-47)."""

    _transition = FFTransition

    def add(self, batch, priority):
        """add(const FFTransition& batch, priority): row i of every tensor is
        element i (FFTransition::index, rela/types.cc:48-67)."""
        leaves = batch._leaves()
        n = int(priority.numel())
        if self._schema is None:
            self._set_schema([_Leaf(name, t.dtype, t.shape[1:], 1) for name, t in leaves])
        tensors = [t.contiguous() for _, t in leaves]
        self._add_flat(n, tensors, priority)


class RNNPrioritizedReplay(_PrioritizedReplay):
    """rela.RNNPrioritizedReplay (pybind.cc:49-59)."""

    _transition = RNNTransition
    _rnn = True

    def add(self, elements, priority):
        """add(const std::vector<RNNTransition>&, priority): each element holds
        [T, ...] tensors, h0 [layers, hid], seq_len scalar (r2d2_actor.h:109,301)."""
        n = len(elements)
        assert n == int(priority.numel())
        first = elements[0]._leaves()
        if self._schema is None:
            leaves = []
            for name, t in first:
                if name == "seq_len":
                    leaves.append(_Leaf(name, t.dtype, (1,), 1))
                else:
                    leaves.append(_Leaf(name, t.dtype, t.shape, t.shape[0], True))
            self._set_schema(leaves)
        keep = []
        ptrs = [0] * (len(first) * n)
        for e, el in enumerate(elements):
            for f, (_, t) in enumerate(el._leaves()):
                t = t.contiguous()
                keep.append(t)
                ptrs[f * n + e] = t.data_ptr()
        self._add_flat(n, [keep[0]], priority, row_ptrs=ptrs)


# ------------------------------------------------------------- actor-side math
def nstep_return(reward, terminal, gamma, contract_fma=True):
    """MultiStepTransitionBuffer::popTransition arithmetic (rela/dqn_actor.h:58-106).
    reward float32 [n+1, K], terminal bool/uint8 [n+1, K] on CUDA ->
    (return [K], bootstrap [K], terminal [K] bool)."""
    assert reward.is_cuda and reward.dtype == torch.float32 and reward.dim() == 2
    n1, K = reward.shape
    term = terminal.contiguous().view(torch.uint8) if terminal.dtype == torch.bool else terminal.contiguous()
    reward = reward.contiguous()
    out_r = torch.empty(K, dtype=torch.float32, device=reward.device)
    out_b = torch.empty(K, dtype=torch.float32, device=reward.device)
    out_t = torch.empty(K, dtype=torch.uint8, device=reward.device)
    capi.check(capi.lib().rr_nstep_return(n1 - 1, K, float(gamma), reward.data_ptr(), term.data_ptr(),
                                          out_r.data_ptr(), out_b.data_ptr(), out_t.data_ptr(),
                                          1 if contract_fma else 0, _stream_ptr(reward.device)))
    return out_r, out_b, out_t.view(torch.bool)


def td_priority(reward, bootstrap, q_next, q_sa, gamma_n):
    """|reward + bootstrap * gamma^n * q_next - q_sa| (pyrela/apex.py:44-45,77-78)."""
    K = reward.numel()
    args = [t.contiguous() for t in (reward, bootstrap, q_next, q_sa)]
    out = torch.empty(K, dtype=torch.float32, device=reward.device)
    capi.check(capi.lib().rr_td_priority(K, *[t.data_ptr() for t in args], float(gamma_n), out.data_ptr(),
                                         _stream_ptr(reward.device)))
    return out


def aggregate_priority(priority, seq_len, eta, burn_in):
    """eta * max + (1 - eta) * mean per sequence (pyrela/r2d2.py:103-120)."""
    nseq, L = priority.shape
    priority = priority.contiguous()
    seq_len = seq_len.contiguous()
    out = torch.empty(nseq, dtype=torch.float32, device=priority.device)
    capi.check(capi.lib().rr_aggregate_priority(nseq, L, priority.data_ptr(), seq_len.data_ptr(), float(eta),
                                                int(burn_in), out.data_ptr(), _stream_ptr(priority.device)))
    return out
