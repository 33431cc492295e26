"""GPU: rr_update_priority_noalloc (include/rela_b200.h) against rr_update_priority.

The entry point has no counterpart in the reference; it was added
for callers that cannot call cudaMalloc. Its update kernel works in a file-scope
__device__ array (RR_NOALLOC_MAX_BATCH elements) instead of device memory the replay allocates.

Two twin replays receive the same adds and draw the same batches; one applies every batch's new
priorities with rr_update_priority (device-resident priorities), the other with the no-alloc
entry point, through the C ABI. After each of several updates IN A ROW (the array keeps the
previous call's contents) the two replays must be in the same state: stored weights, eviction
flags, sum_, ring state and the selection hierarchy bit for bit (integer / index work and the
same arithmetic: the tolerance of the other parity tests for these quantities is exact
equality), and the next batches drawn (ids exact, importance weights within 1e-6 relative).

The file is named so that pytest collects it last; it passes alone in a fresh process and after
the other tests (the static array is per process and nothing else uses it).
"""
import ctypes as C

import numpy as np
import pytest

from harness import CudaReplay, Field, gen_trace

pytestmark = pytest.mark.gpu


def _twins(cap, seed):
    schema = [Field("action.a", "int64", ()), Field("reward", "float32", (3,))]
    a = CudaReplay(schema, cap, seed, 0.6, 0.4, 0)
    b = CudaReplay(schema, cap, seed, 0.6, 0.4, 0)
    return schema, a, b


def _same_state(a, b, what):
    assert a.sync() == 0 and b.sync() == 0
    assert a.state() == b.state(), what
    wa, ea = a.weights()
    wb, eb = b.weights()
    np.testing.assert_array_equal(wa.view(np.uint32), wb.view(np.uint32), err_msg=what + ": stored weights")
    np.testing.assert_array_equal(ea, eb, err_msg=what + ": eviction flags")
    for level in (0, 1):
        sa, ma, _ = a.tiles(level=level)
        sb, mb, _ = b.tiles(level=level)
        np.testing.assert_array_equal(sa.view(np.uint64), sb.view(np.uint64), err_msg=what + ": hierarchy sums")
        np.testing.assert_array_equal(ma.view(np.uint32), mb.view(np.uint32), err_msg=what + ": hierarchy minima")


@pytest.mark.parametrize("B", [64, 700, 4096])
def test_no_alloc_update_matches_update_priority(B):
    """B = 64: the existing entry point stages the batch in shared memory; 700 / 4096: it works in
    memory the replay allocated with cudaMalloc. The no-alloc variant uses its static array for all."""
    import torch
    from rela_b200 import capi

    cap = 20000
    schema, a, b = _twins(cap, 11)
    L = a.lib
    assert B <= capi.RR_NOALLOC_MAX_BATCH
    ops = gen_trace(schema, cap, 5 + B, 6, 3000, B, prefill=cap - 1000, adds_per_cycle=(1, 2), zero_frac=0.02)
    g = torch.Generator(device="cuda")
    g.manual_seed(B)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    flags = capi.RR_PRIO_DEVICE
    k = 0
    for op in ops:
        if op[0] == "add":
            a.add(op[1], op[2])
            b.add(op[1], op[2])
        elif op[0] == "sample":
            rc_a, _, w_a, ids_a = a.sample(op[1], want_rows=False)
            rc_b, _, w_b, ids_b = b.sample(op[1], want_rows=False)
            assert rc_a == 0 and rc_b == 0
            np.testing.assert_array_equal(ids_a, ids_b, err_msg="batch %d: ids" % k)
            np.testing.assert_allclose(w_a, w_b, rtol=1e-6, atol=0, err_msg="batch %d: importance weights" % k)
        else:
            # raw priorities (the kernel raises them to alpha), some of them zero, device-resident
            p = torch.rand(B, device="cuda", generator=g) + 0.01
            p[torch.rand(B, device="cuda", generator=g) < 0.03] = 0.0
            capi.check(L.rr_update_priority(a.h, B, C.c_void_p(p.data_ptr()), flags, stream))
            capi.check(L.rr_update_priority_noalloc(b.h, B, C.c_void_p(p.data_ptr()), flags, stream))
            k += 1
            _same_state(a, b, "after update %d (B = %d)" % (k, B))
    assert k >= 2  # at least two calls in a row
    sa, sb = a.stats(), b.stats()
    assert sa["device_errors"] == 0 and sb["device_errors"] == 0
    a.close()
    b.close()


def test_no_alloc_update_rejects_what_it_cannot_do():
    import torch
    from rela_b200 import capi

    cap = 40000
    schema, a, _b = _twins(cap, 3)
    _b.close()
    L = a.lib
    ops = gen_trace(schema, cap, 1, 0, 5000, 16, prefill=cap - 2000)
    for op in ops:
        a.add(op[1], op[2])
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    too_many = capi.RR_NOALLOC_MAX_BATCH + 8
    p = torch.rand(too_many, device="cuda") + 0.01
    host = np.ones(64, dtype=np.float32)
    # nothing sampled yet
    assert L.rr_update_priority_noalloc(a.h, 64, C.c_void_p(p.data_ptr()), capi.RR_PRIO_DEVICE, stream) == capi.RR_ERR_STATE
    rc, _, _, _ = a.sample(too_many, want_rows=False)
    assert rc == 0
    # beyond the static array: a clear error, and the batch stays outstanding for rr_update_priority
    assert L.rr_update_priority_noalloc(a.h, too_many, C.c_void_p(p.data_ptr()), capi.RR_PRIO_DEVICE, stream) == capi.RR_ERR_ARG
    assert b"RR_NOALLOC_MAX_BATCH" in L.rr_last_error()
    capi.check(L.rr_update_priority(a.h, too_many, C.c_void_p(p.data_ptr()), capi.RR_PRIO_DEVICE, stream))
    rc, _, _, _ = a.sample(64, want_rows=False)
    assert rc == 0
    # host priorities would need a staging buffer
    assert L.rr_update_priority_noalloc(a.h, 64, host.ctypes.data, 0, stream) == capi.RR_ERR_ARG
    # wrong length (prioritized_replay.h:237)
    assert L.rr_update_priority_noalloc(a.h, 32, C.c_void_p(p.data_ptr()), capi.RR_PRIO_DEVICE, stream) == capi.RR_ERR_ARG
    capi.check(L.rr_update_priority_noalloc(a.h, 64, C.c_void_p(p.data_ptr()), capi.RR_PRIO_DEVICE, stream))
    assert a.sync() == 0
    assert a.stats()["device_errors"] == 0
    a.close()
