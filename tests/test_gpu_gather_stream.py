"""$RR_GATHER_STREAM=1: gathers on their own stream, control plane
on the internal stream. An EXPERIMENT that is off by default, so these tests only run with
RELA_B200_EXPERIMENTAL=1 (the whole GPU suite can also be run under RR_GATHER_STREAM=1: the
switch is read by rr_create). Results must not change: same traces, same oracle, bit for bit."""
import os

import numpy as np
import pytest

from harness import (CudaReplay, CudaReplayHostPrefetch, Oracle, apply_trace, compare_records, ff_schema,
                     gen_trace, rnn_schema)

pytestmark = [
    pytest.mark.gpu,
    pytest.mark.skipif(os.environ.get("RELA_B200_EXPERIMENTAL", "0") in ("", "0"),
                       reason="experimental switch; set RELA_B200_EXPERIMENTAL=1"),
]


@pytest.fixture(autouse=True)
def gather_stream_on():
    old = os.environ.get("RR_GATHER_STREAM")
    os.environ["RR_GATHER_STREAM"] = "1"  # read by rr_create
    yield
    if old is None:
        del os.environ["RR_GATHER_STREAM"]
    else:
        os.environ["RR_GATHER_STREAM"] = old


def _pair(make_cuda, schema, cap, seed, alpha, beta, prefetch, ops):
    orc = Oracle(schema, cap, seed, alpha, beta, prefetch)
    cu = make_cuda()
    ra = apply_trace(orc, ops, alpha, False)
    rb = apply_trace(cu, ops, alpha, False)
    assert cu.sync() == 0
    compare_records(ra, rb, "oracle vs cuda (gather stream)")
    assert orc.state() == cu.state()
    wa, ea = orc.weights()
    wb, eb = cu.weights()
    st = orc.state()
    idx = (st["head"] + np.arange(st["safe_size"])) % orc.S
    np.testing.assert_array_equal(wa[idx], wb[idx])
    np.testing.assert_array_equal(ea[idx], eb[idx])


@pytest.mark.parametrize("prefetch", [0, 1, 2])
@pytest.mark.parametrize("own_stream", [False, True])
def test_ff_trace(prefetch, own_stream):
    """Library-managed prefetch (device-to-device hand-out) and none, from a foreign stream and
    from the replay's own stream (which joins the gather stream at every hand-out)."""
    schema = ff_schema(frame=(4, 12, 12), n_actions=6)
    cap, alpha = 256, 0.6
    ops = gen_trace(schema, cap, 3, 80, 20, 32)
    _pair(lambda: CudaReplay(schema, cap, 7, alpha, 0.4, prefetch, own_stream=own_stream),
          schema, cap, 7, alpha, 0.4, prefetch, ops)


@pytest.mark.parametrize("prefetch", [1, 2, 3])
def test_host_managed_prefetch(prefetch):
    """The pipelined mode the switch is for: rr_prefetch_into / rr_take_prefetched, where the
    selection and update of later batches overlap the gathers of earlier ones."""
    schema = ff_schema(frame=(4, 84, 84))  # full frames: the gather really is the long kernel
    cap, alpha = 2048, 0.6
    ops = gen_trace(schema, cap, 19, 60, 80, 256)
    _pair(lambda: CudaReplayHostPrefetch(schema, cap, 3, alpha, 0.4, prefetch),
          schema, cap, 3, alpha, 0.4, prefetch, ops)


def test_eviction_and_slot_reuse_under_adds():
    """Small ring, adds between every sample: evicted rows are re-reserved while earlier gathers
    may still be queued on the gather stream (gate2), and id buffers are recycled (evSlot)."""
    schema = ff_schema(frame=(4, 84, 84))
    cap, alpha = 300, 0.6
    ops = gen_trace(schema, cap, 5, 150, 60, 128, adds_per_cycle=(1, 3))
    _pair(lambda: CudaReplayHostPrefetch(schema, cap, 11, alpha, 0.4, 2),
          schema, cap, 11, alpha, 0.4, 2, ops)


def test_rnn_trace():
    schema = rnn_schema(T=12, frame=(1, 16, 16), n_actions=4, hid=32)
    cap, alpha = 200, 0.9
    ops = gen_trace(schema, cap, 23, 40, 10, 16)
    _pair(lambda: CudaReplay(schema, cap, 9, alpha, 0.6, 0), schema, cap, 9, alpha, 0.6, 0, ops)
