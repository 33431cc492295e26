"""Learner-side fusion (rr_set_output_format): a u8 field gathered as
float32 / bfloat16 = byte / 255 is bit-identical to the learner's own `s.float() / 255`
(pyrela/net.py:46,104) applied ON THE GPU to the reference's batch."""
import numpy as np
import pytest

from harness import CudaReplay, Oracle, apply_trace, ff_schema, gen_trace, rnn_schema

pytestmark = pytest.mark.gpu


def _expect(rows_u8, fmt):
    import torch

    # torch's CUDA `x.float() / 255` multiplies by the float reciprocal of the scalar divisor
    f32 = rows_u8.astype(np.float32) * (np.float32(1.0) / np.float32(255.0))
    if fmt == 1:
        return f32
    return torch.from_numpy(f32).bfloat16().view(torch.int16).numpy().view(np.uint16)


@pytest.mark.parametrize("kind,fmt", [("ff", 1), ("ff", 2), ("rnn", 1), ("rnn", 2)])
def test_converted_gather_matches_torch_expression(kind, fmt):
    from rela_b200 import capi

    if kind == "ff":
        schema, cap, K, B = ff_schema(frame=(4, 84, 84), n_actions=6), 400, 40, 96
        conv = [i for i, f in enumerate(schema) if f.name in ("obs.s", "next_obs.s")]
    else:
        schema, cap, K, B = rnn_schema(T=9, frame=(2, 10, 10), n_actions=4, hid=16), 80, 8, 12
        conv = [0]                   # obs.s, gathered time-major [T, B, ...]
    assert all(schema[i].dtype == np.uint8 for i in conv)
    ops = gen_trace(schema, cap, 41, 25, K, B)
    orc = Oracle(schema, cap, 9, 0.6, 0.4, 0)
    cu = CudaReplay(schema, cap, 9, 0.6, 0.4, 0)
    for i in conv:
        assert cu.set_output_format(i, fmt) == capi.RR_OK
    ra = apply_trace(orc, ops, 0.6, False)
    rb = apply_trace(cu, ops, 0.6, False)
    assert cu.sync() == 0
    assert len(ra) == len(rb) > 0
    for a, b in zip(ra, rb):
        np.testing.assert_array_equal(a["ids"], b["ids"])
        for i, f in enumerate(schema):
            if i in conv:
                np.testing.assert_array_equal(_expect(a["rows"][i], fmt), b["rows"][i], err_msg=f.name)
            else:
                np.testing.assert_array_equal(a["rows"][i], b["rows"][i], err_msg=f.name)


def test_output_format_refusals():
    from harness import Field
    from rela_b200 import capi

    schema = [Field("obs.s", "uint8", (3, 7, 7)), Field("obs.t", "uint8", (8,)), Field("reward", "float32", ())]
    cu = CudaReplay(schema, 32, 1, 0.6, 0.4, 0)
    assert cu.set_output_format(0, 1) == capi.RR_ERR_ARG and b"multiple of 4" in cu.lib.rr_last_error()  # 147 B rows
    assert cu.set_output_format(1, 1) == capi.RR_OK
    assert cu.set_output_format(1, 0) == capi.RR_OK
    assert cu.set_output_format(7, 1) == capi.RR_ERR_ARG
    assert cu.set_output_format(1, 9) == capi.RR_ERR_ARG
    pre = CudaReplay(schema, 32, 1, 0.6, 0.4, 2)
    assert pre.set_output_format(1, 1) == capi.RR_ERR_ARG and b"prefetch" in pre.lib.rr_last_error()


def test_python_layer_output_format():
    import torch

    import rela_b200.rela as rela

    def transition(n, g):
        obs = {"s": torch.randint(0, 256, (n, 4, 12, 12), dtype=torch.uint8, generator=g)}
        act = {"a": torch.randint(0, 4, (n,), dtype=torch.int64, generator=g)}
        nxt = {"s": torch.randint(0, 256, (n, 4, 12, 12), dtype=torch.uint8, generator=g)}
        return rela.FFTransition(obs, act, torch.randn(n, generator=g), torch.zeros(n, dtype=torch.bool),
                                 torch.ones(n), nxt)

    plain = rela.FFPrioritizedReplay(200, 3, 0.6, 0.4, 0)
    fused = rela.FFPrioritizedReplay(200, 3, 0.6, 0.4, 0)
    g = torch.Generator().manual_seed(8)
    for _ in range(10):
        tr, p = transition(16, g), torch.rand(16, generator=g) + 0.01
        plain.add(tr, p)
        fused.add(tr, p)
    fused.set_output_format("obs.s", torch.float32)
    fused.set_output_format("next_obs.s", torch.bfloat16)
    with pytest.raises(RuntimeError):
        fused.set_output_format("reward", torch.float32)   # not a uint8 leaf
    with pytest.raises(RuntimeError):
        fused.set_output_format("obs.nope", torch.float32)
    for _ in range(3):  # the second and third batches come out of the (re-typed) batch-set pool
        a, wa = plain.sample(48, "cuda:0")
        b, wb = fused.sample(48, "cuda:0")
        assert torch.equal(wa, wb) and torch.equal(a.action["a"], b.action["a"])
        assert b.obs["s"].dtype == torch.float32 and b.next_obs["s"].dtype == torch.bfloat16
        assert torch.equal(a.obs["s"].float() / 255, b.obs["s"])
        assert torch.equal((a.next_obs["s"].float() / 255).bfloat16(), b.next_obs["s"])
        prio = torch.rand(48, generator=g) + 0.01
        plain.update_priority(prio)
        fused.update_priority(prio)
        del a, b
    fused.set_output_format("obs.s", torch.uint8)
    b, _ = fused.sample(48, "cuda:0")
    assert b.obs["s"].dtype == torch.uint8
