"""CPU: the C-ABI library builds for sm_100a, loads, exports every symbol that
include/rela_b200.h declares, and fails loudly (no CPU fallback) without a GPU."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "rela_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(rr_[a-z_0-9]+)\s*\(", text)))


def test_header_symbols_exported():
    from rela_b200 import build, capi

    build.build_kernels()
    lib = capi.lib()
    names = _declared()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/rela_b200.h but not exported"
    assert sorted(capi.SYMBOLS) == names, (sorted(set(names) - set(capi.SYMBOLS)), sorted(set(capi.SYMBOLS) - set(names)))


def test_no_cpu_fallback():
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from rela_b200 import capi

    lib = capi.lib()
    h = C.c_void_p()
    rc = lib.rr_create(C.byref(h), 128, 1, 0.6, 0.4, 0, 0)
    assert rc == capi.RR_ERR_CUDA
    assert b"no CPU path" in lib.rr_last_error()
    import rela_b200

    with pytest.raises(RuntimeError, match="no CPU path"):
        rela_b200.FFPrioritizedReplay(128, 1, 0.6, 0.4, 0)


def test_product_does_not_touch_oracle():
    """The oracle is test infrastructure: nothing under rela_b200/ may name it."""
    for dirpath, _, files in os.walk(os.path.join(ROOT, "rela_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cc")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "oracle" not in text.lower() or f == "build.py", f"{f} mentions the oracle"
