"""rela_b200: B200-native replay hot path behind rela's replay API.

Public surface (mirrors reference rela/pybind.cc:19-59):
    FFTransition, RNNTransition, FFPrioritizedReplay, RNNPrioritizedReplay
plus the actor-side arithmetic kernels nstep_return / td_priority /
aggregate_priority. Everything runs in librela_b200.so (CUDA, sm_100a).

Two host layers sit on the C ABI (include/rela_b200.h):
  rela_b200.rela   the C++ host layer + pybind module (rela_b200/cpp/, built as
                   rela_b200/_C.so): same class and method names as the
                   reference's `rela` module; `import rela_b200.rela as rela`.
  rela_b200.*      a ctypes mirror of the same classes (replay.py), used by the
                   parity tests and the bench because it exposes the raw handle.
"""
from .capi import ReplayError, lib  # noqa: F401
from .replay import (  # noqa: F401
    FFPrioritizedReplay,
    FFTransition,
    RNNPrioritizedReplay,
    RNNTransition,
    aggregate_priority,
    nstep_return,
    td_priority,
)

