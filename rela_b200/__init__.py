"""rela_b200: B200-native replay hot path behind rela's replay API.

Public surface (mirrors reference rela/pybind.cc:19-59):
    FFTransition, RNNTransition, FFPrioritizedReplay, RNNPrioritizedReplay
plus the actor-side arithmetic kernels nstep_return / td_priority /
aggregate_priority and SequenceAssembler. Everything runs in librela_b200.so
(CUDA, sm_100a) behind the C ABI of include/rela_b200.h; the classes are the
C++ host layer (rela_b200/cpp/, built in-tree as rela_b200/_C.so). There is no
fallback: importing this package fails if the extension has not been built, and
constructing a replay fails without a CUDA device.

`import rela_b200.rela as rela` gives the same classes under the reference's
module name.
"""
from .capi import ReplayError, lib  # noqa: F401
from ._C import (  # noqa: F401
    FFPrioritizedReplay,
    FFTransition,
    RNNPrioritizedReplay,
    RNNTransition,
    SequenceAssembler,
    aggregate_priority,
    nstep_return,
    td_priority,
)
