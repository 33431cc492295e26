"""rela_b200: B200-native replay hot path behind rela's replay API.

Public surface (mirrors reference rela/pybind.cc:19-59):
    FFTransition, RNNTransition, FFPrioritizedReplay, RNNPrioritizedReplay
plus the actor-side arithmetic kernels nstep_return / td_priority /
aggregate_priority. Everything runs in librela_b200.so (CUDA, sm_100a).
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
