"""rela_b200: B200-native replay hot path behind rela's replay API.

Public surface (mirrors reference rela/pybind.cc:19-59):
    FFTransition, RNNTransition, FFPrioritizedReplay, RNNPrioritizedReplay
plus the actor-side arithmetic kernels nstep_return / td_priority /
aggregate_priority and SequenceAssembler. Everything runs in librela_b200.so
(CUDA, sm_100a) behind the C ABI of include/rela_b200.h; the classes are the
C++ host layer (rela_b200/cpp/, built in-tree as rela_b200/_C.so). There is no
fallback: using any of the classes fails if the extension has not been built, and
constructing a replay fails without a CUDA device.

`import rela_b200.rela as rela` gives the same classes under the reference's
module name.
"""
from .capi import ReplayError, lib  # noqa: F401

_FROM_C = ("FFPrioritizedReplay", "FFTransition", "RNNPrioritizedReplay", "RNNTransition", "SequenceAssembler",
           "aggregate_priority", "nstep_return", "td_priority", "raise_at_position")
__all__ = ("ReplayError", "lib") + _FROM_C


def __getattr__(name):
    # The classes live in the compiled host layer. It is imported on first use, not at package
    # import, so that `rela_b200.build` (which compiles it) can be imported from a fresh checkout.
    if name in _FROM_C:
        try:
            from . import _C
        except ImportError as e:
            raise ImportError("rela_b200._C is not built (python -c 'import __graft_entry__ as g; g.build()'): "
                              "there is no Python or CPU fallback") from e
        return getattr(_C, name)
    raise AttributeError("module 'rela_b200' has no attribute %r" % name)
