"""`import rela_b200.rela as rela`: the replay classes of the reference's `rela`
module (rela/pybind.cc:19-59) served by the C++ host layer (rela_b200/cpp/,
built in-tree as rela_b200/_C.so by rela_b200.build.build_shim). No fallback:
importing this module fails if the extension has not been built."""
from ._C import (FFPrioritizedReplay, FFTransition, RNNPrioritizedReplay, RNNTransition,  # noqa: F401
                 SequenceAssembler)
