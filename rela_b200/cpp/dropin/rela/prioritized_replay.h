// rela/prioritized_replay.h -- DROP-IN replacement for the reference header of
// the same path (rela/prioritized_replay.h:173-347 of facebookresearch/rela).
//
// Put this directory BEFORE the reference tree on the include path:
//
//   g++ -I<rela_b200>/cpp/dropin -I<rela_b200>/cpp -I<rela_b200>/../include \
//       -I<reference> ... rela/pybind.cc rela/types.cc <rela_b200>/cpp/replay_core.cc \
//       -L<rela_b200> -lrela_b200
//
// `#include "rela/prioritized_replay.h"` in the reference's dqn_actor.h:10,
// r2d2_actor.h and pybind.cc:11 then resolves to this file, while
// "rela/types.h" and every other header still come from the reference. The
// reference's actors, thread loops and pybind module compile UNMODIFIED on top:
// DQNActor / R2D2Actor keep calling replayBuffer_->add(...) from their threads
// (dqn_actor.h:189, r2d2_actor.h:301) and Python keeps calling
// rela.FFPrioritizedReplay(...).sample / update_priority (pybind.cc:37-59).
#pragma once

#include <future>
#include <random>
#include <vector>

#include <torch/extension.h>

#include "rela/types.h"  // the reference's records (resolved through -I<reference>)

#include "rela_b200/transition_traits.h"

namespace rela_b200 {
template <>
struct TransitionTraits<rela::FFTransition> : FFTraits<rela::FFTransition> {};
template <>
struct TransitionTraits<rela::RNNTransition> : RNNTraits<rela::RNNTransition> {};
}  // namespace rela_b200

namespace rela {

template <class DataType>
using PrioritizedReplay = rela_b200::PrioritizedReplay<DataType>;

using FFPrioritizedReplay = PrioritizedReplay<FFTransition>;
using RNNPrioritizedReplay = PrioritizedReplay<RNNTransition>;

}  // namespace rela
