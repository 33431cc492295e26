// types.h -- rela_b200's own record types for the stand-alone Python module
// (rela_b200/_C). Same members as the reference's records (rela/types.h:18-73,
// bound at rela/pybind.cc:20-35); the reference's index / makeBatch / padLike
// methods are not needed here because splitting and batching happen on the
// device (row insert and gather kernels).
#pragma once

#include <string>
#include <unordered_map>

#include <torch/extension.h>

#include "rela_b200/transition_traits.h"

namespace rela_b200 {

using TensorDict = std::unordered_map<std::string, torch::Tensor>;

struct FFTransition {
  TensorDict obs;
  TensorDict action;
  torch::Tensor reward;
  torch::Tensor terminal;
  torch::Tensor bootstrap;
  TensorDict nextObs;
};

struct RNNTransition {
  TensorDict obs;
  TensorDict h0;
  TensorDict action;
  torch::Tensor reward;
  torch::Tensor terminal;
  torch::Tensor bootstrap;
  torch::Tensor seqLen;

  // element i of a batch of sequences, with the reference's conventions
  // (RNNTransition::index, rela/types.cc:119-138): dim 0 of every tensor is the
  // element, except h0 whose element dim is 1 ([layers, K, hid]).
  RNNTransition index(int i) const {
    RNNTransition e;
    for (const auto& kv : obs) e.obs.insert({kv.first, kv.second.select(0, i)});
    for (const auto& kv : h0) e.h0.insert({kv.first, kv.second.select(1, i)});
    for (const auto& kv : action) e.action.insert({kv.first, kv.second.select(0, i)});
    e.reward = reward.select(0, i);
    e.terminal = terminal.select(0, i);
    e.bootstrap = bootstrap.select(0, i);
    e.seqLen = seqLen.select(0, i);
    return e;
  }
};

template <>
struct TransitionTraits<FFTransition> : FFTraits<FFTransition> {};
template <>
struct TransitionTraits<RNNTransition> : RNNTraits<RNNTransition> {};

using FFPrioritizedReplay = PrioritizedReplay<FFTransition>;
using RNNPrioritizedReplay = PrioritizedReplay<RNNTransition>;

}  // namespace rela_b200
